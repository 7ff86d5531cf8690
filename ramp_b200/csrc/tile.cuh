// tile.cuh -- Model tables in shared memory, the per-warp slab, row loads / TMA prefetch, q expansion and
// forward kinematics of a tile of 32 configurations (lane = configuration).
// Part of the single translation unit ramp_b200.cu (included there, in this order).
#pragma once

namespace ramp {

struct Tables {
  const ImgHeader* h;
  const DJoint* joints;
  const DGeom* geoms;
  const DPair* pairs;
  const float4* verts;
  const float* stat;  // [n_static][12]: x axis, z axis, translation, y axis
  const int32_t* dyn; // slot -> geometry index
  const int32_t* act;
  const DMimic* mimic;
  const int32_t* cont;
  const float* refq;
  const int32_t* plan;  // FK plans for teams of warps (see build_image)
};

__device__ __forceinline__ Tables make_tables(const unsigned char* base) {
  Tables t;
  t.h = reinterpret_cast<const ImgHeader*>(base);
  t.joints = reinterpret_cast<const DJoint*>(base + t.h->off_joints);
  t.geoms = reinterpret_cast<const DGeom*>(base + t.h->off_geoms);
  t.pairs = reinterpret_cast<const DPair*>(base + t.h->off_pairs);
  t.verts = reinterpret_cast<const float4*>(base + t.h->off_verts);
  t.stat = reinterpret_cast<const float*>(base + t.h->off_static);
  t.dyn = reinterpret_cast<const int32_t*>(base + t.h->off_dyn);
  t.act = reinterpret_cast<const int32_t*>(base + t.h->off_act);
  t.mimic = reinterpret_cast<const DMimic*>(base + t.h->off_mimic);
  t.cont = reinterpret_cast<const int32_t*>(base + t.h->off_cont);
  t.refq = reinterpret_cast<const float*>(base + t.h->off_refq);
  t.plan = reinterpret_cast<const int32_t*>(base + t.h->off_plan);
  return t;
}

extern __shared__ __align__(128) unsigned char g_smem[];

// Per-warp shared-memory slab (offsets in 4-byte words from the slab base).
struct SlabLayout {
  int32_t words;      // slab size
  int32_t off_q;      // [nq][32]   Pinocchio q, slot-major
  int32_t off_act;    // [n_cols][32] raw actuated input
  int32_t off_geom;   // [n_dyn][12][32] geometry placements
  int32_t off_save;   // [n_save][12][32] transforms of branching joints
  int32_t off_queue;  // uint16[kQueueCap]
  int32_t off_mask;   // uint32[mask_words][32]
  int32_t off_first;  // int32[32] first colliding pair per configuration
  int32_t off_stage;  // [32][stage_stride] output staging of the FK kernels
  int32_t stage_stride;
  int32_t table_bytes;  // image size rounded up to 16
  int32_t team;         // warps sharing one slab (1, 2 or 4): they split the pair list of a tile
  int32_t chunk_pairs;  // pairs per queue round of one warp
  int32_t queue_words;  // per-warp queue size
  int32_t raw_cols;     // validity kernels: widest input row (floats) the TMA row prefetch takes; 0 = off
};

__device__ __forceinline__ int64_t imin64(int64_t a, int64_t b) { return a < b ? a : b; }

// ---- TMA bulk copy of the model image into shared memory
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void stage_tables(unsigned char* dst, const unsigned char* __restrict__ src, int bytes,
                                             uint64_t* mbar) {
  const uint32_t bar = smem_u32(mbar);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
  }
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar)
        : "memory");
  }
}

struct Warp {
  Tables t;
  float* q;
  float* act;
  float* geom;
  float* save;
  uint16_t* queue;
  uint32_t* mask;
  int32_t* first;
  float* stage;
  int stage_stride;
  int lane;
  // 32-bit word offsets into the CTA's shared memory (g_smem) of this warp's geometry slab and of
  // the static placements: placement addressing in the pair loops is plain integer arithmetic
  int geom_w, stat_w;
};

__device__ __forceinline__ Warp make_warp(const Tables& t, float* slab, const SlabLayout& L) {
  Warp w;
  w.t = t;
  w.q = slab + L.off_q;
  w.act = slab + L.off_act;
  w.geom = slab + L.off_geom;
  w.save = slab + L.off_save;
  w.queue = reinterpret_cast<uint16_t*>(slab + L.off_queue + ((threadIdx.x >> 5) % L.team) * L.queue_words);
  w.mask = reinterpret_cast<uint32_t*>(slab + L.off_mask);
  w.first = reinterpret_cast<int32_t*>(slab + L.off_first);
  w.stage = slab + L.off_stage;
  w.stage_stride = L.stage_stride;
  w.lane = threadIdx.x & 31;
  w.geom_w = static_cast<int>(w.geom - reinterpret_cast<const float*>(g_smem));
  w.stat_w = static_cast<int>(t.stat - reinterpret_cast<const float*>(g_smem));
  return w;
}

// ---- coalesced load of a [rows<=32][cols] float tile into slot-major shared memory
// dst[col*32 + row].  Vectorised float4 path when the tile is 16-byte aligned.
__device__ __forceinline__ void load_rows(float* dst, const float* __restrict__ src, int64_t row0, int64_t n_rows,
                                          int cols, int ld, int lane) {
  const int rows = static_cast<int>(imin64(32, n_rows - row0));
  const float* base = src + row0 * ld;
  if (ld == cols && ((rows * cols) & 3) == 0 && (reinterpret_cast<uintptr_t>(base) & 15) == 0) {
    // contiguous rows: the tile is rows * cols / 4 aligned float4 (one may straddle two rows)
    const int n4 = rows * cols / 4;
    const float4* b4 = reinterpret_cast<const float4*>(base);
    const float inv = 1.f / cols;  // floor(i / cols) == floor((i + 0.5) * inv) for i < 4096, cols <= 128
#pragma unroll 2
    for (int e = lane; e < n4; e += 32) {
      const float4 q4 = __ldg(b4 + e);
      int r = __float2int_rd((4 * e + 0.5f) * inv), c = 4 * e - r * cols;
      const float v[4] = {q4.x, q4.y, q4.z, q4.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        dst[c * 32 + r] = v[j];
        if (++c == cols) { c = 0; ++r; }
      }
    }
  } else {
    const int n = rows * cols;
#pragma unroll 1
    for (int e = lane; e < n; e += 32) {
      const int r = e / cols, c = e - r * cols;
      dst[c * 32 + r] = __ldg(base + static_cast<int64_t>(r) * ld + c);
    }
  }
  // lanes past the end of the batch replay row 0 so every lane does well-defined work
  if (rows < 32) {
    __syncwarp();
    if (lane >= rows) {
#pragma unroll 1
      for (int c = 0; c < cols; ++c) dst[c * 32 + lane] = dst[c * 32];
    }
  }
  __syncwarp();
}

// ---- software prefetch of the NEXT tile's rows (FK kernels; the validity kernels use the TMA
// prefetch below, which costs no registers).  Each warp works through load -> FK -> store on
// its own, so without this the global-load latency of every tile is exposed (long_scoreboard
// was the top stall of the FK kernel).  The next tile's float4s are fetched into registers
// before the current tile's arithmetic and committed to shared memory after it.  kVec float4
// per lane cover rows of up to 4 * kVec floats; wider rows take the plain load.  Rows must be
// contiguous (ld == cols, any width: a tile of 32 rows is then 8 * cols aligned float4, and one
// float4 may straddle two rows).
template <int kVec>
struct RowPrefetch {
  float4 v[kVec];
  bool valid;  // the vector path applied (else commit falls back to load_rows)
};

template <int kVec>
__device__ __forceinline__ bool rows_vectorisable(const float* base, int cols, int ld) {
  return ld == cols && cols <= 4 * kVec && (reinterpret_cast<uintptr_t>(base) & 15) == 0;
}

template <int kVec>
__device__ __forceinline__ void prefetch_rows(RowPrefetch<kVec>& pf, const float* __restrict__ src, int64_t row0, int64_t n_rows,
                                              int cols, int ld, int lane) {
  pf.valid = false;
  if (row0 >= n_rows) return;
  const float* base = src + row0 * ld;
  if (!rows_vectorisable<kVec>(base, cols, ld)) return;
  const int n_el = static_cast<int>(imin64(32, n_rows - row0)) * cols;
  if (n_el & 3) return;  // a ragged last tile: plain load
  pf.valid = true;
  const int n4 = n_el / 4;
  const float4* b4 = reinterpret_cast<const float4*>(base);
#pragma unroll
  for (int k = 0; k < kVec; ++k) {
    const int e = lane + 32 * k;
    if (e < n4) pf.v[k] = __ldg(b4 + e);
  }
}

template <int kVec>
__device__ __forceinline__ void commit_rows(const RowPrefetch<kVec>& pf, float* dst, const float* __restrict__ src, int64_t row0,
                                            int64_t n_rows, int cols, int ld, int lane) {
  if (!pf.valid) {
    load_rows(dst, src, row0, n_rows, cols, ld, lane);
    return;
  }
  const int rows = static_cast<int>(imin64(32, n_rows - row0));
  const int n4 = rows * cols / 4;
  const float inv = 1.f / cols;  // floor(i / cols) == floor((i + 0.5) * inv) for i < 4096, cols <= 128
#pragma unroll
  for (int k = 0; k < kVec; ++k) {
    const int e = lane + 32 * k;
    if (e < n4) {
      int r = __float2int_rd((4 * e + 0.5f) * inv), c = 4 * e - r * cols;
      const float v[4] = {pf.v[k].x, pf.v[k].y, pf.v[k].z, pf.v[k].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        dst[c * 32 + r] = v[j];
        if (++c == cols) { c = 0; ++r; }
      }
    }
  }
  if (rows < 32) {
    __syncwarp();
    if (lane >= rows) {
#pragma unroll 1
      for (int c = 0; c < cols; ++c) dst[c * 32 + lane] = dst[c * 32];
    }
  }
  __syncwarp();
}

// ---- TMA prefetch of the NEXT tile's rows (validity kernels).  The 32 rows of a tile are one
// contiguous, 16-byte aligned block when ld == cols; a single cp.async.bulk lands it in the
// slab's `save` region -- dead between the end of FK and the next tile's FK -- behind a
// per-slab mbarrier, so the copy costs no registers and runs under the whole narrow phase.
// The mbarrier sits in the 16 bytes in front of the region (never written by FK).
struct TileFetch {
  uint32_t bar, raw;  // shared-space addresses
  uint32_t phase;
  bool pending;       // a copy for the tile about to be committed is in flight
};

__device__ __forceinline__ void tile_fetch_init(TileFetch& tf, float* region, int lane) {
  tf.bar = smem_u32(region - 4);
  tf.raw = smem_u32(region);
  tf.phase = 0;
  tf.pending = false;
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tf.bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
}

// Start the copy of tile `row0` if it is a full, contiguous, aligned tile that fits; otherwise
// leave it to the plain load at commit time.  Warp-uniform.
__device__ __forceinline__ void tile_fetch_issue(TileFetch& tf, const float* __restrict__ src, int64_t row0, int64_t n_rows,
                                                 int cols, int ld, int raw_cols, int lane) {
  tf.pending = false;
  if (row0 + 32 > n_rows || ld != cols || cols > raw_cols) return;
  const float* base = src + row0 * ld;
  if (reinterpret_cast<uintptr_t>(base) & 15) return;
  tf.pending = true;
  __syncwarp();  // every lane is done reading the region (previous commit / FK saves)
  if (lane == 0) {
    const uint32_t bytes = 128u * cols;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tf.bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(tf.raw),
                 "l"(base), "r"(bytes), "r"(tf.bar)
                 : "memory");
  }
}

// Rows of the tile into slot-major dst[col * 32 + row]: from the landed copy (lane = row), or
// with the plain load.
__device__ __forceinline__ void tile_fetch_commit(TileFetch& tf, float* dst, const float* region, const float* __restrict__ src,
                                                  int64_t row0, int64_t n_rows, int cols, int ld, int lane) {
  if (!tf.pending) {
    load_rows(dst, src, row0, n_rows, cols, ld, lane);
    return;
  }
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(tf.bar), "r"(tf.phase)
        : "memory");
  }
  tf.phase ^= 1;
  const float* raw = region + lane * cols;
#pragma unroll 4
  for (int c = 0; c < cols; ++c) dst[c * 32 + lane] = raw[c];
  __syncwarp();
}

// ---- actuated values -> Pinocchio q (reference robot_state.py:117-119: scatter, mimic,
// then continuous).  Lane-private columns of shared memory, no cross-lane traffic.
__device__ __forceinline__ void expand_q(const Warp& w) {
  const ImgHeader* h = w.t.h;
  const int l = w.lane;
#pragma unroll 1
  for (int s = 0; s < h->nq; ++s) w.q[s * 32 + l] = w.t.refq[s];
#pragma unroll 1
  for (int k = 0; k < h->n_act; ++k) w.q[w.t.act[k] * 32 + l] = w.act[k * 32 + l];
#pragma unroll 1
  for (int k = 0; k < h->n_mimic; ++k) {
    const DMimic m = w.t.mimic[k];
    w.q[m.dst * 32 + l] = fmaf(w.q[m.src * 32 + l], m.mult, m.off);
  }
#pragma unroll 1
  for (int k = 0; k < h->n_cont; ++k) {
    const int c = w.t.cont[k];
    float s, co;
    sincos_small(w.q[c * 32 + l], &s, &co);
    w.q[c * 32 + l] = co;
    w.q[(c + 1) * 32 + l] = s;
  }
}

// a 3x4 placement of the model image (DJoint::pl, DGeom::pl / opl: all 16-byte aligned)
__device__ __forceinline__ Xf load_table_xf(const float* p) {
  const float4* p4 = reinterpret_cast<const float4*>(p);
  const float4 r0 = p4[0], r1 = p4[1], r2 = p4[2];
  Xf o;
  o.m[0] = r0.x; o.m[1] = r0.y; o.m[2] = r0.z; o.m[3] = r0.w;
  o.m[4] = r1.x; o.m[5] = r1.y; o.m[6] = r1.z; o.m[7] = r1.w;
  o.m[8] = r2.x; o.m[9] = r2.y; o.m[10] = r2.z; o.m[11] = r2.w;
  return o;
}

__device__ __forceinline__ void store_slab_xf(float* slab, int slot, int lane, const Xf& x) {
#pragma unroll
  for (int k = 0; k < 12; ++k) slab[(slot * 12 + k) * 32 + lane] = x.m[k];
}

__device__ __forceinline__ Xf load_slab_xf(const float* slab, int slot, int lane) {
  Xf o;
#pragma unroll
  for (int k = 0; k < 12; ++k) o.m[k] = slab[(slot * 12 + k) * 32 + lane];
  return o;
}

// Geometry placements are kept as 9 floats -- x axis, z axis, translation -- in the slab
// (and in the static table): the y axis is z cross x.  A quarter less shared memory per warp
// means more resident warps, and stage 1 only ever reads z and t.
constexpr int kGeomFloats = 9;
// static placements keep their own y axis too (x, z, t, y): a scene object may carry a rotation
// that is not orthonormal (the reference builds poses from un-normalised quaternions)
constexpr int kStatFloats = 12;

__device__ __forceinline__ void store_geom_xf(float* slab, int slot, int lane, const Xf& x) {
  float* p = slab + slot * kGeomFloats * 32 + lane;
  p[0 * 32] = x.m[0]; p[1 * 32] = x.m[4]; p[2 * 32] = x.m[8];
  p[3 * 32] = x.m[2]; p[4 * 32] = x.m[6]; p[5 * 32] = x.m[10];
  p[6 * 32] = x.m[3]; p[7 * 32] = x.m[7]; p[8 * 32] = x.m[11];
}

// World placement of geometry g for configuration c (lane-divergent g and c allowed).
__device__ __forceinline__ Xf geom_xf(const Warp& w, int gslot, int c) {
  const float* S = reinterpret_cast<const float*>(g_smem);
  const float* p = S + (gslot >= 0 ? w.geom_w + gslot * (kGeomFloats * 32) + c : w.stat_w - (gslot + 1) * kStatFloats);
  const int st = gslot >= 0 ? 32 : 1;
  const V3 x = mk(p[0], p[st], p[2 * st]), z = mk(p[3 * st], p[4 * st], p[5 * st]);
  const V3 y = gslot >= 0 ? cross(z, x) : mk(p[9], p[10], p[11]);
  Xf o;
  o.m[0] = x.x; o.m[1] = y.x; o.m[2] = z.x; o.m[3] = p[6 * st];
  o.m[4] = x.y; o.m[5] = y.y; o.m[6] = z.y; o.m[7] = p[7 * st];
  o.m[8] = x.z; o.m[9] = y.z; o.m[10] = z.z; o.m[11] = p[8 * st];
  return o;
}

// ---- forward kinematics for this lane's configuration (pinocchio::forwardKinematics +
// updateGeometryPlacements).  kStage: 0 = placements to the slab only, 1 = also stage oMi
// rows, 2 = also stage oMg rows (output kernels).
// liMi = placement * M(q).  The aligned revolute / prismatic joints -- all but a handful in
// practice -- only mix two columns of the placement's rotation (or slide along one), a third of
// the arithmetic of a general 3x4 product.
template <int A, int B>
__device__ __forceinline__ void mix_columns(Xf& o, float c, float s) {  // columns A, B of R times a plane rotation
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float ua = o.m[4 * r + A], ub = o.m[4 * r + B];
    o.m[4 * r + A] = fmaf(c, ua, s * ub);
    o.m[4 * r + B] = fmaf(c, ub, -s * ua);
  }
}

template <int A>
__device__ __forceinline__ void slide_column(Xf& o, float d) {
#pragma unroll
  for (int r = 0; r < 3; ++r) o.m[4 * r + 3] = fmaf(d, o.m[4 * r + A], o.m[4 * r + 3]);
}

// kFast: the validity kernels take the hardware sine / cosine (sincos_fast); every kernel that
// REPORTS poses or distances keeps the 1e-7 polynomial.
template <bool kFast = false>
__device__ __forceinline__ Xf local_xf(const DJoint* j, const float* qv) {
  Xf o = load_table_xf(j->pl);
  const int type = j->type;
  if (type <= 3) {  // RX / RY / RZ
    float s, c;
    if (kFast) sincos_fast(qv[0], &s, &c);
    else sincos_small(qv[0], &s, &c);
    if (type == 1) mix_columns<1, 2>(o, c, s);
    else if (type == 2) mix_columns<2, 0>(o, c, s);
    else mix_columns<0, 1>(o, c, s);
    return o;
  }
  if (type >= 9 && type <= 11) {  // PX / PY / PZ
    if (type == 9) slide_column<0>(o, qv[0]);
    else if (type == 10) slide_column<1>(o, qv[0]);
    else slide_column<2>(o, qv[0]);
    return o;
  }
  return xf_mul(o, joint_xf(type, j->ax, j->ay, j->az, qv));
}

// Staging of output rows for the FK kernels: row l (this lane's configuration) holds its 3x4
// matrices as float4 triples.  The row pitch (in float4) is odd, so the 32 lanes' STS.128 of
// one matrix row land in distinct bank groups; an odd pitch that equals the row length makes
// the tile in shared memory the exact image of the tile in global memory.
__device__ __forceinline__ void stage_xf(const Warp& w, int slot, const Xf& x) {
  float4* p = reinterpret_cast<float4*>(w.stage) + w.lane * w.stage_stride + slot * 3;
  p[0] = make_float4(x.m[0], x.m[1], x.m[2], x.m[3]);
  p[1] = make_float4(x.m[4], x.m[5], x.m[6], x.m[7]);
  p[2] = make_float4(x.m[8], x.m[9], x.m[10], x.m[11]);
}

// parent * placement for a placement whose rotation is a quarter / half turn about one axis
// (PlacementKind): the parent's columns, shuffled and negated, and the translation.
template <int A, int B>
__device__ __forceinline__ void turn_columns(Xf& o, const Xf& a, int how) {  // how: 0 = +90, 1 = -90, 2 = 180 degrees in the (A, B) plane
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float ua = a.m[4 * r + A], ub = a.m[4 * r + B];
    o.m[4 * r + A] = how == 0 ? ub : (how == 1 ? -ub : -ua);
    o.m[4 * r + B] = how == 0 ? -ua : (how == 1 ? ua : -ub);
  }
}

// kShuffle = false: always the full product (the pose-output kernels are bound by HBM and by issue
// slots about equally; the selects of the shuffle cost them more than the FMAs they replace)
template <bool kShuffle>
__device__ __forceinline__ Xf mul_placement(const Xf& a, const DJoint* j) {
  const Xf p = load_table_xf(j->pl);
  const int kind = kShuffle ? j->pkind : kPlGeneral;
  if (kind == kPlGeneral) return xf_mul(a, p);
  Xf o = a;
#pragma unroll
  for (int r = 0; r < 3; ++r) o.m[4 * r + 3] = fmaf(a.m[4 * r], p.m[3], fmaf(a.m[4 * r + 1], p.m[7], fmaf(a.m[4 * r + 2], p.m[11], a.m[4 * r + 3])));
  if (kind >= kPlZp) turn_columns<0, 1>(o, a, kind - kPlZp);       // about z: columns x, y
  else if (kind >= kPlYp) turn_columns<2, 0>(o, a, kind - kPlYp);  // about y: columns z, x
  else if (kind >= kPlXp) turn_columns<1, 2>(o, a, kind - kPlXp);  // about x: columns y, z
  return o;
}

// ---- forward kinematics for this lane's configuration (pinocchio::forwardKinematics +
// updateGeometryPlacements).  kStage: 0 = geometry placements to the slab (validity path),
// 1 = stage the oMi rows, 2 = stage the oMg rows (output kernels).
// oMi = oMi[parent] * placement * M(q): the placement is applied first -- for the common quarter-turn
// and pure-translation placements that costs no multiplications for the rotation -- and the joint's
// own motion then mixes two columns (revolute about x / y / z) or slides along one (prismatic).
template <int kStage, bool kFast = (kStage == 0)>
__device__ __forceinline__ void fk_joint(const Warp& w, int i, Xf& cur, bool emit) {
  const int l = w.lane;
  const DJoint* j = &w.t.joints[i];
  float qv[7];
  const float* qj = w.q + j->idx_q * 32 + l;
  qv[0] = qj[0];
  const int n = j->nq;  // 1 for all but a few joints of a robot: skip the other six loads
#pragma unroll
  for (int k = 1; k < 7; ++k) qv[k] = 0.f;
  if (n > 1) {
#pragma unroll
    for (int k = 1; k < 7; ++k) qv[k] = k < n ? qj[k * 32] : 0.f;
  }
  const int pc = j->parent_code;
  if (pc == kParentUniverse) cur = load_table_xf(j->pl);
  else cur = mul_placement<kStage == 0>(pc == kParentRunning ? cur : load_slab_xf(w.save, pc, l), j);
  const int type = j->type;
  if (type <= 3) {  // RX / RY / RZ
    float s, c;
    if (kFast) sincos_fast(qv[0], &s, &c);  // validity path: see local_xf
    else sincos_small(qv[0], &s, &c);
    if (type == 1) mix_columns<1, 2>(cur, c, s);
    else if (type == 2) mix_columns<2, 0>(cur, c, s);
    else mix_columns<0, 1>(cur, c, s);
  } else if (type >= 9 && type <= 11) {  // PX / PY / PZ
    if (type == 9) slide_column<0>(cur, qv[0]);
    else if (type == 10) slide_column<1>(cur, qv[0]);
    else slide_column<2>(cur, qv[0]);
  } else {
    cur = xf_mul(cur, joint_xf(type, j->ax, j->ay, j->az, qv));
  }
  if (j->save_slot >= 0) store_slab_xf(w.save, j->save_slot, l, cur);
  if (kStage == 1) stage_xf(w, i, cur);
  const int gend = (kStage == 1 || !emit) ? j->gfirst : j->gfirst + j->gcount;
#pragma unroll 1
  for (int s = j->gfirst; s < gend; ++s) {
    const int gi = w.t.dyn[s];
    const Xf og = xf_mul(cur, load_table_xf(kStage == 2 ? w.t.geoms[gi].opl : w.t.geoms[gi].pl));
    if (kStage == 2) stage_xf(w, gi, og);
    else store_geom_xf(w.geom, s, l, og);
  }
}

template <int kStage, bool kFast = (kStage == 0)>
__device__ __forceinline__ void fk_lane(const Warp& w) {
  Xf cur = xf_identity();
#pragma unroll 1
  for (int i = 0; i < w.t.h->n_joints; ++i) fk_joint<kStage, kFast>(w, i, cur, true);
}

// The same for a team of kTeam warps sharing one slab: warp `tw` walks its FK plan -- a
// contiguous run of the joint order whose geometries it places, preceded by the ancestors of
// that run (recomputed, not placed) -- so the kinematic tree is split between the warps
// instead of one of them doing all of it while the others wait.
template <int kTeam>
__device__ __forceinline__ void fk_team(const Warp& w, int tw) {
  const int32_t* plan = w.t.plan;
  const int base = kTeam == 2 ? kPlanTeam2 : kPlanTeam4;
  const int e = plan[base + tw + 1];
  Xf cur = xf_identity();
#pragma unroll 1
  for (int k = plan[base + tw]; k < e; ++k) {
    const int ent = plan[kPlanEntries + k];
    fk_joint<0>(w, ent & 0xffff, cur, (ent >> 16) != 0);
  }
}

}  // namespace ramp
