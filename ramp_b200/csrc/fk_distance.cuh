// fk_distance.cuh -- fk_kernel (poses out) and distance_kernel (signed distances).
// Part of the single translation unit ramp_b200.cu (included there, in this order).
#pragma once

namespace ramp {

// --------------------------------------------------------------- kernel: FK / placements
struct FkArgs {
  const unsigned char* img;
  const float* q;
  int64_t n;
  int32_t ld;
  uint32_t flags;
  float* out;
  int32_t row_floats;  // n_joints*12 or n_geoms*12
  SlabLayout L;
};

template <int kWhat>  // 1 = oMi, 2 = oMg
__global__ void __launch_bounds__(512, 1) fk_kernel(const __grid_constant__ FkArgs a) {
  uint64_t* mbar = reinterpret_cast<uint64_t*>(g_smem);
  unsigned char* tbl = g_smem + 16;
  stage_tables(tbl, a.img, a.L.table_bytes, mbar);
  const Tables t = make_tables(tbl);
  const int nwarps = blockDim.x >> 5, wid = threadIdx.x >> 5;
  float* slab = reinterpret_cast<float*>(tbl + a.L.table_bytes) + static_cast<size_t>(wid) * a.L.words;
  const Warp w = make_warp(t, slab, a.L);
  const bool actuated = a.flags & RAMP_Q_ACTUATED;
  int cols = actuated ? t.h->n_act : t.h->nq;
  if ((a.ld & 3) == 0 && a.ld <= cols + 3) cols = a.ld;  // padded rows: take the float4 path
  const int64_t ntiles = (a.n + 31) >> 5;
  const int64_t tile_step = static_cast<int64_t>(gridDim.x) * nwarps;
  RowPrefetch<8> pf;
  prefetch_rows(pf, a.q, (static_cast<int64_t>(blockIdx.x) * nwarps + wid) << 5, a.n, cols, a.ld, w.lane);
  for (int64_t tile = static_cast<int64_t>(blockIdx.x) * nwarps + wid; tile < ntiles; tile += tile_step) {
    const int64_t row0 = tile << 5;
    commit_rows(pf, actuated ? w.act : w.q, a.q, row0, a.n, cols, a.ld, w.lane);
    prefetch_rows(pf, a.q, (tile + tile_step) << 5, a.n, cols, a.ld, w.lane);
    if (actuated) expand_q(w);
    // the bulk stores of the previous tile must have finished READING the staging rows
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    __syncwarp();
    fk_lane<kWhat>(w);
    if (kWhat == 2) {  // universe-attached geometries have constant placements
      for (int g = 0; g < t.h->n_geoms; ++g)
        if (t.geoms[g].gslot < 0) stage_xf(w, g, load_table_xf(t.geoms[g].opl));
    }
    // make the generic-proxy writes to shared memory visible to the async (TMA) proxy
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    // the tile's rows are contiguous in the output: one TMA bulk store for the whole tile when
    // the staging pitch equals the row length, else one per row (cp.async.bulk, SASS UBLKCP)
    const int rows = static_cast<int>(imin64(32, a.n - row0));
    float* dst = a.out + row0 * a.row_floats;
    const uint32_t row_bytes = static_cast<uint32_t>(a.row_floats) * 4u;
    if (w.stage_stride * 4 == a.row_floats) {
      if (w.lane == 0)
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(w.stage)),
                     "r"(row_bytes * rows)
                     : "memory");
    } else if (w.lane < rows) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + static_cast<size_t>(w.lane) * a.row_floats),
                   "r"(smem_u32(reinterpret_cast<float4*>(w.stage) + w.lane * w.stage_stride)), "r"(row_bytes)
                   : "memory");
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// ------------------------------------------------------------- kernel: signed distances
struct DistArgs {
  const unsigned char* img;
  const float* q;
  int64_t n;
  int32_t ld;
  uint32_t flags;
  const int32_t* sel;  // [n_sel] pair indices (device); nullptr = every pair
  int32_t n_sel;
  float* out;          // [n][n_sel]
  float* min_out;      // when set: [n] smallest distance over the ACTIVE pairs instead of `out`
  SlabLayout L;
};

// pinocchio::computeDistance for selected pairs: FK per configuration, then lane l walks the
// selected pairs of ITS configuration (closed forms, or GJK to convergence + EPA).  Not a
// throughput kernel: lanes diverge in GJK/EPA and the EPA polytope sits in local memory.
__global__ void __launch_bounds__(256, 1) distance_kernel(const __grid_constant__ DistArgs a) {
  uint64_t* mbar = reinterpret_cast<uint64_t*>(g_smem);
  unsigned char* tbl = g_smem + 16;
  stage_tables(tbl, a.img, a.L.table_bytes, mbar);
  const Tables t = make_tables(tbl);
  const int nwarps = blockDim.x >> 5, wid = threadIdx.x >> 5;
  float* slab = reinterpret_cast<float*>(tbl + a.L.table_bytes) + static_cast<size_t>(wid) * a.L.words;
  const Warp w = make_warp(t, slab, a.L);
  const bool actuated = a.flags & RAMP_Q_ACTUATED;
  int cols = actuated ? t.h->n_act : t.h->nq;
  if ((a.ld & 3) == 0 && a.ld <= cols + 3) cols = a.ld;
  const int64_t ntiles = (a.n + 31) >> 5;
  for (int64_t tile = static_cast<int64_t>(blockIdx.x) * nwarps + wid; tile < ntiles;
       tile += static_cast<int64_t>(gridDim.x) * nwarps) {
    const int64_t row0 = tile << 5;
    load_rows(actuated ? w.act : w.q, a.q, row0, a.n, cols, a.ld, w.lane);
    if (actuated) expand_q(w);
    fk_lane<0, false>(w);  // placements to the slab, with the accurate sine / cosine: distances are reported
    __syncwarp();
    const int64_t row = row0 + w.lane;
    float dmin = 3.0e38f;
    for (int k = 0; k < a.n_sel; ++k) {
      const DPair pr = t.pairs[a.sel ? a.sel[k] : k];  // sel holds processing indices
      const int col = a.sel ? k : pr.orig;
      if (a.min_out && pr.cls == kClsInactive) continue;
      const DGeom* ga = &t.geoms[pr.a];
      const DGeom* gb = &t.geoms[pr.b];
      const Xf rel = xf_inv_mul(geom_xf(w, ga->gslot, w.lane), geom_xf(w, gb->gslot, w.lane));
      const int cls = pair_class_of(ga->type, gb->type);  // also for inactive pairs
      float d = cls < kClsBoxBox ? closed_form_dist(cls, ga, gb, rel)
                                 : gjk_signed_distance(load_shape(ga, t.verts), load_shape(gb, t.verts), rel,
                                                       swept_radius(ga) + swept_radius(gb));
      if (a.min_out) dmin = fminf(dmin, d);
      else if (row < a.n) a.out[row * a.n_sel + col] = d;
    }
    if (a.min_out && row < a.n) a.min_out[row] = dmin;
    __syncwarp();
  }
}

// ------------------------------------------------------------------ kernel: flag packing
// uint8 flags -> one bit per configuration (bit r & 31 of word r >> 5): what the result gather
// between GPUs ships (1/8 of the bytes).  A warp packs 32 consecutive flags with one ballot.
__global__ void pack_flags_kernel(const uint8_t* __restrict__ hit, int64_t n, uint32_t* __restrict__ words) {
  const int64_t nwords = (n + 31) >> 5;
  for (int64_t w = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; w < nwords;
       w += (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5) {
    const int64_t r = (w << 5) + (threadIdx.x & 31);
    const unsigned b = __ballot_sync(0xffffffffu, r < n && hit[r] != 0);
    if ((threadIdx.x & 31) == 0) words[w] = b;
  }
}

}  // namespace ramp
