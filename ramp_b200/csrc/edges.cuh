// edges.cuh -- Motion validity: edge discretisation kernels (OMPL DiscreteMotionValidator).
// Part of the single translation unit ramp_b200.cu (included there, in this order).
#pragma once

namespace ramp {

// ------------------------------------------------------------------ kernels: edges
// OMPL state-space maths in FP32 (reference: OMPL RealVector/SO2/SO3 spaces).
#define RAMP_PI_F 3.14159265358979323846f

__device__ __forceinline__ float so2_dist(float a, float b) {
  const float d = fabsf(a - b);
  return d > RAMP_PI_F ? 2.f * RAMP_PI_F - d : d;
}

__device__ __forceinline__ float so2_interp(float from, float to, float t) {
  float diff = to - from, v;
  if (fabsf(diff) <= RAMP_PI_F) v = fmaf(diff, t, from);
  else {
    diff = diff > 0.f ? 2.f * RAMP_PI_F - diff : -2.f * RAMP_PI_F - diff;
    v = from - diff * t;
    if (v > RAMP_PI_F) v -= 2.f * RAMP_PI_F;
    else if (v < -RAMP_PI_F) v += 2.f * RAMP_PI_F;
  }
  return v;
}

__device__ __forceinline__ float so3_arc(const float* a, const float* b, int st) {
  const float dq = fabsf(a[0] * b[0] + a[st] * b[st] + a[2 * st] * b[2 * st] + a[3 * st] * b[3 * st]);
  return dq > 1.f - 1e-7f ? 0.f : acosf(dq);
}

__device__ __forceinline__ int seg_count(float dist, float inv_lvs) { return static_cast<int>(ceilf(dist * inv_lvs)); }

// validSegmentCount of one edge; q0/q1 are the edge's rows in global memory
__device__ __forceinline__ int edge_nd(const DSpace& sp, const float* __restrict__ q0, const float* __restrict__ q1) {
  int nd = 0;
  for (int i = 0; i < sp.n_sub; ++i) {
    const DSub& s = sp.sub[i];
    const float* a = q0 + s.col;
    const float* b = q1 + s.col;
    int c = 0;
    const int nreal = s.type == RAMP_SPACE_REAL ? s.dim : (s.type == RAMP_SPACE_SE2 ? 2 : (s.type == RAMP_SPACE_SE3 ? 3 : 0));
    if (nreal) {
      float e = 0.f;
      for (int k = 0; k < nreal; ++k) { const float d = a[k] - b[k]; e = fmaf(d, d, e); }
      c = seg_count(sqrtf(e), s.inv_lvs_real);
    }
    if (s.type == RAMP_SPACE_SO2) c = seg_count(so2_dist(a[0], b[0]), s.inv_lvs_rot);
    else if (s.type == RAMP_SPACE_SE2) c = max(c, seg_count(so2_dist(a[2], b[2]), s.inv_lvs_rot));
    else if (s.type == RAMP_SPACE_SE3) c = max(c, seg_count(so3_arc(a + 3, b + 3, 1), s.inv_lvs_rot));
    else if (s.type == RAMP_SPACE_SO3) c = seg_count(so3_arc(a, b, 1), s.inv_lvs_rot);
    nd = max(nd, c);
  }
  return nd;
}

struct EdgeArgs {
  const unsigned char* img;
  const float* q0;
  const float* q1;
  int64_t n_edges;
  int32_t ld;
  uint8_t* valid;
  int32_t* first_invalid;  // always non-null inside the library (scratch when the caller passes none)
  int32_t* nd;             // idem
  int64_t* offs;           // [n_edges+1] exclusive prefix of per-edge state counts (scratch)
  GjkQueue gq;
  SlabLayout L;
  DSpace sp;
};

// pass 1: nd per edge and the number of states to check per edge (nd interior + end states, >= 1),
// scanned within the block of 256 edges: offs[e + 1] = inclusive prefix inside the block,
// blk[block] = the block's total
__global__ void __launch_bounds__(256) edge_count_kernel(const __grid_constant__ EdgeArgs a, int64_t* __restrict__ blk) {
  __shared__ int64_t wsum[8];
  const int64_t e = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x;
  const int l = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int64_t v = 0;
  if (e < a.n_edges) {
    const int nd = edge_nd(a.sp, a.q0 + e * a.ld, a.q1 + e * a.ld);
    a.nd[e] = nd;
    a.first_invalid[e] = INT32_MAX;
    v = max(nd, 1);
  }
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int64_t u = __shfl_up_sync(0xffffffffu, v, d);
    if (l >= d) v += u;
  }
  if (l == 31) wsum[wid] = v;
  __syncthreads();
  int64_t before = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) before += k < wid ? wsum[k] : 0;
  v += before;
  if (e < a.n_edges) a.offs[e + 1] = v;
  if (threadIdx.x == 255) blk[blockIdx.x] = v;
}

// pass 2a: exclusive scan of the block totals by one CTA, 1024 at a time with a running carry
// (a slice of 2^18 edges has 1024 blocks)
__global__ void __launch_bounds__(1024) edge_scan_kernel(int64_t* __restrict__ blk, int64_t nb, int64_t* __restrict__ offs) {
  __shared__ int64_t wsum[32];
  __shared__ int64_t carry_s;
  const int t = threadIdx.x, l = t & 31, wid = t >> 5;
  if (t == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < nb; base += 1024) {
    const int64_t i = base + t;
    const int64_t own = i < nb ? blk[i] : 0;
    int64_t v = own;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int64_t u = __shfl_up_sync(0xffffffffu, v, d);
      if (l >= d) v += u;
    }
    if (l == 31) wsum[wid] = v;
    __syncthreads();
    if (wid == 0) {
      int64_t ws = wsum[l];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int64_t u = __shfl_up_sync(0xffffffffu, ws, d);
        if (l >= d) ws += u;
      }
      wsum[l] = ws;  // inclusive over the warps
    }
    __syncthreads();
    const int64_t carry = carry_s;
    const int64_t incl = v + (wid ? wsum[wid - 1] : 0) + carry;
    if (i < nb) blk[i] = incl - own;  // exclusive
    __syncthreads();
    if (t == 1023) carry_s = incl;
    __syncthreads();
  }
  // more states than a 32-bit flat index can name: offs[0] = -1 tells the check kernel to do
  // nothing and the finish kernel to report every edge of the slice as unchecked (-2)
  if (t == 0) offs[0] = carry_s > 0xffffffffll ? -1 : 0;
}

// pass 2b: the blocks' offsets added to the prefixes inside the blocks
__global__ void __launch_bounds__(256) edge_offsets_kernel(int64_t* __restrict__ offs, const int64_t* __restrict__ blk, int64_t n) {
  const int64_t e = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x;
  if (blockIdx.x > 0 && e < n) offs[e + 1] += blk[blockIdx.x];
}

// pass 3: every warp tile is 32 consecutive (edge, j) states of the flattened list
__global__ void __launch_bounds__(512, 1) edge_check_kernel(const __grid_constant__ EdgeArgs a) {
  uint64_t* mbar = reinterpret_cast<uint64_t*>(g_smem);
  unsigned char* tbl = g_smem + 16;
  stage_tables(tbl, a.img, a.L.table_bytes, mbar);
  Tables t = make_tables(tbl);
  t.verts = reinterpret_cast<const float4*>(a.img + t.h->off_verts);  // not staged: global memory (overflow path only)
  const int nwarps = blockDim.x >> 5, wid = threadIdx.x >> 5;
  float* slab = reinterpret_cast<float*>(tbl + a.L.table_bytes) + static_cast<size_t>(wid) * a.L.words;
  const Warp w = make_warp(t, slab, a.L);
  const int l = w.lane;
  const int64_t total = a.offs[0] < 0 ? 0 : a.offs[a.n_edges];
  const int64_t ntiles = (total + 31) >> 5;
  for (int64_t tile = static_cast<int64_t>(blockIdx.x) * nwarps + wid; tile < ntiles;
       tile += static_cast<int64_t>(gridDim.x) * nwarps) {
    const int64_t s = (tile << 5) + l;
    const bool live = s < total;
    // edge of state s: last e with offs[e] <= s
    int64_t lo = 0, hi = a.n_edges - 1;
    const int64_t ss = live ? s : total - 1;
    while (lo < hi) {
      const int64_t mid = (lo + hi + 1) >> 1;
      if (a.offs[mid] <= ss) lo = mid; else hi = mid - 1;
    }
    const int64_t e = lo;
    const int nd = a.nd[e];
    const int j = static_cast<int>(ss - a.offs[e]) + 1;  // 1..max(nd,1); j >= nd means q1 itself
    const float* r0 = a.q0 + e * a.ld;
    const float* r1 = a.q1 + e * a.ld;
    const bool is_end = j >= nd;
    const float tt = is_end ? 1.f : static_cast<float>(j) / static_cast<float>(nd);
    // an edge already known to be invalid at a smaller j needs no further states
    bool skip = !live || a.first_invalid[e] < min(j, nd);
    bool in_bounds = true;
    for (int i = 0; i < a.sp.n_sub; ++i) {
      const DSub& sb = a.sp.sub[i];
      const float* p0 = r0 + sb.col;
      const float* p1 = r1 + sb.col;
      float* o = w.act + sb.col * 32 + l;
      const int nreal = sb.type == RAMP_SPACE_REAL ? sb.dim : (sb.type == RAMP_SPACE_SE2 ? 2 : (sb.type == RAMP_SPACE_SE3 ? 3 : 0));
      for (int k = 0; k < nreal; ++k) {
        const float v = is_end ? p1[k] : fmaf(p1[k] - p0[k], tt, p0[k]);
        o[k * 32] = v;
        in_bounds = in_bounds && v <= sb.hi[k] && v >= sb.lo[k];
      }
      if (sb.type == RAMP_SPACE_SO2 || sb.type == RAMP_SPACE_SE2) {
        const int k = sb.type == RAMP_SPACE_SE2 ? 2 : 0;
        const float v = is_end ? p1[k] : so2_interp(p0[k], p1[k], tt);
        o[k * 32] = v;
        in_bounds = in_bounds && v <= RAMP_PI_F && v >= -RAMP_PI_F;
      } else if (sb.type == RAMP_SPACE_SE3 || sb.type == RAMP_SPACE_SO3) {
        const int k0 = sb.type == RAMP_SPACE_SE3 ? 3 : 0;
        const float* qa = p0 + k0;
        const float* qb = p1 + k0;
        float out[4] = {qb[0], qb[1], qb[2], qb[3]};
        if (!is_end) {
          const float th = so3_arc(qa, qb, 1);
          if (th > 1e-7f) {
            const float d = 1.f / sinf(th), s0 = sinf((1.f - tt) * th);
            float s1 = sinf(tt * th);
            if (qa[0] * qb[0] + qa[1] * qb[1] + qa[2] * qb[2] + qa[3] * qb[3] < 0.f) s1 = -s1;
            for (int k = 0; k < 4; ++k) out[k] = (qa[k] * s0 + qb[k] * s1) * d;
          } else {
            for (int k = 0; k < 4; ++k) out[k] = qa[k];
          }
        }
        for (int k = 0; k < 4; ++k) o[(k0 + k) * 32] = out[k];
      }
    }
    __syncwarp();
    expand_q(w);
    reset_results(w, 0);
    fk_lane<0>(w);
    __syncwarp();
    narrow_phase(w, true, true, false, a.gq, static_cast<uint32_t>(tile << 5), static_cast<int>(imin64(32, total - (tile << 5))));
    const bool invalid = w.first[l] != INT32_MAX || (a.sp.check_bounds && !in_bounds);
    if (!skip && invalid) atomicMin(&a.first_invalid[e], min(j, nd));
    __syncwarp();
  }
}

// ---- latency mode for a handful of edges (one planner query calls checkMotion edge by
// edge): ONE launch, one warp per interpolated state.  Every warp recomputes the (tiny) segment
// counts of all edges, picks its states, and the last warp to finish folds the per-edge minima
// into the outputs.  `work` = 33 ints pre-set to 0x7f7f7f7f: [e] smallest invalid j, [32] a
// counter of finished states.
constexpr int kSmallEdges = 32;
constexpr int32_t kWorkInit = 0x7f7f7f7f;

struct EdgeSmallArgs {
  const unsigned char* img;
  int32_t table_bytes;
  const float* q0;
  const float* q1;
  int32_t n_edges;
  int32_t ld;
  uint8_t* valid;
  int32_t* first_invalid;  // nullable
  int32_t* nd;             // nullable
  int32_t* work;
  WarpSlab L;
  DSpace sp;
};

__global__ void __launch_bounds__(128) edge_warp_kernel(const __grid_constant__ EdgeSmallArgs a) {
  uint64_t* mbar = reinterpret_cast<uint64_t*>(g_smem);
  unsigned char* tbl = g_smem + 16;
  stage_tables(tbl, a.img, a.table_bytes, mbar);
  const Tables t = make_tables(tbl);
  const int nwarps = blockDim.x >> 5, wid = threadIdx.x >> 5;
  float* slab = reinterpret_cast<float*>(tbl + a.table_bytes) + static_cast<size_t>(wid) * a.L.words;
  const SmallWarp w = make_small_warp(t, slab, a.L);
  const int l = w.lane;
  const unsigned FULL = 0xffffffffu;
  // lane e: segment count of edge e, then an inclusive scan of the states per edge
  const int my_nd = l < a.n_edges ? edge_nd(a.sp, a.q0 + static_cast<int64_t>(l) * a.ld, a.q1 + static_cast<int64_t>(l) * a.ld) : 0;
  int incl = l < a.n_edges ? max(my_nd, 1) : 0;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_up_sync(FULL, incl, d);
    if (l >= d) incl += v;
  }
  const int total = __shfl_sync(FULL, incl, 31);
  const int gw = blockIdx.x * nwarps + wid, gstep = gridDim.x * nwarps;
  int done = 0;
  for (int s = gw; s < total; s += gstep, ++done) {
    const int e = __popc(__ballot_sync(FULL, l < a.n_edges && incl <= s));
    const int excl = __shfl_sync(FULL, incl, e) - max(__shfl_sync(FULL, my_nd, e), 1);
    const int nd = __shfl_sync(FULL, my_nd, e);
    const int j = s - excl + 1;  // 1 .. max(nd, 1); j >= nd is the end state itself
    const int jj = min(j, nd);
    if (*reinterpret_cast<volatile int32_t*>(a.work + e) < jj) continue;  // an earlier state already failed
    const float* r0 = a.q0 + static_cast<int64_t>(e) * a.ld;
    const float* r1 = a.q1 + static_cast<int64_t>(e) * a.ld;
    const bool is_end = j >= nd;
    const float tt = is_end ? 1.f : static_cast<float>(j) / static_cast<float>(nd);
    bool in_bounds = true;
    for (int i = l; i < a.sp.n_sub; i += 32) {  // lane i interpolates subspace i
      const DSub& sb = a.sp.sub[i];
      const float* p0 = r0 + sb.col;
      const float* p1 = r1 + sb.col;
      float* o = w.act + sb.col;
      const int nreal = sb.type == RAMP_SPACE_REAL ? sb.dim : (sb.type == RAMP_SPACE_SE2 ? 2 : (sb.type == RAMP_SPACE_SE3 ? 3 : 0));
      for (int k = 0; k < nreal; ++k) {
        const float v = is_end ? p1[k] : fmaf(p1[k] - p0[k], tt, p0[k]);
        o[k] = v;
        in_bounds = in_bounds && v <= sb.hi[k] && v >= sb.lo[k];
      }
      if (sb.type == RAMP_SPACE_SO2 || sb.type == RAMP_SPACE_SE2) {
        const int k = sb.type == RAMP_SPACE_SE2 ? 2 : 0;
        const float v = is_end ? p1[k] : so2_interp(p0[k], p1[k], tt);
        o[k] = v;
        in_bounds = in_bounds && v <= RAMP_PI_F && v >= -RAMP_PI_F;
      } else if (sb.type == RAMP_SPACE_SE3 || sb.type == RAMP_SPACE_SO3) {
        const int k0 = sb.type == RAMP_SPACE_SE3 ? 3 : 0;
        const float* qa = p0 + k0;
        const float* qb = p1 + k0;
        float out[4] = {qb[0], qb[1], qb[2], qb[3]};
        if (!is_end) {
          const float th = so3_arc(qa, qb, 1);
          if (th > 1e-7f) {
            const float dn = 1.f / sinf(th), s0 = sinf((1.f - tt) * th);
            float s1 = sinf(tt * th);
            if (qa[0] * qb[0] + qa[1] * qb[1] + qa[2] * qb[2] + qa[3] * qb[3] < 0.f) s1 = -s1;
            for (int k = 0; k < 4; ++k) out[k] = (qa[k] * s0 + qb[k] * s1) * dn;
          } else {
            for (int k = 0; k < 4; ++k) out[k] = qa[k];
          }
        }
        for (int k = 0; k < 4; ++k) o[k0 + k] = out[k];
      }
    }
    in_bounds = __all_sync(FULL, in_bounds);
    __syncwarp();
    expand_q_warp(w);
    bool invalid = a.sp.check_bounds && !in_bounds;
    if (!invalid) invalid = check_one_config(w, true, false) != INT32_MAX;
    if (invalid && l == 0) atomicMin(a.work + e, jj);
    __syncwarp();
  }
  // the last warp to hand in its states folds the minima into the outputs
  __threadfence();
  unsigned fin = 0;
  if (l == 0) fin = atomicAdd(reinterpret_cast<unsigned*>(a.work + kSmallEdges), static_cast<unsigned>(done)) + done;
  fin = __shfl_sync(FULL, fin, 0);
  if (fin == static_cast<unsigned>(kWorkInit) + static_cast<unsigned>(total) && done > 0 || (total == 0 && gw == 0)) {
    __threadfence();
    if (l < a.n_edges) {
      const int32_t f = *reinterpret_cast<volatile int32_t*>(a.work + l);
      a.valid[l] = f == kWorkInit;
      if (a.first_invalid) a.first_invalid[l] = f == kWorkInit ? -1 : f;
      if (a.nd) a.nd[l] = my_nd;
    }
  }
}

// pass 4: fold the per-edge minimum into the caller's outputs
__global__ void edge_finish_kernel(const __grid_constant__ EdgeArgs a, int32_t* out_first, int32_t* out_nd) {
  const int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= a.n_edges) return;
  const bool unchecked = a.offs[0] < 0;  // the slice held more than 2^32 - 1 states
  const int f = unchecked ? -2 : a.first_invalid[e];
  a.valid[e] = f == INT32_MAX;
  if (out_first) out_first[e] = f == INT32_MAX ? -1 : f;
  if (out_nd && out_nd != a.nd) out_nd[e] = a.nd[e];
}

}  // namespace ramp
