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
    if (a.min_out) {
      // Smallest distance only: a pair whose OUTER capsules (or bounding spheres) are farther apart
      // than the smallest distance known so far cannot lower it and skips GJK / EPA.  A first pass
      // over the pairs gives that bound a good start: closed-form pairs exactly, the others by
      // their INNER capsules, which lie inside the shapes and are therefore at least as far apart.
      // The pair that set the bound always passes its own test (inner >= exact >= outer), so the
      // minimum of the distances that are evaluated is the minimum over all pairs.
      float bound = 3.0e38f;
      auto pair_at = [&](int k) { return t.pairs[a.sel ? a.sel[k] : k]; };
      auto rel_of = [&](const DGeom* ga, const DGeom* gb) { return xf_inv_mul(geom_xf(w, ga->gslot, w.lane), geom_xf(w, gb->gslot, w.lane)); };
      // distance of the two capsules (half length s, radius r) around the local z axes of a pair
      auto capsules = [&](const Xf& rel, float sa, float sb, float rsum) {
        const V3 zb = mk(rel.m[2], rel.m[6], rel.m[10]), tb = trans(rel);
        return seg_seg_dist(mk(0.f, 0.f, -sa), mk(0.f, 0.f, 2.f * sa), tb - zb * sb, zb * (2.f * sb)) - rsum;
      };
      auto lower_of = [&](const DPair& pr, const DGeom* ga, const DGeom* gb, const Xf& rel) {
        const V3 tb = trans(rel);
        float lower = sqrtf(dot(tb, tb)) - ga->bound_r - gb->bound_r;
        if (pr.cls == kClsGjk) lower = fmaxf(lower, capsules(rel, ga->s_out, gb->s_out, ga->r_out + gb->r_out));
        return lower - pr.slack;  // (slack: the pair's share of non-rigid placements)
      };
      for (int k = 0; k < a.n_sel; ++k) {
        const DPair pr = pair_at(k);
        if (pr.cls == kClsInactive) continue;
        const DGeom* ga = &t.geoms[pr.a];
        const DGeom* gb = &t.geoms[pr.b];
        const Xf rel = rel_of(ga, gb);
        if (pr.cls < kClsBoxBox) dmin = fminf(dmin, closed_form_dist(pr.cls, ga, gb, rel));
        else if (pr.cls == kClsGjk) bound = fminf(bound, capsules(rel, ga->s_in, gb->s_in, ga->r_in + gb->r_in) + pr.slack);
      }
      bound = fminf(bound, dmin) + 2e-5f;
      // The survivors, 32 pairs at a time: every lane first marks the pairs of ITS configuration
      // that pass the bound, then the lanes work through their marks together -- each on a pair of
      // its own, so one trip of the loop is one GJK per lane, not one pair index for the warp.
      for (int base = 0; base < a.n_sel; base += 32) {
        unsigned todo = 0u;
        const int end = min(a.n_sel, base + 32);
        for (int k = base; k < end; ++k) {
          const DPair pr = pair_at(k);
          if (pr.cls < kClsBoxBox) continue;  // inactive, or a closed form (done above)
          const DGeom* ga = &t.geoms[pr.a];
          const DGeom* gb = &t.geoms[pr.b];
          if (lower_of(pr, ga, gb, rel_of(ga, gb)) <= bound) todo |= 1u << (k - base);
        }
        while (__any_sync(0xffffffffu, todo != 0u)) {
          if (todo != 0u) {
            const DPair pr = pair_at(base + __ffs(todo) - 1);
            todo &= todo - 1u;
            const DGeom* ga = &t.geoms[pr.a];
            const DGeom* gb = &t.geoms[pr.b];
            const Xf rel = rel_of(ga, gb);
            if (lower_of(pr, ga, gb, rel) <= bound) {  // (the bound may have dropped since the pair was marked)
              const float d = gjk_signed_distance(load_shape(ga, t.verts), load_shape(gb, t.verts), rel, swept_radius(ga) + swept_radius(gb));
              dmin = fminf(dmin, d);
              bound = fminf(bound, d + 2e-5f);
            }
          }
        }
      }
    } else {
      for (int k = 0; k < a.n_sel; ++k) {
        const DPair pr = t.pairs[a.sel ? a.sel[k] : k];  // sel holds processing indices
        const int col = a.sel ? k : pr.orig;
        const DGeom* ga = &t.geoms[pr.a];
        const DGeom* gb = &t.geoms[pr.b];
        const Xf rel = xf_inv_mul(geom_xf(w, ga->gslot, w.lane), geom_xf(w, gb->gslot, w.lane));
        const int cls = pair_class_of(ga->type, gb->type);  // also for inactive pairs
        const float d = cls < kClsBoxBox ? closed_form_dist(cls, ga, gb, rel)
                                         : gjk_signed_distance(load_shape(ga, t.verts), load_shape(gb, t.verts), rel,
                                                               swept_radius(ga) + swept_radius(gb));
        if (row < a.n) a.out[row * a.n_sel + col] = d;
      }
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
