// small_batch.cuh -- Cooperative support scans and the one-warp-per-configuration kernel (latency mode).
// Part of the single translation unit ramp_b200.cu (included there, in this order).
#pragma once

namespace ramp {

// Support scan of the few items left in a draining warp, shared between lanes: the k-th active
// lane owns item k; lanes [k * kG, (k + 1) * kG) each scan 1/kG of its vertex list and a
// butterfly picks the best.  Closed-form shapes (no vertex list) are left to their owner.
template <int kG>
__device__ __forceinline__ V3 coop_support(unsigned am, bool active, const GjkShape& S, V3 d, const float4* verts0) {
  const unsigned FULL = 0xffffffffu;
  const int l = threadIdx.x & 31;
  const int grp = l / kG, sub = l % kG;
  const int owner = __fns(am, 0, grp + 1) & 31;  // lane of the grp-th active item (any lane if there is none)
  const bool has = grp < __popc(am);
  const bool listed = S.type != SH_CYLINDER && S.type != SH_ELLIPSOID;
  const int o_cnt = __shfl_sync(FULL, active && listed ? S.vcnt : 0, owner);
  const int o_off = __shfl_sync(FULL, static_cast<int>(S.verts - verts0), owner);
  const V3 od = mk(__shfl_sync(FULL, d.x, owner), __shfl_sync(FULL, d.y, owner), __shfl_sync(FULL, d.z, owner));
  const int per = (o_cnt + kG - 1) / kG;
  const int begin = sub * per, cnt = has ? max(min(per, o_cnt - begin), 0) : 0;
  V3 r = support(SH_CONVEX, 0.f, 0.f, 0.f, verts0 + o_off + begin, cnt, od);
  float val = cnt > 0 ? dot(r, od) : -3.0e38f;
  if constexpr (kG == 32) {
    // whole warp on one item: one integer max-reduction on an order-preserving key, then the
    // winning lane hands out its vertex (5 dependent steps instead of a 20-shuffle butterfly)
    int key = __float_as_int(val);
    key ^= (key >> 31) & 0x7fffffff;
    const int best = __reduce_max_sync(FULL, key);
    const int win = __ffs(__ballot_sync(FULL, key == best)) - 1;
    V3 res = mk(__shfl_sync(FULL, r.x, win), __shfl_sync(FULL, r.y, win), __shfl_sync(FULL, r.z, win));
    if (active && !listed) res = support(S.type, S.p0, S.p1, S.p2, S.verts, S.vcnt, d);
    return res;
  } else {
#pragma unroll
    for (int m = 1; m < kG; m <<= 1) {
      const float ov = __shfl_xor_sync(FULL, val, m);
      const V3 orr = mk(__shfl_xor_sync(FULL, r.x, m), __shfl_xor_sync(FULL, r.y, m), __shfl_xor_sync(FULL, r.z, m));
      if (ov > val) { val = ov; r = orr; }
    }
    // back to the owners: active lane with rank k reads the result of group k
    const int src = (__popc(am & ((1u << l) - 1)) * kG) & 31;
    V3 res = mk(__shfl_sync(FULL, r.x, src), __shfl_sync(FULL, r.y, src), __shfl_sync(FULL, r.z, src));
    if (active && !listed) res = support(S.type, S.p0, S.p1, S.p2, S.verts, S.vcnt, d);
    return res;
  }
}

// ------------------------------------------------- kernel: validity, one warp per configuration
// Latency mode for small batches (a few thousand configurations, the states of one planner
// edge): the tile kernel above needs 32 configurations per warp and tens of thousands of them
// to fill the machine; here every configuration gets its own warp.  All lanes run the (cheap)
// forward kinematics redundantly, lane g places geometry g, lane p tests pair p -- spheres,
// capsules, closed forms / SAT, and GJK in the lane -- and the warp reduces the first hit.
struct WarpSlab {  // per-warp shared memory, in words
  int32_t words, off_q, off_act, off_omi, off_geom, off_mask;
};

struct SmallArgs {
  const unsigned char* img;
  int32_t table_bytes;
  const float* q;
  int64_t n;
  int32_t ld;
  uint32_t flags;
  uint8_t* hit;
  int32_t* first_pair;
  uint32_t* pair_mask;
  WarpSlab L;
};

struct SmallWarp {
  Tables t;
  float *q, *act, *omi, *geom;
  uint32_t* mask;
  int lane;
};

// actuated values (already in w.act) -> Pinocchio q, cooperatively
__device__ __forceinline__ void expand_q_warp(const SmallWarp& w) {
  const ImgHeader* h = w.t.h;
  for (int s = w.lane; s < h->nq; s += 32) w.q[s] = w.t.refq[s];
  __syncwarp();
  for (int k = w.lane; k < h->n_act; k += 32) w.q[w.t.act[k]] = w.act[k];
  __syncwarp();
  if (w.lane == 0)  // in list order: a mimic source may itself be a mimic target
    for (int k = 0; k < h->n_mimic; ++k) {
      const DMimic m = w.t.mimic[k];
      w.q[m.dst] = fmaf(w.q[m.src], m.mult, m.off);
    }
  __syncwarp();
  for (int k = w.lane; k < h->n_cont; k += 32) {
    const int c = w.t.cont[k];
    float sn, cs;
    sincos_small(w.q[c], &sn, &cs);
    w.q[c] = cs;
    w.q[c + 1] = sn;
  }
  __syncwarp();
}

__device__ __forceinline__ Xf small_geom_xf(const SmallWarp& w, int gslot) {
  const float* p = gslot >= 0 ? w.geom + gslot * kStatFloats : w.t.stat + (-1 - gslot) * kStatFloats;
  Xf o;
  o.m[0] = p[0]; o.m[4] = p[1]; o.m[8] = p[2];
  o.m[2] = p[3]; o.m[6] = p[4]; o.m[10] = p[5];
  o.m[3] = p[6]; o.m[7] = p[7]; o.m[11] = p[8];
  o.m[1] = p[9]; o.m[5] = p[10]; o.m[9] = p[11];
  return o;
}

// FK + placements + every pair for ONE configuration whose q is in w.q.  Returns the smallest
// colliding pair index (caller's numbering) or INT32_MAX; sets mask bits when asked.
__device__ __forceinline__ int check_one_config(const SmallWarp& w, bool stop_at_first, bool want_mask) {
  const ImgHeader* h = w.t.h;
  const int l = w.lane;
  const unsigned FULL = 0xffffffffu;
  // every lane walks the whole chain (identical values); lane i % 32 publishes joint i
  {
    Xf cur = xf_identity();
#pragma unroll 1
    for (int i = 0; i < h->n_joints; ++i) {
      const DJoint* j = &w.t.joints[i];
      float qv[7];
      const int n = joint_nq(j->type);
#pragma unroll
      for (int k = 0; k < 7; ++k) qv[k] = k < n ? w.q[j->idx_q + k] : 0.f;
      const Xf loc = local_xf(j, qv);
      const int pc = j->parent_code;
      if (pc == kParentUniverse) cur = loc;
      else if (pc == kParentRunning) cur = xf_mul(cur, loc);
      else {  // parent published earlier (its index is recoverable from the geometry-free table: stored per joint)
        Xf par;
#pragma unroll
        for (int k = 0; k < 12; ++k) par.m[k] = w.omi[j->parent * 12 + k];
        cur = xf_mul(par, loc);
      }
      if ((i & 31) == l) {
#pragma unroll
        for (int k = 0; k < 12; ++k) w.omi[i * 12 + k] = cur.m[k];
      }
      __syncwarp();
    }
  }
  // lane s places geometry s (x, z, t, y layout of the static table)
  for (int s = l; s < h->n_dyn; s += 32) {
    const DGeom* g = &w.t.geoms[w.t.dyn[s]];
    Xf par;
#pragma unroll
    for (int k = 0; k < 12; ++k) par.m[k] = w.omi[g->parent * 12 + k];
    const Xf og = xf_mul(par, load_table_xf(g->pl));
    float* p = w.geom + s * kStatFloats;
    p[0] = og.m[0]; p[1] = og.m[4]; p[2] = og.m[8];
    p[3] = og.m[2]; p[4] = og.m[6]; p[5] = og.m[10];
    p[6] = og.m[3]; p[7] = og.m[7]; p[8] = og.m[11];
    p[9] = og.m[1]; p[10] = og.m[5]; p[11] = og.m[9];
  }
  for (int k = l; k < h->mask_words; k += 32) w.mask[k] = 0u;
  __syncwarp();
  int first = INT32_MAX;
#pragma unroll 1
  for (int base = 0; base < h->n_pairs; base += 32) {
    const int p = base + l;
    bool hit = false, need_gjk = false;
    GjkLane g;
    GjkShape SA, SB;
    SA.type = SB.type = 0; SA.vcnt = SB.vcnt = 0; SA.p0 = SA.p1 = SA.p2 = SB.p0 = SB.p1 = SB.p2 = 0.f; SA.verts = SB.verts = w.t.verts;
    g.s.n = 0; g.it = 0; g.vv = 1.f; g.rsum = 0.f; g.lower = 0.f; g.v = mk(1.f, 0.f, 0.f); g.rel = xf_identity();
    DPair pr;
    pr.cls = kClsInactive; pr.orig = 0;
    if (p < h->n_pairs) pr = w.t.pairs[p];
    if (pr.cls != kClsInactive) {
      const DGeom* ga = &w.t.geoms[pr.a];
      const DGeom* gb = &w.t.geoms[pr.b];
      const Xf A = small_geom_xf(w, pr.sa), B = small_geom_xf(w, pr.sb);
      const V3 dc = trans(A) - trans(B);
      if (dot(dc, dc) <= pr.rs2) {
        {
          const Xf rel = xf_inv_mul(A, B);
          if (pr.cls < kClsBoxBox) hit = closed_form_hit(pr.cls, ga, gb, rel);
          else if (pr.cls == kClsBoxBox) hit = obb_overlap(mk(ga->p0, ga->p1, ga->p2), mk(gb->p0, gb->p1, gb->p2), rel, 0.f);
          else {
            const V3 zb = mk(rel.m[2], rel.m[6], rel.m[10]), tb = trans(rel);
            const V3 wo = seg_seg_witness(mk(0.f, 0.f, -ga->s_out), mk(0.f, 0.f, 2.f * ga->s_out), tb - zb * gb->s_out, zb * (2.f * gb->s_out));
            if (dot(wo, wo) <= pr.rsum_out * pr.rsum_out) {
              const V3 v0 = seg_seg_witness(mk(0.f, 0.f, -ga->s_in), mk(0.f, 0.f, 2.f * ga->s_in), tb - zb * gb->s_in, zb * (2.f * gb->s_in));
              const float rin = ga->r_in + gb->r_in - pr.slack;
              if (rin > 0.f && dot(v0, v0) < rin * rin) hit = true;
              else if (obb_overlap(mk(ga->hx, ga->hy, ga->hz), mk(gb->hx, gb->hy, gb->hz), rel, pr.slack)) {
                gjk_init(g, rel, swept_radius(ga) + swept_radius(gb), v0);
                SA = load_shape(ga, w.t.verts);
                SB = load_shape(gb, w.t.verts);
                need_gjk = true;
              }
            }
          }
        }
      }
    }
    // GJK for the few pairs that got this far, one after the other with the WHOLE warp on each:
    // every lane scans two vertices of a 64-vertex hull and a butterfly picks the support
    // point, the owner lane runs the simplex update (this kernel is about latency, and a
    // configuration rarely has more than one or two such pairs).
    for (unsigned gm = __ballot_sync(FULL, need_gjk); gm; gm &= gm - 1) {
      const int owner = __ffs(gm) - 1;
      const bool mine = l == owner;
      int st = GJK_CONTINUE;
      while (st == GJK_CONTINUE) {
        const V3 sa = coop_support<32>(1u << owner, mine, SA, -g.v, w.t.verts);
        const V3 sb = coop_support<32>(1u << owner, mine, SB, rot_t(g.rel, g.v), w.t.verts);
        float dist;
        if (mine) st = gjk_advance(g, sa, sb, false, dist);
        st = __shfl_sync(FULL, st, owner);
        if (stop_at_first && __any_sync(FULL, hit)) st = GJK_MISS;  // settled already by another pair
      }
      if (mine) hit = hit || st == GJK_HIT;
    }
    if (hit) {
      first = min(first, static_cast<int>(pr.orig));
      if (want_mask) atomicOr(&w.mask[pr.orig >> 5], 1u << (pr.orig & 31));
    }
    // without a first-pair / mask request any hit settles the configuration
    if (stop_at_first && __any_sync(FULL, hit)) break;
  }
  __syncwarp();
  return __reduce_min_sync(FULL, first);
}

__device__ __forceinline__ SmallWarp make_small_warp(const Tables& t, float* slab, const WarpSlab& L) {
  SmallWarp w;
  w.t = t;
  w.q = slab + L.off_q; w.act = slab + L.off_act; w.omi = slab + L.off_omi; w.geom = slab + L.off_geom;
  w.mask = reinterpret_cast<uint32_t*>(slab + L.off_mask);
  w.lane = threadIdx.x & 31;
  return w;
}

// one configuration through the warp: row in, verdict (and first pair / mask) out
__device__ __forceinline__ void check_row(const SmallWarp& w, const SmallArgs& a, int64_t row, int cols, bool actuated, bool stop, bool want_mask) {
  const float* src = a.q + row * a.ld;
  for (int k = w.lane; k < cols; k += 32) (actuated ? w.act : w.q)[k] = __ldg(src + k);
  __syncwarp();
  if (actuated) expand_q_warp(w);
  const int f = check_one_config(w, stop, want_mask);
  if (w.lane == 0) {
    a.hit[row] = f != INT32_MAX;
    if (a.first_pair) a.first_pair[row] = f != INT32_MAX ? f : -1;
  }
  if (want_mask)
    for (int k = w.lane; k < w.t.h->mask_words; k += 32) a.pair_mask[row * w.t.h->mask_words + k] = w.mask[k];
  __syncwarp();
}

__global__ void __launch_bounds__(128, 4) check_warp_kernel(const __grid_constant__ SmallArgs a) {
  uint64_t* mbar = reinterpret_cast<uint64_t*>(g_smem);
  unsigned char* tbl = g_smem + 16;
  stage_tables(tbl, a.img, a.table_bytes, mbar);
  const Tables t = make_tables(tbl);
  const int nwarps = blockDim.x >> 5, wid = threadIdx.x >> 5;
  float* slab = reinterpret_cast<float*>(tbl + a.table_bytes) + static_cast<size_t>(wid) * a.L.words;
  const SmallWarp w = make_small_warp(t, slab, a.L);
  const bool actuated = a.flags & RAMP_Q_ACTUATED;
  const bool want_mask = a.pair_mask != nullptr;
  // early-out only when nobody asked which pair came first
  const bool stop = !(a.flags & RAMP_ALL_PAIRS) && !want_mask && a.first_pair == nullptr;
  const int cols = actuated ? t.h->n_act : t.h->nq;
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * nwarps + wid; row < a.n; row += static_cast<int64_t>(gridDim.x) * nwarps)
    check_row(w, a, row, cols, actuated, stop, want_mask);
}

// The rows of the tiles the batch validity kernel flagged (GjkQueue::tile_flags: a tile whose GJK
// items did not all fit the queue -- a robot that produces more items than the queue was sized
// for, or a call without a queue), redone from scratch with one warp per configuration.  Runs
// between the tile kernel and the queue kernel; when nothing was flagged (count[3] == 0, the
// normal case) every CTA leaves before it has staged anything.
__global__ void __launch_bounds__(128, 4) recheck_kernel(const __grid_constant__ SmallArgs a, const uint8_t* __restrict__ tile_flags,
                                                         const uint32_t* __restrict__ count) {
  if (count[3] == 0u) return;
  uint64_t* mbar = reinterpret_cast<uint64_t*>(g_smem);
  unsigned char* tbl = g_smem + 16;
  stage_tables(tbl, a.img, a.table_bytes, mbar);
  const Tables t = make_tables(tbl);
  const int nwarps = blockDim.x >> 5, wid = threadIdx.x >> 5;
  float* slab = reinterpret_cast<float*>(tbl + a.table_bytes) + static_cast<size_t>(wid) * a.L.words;
  const SmallWarp w = make_small_warp(t, slab, a.L);
  const bool actuated = a.flags & RAMP_Q_ACTUATED;
  const bool want_mask = a.pair_mask != nullptr;
  const bool stop = !(a.flags & RAMP_ALL_PAIRS) && !want_mask && a.first_pair == nullptr;
  const int cols = actuated ? t.h->n_act : t.h->nq;
  // a warp takes a quarter of a flagged tile (8 rows): flagged tiles come in runs
  const int64_t nparts = ((a.n + 31) >> 5) * 4;
  for (int64_t part = static_cast<int64_t>(blockIdx.x) * nwarps + wid; part < nparts; part += static_cast<int64_t>(gridDim.x) * nwarps) {
    if (!tile_flags[part >> 2]) continue;
    const int64_t r0 = part * 8, r1 = imin64(a.n, r0 + 8);
    for (int64_t row = r0; row < r1; ++row) check_row(w, a, row, cols, actuated, stop, want_mask);
  }
}

}  // namespace ramp
