// narrow_phase.cuh -- Narrow phase of a tile (stages 1 / 1.5 / 2, GJK items to the global queue) and check_kernel.
// Part of the single translation unit ramp_b200.cu (included there, in this order).
#pragma once

namespace ramp {

// ---- narrow phase over the pair list for the 32 configurations of the tile.
// A (configuration, pair) work item that survived every cull and needs GJK.  The validity
// kernel appends such items to a global queue; gjk_queue_kernel drains it with full warps and
// far more resident warps than the slab-bound validity kernel can have.
struct __align__(16) GjkItem {
  float rel[12];    // pose of shape b in shape a's frame
  uint32_t dir[2];  // first search direction for GJK -- the witness vector of the two inner capsules,
                    // normalised -- as three halves (only its direction is ever used)
  uint32_t row;     // configuration index inside the launch's slice / flat state index (edges)
  uint32_t p;       // pair index (processing order)
};
static_assert(sizeof(GjkItem) == 64, "GjkItem");

__device__ __forceinline__ void pack_dir(V3 v, uint32_t& a, uint32_t& b) {
  const float vv = dot(v, v);
  v = vv > 1e-24f ? v * rsqrtf(vv) : mk(1.f, 0.f, 0.f);
  const __half2 xy = __floats2half2_rn(v.x, v.y), z0 = __floats2half2_rn(v.z, 0.f);
  a = *reinterpret_cast<const uint32_t*>(&xy);
  b = *reinterpret_cast<const uint32_t*>(&z0);
}

__device__ __forceinline__ V3 unpack_dir(uint32_t a, uint32_t b) {
  const float2 xy = __half22float2(*reinterpret_cast<const __half2*>(&a)), z0 = __half22float2(*reinterpret_cast<const __half2*>(&b));
  return mk(xy.x, xy.y, z0.x);
}

// The queue has two regions.  Items whose inner capsules are still far apart compared with the
// shell between inner and outer capsule almost always end after one GJK iteration ("easy");
// the others take 4-5 ("hard").  Hard items go to the front region and are handed out first,
// so the consumer kernel drains on one-iteration items and its warps finish together.
struct GjkQueue {
  GjkItem* items;     // nullptr: no queue, GJK runs inside the validity kernel
  uint32_t* count;    // [0] hard, [1] easy items reserved so far (may exceed the regions: see tile_flags),
                      // [2] the consumer's cursor, [3] flagged tiles; zeroed by the host
  uint32_t capacity;  // items: hard region [0, cap_hard), easy region [cap_hard, capacity); cap_hard = a quarter
  uint32_t cap_hard;
  // One byte per tile of 32 rows of the result arrays the queue serves (zeroed by the host).  The
  // batch kernel keeps no GJK loop of its own: a tile that could not queue all its items sets its
  // byte (and bumps count[3]), and recheck_kernel redoes the rows of such tiles from scratch.
  uint8_t* tile_flags;
};

// (a quarter: the H1 stand-in produces 0.05 hard and 2.02 easy items per configuration; with a third
// for the hard region its easy region, 2 items per row, overflowed by half a per cent on every call.
// The hard items of the arms stay below 0.5 per row)
__host__ __device__ constexpr uint32_t queue_cap_hard(uint32_t capacity) { return capacity / 4; }
// hard: |witness of the inner capsule axes| below inner radii + 7/8 of the shell
constexpr float kHardShell = 0.875f;

__device__ __forceinline__ void record_hit(const Warp& w, int c, int p, bool want_mask) {
  atomicMin(&w.first[c], p);
  if (want_mask) atomicOr(&w.mask[(p >> 5) * 32 + c], 1u << (p & 31));
}

__device__ __forceinline__ GjkShape load_shape(const DGeom* g, const float4* verts) {
  GjkShape s;
  s.type = g->type; s.vcnt = g->vcnt; s.p0 = g->p0; s.p1 = g->p1; s.p2 = g->p2; s.verts = verts + g->voff;
  return s;
}

__device__ __forceinline__ float swept_radius(const DGeom* g) {
  return (g->type == SH_SPHERE || g->type == SH_CAPSULE) ? g->p0 : 0.f;
}

__device__ __forceinline__ int pair_class_of(int ta, int tb) {  // (sphere classes are stored sphere first)
  if (ta != SH_SPHERE && ta > tb) { const int t = ta; ta = tb; tb = t; }
  if (ta == SH_SPHERE && tb <= SH_CYLINDER) return kClsSphereSphere + tb;
  if (ta == SH_CAPSULE && tb == SH_CAPSULE) return kClsCapsuleCapsule;
  if (ta == SH_BOX && tb == SH_BOX) return kClsBoxBox;
  return kClsGjk;
}

// exact closed-form signed distance of the primitive classes; rel = pose of b in a's frame
__device__ __forceinline__ float closed_form_dist(int cls, const DGeom* ga, const DGeom* gb, const Xf& rel) {
  switch (cls) {
    case kClsSphereSphere: return sqrtf(dot(trans(rel), trans(rel))) - ga->p0 - gb->p0;
    case kClsSphereCapsule: {  // a sphere, b capsule: sphere centre in b's frame
      const V3 c = rot_t(rel, -trans(rel));
      return point_zseg_dist(c, gb->p1) - ga->p0 - gb->p0;
    }
    case kClsSphereBox: {
      const V3 c = rot_t(rel, -trans(rel));
      return sphere_box_dist(c, gb->p0, gb->p1, gb->p2, ga->p0);
    }
    case kClsSphereCylinder: {
      const V3 c = rot_t(rel, -trans(rel));
      return sphere_cyl_dist(c, gb->p0, gb->p1, ga->p0);
    }
    case kClsCapsuleCapsule: {
      const V3 zb = mk(rel.m[2], rel.m[6], rel.m[10]);
      const V3 p2 = trans(rel) - zb * gb->p1;
      return seg_seg_dist(mk(0.f, 0.f, -ga->p1), mk(0.f, 0.f, 2.f * ga->p1), p2, zb * (2.f * gb->p1)) - ga->p0 - gb->p0;
    }
    default: return 3.0e38f;
  }
}

__device__ __forceinline__ bool closed_form_hit(int cls, const DGeom* ga, const DGeom* gb, const Xf& rel) {
  return closed_form_dist(cls, ga, gb, rel) <= 0.f;
}

// Early out (`stop_at_first`): pair p of a configuration is dropped once a pair with a smaller
// list index has hit (the first-pair output needs nothing else), or -- `stop_any`, when only
// the flag is wanted -- once any pair has hit.
// `tw` / kTeam: this warp takes every kTeam-th chunk of `chunk_pairs` pairs (kTeam == 1: all of
// them, in chunks of kChunkPairs -- a compile-time constant on the common path)
// kStage3: items that do not fit the queue are solved here, in the kernel (the planner-edge kernel).
// The batch validity kernel passes false: it flags the tile instead (GjkQueue::tile_flags) and
// carries neither the GJK loop's ~1500 instructions nor its 22 registers.
template <int kTeam = 1, bool kStage3 = true>
__device__ __forceinline__ void narrow_phase(const Warp& w, bool stop_at_first, bool stop_any, bool want_mask,
                                             const GjkQueue& gq, uint32_t row_base, int live_rows, int tw = 0,
                                             int chunk_pairs_rt = kChunkPairs) {
  const int chunk_pairs = kTeam == 1 ? kChunkPairs : chunk_pairs_rt;
  const ImgHeader* h = w.t.h;
  const int l = w.lane;
  const unsigned FULL = 0xffffffffu;
  // decided already?  (first = smallest list index that hit so far, INT32_MAX if none)
  auto dead = [&](int first, int orig) { return stop_at_first && first < (stop_any ? INT32_MAX : orig); };
  // kTeam == 1: the chunks in order.  Teams: every warp takes the next chunk from a counter in
  // the slab (the word behind the prefetch mbarrier, reset with the results), so a warp that
  // drew cheap chunks takes more of them -- chunk costs differ by the share of pairs that
  // survive to stage 2.
  int* next_chunk = reinterpret_cast<int*>(w.save) - 2;
  auto take_chunk = [&]() {
    int c = 0;
    if (l == 0) c = atomicAdd(next_chunk, chunk_pairs);
    return __shfl_sync(FULL, c, 0);
  };
  bool over = false;  // (!kStage3) an item of this tile found no room in the queue
  for (int chunk = kTeam == 1 ? 0 : take_chunk(); chunk < h->n_pairs; chunk = kTeam == 1 ? chunk + chunk_pairs : take_chunk()) {
    const int chunk_end = min(chunk + chunk_pairs, h->n_pairs);
    // Lanes past the end of the batch hold a replayed row (load_rows): they must never produce
    // work, or their results and queued GJK items would land behind the caller's arrays.  A
    // "first hit" of INT32_MIN fails the test in push() in every mode, at no cost in the loop.
    const int my_first = l < live_rows ? w.first[l] : INT32_MIN;
    int qn = 0;
    // ---- stage 1: bounding spheres; lane = configuration, pair index uniform.  The first 16
    // bytes of a DPair are all this stage reads (one vector load): two byte offsets of the
    // placement origins, the squared radius sum (-1: inactive pair, never passes) and the
    // caller's pair index for the early-out test.
    const char* T = reinterpret_cast<const char*>(g_smem) + 4 * (w.geom_w + l + 6 * 32);  // origin rows of this lane's slab column
    const char* TS = reinterpret_cast<const char*>(g_smem) + 4 * (w.stat_w + 6);          // origins of the static placements
    const int4* rec = reinterpret_cast<const int4*>(w.t.pairs);  // 3 per DPair
    const unsigned lt = (1u << l) - 1;
    // dead(my_first, orig) as one LOP3 + compare: pass needs my_first >= (orig & lim_and) | lim_or
    const int lim_and = (stop_at_first && !stop_any) ? 0xffff : 0;
    const int lim_or = stop_at_first ? (stop_any ? INT32_MAX : 0) : INT32_MIN + 1;
    const int item0 = l - (chunk << 5);
    auto push = [&](int p, const int4& r, bool pass) {
      pass = pass && my_first >= ((r.w & lim_and) | lim_or);
      const unsigned b = __ballot_sync(FULL, pass);
      if (pass) w.queue[qn + __popc(b & lt)] = static_cast<uint16_t>((p << 5) + item0);
      qn += __popc(b);
    };
    // squared distance of the two placement origins against the bounding-sphere sum
    auto near = [&](const int4& r, const char* base_a, const char* base_b, int stride_a, int stride_b) {
      const float* a = reinterpret_cast<const float*>(base_a + r.x);
      const float* b = reinterpret_cast<const float*>(base_b + r.y);
      const float dx = a[0] - b[0], dy = a[stride_a] - b[stride_b], dz = a[2 * stride_a] - b[2 * stride_b];
      return fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= __int_as_float(r.z);
    };
    const int dd_end = min(chunk_end, max(chunk, h->n_dd));
    // four pairs per trip: their loads are independent, which hides shared-memory latency
    // when few warps fit next to the slabs (20-geometry robots: 6 warps per SM)
#pragma unroll 1
    for (int p = chunk; p < dd_end; p += 4) {
      const int last = dd_end - 1;
      const int4 r0 = rec[3 * p], r1 = rec[3 * min(p + 1, last)], r2 = rec[3 * min(p + 2, last)], r3 = rec[3 * min(p + 3, last)];
      const bool k0 = near(r0, T, T, 32, 32), k1 = near(r1, T, T, 32, 32), k2 = near(r2, T, T, 32, 32), k3 = near(r3, T, T, 32, 32);
      push(p, r0, k0);
      push(p + 1, r1, k1 && p + 1 <= last);
      push(p + 2, r2, k2 && p + 2 <= last);
      push(p + 3, r3, k3 && p + 3 <= last);
    }
    const int ds_end = min(chunk_end, max(dd_end, h->ss_begin));
#pragma unroll 1
    for (int p = dd_end; p < ds_end; p += 2) {
      const bool two = p + 1 < ds_end;
      const int4 r0 = rec[3 * p], r1 = rec[3 * (two ? p + 1 : p)];
      const bool k0 = near(r0, T, TS, 32, 1), k1 = near(r1, T, TS, 32, 1);
      push(p, r0, k0);
      push(p + 1, r1, k1 && two);
    }
    // two universe-attached geometries (a fixed base link against a scene object): the same
    // answer for every configuration, but it is part of the pair list
#pragma unroll 1
    for (int p = ds_end; p < chunk_end; ++p) {
      const int4 r = rec[3 * p];
      push(p, r, near(r, TS, TS, 1, 1));
    }
    __syncwarp();
    // ---- stage 1.5: OUTER capsules for the GJK class, on dense groups of 32 survivors of the
    // sphere test (only ~20 % of the lanes pass it, so testing capsules per lane in stage 1
    // would waste four lanes in five).  Needs just the z axis and the origin of each placement.
    {
      int kept = 0;
#pragma unroll 1
      for (int base = 0; base < qn; base += 32) {
        const bool live = base + l < qn;
        const int item = live ? w.queue[base + l] : 0;
        bool keep = live;
        if (live) {
          // record 1 of the pair: outer capsules in o1 / o2 order (rsum_out = inf: not the GJK class)
          const int pidx = chunk + (item >> 5);
          const int4 r0 = rec[3 * pidx];
          const float4 r1 = reinterpret_cast<const float4*>(rec)[3 * pidx + 1];
          const int kind = __float_as_int(r1.w);
          const char* col = T + 4 * ((item & 31) - l);  // origin rows of the item's configuration
          const char* pa = (kind < 2 ? col : TS) + r0.x;
          const char* pb = (kind < 1 ? col : TS) + r0.y;
          const int sa = kind < 2 ? 128 : 4, sb = kind < 1 ? 128 : 4;  // row strides in bytes
          auto f = [](const char* p) { return *reinterpret_cast<const float*>(p); };
          const V3 za = mk(f(pa - 3 * sa), f(pa - 2 * sa), f(pa - sa)), ta = mk(f(pa), f(pa + sa), f(pa + 2 * sa));
          const V3 zb = mk(f(pb - 3 * sb), f(pb - 2 * sb), f(pb - sb)), tb = mk(f(pb), f(pb + sb), f(pb + 2 * sb));
          const V3 wv = seg_seg_witness(ta - za * r1.x, za * (2.f * r1.x), tb - zb * r1.y, zb * (2.f * r1.y));
          keep = dot(wv, wv) <= r1.z * r1.z;
#ifdef RAMP_TRACE_PAIR
          if ((r0.w & 0xffff) == RAMP_TRACE_PAIR && row_base + (item & 31) == 0)
            printf("trace stage1.5: |w| %.9g rsum_out %.9g keep %d\n", sqrtf(dot(wv, wv)), r1.z, int(keep));
#endif
        }
        __syncwarp();  // every lane has read its item before the in-place compaction writes
        const unsigned b = __ballot_sync(FULL, keep);
        if (keep) w.queue[kept + __popc(b & ((1u << l) - 1))] = static_cast<uint16_t>(item);
        kept += __popc(b);
        __syncwarp();
      }
      qn = kept;
    }
    // ---- stage 2: closed forms, exact box-box, OBB cull; survivors re-compacted in place
    int gn = 0;
#pragma unroll 1
    for (int base = 0; base < qn; base += 32) {
      const bool live = base + l < qn;
      const int item = live ? w.queue[base + l] : 0;
      bool to_gjk = false, hard = false;
      Xf rel = xf_identity();
      V3 v0 = mk(1.f, 0.f, 0.f);
      int orig = 0, cls = 0, p = 0;
      bool go = live;
      if (live) {  // a configuration settled by an earlier group needs no more work
        p = chunk + (item >> 5);
        const int oc = rec[3 * p].w;  // orig | cls << 16
        orig = oc & 0xffff;
        cls = oc >> 16;
        go = !dead(w.first[item & 31], orig);
      }
      if (go) {
        const int c = item & 31;
        const int4 r2 = rec[3 * p + 2];  // slack, a | b << 16, sa | sb << 16
        const float slack = __int_as_float(r2.x);
        const DGeom* ga = &w.t.geoms[r2.y & 0xffff];
        const DGeom* gb = &w.t.geoms[r2.y >> 16];
        rel = xf_inv_mul(geom_xf(w, static_cast<int16_t>(r2.z & 0xffff), c), geom_xf(w, r2.z >> 16, c));
        if (cls >= kClsBoxBox) {
          // one SAT instance: exact decision on the boxes themselves, or a cull on the
          // local bounding boxes of two GJK shapes (DGeom words 4-7: the box, 8-10: the bounds)
          const bool bb = cls == kClsBoxBox;
          const float4 ha4 = reinterpret_cast<const float4*>(ga)[bb ? 1 : 2], hb4 = reinterpret_cast<const float4*>(gb)[bb ? 1 : 2];
          const V3 ha = mk(ha4.x, ha4.y, ha4.z), hb = mk(hb4.x, hb4.y, hb4.z);
          // GJK class: overlapping INNER capsules prove the collision (the outer ones were
          // tested in stage 1); only the shell in between goes on to the OBB cull and GJK.
          bool decided = false;
          if (!bb) {
            const float4 ca = reinterpret_cast<const float4*>(ga)[6], cb = reinterpret_cast<const float4*>(gb)[6];  // s_out r_out s_in r_in
            const V3 zb = mk(rel.m[2], rel.m[6], rel.m[10]);
            v0 = seg_seg_witness(mk(0.f, 0.f, -ca.z), mk(0.f, 0.f, 2.f * ca.z), trans(rel) - zb * cb.z, zb * (2.f * cb.z));
            const float rin = ca.w + cb.w - slack;
            const float vv0 = dot(v0, v0);
            if (rin > 0.f && vv0 < rin * rin) { record_hit(w, c, orig, want_mask); decided = true; }
            const float thr = fmaf(kHardShell, ca.y + cb.y - (ca.w + cb.w), ca.w + cb.w);  // r_out, r_in sums
            hard = vv0 < thr * thr;
#ifdef RAMP_TRACE_PAIR
            if (orig == RAMP_TRACE_PAIR && row_base + c == 0)
              printf("trace stage2: |v0| %.9g rin %.9g decided %d hard %d slack %.3g\n", sqrtf(vv0), rin, int(decided), int(hard), slack);
#endif
          }
          if (!decided) {
            const bool ov = obb_overlap(ha, hb, rel, bb ? 0.f : slack);
            if (ov && bb) record_hit(w, c, orig, want_mask);
            to_gjk = ov && !bb;
#ifdef RAMP_TRACE_PAIR
            if (orig == RAMP_TRACE_PAIR && row_base + c == 0) printf("trace stage2: obb overlap %d to_gjk %d\n", int(ov), int(to_gjk));
#endif
          }
        } else if (closed_form_hit(cls, ga, gb, rel)) {
          record_hit(w, c, orig, want_mask);
        }
      }
      __syncwarp();  // all lanes have read their item before the in-place compaction writes
      // hits of this group may have settled configurations whose GJK items are still pending
      if (to_gjk && dead(w.first[item & 31], orig)) to_gjk = false;
      if (gq.items != nullptr) {
        // reserve slots of the two queue regions for this group's survivors (one 64-bit atomic
        // per group: hard count in the low word, easy count in the high word)
        const unsigned mh = __ballot_sync(FULL, to_gjk && hard), me = __ballot_sync(FULL, to_gjk && !hard);
        if (mh | me) {
          const int leader = __ffs(mh | me) - 1;
          unsigned long long base = 0;
          if (l == leader)
            base = atomicAdd(reinterpret_cast<unsigned long long*>(gq.count),
                             static_cast<unsigned long long>(__popc(mh)) | (static_cast<unsigned long long>(__popc(me)) << 32));
          base = __shfl_sync(FULL, base, leader);
          const uint32_t lt_l = (1u << l) - 1;
          const uint32_t slot = hard ? static_cast<uint32_t>(base) + __popc(mh & lt_l)
                                     : gq.cap_hard + static_cast<uint32_t>(base >> 32) + __popc(me & lt_l);
          if (to_gjk && slot < (hard ? gq.cap_hard : gq.capacity)) {
            const uint32_t c = item & 31, p = chunk + (item >> 5);
            float4* dst = reinterpret_cast<float4*>(gq.items + slot);
            dst[0] = make_float4(rel.m[0], rel.m[1], rel.m[2], rel.m[3]);
            dst[1] = make_float4(rel.m[4], rel.m[5], rel.m[6], rel.m[7]);
            dst[2] = make_float4(rel.m[8], rel.m[9], rel.m[10], rel.m[11]);
            uint32_t da, db;
            pack_dir(v0, da, db);
            dst[3] = make_float4(__uint_as_float(da), __uint_as_float(db), __uint_as_float(row_base + c), __uint_as_float(p));
            to_gjk = false;  // queued; anything past the capacity stays: in-kernel loop (kStage3) or tile flag
          }
        }
      }
      if (!kStage3) {
        over = over || to_gjk;
      } else {
        const unsigned b = __ballot_sync(FULL, to_gjk);
        if (to_gjk) w.queue[gn + __popc(b & ((1u << l) - 1))] = static_cast<uint16_t>(item);
        gn += __popc(b);
      }
      __syncwarp();
    }
    // ---- stage 3: GJK.  Each lane owns one work item and advances it one iteration per
    // trip; lanes whose item finished fetch the next one from the queue, so lanes stay busy
    // although iteration counts differ per item.
    if (kStage3 && gn > 0) {
      int next = 0, c = 0, p = 0;
      bool active = false;
      GjkLane g;
      GjkShape A, B;
      A.type = B.type = 0; A.vcnt = B.vcnt = 0; A.p0 = A.p1 = A.p2 = B.p0 = B.p1 = B.p2 = 0.f; A.verts = B.verts = w.t.verts;
      g.s.n = 0; g.it = 0; g.vv = 1.f; g.rsum = 0.f; g.lower = 0.f; g.v = mk(1.f, 0.f, 0.f); g.rel = xf_identity();
      while (true) {
        const unsigned need = __ballot_sync(FULL, !active);
        if (!active) {
          const int idx = next + __popc(need & ((1u << l) - 1));
          if (idx < gn) {
            const int item = w.queue[idx];
            c = item & 31;
            p = chunk + (item >> 5);
            // a configuration that was hit earlier (at a smaller list index) needs no later pair
            const DPair pr = w.t.pairs[p];
            p = pr.orig;  // results report the caller's index
            if (!dead(w.first[c], p)) {
              const DGeom* ga = &w.t.geoms[pr.a];
              const DGeom* gb = &w.t.geoms[pr.b];
              A = load_shape(ga, w.t.verts);
              B = load_shape(gb, w.t.verts);
              const Xf rl = xf_inv_mul(geom_xf(w, ga->gslot, c), geom_xf(w, gb->gslot, c));
              // same start as the queued items: the witness of the two inner capsule axes
              const V3 zb = mk(rl.m[2], rl.m[6], rl.m[10]);
              const V3 v0 = seg_seg_witness(mk(0.f, 0.f, -ga->s_in), mk(0.f, 0.f, 2.f * ga->s_in), trans(rl) - zb * gb->s_in,
                                            zb * (2.f * gb->s_in));
              gjk_init(g, rl, swept_radius(ga) + swept_radius(gb), v0);
              active = true;
            }
          }
        }
        next += __popc(need);
        if (!__any_sync(FULL, active)) {
          if (next >= gn) break;
          continue;
        }
        if (active) {
          float dist;
          const int st = gjk_step(g, A, B, false, dist);
          if (st != GJK_CONTINUE) {
            if (st == GJK_HIT) record_hit(w, c, p, want_mask);
            active = false;
          }
        }
      }
    }
    __syncwarp();
  }
  if (!kStage3 && __any_sync(FULL, over) && l == 0) {
    gq.tile_flags[row_base >> 5] = 1;
    atomicAdd(gq.count + 3, 1u);
  }
}

__device__ __forceinline__ void reset_results(const Warp& w, int mask_words) {
  w.first[w.lane] = INT32_MAX;
  if (w.lane == 0) reinterpret_cast<int*>(w.save)[-2] = 0;  // the team's chunk counter
  for (int k = 0; k < mask_words; ++k) w.mask[k * 32 + w.lane] = 0u;
  __syncwarp();
}

struct CheckArgs {
  const unsigned char* img;
  const float* q;
  int64_t n;
  int32_t ld;
  uint32_t flags;
  uint8_t* hit;
  int32_t* first_pair;
  uint32_t* pair_mask;
  GjkQueue gq;
  uint32_t row_offset;  // added to the row index of queued items: several launches can feed one queue
  SlabLayout L;
};

// ------------------------------------------------------------------ kernel: validity
template <int kTeam>
__global__ void __launch_bounds__(576, 1) check_kernel(const __grid_constant__ CheckArgs a) {
  uint64_t* mbar = reinterpret_cast<uint64_t*>(g_smem);
  unsigned char* tbl = g_smem + 16;
  stage_tables(tbl, a.img, a.L.table_bytes, mbar);
  Tables t = make_tables(tbl);
  t.verts = reinterpret_cast<const float4*>(a.img + t.h->off_verts);  // not staged: global memory (overflow path only)
  const int nwarps = blockDim.x >> 5, wid = threadIdx.x >> 5;
  // Teams: `team` warps share one slab and one tile.  Warp 0 of the team loads the rows and
  // runs FK, then all of them split the pair list; a named barrier per team orders the phases.
  // More warps per slab is what lets robots with many geometries (large slabs) and small
  // batches (few tiles) keep the schedulers busy.
  constexpr int team = kTeam;
  const int tid = wid / team, tw = wid - tid * team, nteams = nwarps / team;
  float* slab = reinterpret_cast<float*>(tbl + a.L.table_bytes) + static_cast<size_t>(tid) * a.L.words;
  const Warp w = make_warp(t, slab, a.L);
  auto team_sync = [&]() {
    if (team == 1) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(1 + tid), "r"(team * 32) : "memory");
  };
  const bool actuated = a.flags & RAMP_Q_ACTUATED;
  const bool want_mask = a.pair_mask != nullptr;
  const bool stop = !(a.flags & RAMP_ALL_PAIRS) && !want_mask;
  int cols = actuated ? t.h->n_act : t.h->nq;
  if ((a.ld & 3) == 0 && a.ld <= cols + 3) cols = a.ld;  // padded rows: take the float4 path
  const int64_t ntiles = (a.n + 31) >> 5;
  const int64_t tile_step = static_cast<int64_t>(gridDim.x) * nteams;
  const int64_t tile0 = static_cast<int64_t>(blockIdx.x) * nteams + tid;
  TileFetch tf;
  if (tw == 0) {
    tile_fetch_init(tf, w.save, w.lane);  // the first tile takes the plain load: nothing to hide a copy behind
  }
  for (int64_t tile = tile0; tile < ntiles; tile += tile_step) {
    const int64_t row0 = tile << 5;
    if (tw == 0) {
      tile_fetch_commit(tf, actuated ? w.act : w.q, w.save, a.q, row0, a.n, cols, a.ld, w.lane);
      if (actuated) expand_q(w);
      reset_results(w, want_mask ? t.h->mask_words : 0);
      if (kTeam == 1) fk_lane<0>(w);
    }
    if (kTeam > 1) {  // every warp of the team runs its share of the kinematic tree
      team_sync();
      fk_team<kTeam>(w, tw);
    }
    team_sync();
    // FK is done with the saved transforms: their region takes the next tile's rows
    if (tw == 0) tile_fetch_issue(tf, a.q, (tile + tile_step) << 5, a.n, cols, a.ld, a.L.raw_cols, w.lane);
    narrow_phase<kTeam, false>(w, stop, stop && a.first_pair == nullptr, want_mask, a.gq, static_cast<uint32_t>(row0) + a.row_offset,
                        static_cast<int>(imin64(32, a.n - row0)), tw, a.L.chunk_pairs);
    team_sync();
    const int64_t row = row0 + w.lane;
    if (tw == 0 && row < a.n) {
      const int f = w.first[w.lane];
      a.hit[row] = f != INT32_MAX;
      if (a.first_pair) a.first_pair[row] = f != INT32_MAX ? f : -1;
      if (want_mask)
        for (int k = 0; k < t.h->mask_words; ++k) a.pair_mask[row * t.h->mask_words + k] = w.mask[k * 32 + w.lane];
    }
    // the next trip's rows overwrite the region the team's queues live in
    team_sync();
  }
}

}  // namespace ramp
