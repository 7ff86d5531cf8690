// gjk_queue.cuh -- gjk_queue_kernel: drains the GJK work queue the validity kernels fill.
// Part of the single translation unit ramp_b200.cu (included there, in this order).
#pragma once

namespace ramp {

// ------------------------------------------------------------- kernel: GJK work queue
// Drains the queue the validity kernel filled.  Only the tables live in shared memory, so many
// more warps are resident than in the validity kernel; every lane owns one item and advances
// it one GJK iteration per trip, fetching the next item (from the chunk of 32 its warp last drew
// from the queue's cursor, hard region first) when it finishes.
struct GjkQueueArgs {
  const unsigned char* img;
  int32_t table_bytes;
  const GjkItem* items;
  uint32_t* count;   // [0] hard, [1] easy items produced, [2] the consumers' cursor (zeroed by the host)
  uint32_t capacity, cap_hard;
  uint8_t* hit;
  int32_t* first_pair;
  uint32_t* pair_mask;
  int32_t mask_words;
  int32_t stop;  // early-out semantics: a configuration already hit by an earlier pair is skipped
  uint32_t n_rows;  // rows of the result arrays: an item whose row lies beyond them is dropped (second line of
                    // defence; the producers never queue one).  Edge mode: unused
  // edge mode (offs != nullptr): `row` is a flat state index; a hit lowers first_invalid[edge]
  const int64_t* offs;
  const int32_t* nd;
  int32_t* first_invalid;
  int64_t n_edges;
  unsigned long long* stats;  // per-model counters: [0] items consumed, [1] launches (nullable)
};

constexpr int kQueueChunk = 32;  // items a warp takes from the queue at a time
constexpr int kFetchAttempts = 4;
#ifndef RAMP_GJK_CTAS
#define RAMP_GJK_CTAS 2
#endif
constexpr int kGjkCtasPerSm = RAMP_GJK_CTAS;

#ifdef RAMP_STATS
__device__ unsigned long long g_stats[64];  // [0] items [1] iterations [2] hits [3] skipped, [8+it] histogram
#endif

__global__ void __launch_bounds__(256, 2) gjk_queue_kernel(const __grid_constant__ GjkQueueArgs a) {
  uint64_t* mbar = reinterpret_cast<uint64_t*>(g_smem);
  unsigned char* tbl = g_smem + 16;
  stage_tables(tbl, a.img, a.table_bytes, mbar);
  const Tables t = make_tables(tbl);
  // hard items first (see GjkQueue): virtual index i -> slot i, or cap_hard + (i - n_hard)
  const uint32_t n_hard = min(a.count[0], a.cap_hard);
  const uint32_t total = n_hard + min(a.count[1], a.capacity - a.cap_hard);
  if (a.stats != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(a.stats, static_cast<unsigned long long>(total));
    atomicAdd(a.stats + 1, 1ull);
  }
  // Work is handed out dynamically: a warp takes kQueueChunk consecutive items at a time from a
  // cursor next to the producers' counter, so every warp stays busy until the queue is empty
  // (iteration counts per item vary from 1 to 48: a static split leaves long tails).
  uint32_t* cursor = a.count + 2;
  const int l = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu;
  // A short queue (a batch of a few thousand configurations) is a latency problem: 32 items per
  // warp would leave most of the grid idle while a few warps work through their longest item at
  // one vertex per lane and step.  Then every warp takes only its share of the items (at least
  // 4), does not refill before they are done, and all 32 lanes share the vertex scans of those
  // few items from the first trip on.
  const uint32_t warps_total = gridDim.x * (blockDim.x >> 5);
  const uint32_t chunk = min(static_cast<uint32_t>(kQueueChunk), max(4u, (total + warps_total - 1) / warps_total));
  const bool few = chunk <= 8;
  uint32_t lo = 0, hi = 0, row = 0, p = 0;  // [lo, hi): items this warp holds but has not started
  bool active = false, drained = false;
  GjkLane g;
  GjkShape A, B;
  A.type = B.type = 0; A.vcnt = B.vcnt = 0; A.p0 = A.p1 = A.p2 = B.p0 = B.p1 = B.p2 = 0.f; A.verts = B.verts = t.verts;
  g.s.n = 0; g.it = 0; g.vv = 1.f; g.rsum = 0.f; g.lower = 0.f; g.v = mk(1.f, 0.f, 0.f); g.rel = xf_identity();
  while (true) {
    // Fill idle lanes.  In early-out mode many items are skipped on arrival (their
    // configuration was hit meanwhile), so try a few times before running a trip.
#pragma unroll 1
    for (int attempt = 0; attempt < kFetchAttempts; ++attempt) {
      const unsigned need = __ballot_sync(FULL, !active);
      if (!need || (drained && lo == hi) || (few && ~need != 0u)) break;
      uint32_t gi = 0xffffffffu;
      {
        const uint32_t rank = __popc(need & ((1u << l) - 1)), nneed = __popc(need), avail = hi - lo;
        if (!active && rank < avail) gi = lo + rank;
        if (nneed <= avail) {
          lo += nneed;
        } else {
          lo = hi;
          if (!drained) {
            uint32_t base = 0;
            if (l == 0) base = atomicAdd(cursor, chunk);
            base = __shfl_sync(FULL, base, 0);
            if (base >= total) {
              drained = true;
            } else {
              hi = min(base + chunk, total);
              const uint32_t r2 = rank - avail;
              if (!active && rank >= avail && r2 < hi - base) gi = base + r2;
              lo = min(base + (nneed - avail), hi);
            }
          }
        }
      }
      if (gi != 0xffffffffu) {
        const float4* src = reinterpret_cast<const float4*>(a.items + (gi < n_hard ? gi : a.cap_hard + (gi - n_hard)));
        const float4 r0 = __ldcs(src), r1 = __ldcs(src + 1), r2 = __ldcs(src + 2), r3 = __ldcs(src + 3);
        row = __float_as_uint(r3.z);
        const DPair pr = t.pairs[__float_as_uint(r3.w)];
        p = static_cast<uint32_t>(pr.orig);  // results report the caller's index
        bool skip = !a.offs && row >= a.n_rows;
        if (!skip && a.stop && !a.offs) skip = a.first_pair ? static_cast<uint32_t>(a.first_pair[row]) < p : a.hit[row] != 0;
#ifdef RAMP_STATS
        atomicAdd(&g_stats[skip ? 3 : 4], 1ull);
        if (!skip) { const int va = t.geoms[pr.a].vcnt, vb = t.geoms[pr.b].vcnt; atomicAdd(&g_stats[50 + (va > 16) + 2 * (vb > 16)], 1ull); }
#endif
        if (!skip) {
          const DGeom* ga = &t.geoms[pr.a];
          const DGeom* gb = &t.geoms[pr.b];
          A = load_shape(ga, t.verts);
          B = load_shape(gb, t.verts);
          Xf rel;
          rel.m[0] = r0.x; rel.m[1] = r0.y; rel.m[2] = r0.z; rel.m[3] = r0.w;
          rel.m[4] = r1.x; rel.m[5] = r1.y; rel.m[6] = r1.z; rel.m[7] = r1.w;
          rel.m[8] = r2.x; rel.m[9] = r2.y; rel.m[10] = r2.z; rel.m[11] = r2.w;
          gjk_init(g, rel, swept_radius(ga) + swept_radius(gb), unpack_dir(__float_as_uint(r3.x), __float_as_uint(r3.y)));
          active = true;
        }
      }
    }
    if (!__any_sync(FULL, active)) {
      if (drained && lo == hi) break;
      continue;
    }
    // Both support scans run with the whole warp converged (idle lanes scan zero vertices):
    // left to itself the compiler lets lanes with a short vertex list run ahead into the next
    // scan on their own, and the scans execute twice at partial width.
    // Once the queue is empty a warp is left with a few long-running items (38 % of the trips
    // of a UR5 batch ran with 7 of 32 lanes busy): then the idle lanes help with the scans.
    const unsigned am = __ballot_sync(FULL, active);
    const int na = __popc(am);
    V3 sa, sb;
    if ((few || (drained && lo == hi)) && na <= 8) {
      const V3 da = -g.v, db = rot_t(g.rel, g.v);
      if (na <= 4) { sa = coop_support<8>(am, active, A, da, t.verts); sb = coop_support<8>(am, active, B, db, t.verts); }
      else { sa = coop_support<4>(am, active, A, da, t.verts); sb = coop_support<4>(am, active, B, db, t.verts); }
    } else {
      sa = support(A.type, A.p0, A.p1, A.p2, A.verts, active ? A.vcnt : 0, -g.v);
      __syncwarp();
      sb = support(B.type, B.p0, B.p1, B.p2, B.verts, active ? B.vcnt : 0, rot_t(g.rel, g.v));
      __syncwarp();
    }
    if (active) {
      float dist;
#ifdef RAMP_TRACE_PAIR
      const float t_vv = g.vv; const int t_n = g.s.n;
      const V3 t_w = sa - apply(g.rel, sb);
      const float t_vw = dot(g.v, t_w);
#endif
      const int st = gjk_advance(g, sa, sb, false, dist);
#ifdef RAMP_TRACE_PAIR
      if (p == RAMP_TRACE_PAIR && row == 0)
        printf("trace queue gjk: it %d n %d |v| %.9g vw/|v| %.9g -> status %d n %d |v| %.9g dist %.9g\n", g.it, t_n, sqrtf(t_vv), t_vw * rsqrtf(t_vv), st,
               g.s.n, sqrtf(g.vv), dist);
#endif
#ifdef RAMP_STATS
      atomicAdd(&g_stats[1], 1ull);
      atomicAdd(&g_stats[5], static_cast<unsigned long long>(__popc(__activemask())));
      atomicAdd(&g_stats[56 + min(g.s.n, 4)], 1ull);
      if (st != GJK_CONTINUE) { atomicAdd(&g_stats[0], 1ull); atomicAdd(&g_stats[2], st == GJK_HIT ? 1ull : 0ull); atomicAdd(&g_stats[8 + min(g.it, 40)], 1ull); }
#endif
      if (st != GJK_CONTINUE) {
        if (st == GJK_HIT && a.offs) {
          // edge of flat state `row`: last e with offs[e] <= row; its 1-based index j
          int64_t lo = 0, hi = a.n_edges - 1;
          while (lo < hi) {
            const int64_t mid = (lo + hi + 1) >> 1;
            if (a.offs[mid] <= static_cast<int64_t>(row)) lo = mid; else hi = mid - 1;
          }
          const int j = static_cast<int>(static_cast<int64_t>(row) - a.offs[lo]) + 1;
          atomicMin(&a.first_invalid[lo], min(j, a.nd[lo]));
        } else if (st == GJK_HIT) {
          a.hit[row] = 1;
          // -1 ("none") is the largest unsigned value, so an unsigned min keeps list order
          if (a.first_pair) atomicMin(reinterpret_cast<uint32_t*>(a.first_pair) + row, p);
          if (a.pair_mask) atomicOr(a.pair_mask + static_cast<size_t>(row) * a.mask_words + (p >> 5), 1u << (p & 31));
        }
        active = false;
      }
    }
  }
}

}  // namespace ramp
