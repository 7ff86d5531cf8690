// ramp_b200.cu -- the one translation unit of libramp_b200.so (include/ramp_b200.h): this file
// holds the host side (model packing, launch logic, the C ABI); the sm_100a device code lives in
// the headers it includes: device_math.cuh (SE(3), shapes, closed forms, GJK / EPA), tile.cuh
// (tables, slabs, row loads, FK), narrow_phase.cuh (check_kernel), small_batch.cuh
// (check_warp_kernel), gjk_queue.cuh, fk_distance.cuh, edges.cuh.
//
// Execution model (DESIGN.md has the full account):
//   * one persistent CTA per SM; the packed model image is staged into shared memory once
//     per CTA by a TMA bulk copy (cp.async.bulk, mbarrier completion);
//   * a warp owns a TILE of 32 configurations.  Lane l runs forward kinematics for
//     configuration l entirely in FP32 registers (thread-per-configuration, no idle lanes)
//     and leaves the 3x4 placement of every movable geometry in the warp's shared-memory
//     slab, laid out [geom][entry][lane] so every access is conflict free;
//   * narrow phase: the warp walks the pair list; lane l runs the bounding-sphere test of
//     pair p for ITS configuration, survivors are ballot-compacted into a per-warp queue of
//     (lane, pair) work items; the queue is then drained 32 items at a time -- closed-form
//     tests / exact box-box SAT / OBB cull first, survivors re-compacted, GJK last -- so the
//     expensive code always runs with full warps;
//   * no __syncthreads after the table staging: warps are independent.
//
// There is no CPU fallback in this file: every entry point fails with RAMP_ERR_CUDA when
// no device is usable.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <unordered_map>
#include <vector>

#include "../../include/ramp_b200.h"
#include "device_math.cuh"
#include "model_image.h"

// device code, in dependency order
#include "tile.cuh"
#include "narrow_phase.cuh"
#include "small_batch.cuh"
#include "gjk_queue.cuh"
#include "fk_distance.cuh"
#include "edges.cuh"
#include "ik.cuh"

namespace ramp {

// ================================================================================= host
static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}

#define CUDA_TRY(expr)                                                                          \
  do {                                                                                          \
    cudaError_t e_ = (expr);                                                                    \
    if (e_ != cudaSuccess) return fail(RAMP_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)

// bounding / inscribed capsule of a convex vertex set on its principal axis (fit_capsules)
struct CapsuleFit {
  double c[3], u[3], s_out, r_out, s_in, r_in;
};

}  // namespace ramp

struct RampModel {
  int device = 0;
  int sm_count = 0;
  int max_smem = 0;
  int32_t nq = 0;
  std::vector<RampJoint> joints;
  std::vector<RampGeom> robot_geoms, scene_geoms;
  std::vector<int32_t> robot_pairs, scene_pairs;  // flattened [][2]
  std::vector<double> robot_verts, scene_verts;   // [][3]
  std::vector<uint8_t> pair_active;   // by the caller's pair index
  std::vector<int32_t> pair_order;    // caller's pair index -> processing index in the image
  // qmap
  bool has_qmap = false;
  std::vector<int32_t> act_slot, mimic_dst, mimic_src, cont_slot;
  std::vector<double> mimic_mult, mimic_off, ref_q;
  // device image
  std::vector<unsigned char> image;
  unsigned char* d_image = nullptr;
  size_t d_image_cap = 0;
  ramp::ImgHeader hdr{};
  int n_gjk_pairs = 0;            // active pairs of the GJK class
  cudaMemPool_t pool = nullptr;   // stream-ordered scratch for the GJK work queue
  unsigned long long* d_stats = nullptr;  // [0] GJK items the queue kernels consumed, [1] queue kernel launches
  // Observed GJK items per row (hard, easy), learnt from the counters of earlier ramp_check_batch
  // calls: sizes the queue of large batches (see launch_check).  < 0: nothing observed yet.
  mutable std::mutex est_mu;
  mutable float est_hard = -1.f, est_easy = -1.f;
  mutable uint32_t* h_counts = nullptr;    // page-locked: the last slice's hard / easy counts
  mutable cudaEvent_t ev_counts = nullptr;
  mutable bool counts_pending = false;
  mutable int64_t counts_rows = 0;
  // scratch for ramp_check_edges (offsets, nd, first_invalid) and ramp_check_host staging
  mutable std::mutex mu;
  mutable void* d_scratch = nullptr;
  mutable size_t scratch_cap = 0;
  mutable void* h_pinned = nullptr;
  mutable size_t pinned_cap = 0;
  // ramp_check_host pipeline: kHostStreams for chunks, one for the queue kernel + D2H; events: queue
  // ready, tiles done per chunk stream, results copied x2
  mutable cudaStream_t streams[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  mutable cudaEvent_t events[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  mutable void* d_edge_stage = nullptr;  // ramp_check_edges_host staging (device / pinned host)
  mutable void* h_edge_stage = nullptr;
  mutable size_t edge_stage_cap = 0;
};

namespace ramp {

static int pair_class(int ta, int tb) {
  if (ta > tb) std::swap(ta, tb);
  if (ta == RAMP_SHAPE_SPHERE) {
    if (tb == RAMP_SHAPE_SPHERE) return kClsSphereSphere;
    if (tb == RAMP_SHAPE_CAPSULE) return kClsSphereCapsule;
    if (tb == RAMP_SHAPE_BOX) return kClsSphereBox;
    if (tb == RAMP_SHAPE_CYLINDER) return kClsSphereCylinder;
  }
  if (ta == RAMP_SHAPE_CAPSULE && tb == RAMP_SHAPE_CAPSULE) return kClsCapsuleCapsule;
  if (ta == RAMP_SHAPE_BOX && tb == RAMP_SHAPE_BOX) return kClsBoxBox;
  return kClsGjk;
}

static size_t align16(size_t x) { return (x + 15) & ~size_t(15); }

static void mat_mul34(const double* a, const double* b, double* o) {
  double r[12];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) r[4 * i + j] = a[4 * i] * b[j] + a[4 * i + 1] * b[4 + j] + a[4 * i + 2] * b[8 + j];
    r[4 * i + 3] = a[4 * i] * b[3] + a[4 * i + 1] * b[7] + a[4 * i + 2] * b[11] + a[4 * i + 3];
  }
  std::memcpy(o, r, sizeof r);
}


// ---- bounding / inscribed capsules of a vertex set (model-build time, FP64)
static double seg_point_dist(const double* a, const double* b, const double* p) {
  double ab[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, ap[3] = {p[0] - a[0], p[1] - a[1], p[2] - a[2]};
  const double den = ab[0] * ab[0] + ab[1] * ab[1] + ab[2] * ab[2];
  double t = den > 0 ? (ap[0] * ab[0] + ap[1] * ab[1] + ap[2] * ab[2]) / den : 0.0;
  t = std::min(1.0, std::max(0.0, t));
  const double d[3] = {ap[0] - t * ab[0], ap[1] - t * ab[1], ap[2] - t * ab[2]};
  return std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
}

// principal axis of a point cloud (power iteration on the covariance)
static void principal_axis(const std::vector<double>& v, const double* c, double* u) {
  double C[3][3] = {{0}};
  for (size_t i = 0; i < v.size(); i += 3)
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) C[a][b] += (v[i + a] - c[a]) * (v[i + b] - c[b]);
  double best = -1;
  for (int start = 0; start < 3; ++start) {  // three starts: robust against an orthogonal seed
    double x[3] = {start == 0 ? 1.0 : 0.1, start == 1 ? 1.0 : 0.1, start == 2 ? 1.0 : 0.1};
    double lam = 0;
    for (int it = 0; it < 200; ++it) {
      double y[3];
      for (int a = 0; a < 3; ++a) y[a] = C[a][0] * x[0] + C[a][1] * x[1] + C[a][2] * x[2];
      lam = std::sqrt(y[0] * y[0] + y[1] * y[1] + y[2] * y[2]);
      if (lam < 1e-300) break;
      for (int a = 0; a < 3; ++a) x[a] = y[a] / lam;
    }
    if (lam > best) { best = lam; u[0] = x[0]; u[1] = x[1]; u[2] = x[2]; }
  }
  if (best <= 0) { u[0] = 0; u[1] = 0; u[2] = 1; }
}

// Outer capsule: smallest-volume capsule on the principal axis that contains every vertex.
// Inner capsule: largest-volume capsule on the same axis inside every facet plane of the
// hull; facets are enumerated by brute force (V^4/6, fine for V <= 160 at build time).
static CapsuleFit fit_capsules(const std::vector<double>& v) {
  CapsuleFit f{};
  const int n = static_cast<int>(v.size() / 3);
  double c[3] = {0, 0, 0};
  for (int i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) c[k] += v[3 * i + k] / n;
  double u[3];
  principal_axis(v, c, u);
  double tmin = 1e300, tmax = -1e300, rho = 0;
  for (int i = 0; i < n; ++i) {
    const double d[3] = {v[3 * i] - c[0], v[3 * i + 1] - c[1], v[3 * i + 2] - c[2]};
    const double t = d[0] * u[0] + d[1] * u[1] + d[2] * u[2];
    const double r[3] = {d[0] - t * u[0], d[1] - t * u[1], d[2] - t * u[2]};
    tmin = std::min(tmin, t); tmax = std::max(tmax, t);
    rho = std::max(rho, std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]));
  }
  const double mid = 0.5 * (tmin + tmax);
  for (int k = 0; k < 3; ++k) { f.c[k] = c[k] + mid * u[k]; f.u[k] = u[k]; }
  // facet planes n.x <= d
  std::vector<double> planes;
  if (n <= 160) {
    for (int i = 0; i < n; ++i) for (int j = i + 1; j < n; ++j) for (int k = j + 1; k < n; ++k) {
      const double* a = &v[3 * i]; const double* b = &v[3 * j]; const double* cc = &v[3 * k];
      const double e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {cc[0] - a[0], cc[1] - a[1], cc[2] - a[2]};
      double nn[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
      const double l = std::sqrt(nn[0] * nn[0] + nn[1] * nn[1] + nn[2] * nn[2]);
      if (l < 1e-14) continue;
      for (double& x : nn) x /= l;
      const double d0 = nn[0] * a[0] + nn[1] * a[1] + nn[2] * a[2];
      bool pos = true, neg = true;
      for (int m = 0; m < n && (pos || neg); ++m) {
        const double dd = nn[0] * v[3 * m] + nn[1] * v[3 * m + 1] + nn[2] * v[3 * m + 2] - d0;
        if (dd > 1e-10) neg = false;
        if (dd < -1e-10) pos = false;
      }
      if (neg) { planes.insert(planes.end(), {nn[0], nn[1], nn[2], d0}); }
      else if (pos) { planes.insert(planes.end(), {-nn[0], -nn[1], -nn[2], -d0}); }
    }
  }
  const double half = 0.5 * (tmax - tmin);
  double best_out = 1e300, best_in = -1;
  f.s_out = 0; f.r_out = 0; f.s_in = 0; f.r_in = 0;
  for (int step = 0; step <= 40; ++step) {
    const double s = half * step / 40.0;  // candidate half length of the segment
    const double a[3] = {f.c[0] - s * u[0], f.c[1] - s * u[1], f.c[2] - s * u[2]};
    const double b[3] = {f.c[0] + s * u[0], f.c[1] + s * u[1], f.c[2] + s * u[2]};
    double R = 0;
    for (int i = 0; i < n; ++i) R = std::max(R, seg_point_dist(a, b, &v[3 * i]));
    const double vol_out = M_PI * R * R * 2 * s + 4.0 / 3.0 * M_PI * R * R * R;
    if (vol_out < best_out) { best_out = vol_out; f.s_out = s; f.r_out = R; }
    if (!planes.empty()) {
      double r = 1e300;
      for (size_t k = 0; k < planes.size(); k += 4) {
        const double da = planes[k + 3] - (planes[k] * a[0] + planes[k + 1] * a[1] + planes[k + 2] * a[2]);
        const double db = planes[k + 3] - (planes[k] * b[0] + planes[k + 1] * b[1] + planes[k + 2] * b[2]);
        r = std::min(r, std::min(da, db));
      }
      if (r > 0) {
        const double vol_in = M_PI * r * r * 2 * s + 4.0 / 3.0 * M_PI * r * r * r;
        if (vol_in > best_in) { best_in = vol_in; f.s_in = s; f.r_in = r; }
      }
    }
  }
  // FP32 safety: outer grows, inner shrinks by a relative hair
  f.r_out *= 1.0 + 1e-6;
  f.r_in = std::max(0.0, f.r_in * (1.0 - 1e-5) - 1e-7);
  return f;
}

// PlacementKind of a 3x4 placement: which quarter / half turn about one axis its rotation is, if any
static int32_t placement_kind(const double* pl) {
  static const double S[11][9] = {
      {0, 0, 0, 0, 0, 0, 0, 0, 0},                                   // general (unused)
      {1, 0, 0, 0, 1, 0, 0, 0, 1},                                   // identity
      {1, 0, 0, 0, 0, -1, 0, 1, 0}, {1, 0, 0, 0, 0, 1, 0, -1, 0}, {1, 0, 0, 0, -1, 0, 0, 0, -1},   // x: +90, -90, 180
      {0, 0, 1, 0, 1, 0, -1, 0, 0}, {0, 0, -1, 0, 1, 0, 1, 0, 0}, {-1, 0, 0, 0, 1, 0, 0, 0, -1},   // y
      {0, -1, 0, 1, 0, 0, 0, 0, 1}, {0, 1, 0, -1, 0, 0, 0, 0, 1}, {-1, 0, 0, 0, -1, 0, 0, 0, 1}};  // z
  for (int k = 1; k <= 10; ++k) {
    bool same = true;
    for (int r = 0; r < 3 && same; ++r)
      for (int c = 0; c < 3 && same; ++c) same = std::fabs(pl[4 * r + c] - S[k][3 * r + c]) < 1e-9;
    if (same) return k;
  }
  return kPlGeneral;
}

// Capsule fits are a function of the vertex list alone and cost a brute-force facet enumeration
// (tens of milliseconds per 100-vertex hull), while images are rebuilt often -- every set_scene /
// set_qmap / set_pair_active, and a NEW model whenever the host removes collision pairs
// (compute_disable_collisions).  One process-wide cache keyed by the vertex bytes serves them all.
static CapsuleFit cached_fit(const std::vector<double>& core) {
  struct Entry { std::vector<double> v; CapsuleFit f; };
  static std::mutex mu;
  static std::unordered_multimap<uint64_t, Entry> cache;
  uint64_t h = 1469598103934665603ull;
  const unsigned char* b = reinterpret_cast<const unsigned char*>(core.data());
  for (size_t i = 0; i < core.size() * sizeof(double); ++i) { h ^= b[i]; h *= 1099511628211ull; }
  {
    std::lock_guard<std::mutex> lk(mu);
    auto range = cache.equal_range(h);
    for (auto it = range.first; it != range.second; ++it)
      if (it->second.v == core) return it->second.f;
  }
  const CapsuleFit f = fit_capsules(core);
  std::lock_guard<std::mutex> lk(mu);
  if (cache.size() > 4096) cache.clear();  // bounded: a long-lived process that keeps loading new robots
  cache.emplace(h, Entry{core, f});
  return f;
}

// Pack the host description into the device image.  Convex hulls are re-centred on the
// centre of their local bounding box (vertices shifted, placement composed with the shift):
// the placed point set is unchanged, the bounding sphere / OBB get tighter.
static int build_image(RampModel* m) {
  const int nj = static_cast<int>(m->joints.size());
  const int n_robot = static_cast<int>(m->robot_geoms.size());
  const int ng = n_robot + static_cast<int>(m->scene_geoms.size());
  const int n_rp = static_cast<int>(m->robot_pairs.size() / 2);
  const int np = n_rp + static_cast<int>(m->scene_pairs.size() / 2);
  if (nj > kMaxJoints) return fail(RAMP_ERR_LIMIT, "%d joints > limit %d", nj, kMaxJoints);
  if (ng > kMaxGeoms) return fail(RAMP_ERR_LIMIT, "%d geometries > limit %d", ng, kMaxGeoms);
  if (np > kMaxPairs) return fail(RAMP_ERR_LIMIT, "%d pairs > limit %d", np, kMaxPairs);
  if (m->nq > kMaxNq) return fail(RAMP_ERR_LIMIT, "nq %d > limit %d", m->nq, kMaxNq);

  std::vector<DJoint> dj(nj);
  std::vector<DGeom> dg(ng);
  std::vector<float> verts4;  // float4
  std::vector<float> stat;
  std::vector<int32_t> dyn;

  // joints: running / saved parents
  std::vector<int> save_slot(nj, -1);
  int n_save = 0;
  for (int i = 0; i < nj; ++i) {
    const RampJoint& j = m->joints[i];
    if (j.parent >= i || j.parent < -1) return fail(RAMP_ERR_INVALID, "joint %d: parent %d must precede it", i, j.parent);
    if (j.type < RAMP_JOINT_RX || j.type > RAMP_JOINT_TRANSLATION) return fail(RAMP_ERR_INVALID, "joint %d: unknown type %d", i, j.type);
    if (j.parent >= 0 && j.parent != i - 1 && save_slot[j.parent] < 0) save_slot[j.parent] = n_save++;
  }
  for (int i = 0; i < nj; ++i) {
    const RampJoint& j = m->joints[i];
    DJoint& d = dj[i];
    std::memset(&d, 0, sizeof d);
    d.type = j.type;
    d.idx_q = j.idx_q;
    d.parent_code = j.parent < 0 ? kParentUniverse : (j.parent == i - 1 ? kParentRunning : save_slot[j.parent]);
    d.save_slot = save_slot[i];
    d.parent = j.parent;  // plain parent index, for the one-warp-per-configuration kernel
    d.nq = joint_nq(j.type);
    d.pkind = placement_kind(j.placement);
    d.ax = static_cast<float>(j.axis[0]); d.ay = static_cast<float>(j.axis[1]); d.az = static_cast<float>(j.axis[2]);
    for (int k = 0; k < 12; ++k) d.pl[k] = static_cast<float>(j.placement[k]);
  }
  // geometries
  int n_static = 0, n_lists = 0;
  std::vector<double> nonrigid(ng, 0.0);  // max |R^T R - I| of each placement
  std::vector<std::vector<int>> by_joint(nj);
  for (int g = 0; g < ng; ++g) {
    const bool scene = g >= n_robot;
    const RampGeom& src = scene ? m->scene_geoms[g - n_robot] : m->robot_geoms[g];
    const std::vector<double>& vsrc = scene ? m->scene_verts : m->robot_verts;
    DGeom& d = dg[g];
    std::memset(&d, 0, sizeof d);
    if (src.type < RAMP_SHAPE_SPHERE || src.type > RAMP_SHAPE_ELLIPSOID) return fail(RAMP_ERR_INVALID, "geometry %d: unknown shape %d", g, src.type);
    if (src.parent_joint >= nj || src.parent_joint < -1) return fail(RAMP_ERR_INVALID, "geometry %d: bad parent joint %d", g, src.parent_joint);
    d.type = src.type;
    d.parent = src.parent_joint;
    d.p0 = static_cast<float>(src.params[0]); d.p1 = static_cast<float>(src.params[1]); d.p2 = static_cast<float>(src.params[2]);
    double pl[12];
    std::memcpy(pl, src.placement, sizeof pl);
    double hx = 0, hy = 0, hz = 0, br = 0;
    std::vector<double> core;  // vertices of the shape core, geometry-local
    switch (src.type) {
      case RAMP_SHAPE_SPHERE: hx = hy = hz = br = src.params[0]; break;
      case RAMP_SHAPE_CAPSULE: hx = hy = src.params[0]; hz = br = src.params[0] + src.params[1]; break;
      case RAMP_SHAPE_BOX: hx = src.params[0]; hy = src.params[1]; hz = src.params[2]; br = std::sqrt(hx * hx + hy * hy + hz * hz); break;
      case RAMP_SHAPE_CYLINDER: hx = hy = src.params[0]; hz = src.params[1]; br = std::sqrt(hx * hx + hz * hz); break;
      case RAMP_SHAPE_ELLIPSOID: hx = src.params[0]; hy = src.params[1]; hz = src.params[2]; br = std::max(hx, std::max(hy, hz)); break;
      case RAMP_SHAPE_CONVEX: {
        if (src.vert_count <= 0 || src.vert_offset < 0 || static_cast<size_t>(src.vert_offset + src.vert_count) * 3 > vsrc.size())
          return fail(RAMP_ERR_INVALID, "geometry %d: vertex range out of bounds", g);
        const double* v = vsrc.data() + 3 * static_cast<size_t>(src.vert_offset);
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        for (int i = 0; i < src.vert_count; ++i)
          for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], v[3 * i + k]); hi[k] = std::max(hi[k], v[3 * i + k]); }
        const double c[3] = {0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1]), 0.5 * (lo[2] + hi[2])};
        hx = 0.5 * (hi[0] - lo[0]); hy = 0.5 * (hi[1] - lo[1]); hz = 0.5 * (hi[2] - lo[2]);
        for (int i = 0; i < src.vert_count; ++i) {
          const double x = v[3 * i] - c[0], y = v[3 * i + 1] - c[1], z = v[3 * i + 2] - c[2];
          br = std::max(br, std::sqrt(x * x + y * y + z * z));
          core.push_back(x); core.push_back(y); core.push_back(z);
        }
        const double shift[12] = {1, 0, 0, c[0], 0, 1, 0, c[1], 0, 0, 1, c[2]};
        mat_mul34(src.placement, shift, pl);
      } break;
    }
    {  // capsules for the mid-phase of the GJK class, and the re-framing that puts them on the
       // local z axis through the local origin
      CapsuleFit f{};
      f.u[2] = 1.0;
      const double* P = src.params;
      int perm = 2;  // index of the local axis that becomes z (boxes / ellipsoids)
      switch (src.type) {
        case RAMP_SHAPE_SPHERE: f.r_out = f.r_in = P[0]; break;
        case RAMP_SHAPE_CAPSULE: f.s_out = f.s_in = P[1]; f.r_out = f.r_in = P[0]; break;
        case RAMP_SHAPE_BOX: {
          const int L = (P[0] >= P[1] && P[0] >= P[2]) ? 0 : (P[1] >= P[2] ? 1 : 2), a = (L + 1) % 3, b = (L + 2) % 3;
          perm = L;
          f.s_out = P[L]; f.r_out = std::sqrt(P[a] * P[a] + P[b] * P[b]);
          f.r_in = std::min(P[a], P[b]); f.s_in = P[L] - f.r_in;
        } break;
        case RAMP_SHAPE_CYLINDER:
          f.s_out = P[1]; f.r_out = P[0]; f.r_in = std::min(P[0], P[1]); f.s_in = std::max(0.0, P[1] - P[0]); break;
        case RAMP_SHAPE_ELLIPSOID: {
          const int L = (P[0] >= P[1] && P[0] >= P[2]) ? 0 : (P[1] >= P[2] ? 1 : 2), a = (L + 1) % 3, b = (L + 2) % 3;
          perm = L;
          f.r_out = std::max(P[a], P[b]); f.s_out = std::max(0.0, P[L] - f.r_out);  // contains the ellipsoid
          f.r_in = std::min(P[L], std::min(P[a], P[b]));
        } break;
        default:  // convex: cached (set_qmap / set_pair_active / set_scene rebuild the image)
          f = cached_fit(core);
          break;
      }
      // F = (Rf, c): new local frame expressed in the old one; z' = capsule axis
      double Rf[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
      if (src.type == RAMP_SHAPE_BOX || src.type == RAMP_SHAPE_ELLIPSOID) {
        if (perm != 2) {  // cyclic permutation keeps the frame right handed: (x',y',z') = (a, b, L)
          const int a = (perm + 1) % 3, b = (perm + 2) % 3;
          const double hp[3] = {P[a], P[b], P[perm]};
          d.p0 = static_cast<float>(hp[0]); d.p1 = static_cast<float>(hp[1]); d.p2 = static_cast<float>(hp[2]);
          hx = hp[0]; hy = hp[1]; hz = hp[2];
          for (int r = 0; r < 3; ++r) { Rf[r][0] = r == a; Rf[r][1] = r == b; Rf[r][2] = r == perm; }
        }
      } else if (src.type == RAMP_SHAPE_CONVEX) {
        double z[3] = {f.u[0], f.u[1], f.u[2]};
        const double t[3] = {std::fabs(z[0]) < 0.9 ? 1.0 : 0.0, std::fabs(z[0]) < 0.9 ? 0.0 : 1.0, 0.0};
        double x[3] = {t[1] * z[2] - t[2] * z[1], t[2] * z[0] - t[0] * z[2], t[0] * z[1] - t[1] * z[0]};
        const double xl = std::sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
        for (double& v : x) v /= xl;
        const double y[3] = {z[1] * x[2] - z[2] * x[1], z[2] * x[0] - z[0] * x[2], z[0] * x[1] - z[1] * x[0]};
        for (int r = 0; r < 3; ++r) { Rf[r][0] = x[r]; Rf[r][1] = y[r]; Rf[r][2] = z[r]; }
        // vertices into the new frame; bounding box / radius follow
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        br = 0;
        for (size_t i = 0; i < core.size(); i += 3) {
          const double w[3] = {core[i] - f.c[0], core[i + 1] - f.c[1], core[i + 2] - f.c[2]};
          for (int k = 0; k < 3; ++k) {
            core[i + k] = Rf[0][k] * w[0] + Rf[1][k] * w[1] + Rf[2][k] * w[2];
            lo[k] = std::min(lo[k], core[i + k]); hi[k] = std::max(hi[k], core[i + k]);
          }
          br = std::max(br, std::sqrt(core[i] * core[i] + core[i + 1] * core[i + 1] + core[i + 2] * core[i + 2]));
        }
        // the OBB is centred on the origin: take the larger side of each axis
        hx = std::max(-lo[0], hi[0]); hy = std::max(-lo[1], hi[1]); hz = std::max(-lo[2], hi[2]);
        const double F[12] = {Rf[0][0], Rf[0][1], Rf[0][2], f.c[0], Rf[1][0], Rf[1][1], Rf[1][2], f.c[1],
                              Rf[2][0], Rf[2][1], Rf[2][2], f.c[2]};
        double npl[12];
        mat_mul34(pl, F, npl);
        std::memcpy(pl, npl, sizeof pl);
      }
      if ((src.type == RAMP_SHAPE_BOX || src.type == RAMP_SHAPE_ELLIPSOID) && perm != 2) {
        const double F[12] = {Rf[0][0], Rf[0][1], Rf[0][2], 0, Rf[1][0], Rf[1][1], Rf[1][2], 0, Rf[2][0], Rf[2][1], Rf[2][2], 0};
        double npl[12];
        mat_mul34(pl, F, npl);
        std::memcpy(pl, npl, sizeof pl);
      }
      d.s_out = static_cast<float>(f.s_out); d.r_out = static_cast<float>(f.r_out);
      d.s_in = static_cast<float>(f.s_in); d.r_in = static_cast<float>(f.r_in);
    }
    // polytope cores as vertex lists: sphere = its centre, capsule = the segment ends,
    // box = the 8 corners (device_math.cuh support())
    if (src.type == RAMP_SHAPE_SPHERE) core = {0, 0, 0};
    if (src.type == RAMP_SHAPE_CAPSULE) core = {0, 0, -src.params[1], 0, 0, src.params[1]};
    if (src.type == RAMP_SHAPE_BOX)
      for (int k = 0; k < 8; ++k) {
        core.push_back((k & 1) ? hx : -hx); core.push_back((k & 2) ? hy : -hy); core.push_back((k & 4) ? hz : -hz);
      }
    if (!core.empty()) {
      // skew the start of every vertex list by one float4 (4 banks) relative to the previous
      // one, so lanes scanning different lists in lock step hit different banks
      while ((verts4.size() / 4) % 8 != static_cast<size_t>(n_lists % 8)) verts4.insert(verts4.end(), 4, 0.f);
      ++n_lists;
      d.voff = static_cast<int32_t>(verts4.size() / 4);
      d.vcnt = static_cast<int32_t>(core.size() / 3);
      for (size_t i = 0; i < core.size(); i += 3) {
        verts4.push_back(static_cast<float>(core[i])); verts4.push_back(static_cast<float>(core[i + 1]));
        verts4.push_back(static_cast<float>(core[i + 2])); verts4.push_back(0.f);
      }
    }
    d.hx = static_cast<float>(hx); d.hy = static_cast<float>(hy); d.hz = static_cast<float>(hz);
    d.bound_r = static_cast<float>(br);
    if (src.type == RAMP_SHAPE_SPHERE) {  // a sphere has no orientation: its frame is the rigid one of every pair it is in
      const double t3[3] = {pl[3], pl[7], pl[11]};
      for (int k = 0; k < 12; ++k) pl[k] = (k % 5 == 0) ? 1.0 : 0.0;
      pl[3] = t3[0]; pl[7] = t3[1]; pl[11] = t3[2];
    }
    for (int k = 0; k < 12; ++k) { d.pl[k] = static_cast<float>(pl[k]); d.opl[k] = static_cast<float>(src.placement[k]); }
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        const double gram = pl[i] * pl[j] + pl[4 + i] * pl[4 + j] + pl[8 + i] * pl[8 + j];
        nonrigid[g] = std::max(nonrigid[g], std::fabs(gram - (i == j ? 1.0 : 0.0)));
      }
    if (src.parent_joint < 0) {
      d.gslot = -1 - n_static++;
      // same 9-float form as the slab: x axis, z axis, translation
      for (int k : {0, 4, 8, 2, 6, 10, 3, 7, 11, 1, 5, 9}) stat.push_back(d.pl[k]);
    } else {
      by_joint[src.parent_joint].push_back(g);
    }
  }
  for (int i = 0; i < nj; ++i) {
    dj[i].gfirst = static_cast<int32_t>(dyn.size());
    dj[i].gcount = static_cast<int32_t>(by_joint[i].size());
    for (int g : by_joint[i]) { dg[g].gslot = static_cast<int32_t>(dyn.size()); dyn.push_back(g); }
  }
  // pairs, in processing order: DD (both joint-attached) first, then DS, then the rest
  // (two universe-attached geometries cannot collide differently per configuration; such a
  // pair is kept for index bookkeeping but handled like DS with its first member static)
  m->pair_active.resize(np, 1);
  std::vector<int> order;
  order.reserve(np);
  int n_dd = 0, ss_begin = 0;
  auto geoms_of = [&](int p, int& a, int& b) {
    const int32_t* pr = p < n_rp ? &m->robot_pairs[2 * p] : &m->scene_pairs[2 * (p - n_rp)];
    a = pr[0]; b = pr[1];
  };
  for (int pass = 0; pass < 3; ++pass) {
    if (pass == 2) ss_begin = static_cast<int>(order.size());
    for (int p = 0; p < np; ++p) {
      int a, b;
      geoms_of(p, a, b);
      if (a < 0 || b < 0 || a >= ng || b >= ng || a == b) return fail(RAMP_ERR_INVALID, "pair %d: bad geometry indices (%d, %d)", p, a, b);
      const int kind = (dg[a].gslot < 0) + (dg[b].gslot < 0);  // 0 DD, 1 DS, 2 SS
      if (kind == pass) order.push_back(p);
      n_dd += kind == 0 && pass == 0;
    }
  }
  std::vector<DPair> dp(std::max(np, 1));
  m->pair_order.assign(np, 0);
  int n_gjk = 0;
  for (int k = 0; k < np; ++k) {
    const int p = order[k];
    m->pair_order[p] = k;
    int a, b;
    geoms_of(p, a, b);
    // Which geometry is "a", the one whose frame the pair is solved in (rel = a^-1 b, the inverse
    // taken as the transpose): it only matters when a placement is not a rigid motion -- the
    // reference's own scene capsule is built from an un-normalised quaternion, |R^T R - I| = 6e-4
    // -- and then it has to be the operand the reference's arithmetic transposes, or distances come
    // out up to 2e-4 m different.  Sphere classes: the sphere (its rotation is dropped, below), as
    // coal's specialisations do.  Capsule / capsule is solved in world coordinates
    // there: here in the frame of the (more) rigid one, which is the same thing.  Box / box and
    // the GJK class: the pair's first geometry (coal's MinkowskiDiff::set).
    {
      const int cls_ab = pair_class(dg[a].type, dg[b].type);
      if (cls_ab < kClsCapsuleCapsule) { if (dg[a].type > dg[b].type) std::swap(a, b); }
      else if (cls_ab == kClsCapsuleCapsule) { if (nonrigid[a] > nonrigid[b]) std::swap(a, b); }
    }
    DPair& d = dp[k];
    std::memset(&d, 0, sizeof d);
    d.a = static_cast<int16_t>(a); d.b = static_cast<int16_t>(b);
    d.sa = static_cast<int16_t>(dg[a].gslot); d.sb = static_cast<int16_t>(dg[b].gslot);
    d.orig = static_cast<int16_t>(p);
    d.cls = static_cast<int16_t>(m->pair_active[p] ? pair_class(dg[a].type, dg[b].type) : kClsInactive);
    // stage-1 addressing: DD two slab slots; DS slab slot + static index; SS two static indices
    const auto dyn_off = [](int gslot) { return gslot * kGeomFloats * 32 * 4; };     // bytes into a lane's slab column
    const auto stat_off = [](int gslot) { return (-1 - gslot) * kStatFloats * 4; };  // bytes into the static table
    int g1 = a, g2 = b;  // the geometries o1 / o2 address
    if (k < n_dd) { d.o1 = dyn_off(d.sa); d.o2 = dyn_off(d.sb); d.kind = 0; }
    else if (k >= ss_begin) { d.o1 = stat_off(d.sa); d.o2 = stat_off(d.sb); d.kind = 2; }
    else if (dg[a].gslot >= 0) { d.o1 = dyn_off(d.sa); d.o2 = stat_off(d.sb); d.kind = 1; }
    else { d.o1 = dyn_off(d.sb); d.o2 = stat_off(d.sa); d.kind = 1; g1 = b; g2 = a; }
    n_gjk += d.cls == kClsGjk;
    const double rs = static_cast<double>(dg[a].bound_r) + dg[b].bound_r;
    // stage 1 is a cull for every class (slightly inflated); an inactive pair never passes it
    d.rs2 = d.cls == kClsInactive ? -1.f : static_cast<float>(rs * rs * (1.0 + 1e-5) + 1e-9);
    d.h1_out = dg[g1].s_out; d.h2_out = dg[g2].s_out;
    // Conservative tests are widened by 1e-5 m, plus -- when a placement is not a rigid
    // motion -- the amount by which lengths measured in its frame can be off
    const double nr = nonrigid[a] + nonrigid[b];
    d.slack = static_cast<float>(1e-5 + (nr > 1e-6 ? 2.0 * nr * rs : 0.0));
    d.rsum_out = d.cls == kClsGjk ? dg[a].r_out + dg[b].r_out + d.slack : INFINITY;
  }
  // q map
  std::vector<DMimic> dm(m->mimic_dst.size());
  for (size_t k = 0; k < dm.size(); ++k)
    dm[k] = DMimic{m->mimic_dst[k], m->mimic_src[k], static_cast<float>(m->mimic_mult[k]), static_cast<float>(m->mimic_off[k])};
  std::vector<float> refq(m->nq, 0.f);
  if (m->has_qmap) for (int k = 0; k < m->nq; ++k) refq[k] = static_cast<float>(m->ref_q[k]);

  // FK plans for teams of 2 and 4 warps.  Warp k of a team owns a contiguous run [c_k, c_k+1)
  // of the joint order (it places their geometries) and first recomputes the ancestors of that
  // run.  A plan is a sub-sequence of the joint order closed under "parent of", so the
  // running / saved parent codes of the joints stay valid inside it.  The cuts minimise the
  // largest per-warp cost (1 per joint computed, 0.75 per geometry placed).
  std::vector<int32_t> plan(kPlanEntries, 0);
  {
    std::vector<std::vector<double>> cost(nj + 1, std::vector<double>(nj + 1, 0.0));
    std::vector<char> anc(std::max(nj, 1));
    auto mark_ancestors = [&](int s0, int e0) {
      std::fill(anc.begin(), anc.end(), 0);
      int na = 0;
      for (int i = s0; i < e0; ++i)
        for (int p = m->joints[i].parent; p >= 0 && !anc[p]; p = m->joints[p].parent) { anc[p] = 1; na += p < s0; }
      return na;
    };
    for (int s0 = 0; s0 <= nj; ++s0)
      for (int e0 = s0; e0 <= nj; ++e0) {
        if (e0 == s0) continue;
        double c = mark_ancestors(s0, e0) + (e0 - s0);
        for (int i = s0; i < e0; ++i) c += 0.75 * dj[i].gcount;
        cost[s0][e0] = c;
      }
    auto emit_plan = [&](int base, const std::vector<int>& cuts) {  // cuts: T+1 joint indices
      for (size_t k = 0; k + 1 < cuts.size(); ++k) {
        plan[base + k] = static_cast<int32_t>(plan.size()) - kPlanEntries;
        mark_ancestors(cuts[k], cuts[k + 1]);
        for (int i = 0; i < cuts[k]; ++i)
          if (anc[i]) plan.push_back(i);
        for (int i = cuts[k]; i < cuts[k + 1]; ++i) plan.push_back(i | (1 << 16));
      }
      plan[base + cuts.size() - 1] = static_cast<int32_t>(plan.size()) - kPlanEntries;
    };
    int best2 = 0;
    double b2 = 1e300;
    for (int c = 0; c <= nj; ++c) {
      const double v = std::max(cost[0][c], cost[c][nj]);
      if (v < b2) { b2 = v; best2 = c; }
    }
    emit_plan(kPlanTeam2, {0, best2, nj});
    std::vector<int> best4{0, 0, 0, 0, nj};
    double b4 = 1e300;
    for (int c1 = 0; c1 <= nj; ++c1)
      for (int c2 = c1; c2 <= nj; ++c2)
        for (int c3 = c2; c3 <= nj; ++c3) {
          const double v = std::max(std::max(cost[0][c1], cost[c1][c2]), std::max(cost[c2][c3], cost[c3][nj]));
          if (v < b4) { b4 = v; best4 = {0, c1, c2, c3, nj}; }
        }
    emit_plan(kPlanTeam4, best4);
  }

  ImgHeader h;
  std::memset(&h, 0, sizeof h);
  size_t off = sizeof(ImgHeader);
  auto place = [&](size_t bytes) { const size_t o = off; off = align16(off + bytes); return static_cast<int32_t>(o); };
  h.off_joints = place(dj.size() * sizeof(DJoint));
  h.off_geoms = place(dg.size() * sizeof(DGeom));
  h.off_pairs = place(dp.size() * sizeof(DPair));
  h.off_static = place(stat.size() * sizeof(float));
  h.off_dyn = place(dyn.size() * sizeof(int32_t));
  h.off_act = place(m->act_slot.size() * sizeof(int32_t));
  h.off_mimic = place(dm.size() * sizeof(DMimic));
  h.off_cont = place(m->cont_slot.size() * sizeof(int32_t));
  h.off_refq = place(refq.size() * sizeof(float));
  h.off_plan = place(plan.size() * sizeof(int32_t));
  // the vertex lists come last: the validity and FK kernels stage the image without them
  h.off_verts = place(verts4.size() * sizeof(float));
  h.total_bytes = static_cast<int32_t>(off);
  h.nq = m->nq; h.n_joints = nj; h.n_geoms = ng; h.n_dyn = static_cast<int32_t>(dyn.size());
  h.n_static = n_static; h.n_pairs = np; h.n_verts = static_cast<int32_t>(verts4.size() / 4);
  h.n_act = static_cast<int32_t>(m->act_slot.size()); h.n_mimic = static_cast<int32_t>(dm.size());
  h.n_cont = static_cast<int32_t>(m->cont_slot.size()); h.n_save = n_save;
  h.mask_words = (np + 31) / 32; h.n_robot_geoms = n_robot; h.n_robot_pairs = n_rp; h.n_dd = n_dd; h.ss_begin = ss_begin;

  m->image.assign(off, 0);
  unsigned char* B = m->image.data();
  std::memcpy(B, &h, sizeof h);
  auto put = [&](int32_t o, const void* src, size_t bytes) { if (bytes) std::memcpy(B + o, src, bytes); };
  put(h.off_joints, dj.data(), dj.size() * sizeof(DJoint));
  put(h.off_geoms, dg.data(), dg.size() * sizeof(DGeom));
  put(h.off_pairs, dp.data(), static_cast<size_t>(np) * sizeof(DPair));
  put(h.off_verts, verts4.data(), verts4.size() * sizeof(float));
  put(h.off_static, stat.data(), stat.size() * sizeof(float));
  put(h.off_dyn, dyn.data(), dyn.size() * sizeof(int32_t));
  put(h.off_act, m->act_slot.data(), m->act_slot.size() * sizeof(int32_t));
  put(h.off_mimic, dm.data(), dm.size() * sizeof(DMimic));
  put(h.off_cont, m->cont_slot.data(), m->cont_slot.size() * sizeof(int32_t));
  put(h.off_refq, refq.data(), refq.size() * sizeof(float));
  put(h.off_plan, plan.data(), plan.size() * sizeof(int32_t));
  m->hdr = h;
  m->n_gjk_pairs = n_gjk;

  CUDA_TRY(cudaSetDevice(m->device));
  if (off > m->d_image_cap) {
    if (m->d_image) cudaFree(m->d_image);
    m->d_image = nullptr;
    m->d_image_cap = 0;
    CUDA_TRY(cudaMalloc(&m->d_image, off));
    m->d_image_cap = off;
  }
  // the image may be in use by kernels queued on caller streams
  CUDA_TRY(cudaDeviceSynchronize());
  CUDA_TRY(cudaMemcpy(m->d_image, B, off, cudaMemcpyHostToDevice));
  return RAMP_OK;
}

// Shared-memory plan: tables + as many warp slabs as fit (<= max_warps; the batch validity kernel runs at
// 106-110 registers and takes up to kCheckWarps = 18 warps per SM).
constexpr int kCheckWarps = 18;
static int plan_slabs(const RampModel* m, int stage_row_floats, int act_cols, SlabLayout* L, int* nwarps,
                      size_t* smem_bytes, int ld = 0, int team = 1, bool with_verts = false, int max_warps = 16) {
  const ImgHeader& h = m->hdr;
  int w = 0;
  auto take = [&](int words) { const int o = w; w += (words + 3) & ~3; return o; };
  // up to 3 padding columns of the input rows are loaded too, so whole rows move as float4
  // q / actuated rows are dead once FK has run; the work queue of the narrow phase reuses them
  const int q_words = std::max(std::max(h.nq, ld <= h.nq + 3 ? ld : 0), 1) * 32;
  const int act_words = std::max(std::max(act_cols, ld <= act_cols + 3 ? ld : 0), 1) * 32;
  // pairs per queue round: the team's warps take chunks in turn, so aim at an even split
  const int rounds = std::max(1, (h.n_pairs + team * kChunkPairs - 1) / (team * kChunkPairs));
  L->team = team;
  L->chunk_pairs = team == 1 ? kChunkPairs : std::min(kChunkPairs, std::max(4, ((h.n_pairs + team * rounds - 1) / (team * rounds) + 3) & ~3));
  if (const char* e = std::getenv("RAMP_B200_TEAM_CHUNK")) if (team > 1) L->chunk_pairs = std::min(kChunkPairs, std::max(4, std::atoi(e) & ~3));  // experiments
  L->queue_words = L->chunk_pairs * kLanes / 2;
  const int region0 = take(std::max(q_words + act_words, stage_row_floats > 0 ? 0 : team * L->queue_words));
  L->off_q = region0;
  L->off_act = region0 + q_words;
  const bool fk_only = stage_row_floats > 0;  // the FK kernels need no slab, queue or result rows
  L->off_geom = take(fk_only ? 0 : std::max(h.n_dyn, 1) * kGeomFloats * 32);
  // the saved transforms share their region with the TMA row prefetch of the validity kernels
  // (mbarrier + 32 raw rows); rows wider than 35 floats are loaded the plain way
  const int in_cols = std::max(q_words, act_words) / 32;
  L->raw_cols = (!fk_only && in_cols <= 35) ? in_cols : 0;
  take(4);  // the prefetch mbarrier, in front of the shared region
  L->off_save = take(std::max(std::max(h.n_save, 1) * 12 * 32, 32 * L->raw_cols));
  L->off_queue = region0;
  L->off_mask = take(fk_only ? 0 : std::max(h.mask_words, 1) * 32);
  L->off_first = take(fk_only ? 0 : 32);
  // staging pitch in float4 units, odd (see stage_xf); 16-byte aligned region
  L->stage_stride = (stage_row_floats / 4) | 1;
  w = (w + 3) & ~3;
  L->off_stage = take(stage_row_floats ? 32 * L->stage_stride * 4 : 0);
  L->words = w;
  // Only kernels that run GJK on every item keep the vertex lists in shared memory.  The validity
  // kernels hand their GJK items to the queue kernel and touch vertices only on the rare overflow
  // path, where they read them from global memory: the 8 - 26 KB this frees is one more slab
  // (warp) per CTA on Panda- and H1-sized robots.
  L->table_bytes = static_cast<int32_t>(align16(with_verts ? h.total_bytes : h.off_verts));
  const size_t fixed = 16 + L->table_bytes;
  const size_t per_warp = static_cast<size_t>(w) * 4;
  if (fixed + per_warp > static_cast<size_t>(m->max_smem))
    return fail(RAMP_ERR_LIMIT, "model needs %zu B of shared memory per CTA, device allows %d", fixed + per_warp, m->max_smem);
  const int fit = static_cast<int>((m->max_smem - fixed) / per_warp);  // slabs (= teams) that fit
  const int teams = std::min(max_warps / team, fit);
  *nwarps = teams * team;
  *smem_bytes = fixed + per_warp * teams;
  return RAMP_OK;
}

// Every kernel that takes dynamic shared memory is opted in to the device maximum ONCE per device,
// when the first model is created there: launches then only pass their own size, and host threads
// launching with different sizes cannot race on the attribute.
constexpr int kMaxDevices = 64;
static std::mutex g_attr_mu;
static bool g_attr_done[kMaxDevices] = {};

template <typename K>
static cudaError_t opt_in_smem(K kernel, int bytes) {
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

static int init_kernel_attrs(int device, int max_smem) {
  if (device < 0 || device >= kMaxDevices) return fail(RAMP_ERR_LIMIT, "device %d: only %d devices supported", device, kMaxDevices);
  std::lock_guard<std::mutex> lk(g_attr_mu);
  if (g_attr_done[device]) return RAMP_OK;
  CUDA_TRY(cudaSetDevice(device));
  CUDA_TRY(opt_in_smem(check_kernel<1>, max_smem));
  CUDA_TRY(opt_in_smem(check_kernel<2>, max_smem));
  CUDA_TRY(opt_in_smem(check_kernel<4>, max_smem));
  CUDA_TRY(opt_in_smem(gjk_queue_kernel, max_smem));
  CUDA_TRY(opt_in_smem(check_warp_kernel, max_smem));
  CUDA_TRY(opt_in_smem(recheck_kernel, max_smem));
  CUDA_TRY(opt_in_smem(fk_kernel<1>, max_smem));
  CUDA_TRY(opt_in_smem(fk_kernel<2>, max_smem));
  CUDA_TRY(opt_in_smem(distance_kernel, max_smem));
  CUDA_TRY(opt_in_smem(edge_check_kernel, max_smem));
  CUDA_TRY(opt_in_smem(edge_warp_kernel, max_smem));
  CUDA_TRY(opt_in_smem(ik_kernel, max_smem));
  g_attr_done[device] = true;
  return RAMP_OK;
}

static int grid_for(const RampModel* m, int64_t tiles, int nwarps) {
  const int64_t ctas = (tiles + nwarps - 1) / nwarps;
  return static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(m->sm_count, ctas)));
}

// Launch shape of a tile kernel: one CTA per SM with every slab that fits -- unless there are
// fewer tiles than that many slabs.  Then each CTA gets only its share of slabs (warps), so the
// tiles of a mid-size batch spread over all SMs instead of filling the first few: 16 384 Panda
// rows are 512 tiles = 37 full CTAs, or 148 CTAs of 4 warps that finish in a third of the time.
// The kernels take their warp count from blockDim.
static int spread_tiles(const RampModel* m, int64_t tiles, const SlabLayout& L, int* nwarps, size_t* smem) {
  const int team = std::max<int>(L.team, 1);
  const int slabs = *nwarps / team;
  const int share = static_cast<int>(std::max<int64_t>(1, (tiles + m->sm_count - 1) / m->sm_count));
  if (share < slabs) {
    *smem -= static_cast<size_t>(slabs - share) * L.words * 4;
    *nwarps = share * team;
  }
  return grid_for(m, tiles, *nwarps / team);
}

static int check_q_args(const RampModel* m, const void* d_q, int64_t n, int32_t ld, uint32_t flags) {
  if (!m) return fail(RAMP_ERR_INVALID, "null model");
  if (n < 0) return fail(RAMP_ERR_INVALID, "negative batch size");
  if (n > 0 && !d_q) return fail(RAMP_ERR_INVALID, "null q");
  const int cols = (flags & RAMP_Q_ACTUATED) ? m->hdr.n_act : m->hdr.nq;
  if ((flags & RAMP_Q_ACTUATED) && !m->has_qmap) return fail(RAMP_ERR_INVALID, "RAMP_Q_ACTUATED without ramp_model_set_qmap");
  if (ld < cols) return fail(RAMP_ERR_INVALID, "ld_q %d < %d columns", ld, cols);
  return RAMP_OK;
}

// Rows are processed in slices so the GJK work queue has a bounded size.
constexpr int64_t kSliceRows = 1 << 21;
// Queue capacity: 3 items per row (0.75 hard + 2.25 easy; 0.1 - 2.1 per row on the robots measured) plus a
// floor; a tile whose items do not all fit is flagged and redone by recheck_kernel.
// ramp_check_batch_noalloc sizes its queue by the same rule, so both entry points queue (and
// recheck) exactly the same items.
constexpr int64_t kQueueItemsPerRow = 3;
constexpr uint32_t queue_items_for(int64_t rows) { return static_cast<uint32_t>(rows * kQueueItemsPerRow + 4096); }
constexpr int64_t kQueueMinRows = 1024;    // below this a second launch costs more than it saves
// Queue scratch: [0, 256) counters | one flag byte per tile of 32 rows (GjkQueue::tile_flags) | items
static size_t queue_flag_bytes(int64_t rows) { return (static_cast<size_t>((rows + 31) / 32) + 255) & ~size_t{255}; }
static size_t queue_bytes(int64_t rows, uint32_t cap_items) { return 256 + queue_flag_bytes(rows) + static_cast<size_t>(cap_items) * sizeof(GjkItem); }
static GjkQueue queue_at(unsigned char* scratch, int64_t rows, uint32_t cap_items, uint32_t cap_hard) {
  return GjkQueue{cap_items ? reinterpret_cast<GjkItem*>(scratch + 256 + queue_flag_bytes(rows)) : nullptr,
                  reinterpret_cast<uint32_t*>(scratch), cap_items, cap_hard, scratch + 256};
}
// counters and flags start from zero for every launch group that uses the queue
static cudaError_t queue_reset(unsigned char* scratch, int64_t rows, cudaStream_t st) {
  return cudaMemsetAsync(scratch, 0, 256 + queue_flag_bytes(rows), st);
}

constexpr int64_t kEdgeSliceEdges = 1 << 18;  // edges per launch group of ramp_check_edges (see there)
// Up to here one warp per configuration is faster than 32 configurations per warp: 8192 rows for
// an arm (Panda 32 pairs, UR5 45: 35 us against 44 us at 4096 rows, level at 8192), fewer for a
// robot with many pairs, whose pair list takes a warp several rounds (H1, 177 pairs: 61 us
// against 56 us at 4096 rows, 111 against 58 at 8192).
static int64_t small_rows(const RampModel* m) {
  static const int64_t forced = [] { const char* e = std::getenv("RAMP_B200_SMALL_ROWS"); return e ? std::atoll(e) : int64_t{-1}; }();  // experiments
  if (forced >= 0) return forced;
  return std::max<int64_t>(1024, 8192 * 64 / std::max(64, m->hdr.n_pairs));
}

static void plan_warp_slab(const RampModel* m, int cols, WarpSlab* L) {
  const ImgHeader& h = m->hdr;
  int w = 0;
  auto take = [&](int words) { const int o = w; w += (words + 3) & ~3; return o; };
  L->off_q = take(std::max(h.nq, 1));
  L->off_act = take(std::max(std::max(h.n_act, cols), 1));
  L->off_omi = take(std::max(h.n_joints, 1) * 12);
  L->off_geom = take(std::max(h.n_dyn, 1) * kStatFloats);
  L->off_mask = take(std::max(h.mask_words, 1));
  L->words = w;
}

static int launch_check_small(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, uint8_t* d_hit,
                              int32_t* d_first, uint32_t* d_mask, cudaStream_t st) {
  SmallArgs a;
  a.img = m->d_image; a.table_bytes = static_cast<int32_t>(align16(m->hdr.total_bytes));
  a.q = d_q; a.n = n; a.ld = ld; a.flags = flags; a.hit = d_hit; a.first_pair = d_first; a.pair_mask = d_mask;
  plan_warp_slab(m, (flags & RAMP_Q_ACTUATED) ? m->hdr.n_act : m->hdr.nq, &a.L);
  const int nwarps = 4;
  const size_t smem = 16 + a.table_bytes + static_cast<size_t>(a.L.words) * 4 * nwarps;
  if (smem > static_cast<size_t>(m->max_smem)) return fail(RAMP_ERR_LIMIT, "model needs %zu B of shared memory per CTA", smem);
  const int per_sm = static_cast<int>(std::max<size_t>(1, std::min<size_t>(8, (static_cast<size_t>(m->max_smem) + 1024) / (smem + 1024))));
  const int64_t blocks = std::min<int64_t>((n + nwarps - 1) / nwarps, static_cast<int64_t>(m->sm_count) * per_sm);
  check_warp_kernel<<<static_cast<int>(blocks), nwarps * 32, smem, st>>>(a);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  CUDA_TRY(cudaGetLastError());
  return RAMP_OK;
}

// recheck_kernel behind a tile kernel: the rows of the tiles it flagged, one warp per configuration
// (the shape of launch_check_small; every CTA leaves at once when nothing was flagged)
static void launch_recheck(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, uint8_t* d_hit, int32_t* d_first,
                           uint32_t* d_mask, const uint8_t* tile_flags, const uint32_t* count, cudaStream_t st) {
  SmallArgs a;
  a.img = m->d_image; a.table_bytes = static_cast<int32_t>(align16(m->hdr.total_bytes));
  a.q = d_q; a.n = n; a.ld = ld; a.flags = flags; a.hit = d_hit; a.first_pair = d_first; a.pair_mask = d_mask;
  plan_warp_slab(m, (flags & RAMP_Q_ACTUATED) ? m->hdr.n_act : m->hdr.nq, &a.L);
  const int nwarps = 4;
  const size_t smem = 16 + a.table_bytes + static_cast<size_t>(a.L.words) * 4 * nwarps;
  if (smem > static_cast<size_t>(m->max_smem)) return;  // (such a model has no small-batch path either: create fails earlier)
  const int per_sm = static_cast<int>(std::max<size_t>(1, std::min<size_t>(8, (static_cast<size_t>(m->max_smem) + 1024) / (smem + 1024))));
  const int64_t parts = ((n + 31) / 32) * 4;
  const int64_t blocks = std::min<int64_t>((parts + nwarps - 1) / nwarps, static_cast<int64_t>(m->sm_count) * per_sm);
  recheck_kernel<<<static_cast<int>(blocks), nwarps * 32, smem, st>>>(a, tile_flags, count);
  g_launches.fetch_add(1, std::memory_order_relaxed);
}

// One ramp_check_batch is: plan (slab layout, team size, shared-memory opt-in), a GJK queue from
// the model's pool, then per slice the tile kernel and the queue kernel.  The pieces are
// separate because ramp_check_host feeds ONE queue from the tile kernels of several chunks and
// runs the queue kernel once behind them.
struct CheckPlan {
  CheckArgs a;  // img, ld, flags, L; the per-launch fields are filled by launch_tiles
  int nwarps, team;
  size_t smem, smem_b;
  bool use_queue;
};

static int plan_check(const RampModel* m, int32_t ld, uint32_t flags, int64_t n, CheckPlan* P) {
  CheckArgs& a = P->a;
  a.img = m->d_image; a.ld = ld; a.flags = flags; a.row_offset = 0;
  a.gq = GjkQueue{nullptr, nullptr, 0u, 0u, nullptr};
  // no actuated rows in this call: their part of the slab is not reserved (one more warp on Panda-sized robots)
  const int act_cols = (flags & RAMP_Q_ACTUATED) ? m->hdr.n_act : 0;
  int nwarps = 0; size_t smem = 0;
  if (int rc = plan_slabs(m, 0, act_cols, &a.L, &nwarps, &smem, ld, 1, false, kCheckWarps)) return rc;
  // Few slabs fit (a robot with many geometries): let teams of 2 or 4 warps share each slab and
  // split the kinematic tree and the pair list between them -- whichever team size puts the most
  // warps on an SM (H1 stand-in, 20 hulls: 6 slabs -> 6 / 12 / 16 warps: 4.6e8 / 7.6e8 / 8.1e8
  // configurations/s).  Robots whose slabs fit 12+ times are ~20 % faster without teams.
  int team = 1;
  if (nwarps < 12) {
    SlabLayout L2, L4;
    int w2 = 0, w4 = 0;
    size_t s2 = 0, s4 = 0;
    const bool ok2 = plan_slabs(m, 0, act_cols, &L2, &w2, &s2, ld, 2, false, kCheckWarps) == RAMP_OK;
    const bool ok4 = plan_slabs(m, 0, act_cols, &L4, &w4, &s4, ld, 4, false, kCheckWarps) == RAMP_OK;
    if (ok2 && w2 > nwarps) team = 2;
    if (ok4 && w4 > (team == 2 ? w2 : nwarps)) team = 4;
  }
  if (const char* e = std::getenv("RAMP_B200_TEAM")) { team = std::atoi(e); team = team >= 4 ? 4 : (team >= 2 ? 2 : 1); }  // experiments
  if (team > 1 && plan_slabs(m, 0, act_cols, &a.L, &nwarps, &smem, ld, team, false, kCheckWarps) != RAMP_OK) return RAMP_ERR_LIMIT;
  P->nwarps = nwarps; P->team = team; P->smem = smem;
  P->use_queue = m->pool != nullptr && m->n_gjk_pairs > 0 && n >= kQueueMinRows;
  P->smem_b = 16 + align16(m->hdr.total_bytes);  // the queue kernel stages the whole image, vertex lists included
  return RAMP_OK;
}

// a queue for up to `rows` configurations, from the model's pool, in stream order
static int queue_alloc(const RampModel* m, int64_t rows, cudaStream_t st, unsigned char** scratch, GjkQueue* gq) {
  uint32_t cap_items = queue_items_for(rows);
  // test hook: a tiny queue forces the overflow path (flagged tiles, recheck_kernel)
  if (const char* e = std::getenv("RAMP_B200_QUEUE_CAP")) cap_items = static_cast<uint32_t>(std::max(1, std::atoi(e)));
  const size_t bytes = queue_bytes(rows, cap_items);
  if (cudaMallocFromPoolAsync(reinterpret_cast<void**>(scratch), bytes, m->pool, st) != cudaSuccess) {
    cudaGetLastError();
    return fail(RAMP_ERR_ALLOC, "cudaMallocFromPoolAsync(%zu) for the GJK queue failed", bytes);
  }
  *gq = queue_at(*scratch, rows, cap_items, queue_cap_hard(cap_items));
  return RAMP_OK;
}

// tile kernel over ns rows; queued items carry row_offset + their row, results go to hit/first/mask
static void launch_tiles(const RampModel* m, const CheckPlan& P, const float* d_q, int64_t ns, uint32_t row_offset, uint8_t* hit,
                         int32_t* first, uint32_t* mask, const GjkQueue& gq, cudaStream_t st) {
  CheckArgs a = P.a;
  a.q = d_q; a.n = ns; a.hit = hit; a.first_pair = first; a.pair_mask = mask; a.gq = gq; a.row_offset = row_offset;
  int nwarps = P.nwarps;
  size_t smem = P.smem;
  const int grid = spread_tiles(m, (ns + 31) / 32, a.L, &nwarps, &smem);
  if (P.team == 2) check_kernel<2><<<grid, nwarps * 32, smem, st>>>(a);
  else if (P.team == 4) check_kernel<4><<<grid, nwarps * 32, smem, st>>>(a);
  else check_kernel<1><<<grid, nwarps * 32, smem, st>>>(a);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (gq.tile_flags != nullptr) launch_recheck(m, d_q, ns, a.ld, a.flags, hit, first, mask, gq.tile_flags + (row_offset >> 5), gq.count, st);
}

// queue kernel; hit/first/mask are the arrays the items' row indices refer to
static void launch_queue(const RampModel* m, const CheckPlan& P, const GjkQueue& gq, uint8_t* hit, int32_t* first, uint32_t* mask,
                         int64_t rows, cudaStream_t st) {
  GjkQueueArgs b{};
  b.img = m->d_image; b.table_bytes = static_cast<int32_t>(align16(m->hdr.total_bytes));
  b.items = gq.items; b.count = gq.count; b.capacity = gq.capacity; b.cap_hard = gq.cap_hard; b.stats = m->d_stats;
  b.mask_words = m->hdr.mask_words;
  b.stop = !(P.a.flags & RAMP_ALL_PAIRS) && mask == nullptr;
  b.hit = hit; b.first_pair = first; b.pair_mask = mask; b.n_rows = static_cast<uint32_t>(rows);
  gjk_queue_kernel<<<m->sm_count * kGjkCtasPerSm, 256, P.smem_b, st>>>(b);
  g_launches.fetch_add(1, std::memory_order_relaxed);
}

// Per-call scratch in stream order: from the model's pool (kept between calls, so no driver
// allocation after the first call of a size), else from the device's default pool.  Calls on
// different streams or host threads get different blocks -- nothing here is shared.
static int pool_alloc(const RampModel* m, size_t bytes, cudaStream_t st, void** out) {
  const cudaError_t e = m->pool ? cudaMallocFromPoolAsync(out, bytes, m->pool, st) : cudaMallocAsync(out, bytes, st);
  if (e != cudaSuccess) {
    cudaGetLastError();
    *out = nullptr;
    return fail(RAMP_ERR_ALLOC, "stream-ordered allocation of %zu B failed: %s", bytes, cudaGetErrorString(e));
  }
  return RAMP_OK;
}

// Slices and queue capacity of one ramp_check_batch call.  Up to kSliceRows rows: one slice, 3 items
// per row (the rule ramp_check_batch_noalloc shares).  Larger batches: every slice ends with the
// drain of its queue kernel (warps finishing their last long items at low occupancy), so fewer,
// larger slices are faster -- Panda 2^24 rows: 3.97e9 configurations/s in 2^21-row slices, 4.31e9 in
// one slice -- but 3 items per row of queue would then be gigabytes.  The items a robot really
// produces per row are learnt from the counters of earlier calls (0.1 for a capsule-like arm, 2 for
// a humanoid); the queue is sized 1.5x that within kQueueBudgetBytes and the slice is as large as
// that allows (at most 2^26 rows).  An underestimate is not an error: the tiles whose items do not
// fit are redone by recheck_kernel (slowly: one warp per configuration).
constexpr size_t kQueueBudgetBytes = size_t{1} << 30;
constexpr int64_t kMaxSliceRows = int64_t{1} << 26;

static void plan_slices(const RampModel* m, int64_t n, int64_t* slice_rows, uint32_t* cap_items, uint32_t* cap_hard) {
  *slice_rows = std::min(n, kSliceRows);
  *cap_items = queue_items_for(*slice_rows);
  *cap_hard = queue_cap_hard(*cap_items);
  if (n <= kSliceRows) return;
  std::lock_guard<std::mutex> lk(m->est_mu);
  if (m->counts_pending && cudaEventQuery(m->ev_counts) == cudaSuccess) {
    const float rows = static_cast<float>(std::max<int64_t>(m->counts_rows, 1));
    m->est_hard = std::max(m->h_counts[0] / rows, 0.9f * m->est_hard);
    m->est_easy = std::max(m->h_counts[1] / rows, 0.9f * m->est_easy);
    m->counts_pending = false;
  } else {
    cudaGetLastError();  // cudaErrorNotReady is not an error
  }
  if (const char* e = std::getenv("RAMP_B200_SLICE_LOG2")) {  // experiments: fixed slice, conservative capacity
    *slice_rows = std::min(n, int64_t{1} << std::atoi(e));
    *cap_items = queue_items_for(*slice_rows);
    *cap_hard = queue_cap_hard(*cap_items);
    return;
  }
  if (m->est_hard < 0.f) return;  // nothing observed yet: the conservative rule
  // (rounded up to sixteenths: repeated calls must ask the pool for the SAME size, or every call
  // would allocate afresh -- the observed rates wobble in the fourth digit)
  auto up16 = [](double x) { return std::ceil(x * 16.0) / 16.0; };
  const double per_hard = std::min(2.0, up16(1.5 * m->est_hard + 0.02)), per_easy = std::min(3.5, up16(1.5 * m->est_easy + 0.04));
  const double bytes_per_row = (per_hard + per_easy) * sizeof(GjkItem);
  int64_t rows = static_cast<int64_t>(kQueueBudgetBytes / bytes_per_row) & ~int64_t{(1 << 18) - 1};
  rows = std::max(kSliceRows, std::min(rows, kMaxSliceRows));
  *slice_rows = std::min(n, rows);
  // (capacity for the slice rounded up to 2^20 rows: batches of slightly different sizes reuse one block)
  const double cap_rows = static_cast<double>((*slice_rows + (1 << 20) - 1) & ~int64_t{(1 << 20) - 1});
  *cap_hard = static_cast<uint32_t>(cap_rows * per_hard) + 2048;
  *cap_items = *cap_hard + static_cast<uint32_t>(cap_rows * per_easy) + 2048;
}

static int launch_check(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, uint8_t* d_hit,
                        int32_t* d_first, uint32_t* d_mask, cudaStream_t st) {
  if (n <= small_rows(m)) return launch_check_small(m, d_q, n, ld, flags, d_hit, d_first, d_mask, st);
  CheckPlan P;
  if (int rc = plan_check(m, ld, flags, n, &P)) return rc;
  // (the tile kernel hands its GJK work to the queue; without one -- no memory pool on this device --
  // a robot that has GJK pairs takes the one-warp-per-configuration kernel whatever the batch size)
  if (!P.use_queue && m->n_gjk_pairs > 0) return launch_check_small(m, d_q, n, ld, flags, d_hit, d_first, d_mask, st);
  unsigned char* scratch = nullptr;
  GjkQueue gq{nullptr, nullptr, 0u, 0u, nullptr};
  int64_t step = n;
  if (P.use_queue) {
    uint32_t cap_items, cap_hard;
    plan_slices(m, n, &step, &cap_items, &cap_hard);
    if (const char* e = std::getenv("RAMP_B200_QUEUE_CAP")) {  // test hook: a tiny queue forces the overflow path
      cap_items = static_cast<uint32_t>(std::max(1, std::atoi(e)));
      cap_hard = queue_cap_hard(cap_items);
    }
    void* p = nullptr;
    if (int rc = pool_alloc(m, queue_bytes(step, cap_items), st, &p)) return rc;
    scratch = static_cast<unsigned char*>(p);
    gq = queue_at(scratch, step, cap_items, cap_hard);
  }
  int64_t last_rows = 0;
  for (int64_t off = 0; off < n; off += step) {
    const int64_t ns = std::min(step, n - off);
    uint8_t* hit = d_hit + off;
    int32_t* first = d_first ? d_first + off : nullptr;
    uint32_t* mask = d_mask ? d_mask + off * m->hdr.mask_words : nullptr;
    if (P.use_queue) CUDA_TRY(queue_reset(scratch, step, st));
    launch_tiles(m, P, d_q + off * ld, ns, 0u, hit, first, mask, gq, st);
    if (P.use_queue) launch_queue(m, P, gq, hit, first, mask, ns, st);
    last_rows = ns;
  }
  if (P.use_queue && n > kSliceRows) {  // remember what this robot produced per row (read by a later call)
    std::lock_guard<std::mutex> lk(m->est_mu);
    if (!m->counts_pending && m->h_counts && m->ev_counts) {
      CUDA_TRY(cudaMemcpyAsync(m->h_counts, scratch, 8, cudaMemcpyDeviceToHost, st));
      CUDA_TRY(cudaEventRecord(m->ev_counts, st));
      m->counts_pending = true;
      m->counts_rows = last_rows;
    }
  }
  if (scratch) CUDA_TRY(cudaFreeAsync(scratch, st));
  CUDA_TRY(cudaGetLastError());
  return RAMP_OK;
}

static int ensure_scratch(const RampModel* m, size_t bytes) {
  if (bytes <= m->scratch_cap) return RAMP_OK;
  if (m->d_scratch) cudaFree(m->d_scratch);
  m->d_scratch = nullptr; m->scratch_cap = 0;
  const size_t cap = std::max<size_t>(bytes, 1 << 20);
  if (cudaMalloc(&m->d_scratch, cap) != cudaSuccess) return fail(RAMP_ERR_ALLOC, "cudaMalloc(%zu) for scratch failed", cap);
  m->scratch_cap = cap;
  return RAMP_OK;
}

static int launch_min_distance(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, float* d_min,
                               cudaStream_t st);

}  // namespace ramp

using namespace ramp;

// ================================================================================ C ABI
extern "C" {

int ramp_abi_version(void) { return RAMP_B200_ABI_VERSION; }
const char* ramp_last_error(void) { return g_err; }
int64_t ramp_launch_count(void) { return g_launches.load(); }
#ifdef RAMP_STATS
int ramp_debug_image(const RampModel* m, unsigned char* out, int cap) {
  const int n = static_cast<int>(m->image.size());
  if (out && cap >= n) std::memcpy(out, m->image.data(), n);
  return n;
}
}  // extern "C"
namespace ramp {
// single-thread trace of the exact-mode GJK for one pair (debug builds only)
__global__ void debug_gjk_kernel(const float4* va, int na, const float4* vb, int nb, Xf rel, float* trace, int max_it,
                                 float rsum, V3 v0, int use_v0, int exact) {
  GjkShape A, B;
  A.type = SH_CONVEX; A.vcnt = na; A.p0 = A.p1 = A.p2 = 0.f; A.verts = va;
  B.type = SH_CONVEX; B.vcnt = nb; B.p0 = B.p1 = B.p2 = 0.f; B.verts = vb;
  GjkLane g;
  gjk_init(g, rel, rsum, use_v0 ? v0 : -trans(rel));
  float dist = 0.f;
  int st = GJK_CONTINUE, it = 0;
  while (st == GJK_CONTINUE && it < max_it) {
    st = gjk_step(g, A, B, exact != 0, dist);
    trace[8 * it + 0] = g.vv; trace[8 * it + 1] = g.s.n; trace[8 * it + 2] = st; trace[8 * it + 3] = dist;
    trace[8 * it + 4] = g.v.x; trace[8 * it + 5] = g.v.y; trace[8 * it + 6] = g.v.z; trace[8 * it + 7] = g.lower;
    ++it;
  }
}
}  // namespace ramp
extern "C" {
int ramp_debug_gjk(const float* h_va, int na, const float* h_vb, int nb, const float* h_rel, float* h_trace, int max_it,
                   float rsum, const float* v0, int exact) {
  float4 *da, *db; float* dt;
  cudaMalloc(&da, 16 * na); cudaMalloc(&db, 16 * nb); cudaMalloc(&dt, 32 * max_it);
  cudaMemcpy(da, h_va, 16 * na, cudaMemcpyHostToDevice); cudaMemcpy(db, h_vb, 16 * nb, cudaMemcpyHostToDevice);
  cudaMemset(dt, 0, 32 * max_it);
  ramp::Xf rel; for (int k = 0; k < 12; ++k) rel.m[k] = h_rel[k];
  ramp::V3 vv0{v0 ? v0[0] : 0.f, v0 ? v0[1] : 0.f, v0 ? v0[2] : 0.f};
  ramp::debug_gjk_kernel<<<1, 1>>>(da, na, db, nb, rel, dt, max_it, rsum, vv0, v0 != nullptr, exact);
  cudaMemcpy(h_trace, dt, 32 * max_it, cudaMemcpyDeviceToHost);
  cudaFree(da); cudaFree(db); cudaFree(dt);
  return 0;
}
#endif

#ifdef RAMP_STATS
int ramp_debug_stats(unsigned long long* out, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, g_stats, sizeof(unsigned long long) * 64);
  if (reset) { unsigned long long z[64] = {0}; cudaMemcpyToSymbol(g_stats, z, sizeof z); }
  return 0;
}
#endif

int ramp_model_create(const RampModelDesc* d, int device, RampModel** out) {
  if (!d || !out) return fail(RAMP_ERR_INVALID, "null argument");
  *out = nullptr;
  if (d->abi_version != RAMP_B200_ABI_VERSION) return fail(RAMP_ERR_INVALID, "ABI version %d, library is %d", d->abi_version, RAMP_B200_ABI_VERSION);
  if (d->n_joints < 0 || d->n_geoms < 0 || d->n_pairs < 0 || d->n_verts < 0 || d->nq < 0) return fail(RAMP_ERR_INVALID, "negative count");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(RAMP_ERR_CUDA, "no CUDA device (there is no CPU fallback)");
  if (device < 0 || device >= ndev) return fail(RAMP_ERR_INVALID, "device %d out of range (%d devices)", device, ndev);
  RampModel* m = new (std::nothrow) RampModel();
  if (!m) return fail(RAMP_ERR_ALLOC, "out of host memory");
  m->device = device;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete m; return fail(RAMP_ERR_CUDA, "cudaGetDeviceProperties failed"); }
  m->sm_count = prop.multiProcessorCount;
  m->max_smem = static_cast<int>(prop.sharedMemPerBlockOptin);
  if (int rc = init_kernel_attrs(device, m->max_smem)) { delete m; return rc; }
  {  // a private stream-ordered pool that keeps its memory between calls
    cudaMemPoolProps pp{};
    pp.allocType = cudaMemAllocationTypePinned;
    pp.handleTypes = cudaMemHandleTypeNone;
    pp.location.type = cudaMemLocationTypeDevice;
    pp.location.id = device;
    cudaSetDevice(device);
    if (cudaMemPoolCreate(&m->pool, &pp) == cudaSuccess) {
      uint64_t keep = UINT64_MAX;
      cudaMemPoolSetAttribute(m->pool, cudaMemPoolAttrReleaseThreshold, &keep);
    } else {
      cudaGetLastError();
      m->pool = nullptr;  // no pool: GJK stays inside the validity kernel
    }
  }
  m->nq = d->nq;
  m->joints.assign(d->joints, d->joints + d->n_joints);
  m->robot_geoms.assign(d->geoms, d->geoms + d->n_geoms);
  m->robot_pairs.assign(d->pairs, d->pairs + 2 * static_cast<size_t>(d->n_pairs));
  m->robot_verts.assign(d->verts, d->verts + 3 * static_cast<size_t>(d->n_verts));
  for (int i = 0; i < d->n_joints; ++i)
    if (d->joints[i].idx_q < 0 || d->joints[i].idx_q >= std::max(d->nq, 1)) { delete m; return fail(RAMP_ERR_INVALID, "joint %d: idx_q out of range", i); }
  if (cudaMalloc(&m->d_stats, 16) != cudaSuccess || cudaMemset(m->d_stats, 0, 16) != cudaSuccess) {
    cudaGetLastError();
    ramp_model_destroy(m);
    return fail(RAMP_ERR_ALLOC, "cudaMalloc for the model counters failed");
  }
  if (cudaHostAlloc(reinterpret_cast<void**>(&m->h_counts), 16, cudaHostAllocPortable) != cudaSuccess ||
      cudaEventCreateWithFlags(&m->ev_counts, cudaEventDisableTiming) != cudaSuccess) {
    cudaGetLastError();
    ramp_model_destroy(m);
    return fail(RAMP_ERR_ALLOC, "allocation of the model's counter read-back failed");
  }
  if (int rc = build_image(m)) { ramp_model_destroy(m); return rc; }
  *out = m;
  return RAMP_OK;
}

int ramp_model_set_scene(RampModel* m, const RampGeom* geoms, int32_t n_geoms, const int32_t* pairs, int32_t n_pairs,
                         const double* verts, int32_t n_verts) {
  if (!m || n_geoms < 0 || n_pairs < 0 || n_verts < 0) return fail(RAMP_ERR_INVALID, "bad argument");
  if ((n_geoms > 0 && !geoms) || (n_pairs > 0 && !pairs) || (n_verts > 0 && !verts)) return fail(RAMP_ERR_INVALID, "null array");
  // validate before touching the model: a rejected scene leaves the previous one in place
  for (int32_t g = 0; g < n_geoms; ++g)
    if (geoms[g].parent_joint != -1) return fail(RAMP_ERR_INVALID, "scene objects must be attached to the universe");
  std::lock_guard<std::mutex> lk(m->mu);
  m->scene_geoms.assign(geoms, geoms + n_geoms);
  m->scene_pairs.assign(pairs, pairs + 2 * static_cast<size_t>(n_pairs));
  m->scene_verts.assign(verts, verts + 3 * static_cast<size_t>(n_verts));
  m->pair_active.resize(m->robot_pairs.size() / 2);  // scene pairs start active
  return build_image(m);
}

int ramp_model_set_qmap(RampModel* m, const RampQMap* q) {
  if (!m || !q) return fail(RAMP_ERR_INVALID, "null argument");
  if (q->n_act < 0 || q->n_act > kMaxNq || q->n_mimic < 0 || q->n_cont < 0) return fail(RAMP_ERR_INVALID, "bad qmap counts");
  for (int k = 0; k < q->n_act; ++k) if (q->act_slot[k] < 0 || q->act_slot[k] >= m->nq) return fail(RAMP_ERR_INVALID, "act_slot[%d] out of range", k);
  for (int k = 0; k < q->n_mimic; ++k)
    if (q->mimic_dst[k] < 0 || q->mimic_dst[k] >= m->nq || q->mimic_src[k] < 0 || q->mimic_src[k] >= m->nq) return fail(RAMP_ERR_INVALID, "mimic[%d] out of range", k);
  for (int k = 0; k < q->n_cont; ++k) if (q->cont_slot[k] < 0 || q->cont_slot[k] + 1 >= m->nq) return fail(RAMP_ERR_INVALID, "cont_slot[%d] out of range", k);
  std::lock_guard<std::mutex> lk(m->mu);
  {  // the same map again (a planner re-installs its reference state before every query): nothing to rebuild
    auto same_i = [](const std::vector<int32_t>& v, const int32_t* p, int n) { return static_cast<int>(v.size()) == n && std::equal(v.begin(), v.end(), p); };
    auto same_d = [](const std::vector<double>& v, const double* p, int n) { return static_cast<int>(v.size()) == n && std::equal(v.begin(), v.end(), p); };
    if (m->has_qmap && same_i(m->act_slot, q->act_slot, q->n_act) && same_i(m->mimic_dst, q->mimic_dst, q->n_mimic) &&
        same_i(m->mimic_src, q->mimic_src, q->n_mimic) && same_d(m->mimic_mult, q->mimic_mult, q->n_mimic) &&
        same_d(m->mimic_off, q->mimic_off, q->n_mimic) && same_i(m->cont_slot, q->cont_slot, q->n_cont) &&
        same_d(m->ref_q, q->ref_q, m->nq))
      return RAMP_OK;
  }
  m->has_qmap = true;
  m->act_slot.assign(q->act_slot, q->act_slot + q->n_act);
  m->mimic_dst.assign(q->mimic_dst, q->mimic_dst + q->n_mimic);
  m->mimic_src.assign(q->mimic_src, q->mimic_src + q->n_mimic);
  m->mimic_mult.assign(q->mimic_mult, q->mimic_mult + q->n_mimic);
  m->mimic_off.assign(q->mimic_off, q->mimic_off + q->n_mimic);
  m->cont_slot.assign(q->cont_slot, q->cont_slot + q->n_cont);
  m->ref_q.assign(q->ref_q, q->ref_q + m->nq);
  return build_image(m);
}

int ramp_model_set_pair_active(RampModel* m, int32_t p, int32_t active) {
  if (!m) return fail(RAMP_ERR_INVALID, "null model");
  std::lock_guard<std::mutex> lk(m->mu);
  if (p < 0 || p >= static_cast<int32_t>(m->pair_active.size())) return fail(RAMP_ERR_INVALID, "pair index %d out of range", p);
  if (m->pair_active[p] == (active ? 1 : 0)) return RAMP_OK;  // already so
  m->pair_active[p] = active ? 1 : 0;
  return build_image(m);
}

void ramp_model_destroy(RampModel* m) {
  if (!m) return;
  cudaSetDevice(m->device);
  cudaDeviceSynchronize();
  if (m->d_image) cudaFree(m->d_image);
  if (m->d_scratch) cudaFree(m->d_scratch);
  if (m->d_stats) cudaFree(m->d_stats);
  if (m->h_counts) cudaFreeHost(m->h_counts);
  if (m->ev_counts) cudaEventDestroy(m->ev_counts);
  if (m->pool) cudaMemPoolDestroy(m->pool);
  for (cudaStream_t st : m->streams) if (st) cudaStreamDestroy(st);
  for (cudaEvent_t ev : m->events) if (ev) cudaEventDestroy(ev);
  if (m->h_pinned) cudaFreeHost(m->h_pinned);
  if (m->d_edge_stage) cudaFree(m->d_edge_stage);
  if (m->h_edge_stage) cudaFreeHost(m->h_edge_stage);
  delete m;
}

int ramp_model_trim(RampModel* m) {
  if (!m) return fail(RAMP_ERR_INVALID, "null model");
  CUDA_TRY(cudaSetDevice(m->device));
  CUDA_TRY(cudaDeviceSynchronize());
  if (m->pool) CUDA_TRY(cudaMemPoolTrimTo(m->pool, 0));
  return RAMP_OK;
}

int ramp_model_counters(const RampModel* m, uint64_t out[2], int32_t reset) {
  if (!m || !out) return fail(RAMP_ERR_INVALID, "null argument");
  CUDA_TRY(cudaSetDevice(m->device));
  CUDA_TRY(cudaDeviceSynchronize());
  unsigned long long v[2] = {0, 0};
  CUDA_TRY(cudaMemcpy(v, m->d_stats, sizeof v, cudaMemcpyDeviceToHost));
  out[0] = v[0]; out[1] = v[1];
  if (reset) CUDA_TRY(cudaMemset(m->d_stats, 0, sizeof v));
  return RAMP_OK;
}

int ramp_model_info(const RampModel* m, int32_t out[6]) {
  if (!m || !out) return fail(RAMP_ERR_INVALID, "null argument");
  out[0] = m->hdr.nq; out[1] = m->hdr.n_joints; out[2] = m->hdr.n_geoms; out[3] = m->hdr.n_pairs;
  out[4] = m->has_qmap ? m->hdr.n_act : 0; out[5] = m->device;
  return RAMP_OK;
}

static int fk_common(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, float* d_out, void* stream,
                     int what) {
  if (int rc = check_q_args(m, d_q, n, ld, flags)) return rc;
  if (n == 0) return RAMP_OK;
  if (!d_out) return fail(RAMP_ERR_INVALID, "null output");
  CUDA_TRY(cudaSetDevice(m->device));
  FkArgs a;
  a.img = m->d_image; a.q = d_q; a.n = n; a.ld = ld; a.flags = flags; a.out = d_out;
  a.row_floats = 12 * (what == 1 ? m->hdr.n_joints : m->hdr.n_geoms);
  if (a.row_floats == 0) return RAMP_OK;
  int nwarps = 0; size_t smem = 0;
  if (int rc = plan_slabs(m, a.row_floats, m->hdr.n_act, &a.L, &nwarps, &smem, ld)) return rc;
  const int grid = spread_tiles(m, (n + 31) / 32, a.L, &nwarps, &smem);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (what == 1) {
    fk_kernel<1><<<grid, nwarps * 32, smem, st>>>(a);
  } else {
    fk_kernel<2><<<grid, nwarps * 32, smem, st>>>(a);
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  CUDA_TRY(cudaGetLastError());
  return RAMP_OK;
}

int ramp_fk_batch(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, float* d_oMi, void* stream) {
  return fk_common(m, d_q, n, ld, flags, d_oMi, stream, 1);
}

int ramp_placements_batch(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, float* d_oMg,
                          void* stream) {
  return fk_common(m, d_q, n, ld, flags, d_oMg, stream, 2);
}

int ramp_check_batch(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, uint8_t* d_hit,
                     int32_t* d_first_pair, float* d_min_dist, uint32_t* d_pair_mask, void* stream) {
  if (int rc = check_q_args(m, d_q, n, ld, flags)) return rc;
  if (n == 0) return RAMP_OK;
  if (!d_hit) return fail(RAMP_ERR_INVALID, "null d_hit");
  CUDA_TRY(cudaSetDevice(m->device));
  if (int rc = launch_check(m, d_q, n, ld, flags, d_hit, d_first_pair, d_pair_mask, static_cast<cudaStream_t>(stream))) return rc;
  return d_min_dist ? launch_min_distance(m, d_q, n, ld, flags, d_min_dist, static_cast<cudaStream_t>(stream)) : RAMP_OK;
}

// ---- ramp_check_batch_noalloc: the validity check for callers that must not allocate.
//
// Some hosts of this library run where device allocation is off limits: a planner thread
// inside a real-time control loop (cudaMalloc / cudaFree can block on every other stream of
// the process and their latency is unbounded), or a process whose device memory was carved
// up once at start-up and has nothing left to hand out.  ramp_check_batch needs a work queue
// for its GJK items and takes it from a stream-ordered pool; this variant takes the queue
// from the fixed array below instead and calls no allocating routine at all.
//
// The array lives in the library image: every GPU that loads libramp_b200.so carries these
// kNoAllocQueueBytes (about 6.5 MB) until the process exits, whether or not the entry point
// is ever used.  That is why the limit is modest.
constexpr int64_t kNoAllocMaxRows = RAMP_NOALLOC_MAX_ROWS;
constexpr uint32_t kNoAllocQueueItems = queue_items_for(kNoAllocMaxRows);
constexpr size_t kNoAllocQueueBytes = 256 + ((static_cast<size_t>((kNoAllocMaxRows + 31) / 32) + 255) & ~size_t{255}) + static_cast<size_t>(kNoAllocQueueItems) * sizeof(GjkItem);  // counters | tile flags | items
static __device__ __align__(256) unsigned char g_noalloc_queue[kNoAllocQueueBytes];

// In host code `g_noalloc_queue` is only a handle for the symbol, not an address a kernel
// could use: the device address differs per GPU and has to be asked for, once per device,
// with that device current.
static unsigned char* g_noalloc_ptr[kMaxDevices] = {};
static std::mutex g_noalloc_mu;

static int noalloc_queue_for(int device, unsigned char** out) {
  if (device < 0 || device >= kMaxDevices) return fail(RAMP_ERR_LIMIT, "device %d: only %d devices supported", device, kMaxDevices);
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    CUDA_TRY(cudaSetDevice(device));
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_queue);
    if (e != cudaSuccess || p == nullptr)
      return fail(RAMP_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_queue) on device %d: %s", device, cudaGetErrorString(e));
    g_noalloc_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_ptr[device];
  return RAMP_OK;
}

int ramp_check_batch_noalloc(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, uint8_t* d_hit,
                             int32_t* d_first_pair, uint32_t* d_pair_mask, void* stream) {
  if (int rc = check_q_args(m, d_q, n, ld, flags)) return rc;
  if (n > kNoAllocMaxRows)
    return fail(RAMP_ERR_LIMIT, "ramp_check_batch_noalloc: %lld configurations exceed the fixed scratch (max %lld per call); "
                "split the batch or use ramp_check_batch", static_cast<long long>(n), static_cast<long long>(kNoAllocMaxRows));
  if (n == 0) return RAMP_OK;
  if (!d_hit) return fail(RAMP_ERR_INVALID, "null d_hit");
  CUDA_TRY(cudaSetDevice(m->device));
  unsigned char* scratch = nullptr;
  if (int rc = noalloc_queue_for(m->device, &scratch)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CheckArgs a;
  a.row_offset = 0;
  a.img = m->d_image; a.q = d_q; a.n = n; a.ld = ld; a.flags = flags;
  a.hit = d_hit; a.first_pair = d_first_pair; a.pair_mask = d_pair_mask;
  int nwarps = 0; size_t smem = 0;
  if (int rc = plan_slabs(m, 0, m->hdr.n_act, &a.L, &nwarps, &smem, ld)) return rc;
  const size_t smem_b = 16 + align16(m->hdr.total_bytes);
  // the array still holds the previous call's counter and items: only the counter has to
  // start from zero, the items below it are rewritten before they are read
  CUDA_TRY(queue_reset(scratch, kNoAllocMaxRows, st));
  // the capacity rule of ramp_check_batch for this many rows (not the whole array): both entry
  // points then settle every item in the same place and give bit-identical results
  const uint32_t cap_items = queue_items_for(n);
  a.gq = queue_at(scratch, kNoAllocMaxRows, cap_items, queue_cap_hard(cap_items));
  const int grid = spread_tiles(m, (n + 31) / 32, a.L, &nwarps, &smem);
  check_kernel<1><<<grid, nwarps * 32, smem, st>>>(a);
  launch_recheck(m, d_q, n, ld, flags, d_hit, d_first_pair, d_pair_mask, a.gq.tile_flags, a.gq.count, st);
  GjkQueueArgs b{};
  b.img = m->d_image; b.table_bytes = static_cast<int32_t>(align16(m->hdr.total_bytes));
  b.items = a.gq.items; b.count = a.gq.count; b.capacity = cap_items; b.cap_hard = a.gq.cap_hard; b.stats = m->d_stats;
  b.hit = d_hit; b.first_pair = d_first_pair; b.pair_mask = d_pair_mask; b.mask_words = m->hdr.mask_words;
  b.stop = !(flags & RAMP_ALL_PAIRS) && d_pair_mask == nullptr;
  b.n_rows = static_cast<uint32_t>(n);
  gjk_queue_kernel<<<m->sm_count * kGjkCtasPerSm, 256, smem_b, st>>>(b);
  g_launches.fetch_add(2, std::memory_order_relaxed);  // (+1 counted by launch_recheck)
  CUDA_TRY(cudaGetLastError());
  return RAMP_OK;
}

}  // extern "C"

namespace ramp {
// smallest signed distance over the active pairs per configuration (the d_min_dist output):
// the distance kernel in reduce mode -- every pair to convergence, so far slower than the flags
static int launch_min_distance(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, float* d_min,
                               cudaStream_t st) {
  DistArgs a;
  a.img = m->d_image; a.q = d_q; a.n = n; a.ld = ld; a.flags = flags; a.n_sel = m->hdr.n_pairs; a.out = nullptr; a.sel = nullptr;
  a.min_out = d_min;
  int nwarps = 0; size_t smem = 0;
  if (int rc = plan_slabs(m, 0, m->hdr.n_act, &a.L, &nwarps, &smem, ld, 1, true)) return rc;
  nwarps = std::min(nwarps, 8);
  smem = 16 + a.L.table_bytes + static_cast<size_t>(a.L.words) * 4 * nwarps;
  const int grid = spread_tiles(m, (n + 31) / 32, a.L, &nwarps, &smem);
  distance_kernel<<<grid, nwarps * 32, smem, st>>>(a);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  CUDA_TRY(cudaGetLastError());
  return RAMP_OK;
}
}  // namespace ramp

extern "C" {

int ramp_pack_flags(const RampModel* m, const uint8_t* d_hit, int64_t n, uint32_t* d_words, void* stream) {
  if (!m) return fail(RAMP_ERR_INVALID, "null model");
  if (n < 0) return fail(RAMP_ERR_INVALID, "negative count");
  if (n == 0) return RAMP_OK;
  if (!d_hit || !d_words) return fail(RAMP_ERR_INVALID, "null device pointer");
  CUDA_TRY(cudaSetDevice(m->device));
  const int64_t warps = (n + 31) / 32;
  const int blocks = static_cast<int>(std::min<int64_t>((warps + 7) / 8, static_cast<int64_t>(m->sm_count) * 8));
  pack_flags_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(d_hit, n, d_words);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  CUDA_TRY(cudaGetLastError());
  return RAMP_OK;
}

int ramp_ik_batch(const RampModel* m, const float* d_seed, int64_t n, int32_t ld, const float* d_target, int32_t one_target,
                  const RampIkParams* p, float* d_q_out, float* d_err, uint8_t* d_ok, int32_t* d_iters, void* stream) {
  if (!m || !p) return fail(RAMP_ERR_INVALID, "null argument");
  if (!m->has_qmap) return fail(RAMP_ERR_INVALID, "ramp_ik_batch needs ramp_model_set_qmap (the group's actuated columns)");
  if (n < 0) return fail(RAMP_ERR_INVALID, "negative batch size");
  if (n == 0) return RAMP_OK;
  if (!d_seed || !d_target || !d_q_out) return fail(RAMP_ERR_INVALID, "null device pointer");
  const int n_act = m->hdr.n_act, nj = m->hdr.n_joints;
  if (ld < n_act) return fail(RAMP_ERR_INVALID, "ld %d < %d actuated columns", ld, n_act);
  if (n_act > kIkMaxCols) return fail(RAMP_ERR_LIMIT, "%d actuated columns > limit %d", n_act, kIkMaxCols);
  if (p->frame_parent_joint < -1 || p->frame_parent_joint >= nj) return fail(RAMP_ERR_INVALID, "frame parent joint %d out of range", p->frame_parent_joint);
  if (p->max_iterations < 0 || !(p->tolerance > 0) || !(p->damping >= 0)) return fail(RAMP_ERR_INVALID, "bad IK parameters");
  IkArgs a;
  std::memset(&a, 0, sizeof a);
  auto joint_of_slot = [&](int slot) {
    for (int j = 0; j < nj; ++j)
      if (m->joints[j].idx_q == slot) return j;
    return -1;
  };
  auto one_dof = [&](int j) {
    const int t = m->joints[j].type;
    return t <= RAMP_JOINT_PRISMATIC_UNALIGNED;  // revolute, unbounded revolute, prismatic; aligned or not
  };
  for (int k = 0; k < n_act; ++k) {
    const int j = joint_of_slot(m->act_slot[k]);
    if (j < 0 || !one_dof(j))
      return fail(RAMP_ERR_INVALID, "actuated column %d does not drive a one-degree-of-freedom joint: not supported by ramp_ik_batch", k);
    a.col_joint[k] = j;
    a.lo[k] = p->lower ? static_cast<float>(p->lower[k]) : -INFINITY;
    a.hi[k] = p->upper ? static_cast<float>(p->upper[k]) : INFINITY;
  }
  for (size_t i = 0; i < m->mimic_dst.size(); ++i) {
    const int jd = joint_of_slot(m->mimic_dst[i]);
    int col = -1;
    for (int k = 0; k < n_act; ++k) if (m->act_slot[k] == m->mimic_src[i]) col = k;
    if (jd < 0 || col < 0) continue;  // the source is not driven by this group
    if (!one_dof(jd)) return fail(RAMP_ERR_INVALID, "mimic joint %d is not a one-degree-of-freedom joint", jd);
    if (a.n_mimic >= kIkMaxMimic) return fail(RAMP_ERR_LIMIT, "more than %d mimic joints", kIkMaxMimic);
    a.mim_col[a.n_mimic] = col; a.mim_joint[a.n_mimic] = jd; a.mim_mult[a.n_mimic] = static_cast<float>(m->mimic_mult[i]);
    ++a.n_mimic;
  }
  for (int j = p->frame_parent_joint; j >= 0; j = m->joints[j].parent) a.chain |= 1ull << j;
  a.img = m->d_image; a.seed = d_seed; a.n = n; a.ld = ld; a.target = d_target; a.target_stride = one_target ? 0 : 12;
  a.q_out = d_q_out; a.err_out = d_err; a.ok = d_ok; a.iters_out = d_iters;
  a.frame_joint = p->frame_parent_joint;
  for (int k = 0; k < 12; ++k) a.frame_pl[k] = static_cast<float>(p->frame_placement[k]);
  a.max_iter = p->max_iterations;
  a.damping2 = static_cast<float>(p->damping * p->damping);
  a.step = static_cast<float>(p->step > 0 ? p->step : 1.0);
  a.tol = static_cast<float>(p->tolerance);
  a.max_dq = static_cast<float>(p->max_step > 0 ? p->max_step : 1e30);
  a.n_cols = n_act;
  CUDA_TRY(cudaSetDevice(m->device));
  int nwarps = 0; size_t smem = 0;
  if (int rc = plan_slabs(m, 12 * nj, n_act, &a.L, &nwarps, &smem, ld)) return rc;
  nwarps = std::min(nwarps, 16);
  // IK batches are small (hundreds to thousands of seeds) and every seed iterates tens of times:
  // spread the seeds over the SMs rather than filling CTAs
  const int64_t tiles = (n + 31) / 32;
  const int warps_per_cta = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(nwarps, (tiles + m->sm_count - 1) / m->sm_count)));
  const int grid = static_cast<int>(std::min<int64_t>((tiles + warps_per_cta - 1) / warps_per_cta, m->sm_count));
  smem = 16 + a.L.table_bytes + static_cast<size_t>(a.L.words) * 4 * warps_per_cta;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void* cursor = nullptr;  // the seed cursor the warps draw from
  if (int rc = pool_alloc(m, 16, st, &cursor)) return rc;
  CUDA_TRY(cudaMemsetAsync(cursor, 0, 16, st));
  ik_kernel<<<grid, warps_per_cta * 32, smem, st>>>(a, static_cast<unsigned long long*>(cursor));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  CUDA_TRY(cudaFreeAsync(cursor, st));
  CUDA_TRY(cudaGetLastError());
  return RAMP_OK;
}

int ramp_distance_batch(const RampModel* m, const float* d_q, int64_t n, int32_t ld, uint32_t flags, const int32_t* h_pair_idx,
                        int32_t n_sel, float* d_dist, void* stream) {
  if (int rc = check_q_args(m, d_q, n, ld, flags)) return rc;
  const int np = m->hdr.n_pairs;
  if (!h_pair_idx) n_sel = np;
  if (n_sel < 0 || n_sel > np) return fail(RAMP_ERR_INVALID, "n_sel %d out of range (%d pairs)", n_sel, np);
  if (n == 0 || n_sel == 0) return RAMP_OK;
  if (!d_dist) return fail(RAMP_ERR_INVALID, "null d_dist");
  if (h_pair_idx)
    for (int k = 0; k < n_sel; ++k)
      if (h_pair_idx[k] < 0 || h_pair_idx[k] >= np) return fail(RAMP_ERR_INVALID, "pair index %d out of range", h_pair_idx[k]);
  CUDA_TRY(cudaSetDevice(m->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DistArgs a;
  a.img = m->d_image; a.q = d_q; a.n = n; a.ld = ld; a.flags = flags; a.n_sel = n_sel; a.out = d_dist; a.sel = nullptr;
  a.min_out = nullptr;
  int32_t* d_sel = nullptr;
  if (h_pair_idx) {
    if (!m->pool) return fail(RAMP_ERR_ALLOC, "no memory pool for the pair selection");
    if (cudaMallocFromPoolAsync(reinterpret_cast<void**>(&d_sel), sizeof(int32_t) * n_sel, m->pool, st) != cudaSuccess) {
      cudaGetLastError();
      return fail(RAMP_ERR_ALLOC, "cudaMallocFromPoolAsync for the pair selection failed");
    }
    std::vector<int32_t> sel(n_sel);
    for (int k = 0; k < n_sel; ++k) sel[k] = m->pair_order[h_pair_idx[k]];
    CUDA_TRY(cudaMemcpyAsync(d_sel, sel.data(), sizeof(int32_t) * n_sel, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaStreamSynchronize(st));  // `sel` is a pageable temporary
    a.sel = d_sel;
  }
  int nwarps = 0; size_t smem = 0;
  if (int rc = plan_slabs(m, 0, m->hdr.n_act, &a.L, &nwarps, &smem, ld, 1, true)) return rc;
  nwarps = std::min(nwarps, 8);
  smem = 16 + a.L.table_bytes + static_cast<size_t>(a.L.words) * 4 * nwarps;
  const int grid = spread_tiles(m, (n + 31) / 32, a.L, &nwarps, &smem);
  distance_kernel<<<grid, nwarps * 32, smem, st>>>(a);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (d_sel) CUDA_TRY(cudaFreeAsync(d_sel, st));
  CUDA_TRY(cudaGetLastError());
  return RAMP_OK;
}

int ramp_check_edges(const RampModel* m, const float* d_q0, const float* d_q1, int64_t n_edges, int32_t ld,
                     const RampSpaceDesc* sp, uint8_t* d_valid, int32_t* d_first_invalid, int32_t* d_nd, void* stream) {
  if (!m || !sp) return fail(RAMP_ERR_INVALID, "null argument");
  if (!m->has_qmap) return fail(RAMP_ERR_INVALID, "ramp_check_edges needs ramp_model_set_qmap");
  if (n_edges < 0) return fail(RAMP_ERR_INVALID, "negative edge count");
  if (n_edges == 0) return RAMP_OK;
  if (!d_q0 || !d_q1 || !d_valid) return fail(RAMP_ERR_INVALID, "null device pointer");
  if (sp->n_sub < 0 || sp->n_sub > kMaxSubspaces) return fail(RAMP_ERR_LIMIT, "%d subspaces > limit %d", sp->n_sub, kMaxSubspaces);
  EdgeArgs a;
  std::memset(&a.sp, 0, sizeof a.sp);
  const double frac = sp->longest_valid_segment_fraction > 0 ? sp->longest_valid_segment_fraction : 0.01;
  int n_cols = 0;
  for (int i = 0; i < sp->n_sub; ++i) {
    const RampSubspace& s = sp->sub[i];
    DSub& d = a.sp.sub[i];
    d.type = s.type; d.col = s.col; d.dim = s.dim;
    int nreal = 0, width = 0;
    double rot_extent = 0;
    switch (s.type) {
      case RAMP_SPACE_REAL: nreal = s.dim; width = s.dim; break;
      case RAMP_SPACE_SO2: width = 1; rot_extent = M_PI; break;
      case RAMP_SPACE_SE2: nreal = 2; width = 3; rot_extent = M_PI; break;
      case RAMP_SPACE_SE3: nreal = 3; width = 7; rot_extent = 0.5 * M_PI; break;
      case RAMP_SPACE_SO3: width = 4; rot_extent = 0.5 * M_PI; break;
      default: return fail(RAMP_ERR_INVALID, "subspace %d: unknown type %d", i, s.type);
    }
    if (nreal < 0 || nreal > 3 || s.col < 0) return fail(RAMP_ERR_INVALID, "subspace %d: bad dim/col", i);
    double ext = 0;
    for (int k = 0; k < nreal; ++k) { d.lo[k] = static_cast<float>(s.lo[k]); d.hi[k] = static_cast<float>(s.hi[k]); ext += (s.hi[k] - s.lo[k]) * (s.hi[k] - s.lo[k]); }
    ext = std::sqrt(ext);
    d.inv_lvs_real = ext > 0 ? static_cast<float>(1.0 / (ext * frac)) : 0.f;
    d.inv_lvs_rot = rot_extent > 0 ? static_cast<float>(1.0 / (rot_extent * frac)) : 0.f;
    n_cols = std::max(n_cols, s.col + width);
  }
  if (n_cols != m->hdr.n_act) return fail(RAMP_ERR_INVALID, "space covers %d columns, qmap has %d actuated values", n_cols, m->hdr.n_act);
  if (ld < n_cols) return fail(RAMP_ERR_INVALID, "ld %d < %d columns", ld, n_cols);
  a.sp.n_sub = sp->n_sub; a.sp.check_bounds = sp->check_bounds; a.sp.n_cols = n_cols;
  CUDA_TRY(cudaSetDevice(m->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (n_edges <= kSmallEdges) {  // a planner's edge-by-edge calls: one launch
    void* work = nullptr;
    if (int rc = pool_alloc(m, 256, st, &work)) return rc;
    EdgeSmallArgs b;
    b.img = m->d_image; b.table_bytes = static_cast<int32_t>(align16(m->hdr.total_bytes));
    b.q0 = d_q0; b.q1 = d_q1; b.n_edges = static_cast<int32_t>(n_edges); b.ld = ld;
    b.valid = d_valid; b.first_invalid = d_first_invalid; b.nd = d_nd;
    b.work = static_cast<int32_t*>(work);
    b.sp = a.sp;
    plan_warp_slab(m, n_cols, &b.L);
    const int nw = 4;
    const size_t smem = 16 + b.table_bytes + static_cast<size_t>(b.L.words) * 4 * nw;
    if (smem > static_cast<size_t>(m->max_smem)) return fail(RAMP_ERR_LIMIT, "model needs %zu B of shared memory per CTA", smem);
    CUDA_TRY(cudaMemsetAsync(b.work, 0x7f, (kSmallEdges + 1) * sizeof(int32_t), st));
    // enough warps for a few hundred states at once; every warp loops over its share
    edge_warp_kernel<<<std::min(m->sm_count, 128), nw * 32, smem, st>>>(b);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CUDA_TRY(cudaFreeAsync(work, st));
    CUDA_TRY(cudaGetLastError());
    return RAMP_OK;
  }
  // Edges are processed in slices of 2^18.  The flat index of an interpolated state travels as 32
  // bits in the GJK work items: a slice stays below 2^32 states for any mean segment count up to
  // 16 383 (edges with both ends inside the bounds have at most ~101); one that still exceeds it is
  // not checked at all -- every one of its edges is reported invalid with first_invalid = -2
  // (edge_scan_kernel / edge_finish_kernel), which no valid result can be.  And the GJK queue of a
  // slice stays a few hundred MB: 2^18 planner edges are ~4M states and ~0.15 items per state on an
  // arm (FR3 stand-in, 2^20 edges: 2.4e9 states/s in one slice whose queue overflowed into the
  // in-kernel path, 4.1e9 in slices of 2^18).
  int nwarps = 0; size_t smem = 0;
  if (int rc = plan_slabs(m, 0, n_cols, &a.L, &nwarps, &smem)) return rc;
  const size_t smem_b = 16 + align16(m->hdr.total_bytes);
  const int64_t slice = std::min<int64_t>(n_edges, kEdgeSliceEdges);
  // scratch: offs int64[n+1] | nd int32[n] | first int32[n] | totals of the blocks of 256 edges int64[n/256]
  const size_t off_nd = align16(sizeof(int64_t) * (slice + 1));
  const size_t off_first = off_nd + align16(sizeof(int32_t) * slice);
  const size_t off_blk = off_first + align16(sizeof(int32_t) * slice);
  void* edge_scratch = nullptr;
  if (int rc = pool_alloc(m, off_blk + align16(sizeof(int64_t) * ((slice + 255) / 256)), st, &edge_scratch)) return rc;
  unsigned char* S = static_cast<unsigned char*>(edge_scratch);
  // GJK work queue (only worth a launch when there are enough edges to fill warps)
  void* qmem = nullptr;
  const bool use_queue = m->n_gjk_pairs > 0 && n_edges >= 64;
  uint32_t cap = 0;
  if (use_queue) {
    cap = static_cast<uint32_t>(std::min<int64_t>(std::max<int64_t>(1 << 16, slice * 64), 1 << 22));
    if (int rc = pool_alloc(m, 256 + static_cast<size_t>(cap) * sizeof(GjkItem), st, &qmem)) {
      cudaFreeAsync(edge_scratch, st);
      return rc;
    }
  }
  a.img = m->d_image; a.ld = ld;
  a.offs = reinterpret_cast<int64_t*>(S);
  a.first_invalid = reinterpret_cast<int32_t*>(S + off_first);
  for (int64_t e0 = 0; e0 < n_edges; e0 += slice) {
    const int64_t ne = std::min(slice, n_edges - e0);
    a.q0 = d_q0 + e0 * ld; a.q1 = d_q1 + e0 * ld; a.n_edges = ne; a.valid = d_valid + e0;
    a.nd = d_nd ? d_nd + e0 : reinterpret_cast<int32_t*>(S + off_nd);
    a.gq = GjkQueue{nullptr, nullptr, 0u, 0u, nullptr};
    GjkQueueArgs b{};
    if (use_queue) {
      unsigned char* Q = static_cast<unsigned char*>(qmem);
      CUDA_TRY(cudaMemsetAsync(Q, 0, 16, st));
      a.gq = GjkQueue{reinterpret_cast<GjkItem*>(Q + 256), reinterpret_cast<uint32_t*>(Q), cap, queue_cap_hard(cap), nullptr};  // (the edge kernel solves what does not fit in place)
      b.img = m->d_image; b.table_bytes = static_cast<int32_t>(align16(m->hdr.total_bytes)); b.items = a.gq.items; b.count = a.gq.count; b.capacity = cap; b.cap_hard = a.gq.cap_hard; b.stats = m->d_stats;
      b.offs = a.offs; b.nd = a.nd; b.first_invalid = a.first_invalid; b.n_edges = ne;
    }
    const int eb = static_cast<int>((ne + 255) / 256);
    int64_t* blk = reinterpret_cast<int64_t*>(S + off_blk);
    edge_count_kernel<<<eb, 256, 0, st>>>(a, blk);
    edge_scan_kernel<<<1, 1024, 0, st>>>(blk, eb, a.offs);
    if (eb > 1) edge_offsets_kernel<<<eb, 256, 0, st>>>(a.offs, blk, ne);
    // the number of states is only known on the device: size the grid for the SM count
    edge_check_kernel<<<m->sm_count, nwarps * 32, smem, st>>>(a);
    if (use_queue) gjk_queue_kernel<<<m->sm_count * kGjkCtasPerSm, 256, smem_b, st>>>(b);
    edge_finish_kernel<<<eb, 256, 0, st>>>(a, d_first_invalid ? d_first_invalid + e0 : nullptr, d_nd ? d_nd + e0 : nullptr);
    g_launches.fetch_add((use_queue ? 5 : 4) + (eb > 1 ? 1 : 0), std::memory_order_relaxed);
  }
  if (qmem) CUDA_TRY(cudaFreeAsync(qmem, st));
  CUDA_TRY(cudaFreeAsync(edge_scratch, st));
  CUDA_TRY(cudaGetLastError());
  return RAMP_OK;
}

// Host-buffer form of ramp_check_edges for a planner's edge-by-edge calls: end points in,
// verdict out, one synchronisation.  Copies go through the model's pinned staging buffer.
int ramp_check_edges_host(const RampModel* m, const float* h_q0, const float* h_q1, int64_t n_edges, int32_t ld,
                          const RampSpaceDesc* sp, uint8_t* h_valid, int32_t* h_first_invalid, int32_t* h_nd) {
  if (!m || !sp) return fail(RAMP_ERR_INVALID, "null argument");
  if (n_edges < 0) return fail(RAMP_ERR_INVALID, "negative edge count");
  if (n_edges == 0) return RAMP_OK;
  if (!h_q0 || !h_q1 || !h_valid) return fail(RAMP_ERR_INVALID, "null host pointer");
  CUDA_TRY(cudaSetDevice(m->device));
  const size_t b_q = align16(sizeof(float) * n_edges * ld), b_v = align16(n_edges), b_i = align16(sizeof(int32_t) * n_edges);
  const size_t total = 2 * b_q + b_v + 2 * b_i;
  // the staging buffers are the model's: one host-buffer call at a time per model
  std::lock_guard<std::mutex> lk(m->mu);
  if (total > m->edge_stage_cap) {
    if (m->d_edge_stage) cudaFree(m->d_edge_stage);
    if (m->h_edge_stage) cudaFreeHost(m->h_edge_stage);
    m->d_edge_stage = nullptr; m->h_edge_stage = nullptr; m->edge_stage_cap = 0;
    const size_t cap = std::max<size_t>(total, 1 << 16);
    if (cudaMalloc(&m->d_edge_stage, cap) != cudaSuccess || cudaMallocHost(&m->h_edge_stage, cap) != cudaSuccess)
      return fail(RAMP_ERR_ALLOC, "staging allocation (%zu B) failed", cap);
    m->edge_stage_cap = cap;
  }
  unsigned char* D = static_cast<unsigned char*>(m->d_edge_stage);
  unsigned char* H = static_cast<unsigned char*>(m->h_edge_stage);
  cudaStream_t st = cudaStreamPerThread;
  std::memcpy(H, h_q0, sizeof(float) * n_edges * ld);
  std::memcpy(H + b_q, h_q1, sizeof(float) * n_edges * ld);
  CUDA_TRY(cudaMemcpyAsync(D, H, 2 * b_q, cudaMemcpyHostToDevice, st));
  uint8_t* d_valid = D + 2 * b_q;
  int32_t* d_first = reinterpret_cast<int32_t*>(D + 2 * b_q + b_v);
  int32_t* d_nd = reinterpret_cast<int32_t*>(D + 2 * b_q + b_v + b_i);
  if (int rc = ramp_check_edges(m, reinterpret_cast<const float*>(D), reinterpret_cast<const float*>(D + b_q), n_edges, ld, sp,
                                d_valid, d_first, d_nd, st))
    return rc;
  CUDA_TRY(cudaMemcpyAsync(H + 2 * b_q, D + 2 * b_q, b_v + 2 * b_i, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  std::memcpy(h_valid, H + 2 * b_q, n_edges);
  if (h_first_invalid) std::memcpy(h_first_invalid, H + 2 * b_q + b_v, sizeof(int32_t) * n_edges);
  if (h_nd) std::memcpy(h_nd, H + 2 * b_q + b_v + b_i, sizeof(int32_t) * n_edges);
  return RAMP_OK;
}

static bool is_pinned(const void* p) {
  cudaPointerAttributes at{};
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost;
}

void* ramp_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, std::max<size_t>(bytes, 1), cudaHostAllocPortable) != cudaSuccess) {
    cudaGetLastError();
    fail(RAMP_ERR_ALLOC, "cudaHostAlloc(%zu) failed", bytes);
    return nullptr;
  }
  return p;
}

void ramp_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

// Host-buffer check, pipelined.  The rows move to the device in chunks on kHostStreams streams (and
// as many q buffers), so the H2D copies of the next chunks run under the tile kernel of chunk k --
// with four of them the copy engine keeps going while a queue kernel holds the SMs (UR5, 2^24 rows:
// 1.85e9 -> 2.05e9 configurations/s against two).  The tile kernels of all chunks
// of a section (up to 2^21 rows) feed ONE GJK queue; a third stream runs the queue kernel once
// behind them and copies the section's results home -- a queue kernel per chunk would pay its
// drain (warps finishing their last long items at low occupancy) every 2^18 rows.  Buffers
// the caller already pinned (ramp_host_alloc / cudaHostAlloc / cudaHostRegister / torch
// pin_memory) are used for DMA directly; pageable ones are staged through the model's pinned
// buffers.  (Reading a pinned q in place from the kernels was measured too: 35-46 GB/s over
// PCIe against 55 GB/s for the copy engine, and slower end to end at every batch size.)
constexpr int kHostStreams = 4;  // chunk streams (and q buffers) of ramp_check_host: slack for the copy engine while a queue kernel holds the SMs

int ramp_check_host(const RampModel* m, const float* h_q, int64_t n, int32_t ld, uint32_t flags, uint8_t* h_hit,
                    int32_t* h_first_pair, float* h_min_dist, uint32_t* h_pair_mask) {
  if (int rc = check_q_args(m, h_q, n, ld, flags)) return rc;
  if (n == 0) return RAMP_OK;
  if (!h_hit) return fail(RAMP_ERR_INVALID, "null h_hit");
  if (h_min_dist && !m->pool) return fail(RAMP_ERR_ALLOC, "no memory pool for the distance output");
  CUDA_TRY(cudaSetDevice(m->device));
  std::lock_guard<std::mutex> lk(m->mu);
  const int mw = std::max(m->hdr.mask_words, 1);
  // A SECTION is what one queue kernel covers (up to kSliceRows rows); its rows arrive in CHUNKS.
  // A batch that fits one slice is still cut into three sections (of at least 2^18 rows): the queue
  // kernels of the first two run under the copies of the next section and only the last third's
  // queue kernel is exposed at the end (UR5, 2^20 rows: 0.726 -> 0.694 ms; more sections lose to
  // their queue kernels' drains).
  int64_t section = std::min<int64_t>(n, kSliceRows);
  std::vector<int64_t> sec_start;  // section boundaries (the last entry is n)
  if (n <= kSliceRows) {
    const int64_t even = std::min<int64_t>(n, std::max<int64_t>(1 << 18, ((n + 2) / 3 + 31) & ~int64_t{31}));
    std::vector<double> frac(static_cast<size_t>((n + even - 1) / even), static_cast<double>(even) / n);
    int64_t min_rows = 1;
    if (const char* e = std::getenv("RAMP_B200_SECTIONS")) {  // experiments: comma-separated shares
      frac.clear();
      min_rows = 1 << 15;
      for (const char* c = e; *c;) { char* end; frac.push_back(std::strtod(c, &end)); c = *end ? end + 1 : end; }
    }
    sec_start.push_back(0);
    double acc = 0;
    for (size_t i = 0; i + 1 < frac.size(); ++i) {
      acc += frac[i];
      const int64_t b = (static_cast<int64_t>(acc * n) + 31) & ~int64_t{31};
      if (b - sec_start.back() >= min_rows && n - b >= min_rows) sec_start.push_back(b);
    }
    sec_start.push_back(n);
    section = 0;
    for (size_t i = 0; i + 1 < sec_start.size(); ++i) section = std::max(section, sec_start[i + 1] - sec_start[i]);
  } else {
    for (int64_t b = 0; b < n; b += section) sec_start.push_back(b);
    sec_start.push_back(n);
  }
  // Chunk size: 2^18 rows, or for batches of several slices about 16 MB of rows (a power of two
  // between 2^17 and 2^19 rows): the copy engine reaches 53 GB/s on copies of that size and 45-50
  // on 6 MB ones (UR5, 24 B per row, 2^24 rows: 2.05e9 -> 2.13e9 configurations/s).
  int64_t chunk = 1 << 18;
  if (n > kSliceRows) {
    const int64_t row_bytes = static_cast<int64_t>(sizeof(float)) * ld;
    chunk = 1 << 17;
    while (chunk < (1 << 19) && 2 * chunk * row_bytes <= (20 << 20)) chunk *= 2;
  }
  chunk = std::min(section, chunk);
  CheckPlan P;
  const bool tiled = n > small_rows(m);
  if (tiled)
    if (int rc = plan_check(m, ld, flags, n, &P)) return rc;
  const bool queued = tiled && P.use_queue;
  if (tiled && !queued && m->n_gjk_pairs > 0) return fail(RAMP_ERR_ALLOC, "ramp_check_host needs the device's memory pool for batches of this size");
  // device: two q chunks, then two sets of result arrays (one per section parity)
  const size_t b_q = align16(sizeof(float) * chunk * ld), b_hit = align16(section), b_first = align16(sizeof(int32_t) * section),
               b_mask = align16(sizeof(uint32_t) * section * mw);
  const size_t b_out = b_hit + b_first + b_mask, total = kHostStreams * b_q + 2 * b_out;
  if (int rc = ensure_scratch(m, total)) return rc;
  const bool pin_q = is_pinned(h_q), pin_hit = is_pinned(h_hit);
  const bool pin_first = !h_first_pair || is_pinned(h_first_pair), pin_mask = !h_pair_mask || is_pinned(h_pair_mask);
  const bool staged = !(pin_q && pin_hit && pin_first && pin_mask);
  if (staged && total > m->pinned_cap) {
    if (m->h_pinned) cudaFreeHost(m->h_pinned);
    m->h_pinned = nullptr; m->pinned_cap = 0;
    const size_t cap = std::max<size_t>(total, 1 << 16);
    if (cudaMallocHost(&m->h_pinned, cap) != cudaSuccess) return fail(RAMP_ERR_ALLOC, "cudaMallocHost(%zu) failed", cap);
    m->pinned_cap = cap;
  }
  for (auto& st : m->streams)
    if (!st) CUDA_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  for (auto& ev : m->events)
    if (!ev) CUDA_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  cudaStream_t s_tail = m->streams[kHostStreams];
  cudaEvent_t ev_queue = m->events[0], ev_out[2] = {m->events[kHostStreams + 1], m->events[kHostStreams + 2]};
  cudaEvent_t* ev_tiles = m->events + 1;
  unsigned char* D = static_cast<unsigned char*>(m->d_scratch);
  unsigned char* H = static_cast<unsigned char*>(m->h_pinned);
  int64_t k = 0;  // chunk counter over the whole call: picks the stream and the q buffer
  int64_t nsec = 0;
  for (; nsec + 1 < static_cast<int64_t>(sec_start.size()); ++nsec) {
    const int64_t s0 = sec_start[nsec], s1 = sec_start[nsec + 1], rows = s1 - s0;
    const int par = static_cast<int>(nsec & 1);
    unsigned char* Dout = D + kHostStreams * b_q + par * b_out;
    uint8_t* d_hit = Dout;
    int32_t* d_first = h_first_pair ? reinterpret_cast<int32_t*>(Dout + b_hit) : nullptr;
    uint32_t* d_mask = h_pair_mask ? reinterpret_cast<uint32_t*>(Dout + b_hit + b_first) : nullptr;
    // the result arrays of this parity were last copied out by section nsec - 2
    if (nsec >= 2)
      for (int i = 0; i < kHostStreams; ++i) CUDA_TRY(cudaStreamWaitEvent(m->streams[i], ev_out[par], 0));
    unsigned char* qmem = nullptr;
    GjkQueue gq{nullptr, nullptr, 0u, 0u, nullptr};
    if (queued) {
      if (int rc = queue_alloc(m, rows, m->streams[0], &qmem, &gq)) return rc;
      CUDA_TRY(queue_reset(qmem, rows, m->streams[0]));
      CUDA_TRY(cudaEventRecord(ev_queue, m->streams[0]));
      for (int i = 1; i < kHostStreams; ++i) CUDA_TRY(cudaStreamWaitEvent(m->streams[i], ev_queue, 0));
    }
    // Chunk boundaries.  The kernels of the very first chunk cannot start before its copy has
    // landed and nothing hides that copy, so it is shorter (n/8, at least 2^16 rows: smaller
    // chunks cost more in kernel tails than they save); the rest are full chunks.
    std::vector<int64_t> start{s0};
    if (s0 == 0) {
      const int64_t first = std::min<int64_t>(chunk, std::max<int64_t>(1 << 16, (n / 8 + 31) & ~int64_t{31}));
      if (first < rows) start.push_back(first);
    }
    while (start.back() < s1) start.push_back(std::min(s1, start.back() + chunk));
    for (size_t c = 0; c + 1 < start.size(); ++c, ++k) {
      const int b = static_cast<int>(k % kHostStreams);
      cudaStream_t st = m->streams[b];
      const int64_t off = start[c], ns = start[c + 1] - off, rel = off - s0;
      float* d_q = reinterpret_cast<float*>(D + b * b_q);
      const float* src = h_q + off * ld;
      if (!pin_q) {  // pageable q: through the pinned staging buffer the copy of chunk k - 2 read
        if (k >= kHostStreams) CUDA_TRY(cudaStreamSynchronize(st));
        std::memcpy(H + b * b_q, src, sizeof(float) * ns * ld);
        src = reinterpret_cast<const float*>(H + b * b_q);
      }
      CUDA_TRY(cudaMemcpyAsync(d_q, src, sizeof(float) * ns * ld, cudaMemcpyHostToDevice, st));
      uint8_t* c_hit = d_hit + rel;
      int32_t* c_first = d_first ? d_first + rel : nullptr;
      uint32_t* c_mask = d_mask ? d_mask + rel * mw : nullptr;
      if (tiled) launch_tiles(m, P, d_q, ns, static_cast<uint32_t>(rel), c_hit, c_first, c_mask, gq, st);
      else if (int rc = launch_check(m, d_q, ns, ld, flags, c_hit, c_first, c_mask, st)) return rc;
    }
    // tail of the section, behind the tile kernels of both chunk streams: ONE queue kernel for
    // all its chunks, then the results go home
    for (int i = 0; i < kHostStreams; ++i) {
      CUDA_TRY(cudaEventRecord(ev_tiles[i], m->streams[i]));
      CUDA_TRY(cudaStreamWaitEvent(s_tail, ev_tiles[i], 0));
    }
    if (queued) {
      launch_queue(m, P, gq, d_hit, d_first, d_mask, rows, s_tail);
      CUDA_TRY(cudaFreeAsync(qmem, s_tail));
    }
    unsigned char* Hout = staged ? H + kHostStreams * b_q + par * b_out : nullptr;
    CUDA_TRY(cudaMemcpyAsync(pin_hit ? h_hit + s0 : Hout, d_hit, rows, cudaMemcpyDeviceToHost, s_tail));
    if (h_first_pair)
      CUDA_TRY(cudaMemcpyAsync(pin_first ? reinterpret_cast<unsigned char*>(h_first_pair + s0) : Hout + b_hit, d_first,
                               sizeof(int32_t) * rows, cudaMemcpyDeviceToHost, s_tail));
    if (h_pair_mask)
      CUDA_TRY(cudaMemcpyAsync(pin_mask ? reinterpret_cast<unsigned char*>(h_pair_mask + s0 * mw) : Hout + b_hit + b_first, d_mask,
                               sizeof(uint32_t) * rows * mw, cudaMemcpyDeviceToHost, s_tail));
    CUDA_TRY(cudaEventRecord(ev_out[par], s_tail));
    if (staged) {  // pageable results: wait and copy them out of the staging buffer
      CUDA_TRY(cudaStreamSynchronize(s_tail));
      if (!pin_hit) std::memcpy(h_hit + s0, Hout, rows);
      if (h_first_pair && !pin_first) std::memcpy(h_first_pair + s0, Hout + b_hit, sizeof(int32_t) * rows);
      if (h_pair_mask && !pin_mask) std::memcpy(h_pair_mask + s0 * mw, Hout + b_hit + b_first, sizeof(uint32_t) * rows * mw);
    }
  }
  CUDA_TRY(cudaStreamSynchronize(s_tail));
  if (h_min_dist) {
    // smallest signed distance per configuration: the (slow, every pair to convergence) distance
    // kernel in reduce mode, chunk by chunk through the first q buffer
    float* d_min = nullptr;
    if (cudaMallocFromPoolAsync(reinterpret_cast<void**>(&d_min), sizeof(float) * chunk, m->pool, s_tail) != cudaSuccess) {
      cudaGetLastError();
      return fail(RAMP_ERR_ALLOC, "cudaMallocFromPoolAsync for the distance output failed");
    }
    for (int64_t off = 0; off < n; off += chunk) {
      const int64_t ns = std::min(chunk, n - off);
      CUDA_TRY(cudaMemcpyAsync(D, h_q + off * ld, sizeof(float) * ns * ld, cudaMemcpyHostToDevice, s_tail));
      if (int rc = launch_min_distance(m, reinterpret_cast<const float*>(D), ns, ld, flags, d_min, s_tail)) return rc;
      CUDA_TRY(cudaMemcpyAsync(h_min_dist + off, d_min, sizeof(float) * ns, cudaMemcpyDeviceToHost, s_tail));
      CUDA_TRY(cudaStreamSynchronize(s_tail));
    }
    CUDA_TRY(cudaFreeAsync(d_min, s_tail));
  }
  CUDA_TRY(cudaGetLastError());
  return RAMP_OK;
}

}  // extern "C"
