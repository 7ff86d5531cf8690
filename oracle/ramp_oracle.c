/*
 * ramp_oracle.c -- CPU FP64 restatement of RAMP's validity-checking hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (ramp_b200/) may
 * import, link or call this file; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs do, and only as the checker
 * or the timed CPU baseline.
 *
 * PARITY UNPINNED (numerically): the reference (JafarAbdi/ramp) contains no
 * arithmetic on this path.  All of it lives in un-vendored third-party
 * wheels that are absent from /root/reference and not installable offline:
 *     pin  3.7.0  (uv.lock:1982-1983)   forwardKinematics,
 *                                       updateGeometryPlacements,
 *                                       computeCollisions
 *     coal 3.0.1  (uv.lock:479-480)     collide / distance, GJK + EPA
 *     ompl 1.8.0  (uv.lock:1771-1786)   DiscreteMotionValidator, state spaces
 * This file restates their published algorithms and is anchored on the
 * reference's own call sites and the known-answer facts its tests pin
 * (tests/test_oracle_kats.py).  The reference's tests hold no numeric golden
 * vectors for FK poses or collision booleans, so beyond those KATs the oracle
 * is validated against analytic geometry and independent brute-force checks
 * (tests/test_oracle_*.py).
 *
 * Reference call sites restated here (paths under the reference tree):
 *   orc_fk_*          src/ramp/robot_state.py:199-205 (framesForwardKinematics)
 *   orc_placements_*  src/ramp/robot_state.py:345-351 (updateGeometryPlacements)
 *   orc_check_*       src/ramp/robot_state.py:376-392 (computeCollisions),
 *                     src/ramp/compute_disable_collisions.py:152-161
 *   orc_qmap_apply    src/ramp/robot_model.py:363-375,406-426
 *   orc_edge_*        OMPL DiscreteMotionValidator implied by
 *                     src/ramp/motion_planner.py:366,469-483,612-625
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/ramp_b200.h"

#define ORC_MAX_JOINTS 64
#define ORC_MAX_GEOMS 256
#define ORC_MAX_NQ 256
#define ORC_PI 3.14159265358979323846

/* ------------------------------------------------------------------ SE(3) */
/* 3x4 row-major [R|t]:  m[4*r+c], c<3 rotation, c==3 translation.          */

static void se3_identity(double *m) {
  memset(m, 0, 12 * sizeof(double));
  m[0] = m[5] = m[10] = 1.0;
}

/* out = a * b  (pinocchio SE3::act: (R1 R2, t1 + R1 t2)) */
static void se3_mul(const double *a, const double *b, double *out) {
  double r[12];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j)
      r[4 * i + j] = a[4 * i] * b[j] + a[4 * i + 1] * b[4 + j] + a[4 * i + 2] * b[8 + j];
    r[4 * i + 3] = a[4 * i] * b[3] + a[4 * i + 1] * b[7] + a[4 * i + 2] * b[11] + a[4 * i + 3];
  }
  memcpy(out, r, sizeof r);
}

/* out = a^-1 * b */
static void se3_inv_mul(const double *a, const double *b, double *out) {
  double r[12];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j)
      r[4 * i + j] = a[i] * b[j] + a[4 + i] * b[4 + j] + a[8 + i] * b[8 + j];
    r[4 * i + 3] = a[i] * (b[3] - a[3]) + a[4 + i] * (b[7] - a[7]) + a[8 + i] * (b[11] - a[11]);
  }
  memcpy(out, r, sizeof r);
}

static void rot_axis_aligned(int axis, double c, double s, double *m) {
  se3_identity(m);
  switch (axis) {
  case 0: m[5] = c; m[6] = -s; m[9] = s; m[10] = c; break;
  case 1: m[0] = c; m[2] = s; m[8] = -s; m[10] = c; break;
  default: m[0] = c; m[1] = -s; m[4] = s; m[5] = c; break;
  }
}

/* Rodrigues, unit axis a */
static void rot_unaligned(const double *a, double c, double s, double *m) {
  se3_identity(m);
  const double v = 1.0 - c;
  m[0] = c + v * a[0] * a[0];
  m[1] = v * a[0] * a[1] - s * a[2];
  m[2] = v * a[0] * a[2] + s * a[1];
  m[4] = v * a[1] * a[0] + s * a[2];
  m[5] = c + v * a[1] * a[1];
  m[6] = v * a[1] * a[2] - s * a[0];
  m[8] = v * a[2] * a[0] - s * a[1];
  m[9] = v * a[2] * a[1] + s * a[0];
  m[10] = c + v * a[2] * a[2];
}

/* Eigen::Quaternion::toRotationMatrix: assumes unit norm, does NOT normalise
 * (free-flyer q is used as given). */
static void rot_quat(double x, double y, double z, double w, double *m) {
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w;
  const double txx = tx * x, txy = ty * x, txz = tz * x;
  const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  m[0] = 1 - (tyy + tzz); m[1] = txy - twz;       m[2] = txz + twy;
  m[4] = txy + twz;       m[5] = 1 - (txx + tzz); m[6] = tyz - twx;
  m[8] = txz - twy;       m[9] = tyz + twx;       m[10] = 1 - (txx + tyy);
}

/* Joint transform M_i(q) per joint type (pinocchio JointModel*::calc). */
static void joint_transform(const RampJoint *j, const double *q, double *m) {
  const double *qq = q + j->idx_q;
  se3_identity(m);
  switch (j->type) {
  case RAMP_JOINT_RX: case RAMP_JOINT_RY: case RAMP_JOINT_RZ:
    rot_axis_aligned(j->type - RAMP_JOINT_RX, cos(qq[0]), sin(qq[0]), m); break;
  case RAMP_JOINT_REVOLUTE_UNALIGNED:
    rot_unaligned(j->axis, cos(qq[0]), sin(qq[0]), m); break;
  case RAMP_JOINT_RUBX: case RAMP_JOINT_RUBY: case RAMP_JOINT_RUBZ:
    rot_axis_aligned(j->type - RAMP_JOINT_RUBX, qq[0], qq[1], m); break;
  case RAMP_JOINT_RUB_UNALIGNED:
    rot_unaligned(j->axis, qq[0], qq[1], m); break;
  case RAMP_JOINT_PX: case RAMP_JOINT_PY: case RAMP_JOINT_PZ:
    m[4 * (j->type - RAMP_JOINT_PX) + 3] = qq[0]; break;
  case RAMP_JOINT_PRISMATIC_UNALIGNED:
    m[3] = qq[0] * j->axis[0]; m[7] = qq[0] * j->axis[1]; m[11] = qq[0] * j->axis[2]; break;
  case RAMP_JOINT_PLANAR:
    rot_axis_aligned(2, qq[2], qq[3], m); m[3] = qq[0]; m[7] = qq[1]; break;
  case RAMP_JOINT_FREEFLYER:
    rot_quat(qq[3], qq[4], qq[5], qq[6], m); m[3] = qq[0]; m[7] = qq[1]; m[11] = qq[2]; break;
  case RAMP_JOINT_SPHERICAL:
    rot_quat(qq[0], qq[1], qq[2], qq[3], m); break;
  case RAMP_JOINT_SPHERICAL_ZYX: {
    double rz[12], ry[12], rx[12], t[12];
    rot_axis_aligned(2, cos(qq[0]), sin(qq[0]), rz);
    rot_axis_aligned(1, cos(qq[1]), sin(qq[1]), ry);
    rot_axis_aligned(0, cos(qq[2]), sin(qq[2]), rx);
    se3_mul(rz, ry, t); se3_mul(t, rx, m);
  } break;
  case RAMP_JOINT_TRANSLATION:
    m[3] = qq[0]; m[7] = qq[1]; m[11] = qq[2]; break;
  default: break;
  }
}

/* pinocchio::forwardKinematics: liMi = placement * M(q); oMi = oMi[parent] * liMi */
static void fk_one(const RampModelDesc *d, const double *q, double *oMi /*[J][12]*/) {
  for (int i = 0; i < d->n_joints; ++i) {
    const RampJoint *j = &d->joints[i];
    double mj[12], li[12];
    joint_transform(j, q, mj);
    se3_mul(j->placement, mj, li);
    if (j->parent < 0) memcpy(oMi + 12 * i, li, sizeof li);
    else se3_mul(oMi + 12 * j->parent, li, oMi + 12 * i);
  }
}

/* pinocchio::updateGeometryPlacements: oMg = oMi[parentJoint] * placement */
static void placements_one(const RampModelDesc *d, const double *oMi, double *oMg) {
  for (int g = 0; g < d->n_geoms; ++g) {
    const RampGeom *ge = &d->geoms[g];
    if (ge->parent_joint < 0) memcpy(oMg + 12 * g, ge->placement, 12 * sizeof(double));
    else se3_mul(oMi + 12 * ge->parent_joint, ge->placement, oMg + 12 * g);
  }
}

/* ----------------------------------------------------------- vector utils */
typedef struct { double x, y, z; } v3;
static inline v3 V(double x, double y, double z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vscale(v3 a, double s) { return V(a.x * s, a.y * s, a.z * s); }
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline double vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 vcross(v3 a, v3 b) {
  return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline double vnorm(v3 a) { return sqrt(vdot(a, a)); }
static inline v3 m_rot(const double *m, v3 a) {
  return V(m[0] * a.x + m[1] * a.y + m[2] * a.z, m[4] * a.x + m[5] * a.y + m[6] * a.z,
           m[8] * a.x + m[9] * a.y + m[10] * a.z);
}
static inline v3 m_rot_t(const double *m, v3 a) {
  return V(m[0] * a.x + m[4] * a.y + m[8] * a.z, m[1] * a.x + m[5] * a.y + m[9] * a.z,
           m[2] * a.x + m[6] * a.y + m[10] * a.z);
}
static inline v3 m_apply(const double *m, v3 a) {
  v3 r = m_rot(m, a);
  return V(r.x + m[3], r.y + m[7], r.z + m[11]);
}
static inline v3 m_trans(const double *m) { return V(m[3], m[7], m[11]); }

/* ------------------------------------------------------ support functions */
/* Support of the shape CORE in its local frame (coal getSupport): sphere and
 * capsule are reduced to point / segment, their radius is swept afterwards. */
static double swept_radius(const RampGeom *g) {
  return (g->type == RAMP_SHAPE_SPHERE || g->type == RAMP_SHAPE_CAPSULE) ? g->params[0] : 0.0;
}

static v3 support_local(const RampGeom *g, const double *verts, v3 d) {
  switch (g->type) {
  case RAMP_SHAPE_SPHERE: return V(0, 0, 0);
  case RAMP_SHAPE_CAPSULE: return V(0, 0, d.z > 0 ? g->params[1] : (d.z < 0 ? -g->params[1] : 0.0));
  case RAMP_SHAPE_BOX:
    return V(d.x > 0 ? g->params[0] : -g->params[0], d.y > 0 ? g->params[1] : -g->params[1],
             d.z > 0 ? g->params[2] : -g->params[2]);
  case RAMP_SHAPE_CYLINDER: {
    const double zz = d.z > 0 ? g->params[1] : (d.z < 0 ? -g->params[1] : 0.0);
    const double n = sqrt(d.x * d.x + d.y * d.y);
    if (n > 0) return V(g->params[0] * d.x / n, g->params[0] * d.y / n, zz);
    return V(0, 0, zz);
  }
  case RAMP_SHAPE_ELLIPSOID: {
    const double a2 = g->params[0] * g->params[0], b2 = g->params[1] * g->params[1],
                 c2 = g->params[2] * g->params[2];
    const v3 v = V(a2 * d.x, b2 * d.y, c2 * d.z);
    const double den = sqrt(vdot(v, d));
    if (den > 0) return vscale(v, 1.0 / den);
    return V(0, 0, 0);
  }
  case RAMP_SHAPE_CONVEX: {
    const double *p = verts + 3 * (size_t)g->vert_offset;
    int best = 0;
    double bv = -INFINITY;
    for (int i = 0; i < g->vert_count; ++i) {
      const double s = p[3 * i] * d.x + p[3 * i + 1] * d.y + p[3 * i + 2] * d.z;
      if (s > bv) { bv = s; best = i; }
    }
    return V(p[3 * best], p[3 * best + 1], p[3 * best + 2]);
  }
  default: return V(0, 0, 0);
  }
}

typedef struct {
  const RampGeom *g1, *g2;
  const double *verts;
  double m12[12]; /* pose of shape 2 in the frame of shape 1 */
} mink_t;

/* support of core1 - core2 (Minkowski difference) in frame 1 */
static v3 mink_support(const mink_t *mk, v3 d) {
  const v3 s1 = support_local(mk->g1, mk->verts, d);
  const v3 d2 = m_rot_t(mk->m12, vneg(d));
  const v3 s2 = m_apply(mk->m12, support_local(mk->g2, mk->verts, d2));
  return vsub(s1, s2);
}

/* ------------------------------------------------------------------- GJK */
/* Closest point of a simplex to the origin (Johnson sub-algorithm written as
 * Voronoi-region tests, the formulation coal's projectLine/Triangle/
 * TetrahedraOrigin use).  The simplex is reduced in place to the supporting
 * sub-simplex; returns 1 when the origin is enclosed by a tetrahedron. */
typedef struct { v3 p[4]; int n; } simplex_t;

static v3 closest_segment(simplex_t *s) {
  const v3 a = s->p[0], b = s->p[1];
  const v3 ab = vsub(b, a);
  const double t = -vdot(a, ab), den = vdot(ab, ab);
  if (t <= 0 || den <= 0) { s->n = 1; return a; }
  if (t >= den) { s->p[0] = b; s->n = 1; return b; }
  return vadd(a, vscale(ab, t / den));
}

/* closest point on triangle (a,b,c) to origin; writes the reduced simplex */
static v3 closest_triangle_pts(v3 a, v3 b, v3 c, simplex_t *out) {
  const v3 ab = vsub(b, a), ac = vsub(c, a);
  const double d1 = -vdot(ab, a), d2 = -vdot(ac, a);
  if (d1 <= 0 && d2 <= 0) { out->p[0] = a; out->n = 1; return a; }
  const double d3 = -vdot(ab, b), d4 = -vdot(ac, b);
  if (d3 >= 0 && d4 <= d3) { out->p[0] = b; out->n = 1; return b; }
  const double vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) {
    const double v = d1 / (d1 - d3);
    out->p[0] = a; out->p[1] = b; out->n = 2;
    return vadd(a, vscale(ab, v));
  }
  const double d5 = -vdot(ab, c), d6 = -vdot(ac, c);
  if (d6 >= 0 && d5 <= d6) { out->p[0] = c; out->n = 1; return c; }
  const double vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) {
    const double w = d2 / (d2 - d6);
    out->p[0] = a; out->p[1] = c; out->n = 2;
    return vadd(a, vscale(ac, w));
  }
  const double va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) {
    const double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    out->p[0] = b; out->p[1] = c; out->n = 2;
    return vadd(b, vscale(vsub(c, b), w));
  }
  const double den = 1.0 / (va + vb + vc);
  const double v = vb * den, w = vc * den;
  out->p[0] = a; out->p[1] = b; out->p[2] = c; out->n = 3;
  return vadd(a, vadd(vscale(ab, v), vscale(ac, w)));
}

static v3 closest_triangle(simplex_t *s) {
  simplex_t o;
  const v3 r = closest_triangle_pts(s->p[0], s->p[1], s->p[2], &o);
  *s = o;
  return r;
}

/* returns 1 if the origin is inside the tetrahedron */
static int closest_tetra(simplex_t *s, v3 *out) {
  static const int F[4][4] = {{0, 1, 2, 3}, {0, 2, 3, 1}, {0, 3, 1, 2}, {1, 3, 2, 0}};
  double best = INFINITY;
  simplex_t bs;
  memset(&bs, 0, sizeof bs);
  v3 bp = V(0, 0, 0);
  int outside_any = 0;
  for (int f = 0; f < 4; ++f) {
    const v3 a = s->p[F[f][0]], b = s->p[F[f][1]], c = s->p[F[f][2]], d = s->p[F[f][3]];
    const v3 n = vcross(vsub(b, a), vsub(c, a));
    const double so = -vdot(a, n);         /* origin side */
    const double sd = vdot(vsub(d, a), n); /* opposite vertex side */
    /* origin outside this face when it is on the other side than d */
    if (so * sd < 0 || sd == 0) {
      outside_any = 1;
      simplex_t o;
      memset(&o, 0, sizeof o);
      const v3 p = closest_triangle_pts(a, b, c, &o);
      const double dd = vdot(p, p);
      if (dd < best) { best = dd; bs = o; bp = p; }
    }
  }
  if (!outside_any) return 1;
  *s = bs;
  *out = bp;
  return 0;
}

typedef struct {
  double tol;       /* coal GJK tolerance, 1e-6 */
  int max_iter;     /* coal GJK max iterations, 128 */
  double epa_tol;   /* 1e-6 */
  int epa_max_iter; /* coal 64 */
} orc_params_t;

static orc_params_t g_params = {1e-6, 128, 1e-6, 64};

void orc_set_params(double gjk_tol, int gjk_max_iter, double epa_tol, int epa_max_iter) {
  g_params.tol = gjk_tol; g_params.max_iter = gjk_max_iter;
  g_params.epa_tol = epa_tol; g_params.epa_max_iter = epa_max_iter;
}

enum { GJK_SEPARATED = 0, GJK_INSIDE = 1, GJK_EARLY_OUT = 2, GJK_FAILED = 3 };

/* GJK on the cores.  `upper` = distance upper bound on the CORE distance:
 * once a separating plane proves core distance > upper the iteration stops
 * (coal's distance_upper_bound early stop used by collide()).
 * Returns status; *dist = |closest point| (core distance, valid for
 * GJK_SEPARATED), final simplex in *sx. */
static int gjk_run(const mink_t *mk, double upper, double *dist, simplex_t *sx, int *iters) {
  const double tol = g_params.tol;
  v3 ray = V(1, 0, 0); /* coal default initial guess */
  double rl = 1.0, alpha = 0.0;
  simplex_t s; s.n = 0;
  int status = GJK_FAILED, it = 0;
  for (; it < g_params.max_iter; ++it) {
    if (rl < tol) { status = GJK_INSIDE; break; }
    const v3 w = mink_support(mk, vneg(ray));
    const double omega = vdot(ray, w) / rl;
    if (omega > upper) { status = GJK_EARLY_OUT; rl = omega; ++it; break; }
    if (omega > alpha) alpha = omega;
    if (it > 0 && rl - alpha <= tol * rl) { status = GJK_SEPARATED; break; }
    /* duplicate vertex => no more progress possible */
    int dup = 0;
    for (int k = 0; k < s.n; ++k) {
      const v3 e = vsub(s.p[k], w);
      if (vdot(e, e) <= 1e-30) dup = 1;
    }
    if (dup) { status = GJK_SEPARATED; break; }
    s.p[s.n++] = w;
    v3 nr;
    switch (s.n) {
    case 1: nr = s.p[0]; break;
    case 2: nr = closest_segment(&s); break;
    case 3: nr = closest_triangle(&s); break;
    default:
      if (closest_tetra(&s, &nr)) { status = GJK_INSIDE; rl = 0; goto done; }
    }
    const double nrl = vnorm(nr);
    if (nrl >= rl && it > 0) { /* no decrease: converged to machine precision */
      status = GJK_SEPARATED; break;
    }
    ray = nr; rl = nrl;
  }
done:
  if (status == GJK_FAILED) status = (rl < tol) ? GJK_INSIDE : GJK_SEPARATED;
  *dist = rl; *sx = s;
  if (iters) *iters += it;
  return status;
}

/* ------------------------------------------------------------------- EPA */
#define EPA_MAX_V 200
#define EPA_MAX_F 400
typedef struct { int a, b, c; v3 n; double d; int alive; } epa_face;

/* Face (a, b, c) of the polytope, its normal turned AWAY from `inside` (a point strictly inside
 * the polytope: the centroid of the start tetrahedron).  Orienting every face by that test, instead
 * of by the winding of the horizon edge it was built on, keeps normals outward when a new vertex
 * lies in the plane of a face next to the horizon -- flat faces of box / cylinder Minkowski sums
 * yield many coplanar support points, and a winding-derived normal then came out inverted (d < 0). */
static int epa_make_face(const v3 *vt, int a, int b, int c, v3 inside, epa_face *f) {
  v3 n = vcross(vsub(vt[b], vt[a]), vsub(vt[c], vt[a]));
  const double l = vnorm(n);
  f->a = a; f->b = b; f->c = c; f->alive = 0;
  const double scale = vnorm(vsub(vt[b], vt[a])) * vnorm(vsub(vt[c], vt[a]));
  if (l <= 1e-12 * scale || l < 1e-300) return 0; /* a sliver: no plane to speak of */
  n = vscale(n, 1.0 / l);
  if (vdot(n, vsub(vt[a], inside)) < 0) { n = vneg(n); f->b = c; f->c = b; }
  f->n = n; f->d = vdot(n, vt[a]); f->alive = 1;
  return 1;
}

/* Expand a degenerate terminal simplex into a tetrahedron that encloses the
 * origin (the role of coal's EPA::evaluate / encloseOrigin pre-step). */
static int epa_enclose(const mink_t *mk, simplex_t *s) {
  static const v3 AX[6] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
  if (s->n == 0) { s->p[0] = mink_support(mk, V(1, 0, 0)); s->n = 1; }
  if (s->n == 1) {
    for (int k = 0; k < 6; ++k) {
      const v3 w = mink_support(mk, AX[k]);
      if (vnorm(vsub(w, s->p[0])) > 1e-12) { s->p[1] = w; s->n = 2; break; }
    }
    if (s->n == 1) return 0;
  }
  if (s->n == 2) {
    const v3 e = vsub(s->p[1], s->p[0]);
    int done = 0;
    for (int k = 0; k < 6 && !done; k += 2) {
      const v3 dir = vcross(e, AX[k]);
      if (vdot(dir, dir) < 1e-24) continue;
      for (int sg = 0; sg < 2 && !done; ++sg) {
        const v3 w = mink_support(mk, sg ? vneg(dir) : dir);
        const v3 n = vcross(e, vsub(w, s->p[0]));
        if (vdot(n, n) > 1e-24) { s->p[2] = w; s->n = 3; done = 1; }
      }
    }
    if (s->n == 2) return 0;
  }
  if (s->n == 3) {
    const v3 n = vcross(vsub(s->p[1], s->p[0]), vsub(s->p[2], s->p[0]));
    for (int sg = 0; sg < 2; ++sg) {
      const v3 w = mink_support(mk, sg ? vneg(n) : n);
      if (fabs(vdot(vsub(w, s->p[0]), n)) > 1e-14 * (1 + vdot(n, n))) { s->p[3] = w; s->n = 4; break; }
    }
    if (s->n == 3) return 0;
  }
  return 1;
}

/* Penetration depth of the cores (>= 0).  Returns the lower bound (distance
 * of the closest polytope face) when the iteration cap is hit. */
static double epa_run(const mink_t *mk, simplex_t s, int *iters) {
  v3 vt[EPA_MAX_V];
  epa_face fc[EPA_MAX_F];
  int nv = 0, nf = 0;
  if (!epa_enclose(mk, &s)) return 0.0;
  for (int i = 0; i < 4; ++i) vt[nv++] = s.p[i];
  const v3 inside = vscale(vadd(vadd(vt[0], vt[1]), vadd(vt[2], vt[3])), 0.25);
  static const int T[4][3] = {{0, 1, 2}, {0, 1, 3}, {0, 2, 3}, {1, 2, 3}};
  for (int i = 0; i < 4; ++i)
    if (!epa_make_face(vt, T[i][0], T[i][1], T[i][2], inside, &fc[nf++])) return 0.0;
  double best = 0.0;
  for (int it = 0; it < g_params.epa_max_iter; ++it) {
    int bi = -1;
    double bd = INFINITY;
    for (int i = 0; i < nf; ++i)
      if (fc[i].alive && fc[i].d < bd) { bd = fc[i].d; bi = i; }
    if (bi < 0) break;
    best = bd;
    if (iters) ++*iters;
    const v3 w = mink_support(mk, fc[bi].n);
    const double gap = vdot(fc[bi].n, w) - bd;
    if (gap <= g_params.epa_tol * (1.0 + bd) || nv >= EPA_MAX_V) break;
    int dup = 0; /* a support point the polytope already has: nothing left to expand */
    for (int k = 0; k < nv; ++k) {
      const v3 e = vsub(vt[k], w);
      if (vdot(e, e) <= 1e-24) dup = 1;
    }
    if (dup) break;
    /* remove the faces that see w; the horizon is the set of edges with exactly ONE removed
     * neighbour -- counted without regard to direction, so it does not depend on windings */
    int eh[EPA_MAX_F][2], ne = 0;
    for (int i = 0; i < nf; ++i) {
      if (!fc[i].alive) continue;
      if (vdot(fc[i].n, w) - fc[i].d > 1e-13) {
        fc[i].alive = 0;
        const int ed[3][2] = {{fc[i].a, fc[i].b}, {fc[i].b, fc[i].c}, {fc[i].c, fc[i].a}};
        for (int k = 0; k < 3; ++k) {
          const int lo = ed[k][0] < ed[k][1] ? ed[k][0] : ed[k][1], hi = ed[k][0] < ed[k][1] ? ed[k][1] : ed[k][0];
          int found = -1;
          for (int e = 0; e < ne; ++e)
            if (eh[e][0] == lo && eh[e][1] == hi) { found = e; break; }
          if (found >= 0) { eh[found][0] = eh[ne - 1][0]; eh[found][1] = eh[ne - 1][1]; --ne; }
          else if (ne < EPA_MAX_F) { eh[ne][0] = lo; eh[ne][1] = hi; ++ne; }
        }
      }
    }
    if (ne == 0) break;
    const int wi = nv;
    vt[nv++] = w;
    int ok = 1;
    for (int e = 0; e < ne; ++e) {
      int slot = -1;
      for (int i = 0; i < nf; ++i) if (!fc[i].alive) { slot = i; break; }
      if (slot < 0) { if (nf >= EPA_MAX_F) { ok = 0; break; } slot = nf++; }
      if (!epa_make_face(vt, eh[e][0], eh[e][1], wi, inside, &fc[slot])) fc[slot].alive = 0;
    }
    if (!ok) break;
  }
  return best;
}

/* --------------------------------------------------- closed-form distances */
static double seg_seg_dist(v3 p1, v3 q1, v3 p2, v3 q2) {
  const v3 d1 = vsub(q1, p1), d2 = vsub(q2, p2), r = vsub(p1, p2);
  const double a = vdot(d1, d1), e = vdot(d2, d2), f = vdot(d2, r);
  double s, t;
  const double EPS = 1e-30;
  if (a <= EPS && e <= EPS) return vnorm(r);
  if (a <= EPS) { s = 0; t = fmin(fmax(f / e, 0), 1); }
  else {
    const double c = vdot(d1, r);
    if (e <= EPS) { t = 0; s = fmin(fmax(-c / a, 0), 1); }
    else {
      const double b = vdot(d1, d2), den = a * e - b * b;
      s = den > EPS ? fmin(fmax((b * f - c * e) / den, 0), 1) : 0.0;
      t = (b * s + f) / e;
      if (t < 0) { t = 0; s = fmin(fmax(-c / a, 0), 1); }
      else if (t > 1) { t = 1; s = fmin(fmax((b - c) / a, 0), 1); }
    }
  }
  return vnorm(vsub(vadd(p1, vscale(d1, s)), vadd(p2, vscale(d2, t))));
}

static double point_seg_dist(v3 p, v3 a, v3 b) {
  const v3 ab = vsub(b, a);
  const double den = vdot(ab, ab);
  double t = den > 0 ? vdot(vsub(p, a), ab) / den : 0.0;
  t = fmin(fmax(t, 0), 1);
  return vnorm(vsub(p, vadd(a, vscale(ab, t))));
}

static double sphere_box_dist(v3 c, const double *h, double r) {
  const double cc[3] = {c.x, c.y, c.z};
  double out2 = 0, inside = INFINITY;
  int is_in = 1;
  for (int i = 0; i < 3; ++i) {
    const double a = fabs(cc[i]) - h[i];
    if (a > 0) { out2 += a * a; is_in = 0; }
    else if (-a < inside) inside = -a;
  }
  return is_in ? -inside - r : sqrt(out2) - r;
}

static double sphere_cyl_dist(v3 c, double R, double h, double r) {
  const double dr = sqrt(c.x * c.x + c.y * c.y) - R, dz = fabs(c.z) - h;
  if (dr <= 0 && dz <= 0) return fmax(dr, dz) - r;
  if (dr <= 0) return dz - r;
  if (dz <= 0) return dr - r;
  return sqrt(dr * dr + dz * dz) - r;
}

/* Signed distance between two placed shapes (coal::distance semantics:
 * positive separation, negative penetration depth).  want_depth = 0 skips EPA
 * and returns -0.0-ish (a non-positive value) for penetrating pairs, which is
 * all coal::collide needs (CollisionRequest(NO_REQUEST,1), margin 0).
 * `upper`: early-out bound on the signed distance (INFINITY = exact). */
static double shape_distance(const RampGeom *g1, const double *M1, const RampGeom *g2,
                             const double *M2, const double *verts, int want_depth,
                             double upper, int *iters) {
  int t1 = g1->type, t2 = g2->type;
  const RampGeom *a = g1, *b = g2;
  const double *Ma = M1, *Mb = M2;
  if (t1 > t2) { a = g2; b = g1; Ma = M2; Mb = M1; t1 = a->type; t2 = b->type; }
  /* closed forms coal specialises (ShapeShapeDistance<...>) */
  if (t1 == RAMP_SHAPE_SPHERE) {
    const v3 ca = m_trans(Ma);
    if (t2 == RAMP_SHAPE_SPHERE) return vnorm(vsub(ca, m_trans(Mb))) - a->params[0] - b->params[0];
    const v3 c = m_rot_t(Mb, vsub(ca, m_trans(Mb))); /* sphere centre in b's frame */
    if (t2 == RAMP_SHAPE_CAPSULE)
      return point_seg_dist(c, V(0, 0, -b->params[1]), V(0, 0, b->params[1])) - a->params[0] - b->params[0];
    if (t2 == RAMP_SHAPE_BOX) return sphere_box_dist(c, b->params, a->params[0]);
    if (t2 == RAMP_SHAPE_CYLINDER) return sphere_cyl_dist(c, b->params[0], b->params[1], a->params[0]);
  }
  if (t1 == RAMP_SHAPE_CAPSULE && t2 == RAMP_SHAPE_CAPSULE) {
    const v3 p1 = m_apply(Ma, V(0, 0, -a->params[1])), q1 = m_apply(Ma, V(0, 0, a->params[1]));
    const v3 p2 = m_apply(Mb, V(0, 0, -b->params[1])), q2 = m_apply(Mb, V(0, 0, b->params[1]));
    return seg_seg_dist(p1, q1, p2, q2) - a->params[0] - b->params[0];
  }
  /* everything else: GJK (+EPA) on the cores, radii swept afterwards */
  mink_t mk;
  mk.g1 = g1; mk.g2 = g2; mk.verts = verts;
  se3_inv_mul(M1, M2, mk.m12);
  const double rsum = swept_radius(g1) + swept_radius(g2);
  double dist;
  simplex_t sx;
  const int st = gjk_run(&mk, upper + rsum, &dist, &sx, iters);
  if (st == GJK_SEPARATED || st == GJK_EARLY_OUT) {
    const double d = dist - rsum;
    if (d > 0 || !want_depth || rsum == 0) return d;
    return d; /* cores apart, swept volumes overlap: depth = rsum - dist exactly */
  }
  if (!want_depth) return -rsum > -1e-300 ? -1e-300 : -rsum;
  return -epa_run(&mk, sx, iters) - rsum;
}

double orc_shape_distance(const RampGeom *g1, const double *M1, const RampGeom *g2,
                          const double *M2, const double *verts, int *iters) {
  int it = 0;
  const double d = shape_distance(g1, M1, g2, M2, verts, 1, INFINITY, &it);
  if (iters) *iters = it;
  return d;
}

/* coal: collision iff distance - security_margin <= collision_distance_threshold
 * with margin 0 and threshold Eigen dummy_precision<double>() = 1e-12 */
#define ORC_COLLISION_THRESHOLD 1e-12

/* --------------------------------------------------------- batch wrappers */
static int check_desc(const RampModelDesc *d) {
  if (!d || d->n_joints > ORC_MAX_JOINTS || d->n_geoms > ORC_MAX_GEOMS) return -1;
  return 0;
}

int orc_fk_batch(const RampModelDesc *d, const double *q, int64_t n, int ld, double *oMi) {
  if (check_desc(d)) return -1;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) fk_one(d, q + i * ld, oMi + i * 12 * d->n_joints);
  return 0;
}

int orc_placements_batch(const RampModelDesc *d, const double *q, int64_t n, int ld, double *oMg) {
  if (check_desc(d)) return -1;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    double oMi[12 * ORC_MAX_JOINTS];
    fk_one(d, q + i * ld, oMi);
    placements_one(d, oMi, oMg + i * 12 * d->n_geoms);
  }
  return 0;
}

/* pinocchio::computeCollisions restated: pairs in list order, first hit index,
 * optional early return.  mode bits: 1 = stop at first collision,
 * 2 = signed depth (EPA) for penetrating pairs, 4 = exact separation
 * distances (otherwise GJK stops at the first separating plane, like
 * coal::collide with distance_upper_bound 0). */
static void check_one(const RampModelDesc *d, const uint8_t *active, const double *q, int mode,
                      uint8_t *hit, int32_t *first_pair, double *min_dist, double *pair_dist,
                      int64_t *iters_out) {
  double oMi[12 * ORC_MAX_JOINTS], oMg[12 * ORC_MAX_GEOMS];
  fk_one(d, q, oMi);
  placements_one(d, oMi, oMg);
  const int stop = mode & 1, depth = (mode & 2) != 0, exact = (mode & 4) != 0;
  int first = -1, iters = 0;
  double md = INFINITY;
  for (int p = 0; p < d->n_pairs; ++p) {
    if (pair_dist) pair_dist[p] = NAN;
    if (active && !active[p]) continue;
    const int a = d->pairs[2 * p], b = d->pairs[2 * p + 1];
    const double dist = shape_distance(&d->geoms[a], oMg + 12 * a, &d->geoms[b], oMg + 12 * b,
                                       d->verts, depth, exact ? INFINITY : 0.0, &iters);
    if (pair_dist) pair_dist[p] = dist;
    if (dist < md) md = dist;
    if (dist <= ORC_COLLISION_THRESHOLD) {
      if (first < 0) first = p;
      if (stop) break;
    }
  }
  *hit = first >= 0;
  if (first_pair) *first_pair = first;
  if (min_dist) *min_dist = md;
  if (iters_out) *iters_out = iters;
}

int orc_check_batch(const RampModelDesc *d, const uint8_t *active, const double *q, int64_t n,
                    int ld, int mode, uint8_t *hit, int32_t *first_pair, double *min_dist,
                    double *pair_dist, int64_t *iters, int nthreads) {
  if (check_desc(d)) return -1;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t i = 0; i < n; ++i)
    check_one(d, active, q + i * ld, mode, hit + i, first_pair ? first_pair + i : NULL,
              min_dist ? min_dist + i : NULL, pair_dist ? pair_dist + i * d->n_pairs : NULL,
              iters ? iters + i : NULL);
  return 0;
}

/* float32 input variant: the timed CPU baseline consumes the same FP32 batch
 * the GPU does. */
int orc_check_batch_f32(const RampModelDesc *d, const uint8_t *active, const float *q, int64_t n,
                        int ld, int mode, uint8_t *hit, int nthreads) {
  if (check_desc(d) || d->nq > 256) return -1;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t i = 0; i < n; ++i) {
    double qq[256];
    for (int k = 0; k < d->nq; ++k) qq[k] = q[i * ld + k];
    check_one(d, active, qq, mode, hit + i, NULL, NULL, NULL, NULL);
  }
  return 0;
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* ------------------------------------------------------------ q expansion */
/* robot_state.py:117-119: scatter, mimic, then continuous (cos, sin) */
static void qmap_one(const RampQMap *m, int nq, const double *act, double *q) {
  memcpy(q, m->ref_q, nq * sizeof(double));
  for (int k = 0; k < m->n_act; ++k) q[m->act_slot[k]] = act[k];
  for (int k = 0; k < m->n_mimic; ++k)
    q[m->mimic_dst[k]] = q[m->mimic_src[k]] * m->mimic_mult[k] + m->mimic_off[k];
  for (int k = 0; k < m->n_cont; ++k) {
    const int c = m->cont_slot[k];
    const double th = q[c];
    q[c] = cos(th); q[c + 1] = sin(th);
  }
}

int orc_qmap_apply(const RampQMap *m, int nq, const double *act, int64_t n, int ld, double *q) {
  for (int64_t i = 0; i < n; ++i) qmap_one(m, nq, act + i * ld, q + i * nq);
  return 0;
}

/* --------------------------------------------------- OMPL state-space maths */
static double so2_distance(double a, double b) {
  const double d = fabs(a - b);
  return d > ORC_PI ? 2.0 * ORC_PI - d : d;
}

static double so2_interp(double from, double to, double t) {
  double diff = to - from, v;
  if (fabs(diff) <= ORC_PI) v = from + diff * t;
  else {
    if (diff > 0) diff = 2.0 * ORC_PI - diff; else diff = -2.0 * ORC_PI - diff;
    v = from - diff * t;
    if (v > ORC_PI) v -= 2.0 * ORC_PI; else if (v < -ORC_PI) v += 2.0 * ORC_PI;
  }
  return v;
}

static double so3_arc(const double *a, const double *b) {
  const double dq = fabs(a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3]);
  if (dq > 1.0 - 1e-9) return 0.0; /* MAX_QUATERNION_NORM_ERROR */
  return acos(dq);
}

static void so3_slerp(const double *a, const double *b, double t, double *o) {
  const double theta = so3_arc(a, b);
  if (theta > 2.220446049250313e-16) {
    const double d = 1.0 / sin(theta), s0 = sin((1.0 - t) * theta);
    double s1 = sin(t * theta);
    const double dq = a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3];
    if (dq < 0) s1 = -s1;
    for (int k = 0; k < 4; ++k) o[k] = (a[k] * s0 + b[k] * s1) * d;
  } else memcpy(o, a, 4 * sizeof(double));
}

static int seg_count(double dist, double extent, double frac) {
  const double lvs = extent * frac;
  if (!(lvs > 0)) return 0;
  return (int)ceil(dist / lvs);
}

static double real_extent(const RampSubspace *s, int dim) {
  double e = 0;
  for (int k = 0; k < dim; ++k) { const double d = s->hi[k] - s->lo[k]; e += d * d; }
  return sqrt(e);
}

static double real_dist(const double *a, const double *b, int dim) {
  double e = 0;
  for (int k = 0; k < dim; ++k) { const double d = a[k] - b[k]; e += d * d; }
  return sqrt(e);
}

/* CompoundStateSpace::validSegmentCount = max over components (recursively) */
int orc_edge_nd(const RampSpaceDesc *sp, const double *q0, const double *q1) {
  int nd = 0;
  const double f = sp->longest_valid_segment_fraction;
  for (int i = 0; i < sp->n_sub; ++i) {
    const RampSubspace *s = &sp->sub[i];
    const double *a = q0 + s->col, *b = q1 + s->col;
    int c = 0, c2 = 0;
    switch (s->type) {
    case RAMP_SPACE_REAL: c = seg_count(real_dist(a, b, s->dim), real_extent(s, s->dim), f); break;
    case RAMP_SPACE_SO2: c = seg_count(so2_distance(a[0], b[0]), ORC_PI, f); break;
    case RAMP_SPACE_SE2:
      c = seg_count(real_dist(a, b, 2), real_extent(s, 2), f);
      c2 = seg_count(so2_distance(a[2], b[2]), ORC_PI, f); break;
    case RAMP_SPACE_SE3:
      c = seg_count(real_dist(a, b, 3), real_extent(s, 3), f);
      c2 = seg_count(so3_arc(a + 3, b + 3), 0.5 * ORC_PI, f); break;
    case RAMP_SPACE_SO3: c = seg_count(so3_arc(a, b), 0.5 * ORC_PI, f); break;
    default: break;
    }
    if (c2 > c) c = c2;
    if (c > nd) nd = c;
  }
  return nd;
}

void orc_edge_interpolate(const RampSpaceDesc *sp, const double *q0, const double *q1, double t,
                          double *out) {
  for (int i = 0; i < sp->n_sub; ++i) {
    const RampSubspace *s = &sp->sub[i];
    const double *a = q0 + s->col, *b = q1 + s->col;
    double *o = out + s->col;
    switch (s->type) {
    case RAMP_SPACE_REAL: for (int k = 0; k < s->dim; ++k) o[k] = a[k] + (b[k] - a[k]) * t; break;
    case RAMP_SPACE_SO2: o[0] = so2_interp(a[0], b[0], t); break;
    case RAMP_SPACE_SE2:
      for (int k = 0; k < 2; ++k) o[k] = a[k] + (b[k] - a[k]) * t;
      o[2] = so2_interp(a[2], b[2], t); break;
    case RAMP_SPACE_SE3:
      for (int k = 0; k < 3; ++k) o[k] = a[k] + (b[k] - a[k]) * t;
      so3_slerp(a + 3, b + 3, t, o + 3); break;
    case RAMP_SPACE_SO3: so3_slerp(a, b, t, o); break;
    default: break;
    }
  }
}

/* SpaceInformation::satisfiesBounds over the compound space */
int orc_satisfies_bounds(const RampSpaceDesc *sp, const double *q) {
  const double eps = 2.220446049250313e-16;
  for (int i = 0; i < sp->n_sub; ++i) {
    const RampSubspace *s = &sp->sub[i];
    const double *a = q + s->col;
    int nreal = 0;
    switch (s->type) {
    case RAMP_SPACE_REAL: nreal = s->dim; break;
    case RAMP_SPACE_SE2: nreal = 2; if (a[2] < -ORC_PI - eps || a[2] > ORC_PI + eps) return 0; break;
    case RAMP_SPACE_SE3: nreal = 3; break;
    case RAMP_SPACE_SO2: if (a[0] < -ORC_PI - eps || a[0] > ORC_PI + eps) return 0; break;
    default: break;
    }
    for (int k = 0; k < nreal; ++k)
      if (a[k] - eps > s->hi[k] || a[k] + eps < s->lo[k]) return 0;
  }
  return 1;
}

/* MotionPlanner.is_state_valid (motion_planner.py:612-625) on actuated values */
static int state_valid(const RampModelDesc *d, const uint8_t *active, const RampQMap *m,
                       const RampSpaceDesc *sp, const double *act) {
  if (sp->check_bounds && !orc_satisfies_bounds(sp, act)) return 0;
  double q[256];
  uint8_t hit;
  qmap_one(m, d->nq, act, q);
  check_one(d, active, q, 1, &hit, NULL, NULL, NULL, NULL);
  return !hit;
}

/* DiscreteMotionValidator::checkMotion restated as "smallest invalid j":
 * interior states j/nd for j in [1, nd-1], and q1 reported as j = nd.  Both
 * OMPL overloads (bisection order / linear order with lastValid) are
 * functions of that minimum. */
int orc_check_edges(const RampModelDesc *d, const uint8_t *active, const RampQMap *m,
                    const RampSpaceDesc *sp, const double *q0, const double *q1, int64_t n_edges,
                    int ld, uint8_t *valid, int32_t *first_invalid, int32_t *nd_out,
                    int nthreads) {
  if (check_desc(d) || d->nq > 256 || m->n_act > 256) return -1;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 4)
  for (int64_t e = 0; e < n_edges; ++e) {
    const double *a = q0 + e * ld, *b = q1 + e * ld;
    const int nd = orc_edge_nd(sp, a, b);
    int first = -1;
    double st[256];
    for (int j = 1; j < nd && first < 0; ++j) {
      orc_edge_interpolate(sp, a, b, (double)j / (double)nd, st);
      if (!state_valid(d, active, m, sp, st)) first = j;
    }
    if (first < 0 && !state_valid(d, active, m, sp, b)) first = nd;
    valid[e] = first < 0;
    if (first_invalid) first_invalid[e] = first;
    if (nd_out) nd_out[e] = nd;
  }
  return 0;
}

/* ---------------------------------------------------- batched differential IK */
/* FP64 statement of ramp_ik_batch (include/ramp_b200.h): damped least squares on
 * the actuated values of a group, pose error of the tool frame seen from that
 * frame (translation, rotation vector), frame Jacobian in the LOCAL frame.
 * The reference's own differential_ik (src/ramp/robot_state.py:237-304) is a
 * pink / quadprog QP per iteration; what both have in common is the contract
 * (tool frame within MAX_ERROR of the target, inside the limits, or failure),
 * and that is what the tests compare -- plus GPU against THIS iteration. */
static void log_so3(const double *r, double *w) {
  const double vee[3] = {0.5 * (r[9] - r[6]), 0.5 * (r[2] - r[8]), 0.5 * (r[4] - r[1])};
  double c = 0.5 * (r[0] + r[5] + r[10] - 1.0);
  c = c > 1 ? 1 : (c < -1 ? -1 : c);
  const double s = sqrt(vee[0] * vee[0] + vee[1] * vee[1] + vee[2] * vee[2]);
  const double th = atan2(s, c);
  if (s > 1e-4) { for (int k = 0; k < 3; ++k) w[k] = vee[k] * th / s; return; }
  if (c > 0) { for (int k = 0; k < 3; ++k) w[k] = vee[k]; return; }
  const double kk = 1.0 / (1.0 - c);
  double a[3] = {sqrt(fmax((r[0] - c) * kk, 0)), sqrt(fmax((r[5] - c) * kk, 0)), sqrt(fmax((r[10] - c) * kk, 0))};
  if (a[0] >= a[1] && a[0] >= a[2]) { a[1] = copysign(a[1], r[1] + r[4]); a[2] = copysign(a[2], r[2] + r[8]); }
  else if (a[1] >= a[2]) { a[0] = copysign(a[0], r[1] + r[4]); a[2] = copysign(a[2], r[6] + r[9]); }
  else { a[0] = copysign(a[0], r[2] + r[8]); a[1] = copysign(a[1], r[6] + r[9]); }
  /* the largest component is positive (see ik.cuh log_so3) */
  for (int k = 0; k < 3; ++k) w[k] = a[k] * th;
}

static void ik_column(const RampModelDesc *d, const double *oMi, int j, const unsigned char *chain, const double *tcp, double *col) {
  for (int k = 0; k < 6; ++k) col[k] = 0;
  if (j < 0 || !chain[j]) return;
  const RampJoint *jt = &d->joints[j];
  const int t = jt->type;
  double ax[3] = {jt->axis[0], jt->axis[1], jt->axis[2]};
  const int al = t <= 3 ? t - 1 : (t >= 5 && t <= 7 ? t - 5 : (t >= 9 && t <= 11 ? t - 9 : -1));
  if (al >= 0) { ax[0] = al == 0; ax[1] = al == 1; ax[2] = al == 2; }
  const double *o = oMi + 12 * j;
  double aw[3], lw[3] = {0, 0, 0};
  for (int r = 0; r < 3; ++r) aw[r] = o[4 * r] * ax[0] + o[4 * r + 1] * ax[1] + o[4 * r + 2] * ax[2];
  if (t >= 9) { for (int r = 0; r < 3; ++r) lw[r] = aw[r]; }
  else {
    const double dx[3] = {tcp[3] - o[3], tcp[7] - o[7], tcp[11] - o[11]};
    lw[0] = aw[1] * dx[2] - aw[2] * dx[1]; lw[1] = aw[2] * dx[0] - aw[0] * dx[2]; lw[2] = aw[0] * dx[1] - aw[1] * dx[0];
  }
  for (int c = 0; c < 3; ++c) {
    col[c] = tcp[c] * lw[0] + tcp[4 + c] * lw[1] + tcp[8 + c] * lw[2];
    if (t < 9) col[3 + c] = tcp[c] * aw[0] + tcp[4 + c] * aw[1] + tcp[8 + c] * aw[2];
  }
}

int orc_ik_batch(const RampModelDesc *d, const RampQMap *qm, const RampIkParams *p, const double *seed, int64_t n, int ld,
                 const double *target, int one_target, double *q_out, double *err_out, uint8_t *ok, int32_t *iters) {
  if (check_desc(d) || !qm || !p || qm->n_act > 16) return -1;
  const int na = qm->n_act, nj = d->n_joints;
  int col_joint[16], mim_col[8], mim_joint[8], n_mim = 0;
  double mim_mult[8];
  unsigned char chain[ORC_MAX_JOINTS];
  memset(chain, 0, sizeof chain);
  for (int j = p->frame_parent_joint; j >= 0; j = d->joints[j].parent) chain[j] = 1;
  for (int k = 0; k < na; ++k) {
    col_joint[k] = -1;
    for (int j = 0; j < nj; ++j) if (d->joints[j].idx_q == qm->act_slot[k]) col_joint[k] = j;
    if (col_joint[k] < 0 || d->joints[col_joint[k]].type > RAMP_JOINT_PRISMATIC_UNALIGNED) return -1;
  }
  for (int i = 0; i < qm->n_mimic && n_mim < 8; ++i) {
    int jd = -1, col = -1;
    for (int j = 0; j < nj; ++j) if (d->joints[j].idx_q == qm->mimic_dst[i]) jd = j;
    for (int k = 0; k < na; ++k) if (qm->act_slot[k] == qm->mimic_src[i]) col = k;
    if (jd < 0 || col < 0) continue;
    mim_col[n_mim] = col; mim_joint[n_mim] = jd; mim_mult[n_mim] = qm->mimic_mult[i]; ++n_mim;
  }
  const double lam2 = p->damping * p->damping, step = p->step > 0 ? p->step : 1.0, max_dq = p->max_step > 0 ? p->max_step : 1e300;
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t i = 0; i < n; ++i) {
    double act[16], q[ORC_MAX_NQ], oMi[12 * ORC_MAX_JOINTS], tcp[12], dm[12], e[6], err = 1e300;
    for (int k = 0; k < na; ++k) act[k] = seed[i * ld + k];
    const double *tg = target + (one_target ? 0 : 12 * i);
    int it = 0, done = 0;
    for (;;) {
      qmap_one(qm, d->nq, act, q);
      fk_one(d, q, oMi);
      if (p->frame_parent_joint < 0) memcpy(tcp, p->frame_placement, sizeof tcp);
      else se3_mul(oMi + 12 * p->frame_parent_joint, p->frame_placement, tcp);
      se3_inv_mul(tcp, tg, dm);
      e[0] = dm[3]; e[1] = dm[7]; e[2] = dm[11];
      log_so3(dm, e + 3);
      err = 0;
      for (int k = 0; k < 6; ++k) err += e[k] * e[k];
      err = sqrt(err);
      if (err < p->tolerance) { done = 1; break; }
      if (it >= p->max_iterations) break;
      double J[16][6], A[6][6], y[6];
      for (int k = 0; k < na; ++k) {
        ik_column(d, oMi, col_joint[k], chain, tcp, J[k]);
        for (int mi = 0; mi < n_mim; ++mi) if (mim_col[mi] == k) {
          double c2[6];
          ik_column(d, oMi, mim_joint[mi], chain, tcp, c2);
          for (int r = 0; r < 6; ++r) J[k][r] += mim_mult[mi] * c2[r];
        }
      }
      for (int r = 0; r < 6; ++r) for (int c = 0; c < 6; ++c) {
        double sacc = r == c ? lam2 : 0.0;
        for (int k = 0; k < na; ++k) sacc += J[k][r] * J[k][c];
        A[r][c] = sacc;
      }
      for (int r = 0; r < 6; ++r) for (int c = 0; c <= r; ++c) {  /* Cholesky, lower triangle in place */
        double sacc = A[r][c];
        for (int k = 0; k < c; ++k) sacc -= A[r][k] * A[c][k];
        A[r][c] = r == c ? sqrt(fmax(sacc, 1e-300)) : sacc / A[c][c];
      }
      for (int r = 0; r < 6; ++r) { double sacc = e[r]; for (int k = 0; k < r; ++k) sacc -= A[r][k] * y[k]; y[r] = sacc / A[r][r]; }
      for (int r = 5; r >= 0; --r) { double sacc = y[r]; for (int k = r + 1; k < 6; ++k) sacc -= A[k][r] * y[k]; y[r] = sacc / A[r][r]; }
      for (int k = 0; k < na; ++k) {
        double dq = 0;
        for (int r = 0; r < 6; ++r) dq += J[k][r] * y[r];
        dq *= step;
        dq = dq > max_dq ? max_dq : (dq < -max_dq ? -max_dq : dq);
        double v = act[k] + dq;
        if (p->lower && v < p->lower[k]) v = p->lower[k];
        if (p->upper && v > p->upper[k]) v = p->upper[k];
        act[k] = v;
      }
      ++it;
    }
    for (int k = 0; k < na; ++k) q_out[i * na + k] = act[k];
    if (err_out) err_out[i] = err;
    if (ok) ok[i] = (uint8_t)done;
    if (iters) iters[i] = it;
  }
  return 0;
}

/* ---------------------------------------------------------- flop accounting */
/* Algorithmic FP32 work of one validity check, SURVEY §8(d): FK (83 per
 * revolute, 69 per prismatic, 150 per free-flyer/planar/spherical joint, minus
 * 63 for a root joint), 63 per movable geometry, and per pair: 11 when the
 * bounding spheres are disjoint, else the closed-form cost of the class or
 * GJK iterations x (6 (V1 + V2) + 2*33 + 150), iterations counted by THIS
 * oracle's GJK in its boolean early-out mode.  With stop_at_first the pair
 * list of a configuration ends at its first colliding pair in list order,
 * which is where the reference stops (src/ramp/robot_state.py:379-386) and
 * what the timed boolean call computes.  Used by bench.py for the roofline
 * numerator only. */
static double bound_radius(const RampGeom *g, const double *verts) {
  switch (g->type) {
  case RAMP_SHAPE_SPHERE: return g->params[0];
  case RAMP_SHAPE_CAPSULE: return g->params[0] + g->params[1];
  case RAMP_SHAPE_BOX: return sqrt(g->params[0] * g->params[0] + g->params[1] * g->params[1] + g->params[2] * g->params[2]);
  case RAMP_SHAPE_CYLINDER: return sqrt(g->params[0] * g->params[0] + g->params[1] * g->params[1]);
  case RAMP_SHAPE_ELLIPSOID: return fmax(g->params[0], fmax(g->params[1], g->params[2]));
  default: {
    double r = 0;
    const double *p = verts + 3 * (size_t)g->vert_offset;
    for (int i = 0; i < g->vert_count; ++i) r = fmax(r, sqrt(p[3 * i] * p[3 * i] + p[3 * i + 1] * p[3 * i + 1] + p[3 * i + 2] * p[3 * i + 2]));
    return r;
  }
  }
}

int orc_flops_batch(const RampModelDesc *d, const double *q, int64_t n, int ld, int stop_at_first, double *flops) {
  if (check_desc(d)) return -1;
  double fk = 0;
  for (int i = 0; i < d->n_joints; ++i) {
    const int t = d->joints[i].type;
    double f = 83;
    if (t >= RAMP_JOINT_PX && t <= RAMP_JOINT_PRISMATIC_UNALIGNED) f = 69;
    else if (t >= RAMP_JOINT_PLANAR) f = 150;
    if (d->joints[i].parent < 0) f -= 63;
    fk += f;
  }
  for (int g = 0; g < d->n_geoms; ++g) if (d->geoms[g].parent_joint >= 0) fk += 63;
  double br[ORC_MAX_GEOMS];
  for (int g = 0; g < d->n_geoms; ++g) br[g] = bound_radius(&d->geoms[g], d->verts);
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t i = 0; i < n; ++i) {
    double oMi[12 * ORC_MAX_JOINTS], oMg[12 * ORC_MAX_GEOMS];
    fk_one(d, q + i * ld, oMi);
    placements_one(d, oMi, oMg);
    double f = fk;
    for (int p = 0; p < d->n_pairs; ++p) {
      const int a = d->pairs[2 * p], b = d->pairs[2 * p + 1];
      const RampGeom *ga = &d->geoms[a], *gb = &d->geoms[b];
      const double dc = vnorm(vsub(m_trans(oMg + 12 * a), m_trans(oMg + 12 * b)));
      f += 11;
      if (dc > br[a] + br[b]) continue;
      const int t1 = ga->type < gb->type ? ga->type : gb->type, t2 = ga->type < gb->type ? gb->type : ga->type;
      int it = 0;
      const double dist = shape_distance(ga, oMg + 12 * a, gb, oMg + 12 * b, d->verts, 0, 0.0, &it);
      if (t1 == RAMP_SHAPE_SPHERE && t2 == RAMP_SHAPE_SPHERE) { /* the sphere test was the whole pair */ }
      else if (t1 == RAMP_SHAPE_SPHERE && t2 <= RAMP_SHAPE_CYLINDER) f += 30;
      else if (t1 == RAMP_SHAPE_CAPSULE && t2 == RAMP_SHAPE_CAPSULE) f += 90;
      else {
        const double v = (ga->type == RAMP_SHAPE_CONVEX ? ga->vert_count : 0) + (gb->type == RAMP_SHAPE_CONVEX ? gb->vert_count : 0);
        f += it * (6.0 * v + 2 * 33 + 150);
      }
      if (stop_at_first && dist <= 1e-12) break;
    }
    flops[i] = f;
  }
  return 0;
}
