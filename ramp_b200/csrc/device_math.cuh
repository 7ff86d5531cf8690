// device_math.cuh -- FP32 SE(3), joint transforms, closed-form shape tests and GJK
// for the sm_100a validity kernels.  Everything here is register math: no global or
// shared memory traffic except the convex-vertex scans, which read the vertex table
// the caller staged in shared memory.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ramp {

// ----------------------------------------------------------------------------- SE(3)
// 3x4 row-major [R|t]; m[4*r+c].
struct Xf {
  float m[12];
};

__device__ __forceinline__ Xf xf_identity() {
  Xf o;
#pragma unroll
  for (int i = 0; i < 12; ++i) o.m[i] = 0.f;
  o.m[0] = o.m[5] = o.m[10] = 1.f;
  return o;
}

// a * b = (Ra Rb, ta + Ra tb)
__device__ __forceinline__ Xf xf_mul(const Xf& a, const Xf& b) {
  Xf o;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j)
      o.m[4 * i + j] = fmaf(a.m[4 * i], b.m[j], fmaf(a.m[4 * i + 1], b.m[4 + j], a.m[4 * i + 2] * b.m[8 + j]));
    o.m[4 * i + 3] =
        fmaf(a.m[4 * i], b.m[3], fmaf(a.m[4 * i + 1], b.m[7], fmaf(a.m[4 * i + 2], b.m[11], a.m[4 * i + 3])));
  }
  return o;
}

// a^-1 * b
__device__ __forceinline__ Xf xf_inv_mul(const Xf& a, const Xf& b) {
  Xf o;
  const float dx = b.m[3] - a.m[3], dy = b.m[7] - a.m[7], dz = b.m[11] - a.m[11];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j)
      o.m[4 * i + j] = fmaf(a.m[i], b.m[j], fmaf(a.m[4 + i], b.m[4 + j], a.m[8 + i] * b.m[8 + j]));
    o.m[4 * i + 3] = fmaf(a.m[i], dx, fmaf(a.m[4 + i], dy, a.m[8 + i] * dz));
  }
  return o;
}

struct V3 {
  float x, y, z;
};
__device__ __forceinline__ V3 mk(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ V3 rot(const Xf& m, V3 a) {
  return mk(fmaf(m.m[0], a.x, fmaf(m.m[1], a.y, m.m[2] * a.z)), fmaf(m.m[4], a.x, fmaf(m.m[5], a.y, m.m[6] * a.z)),
            fmaf(m.m[8], a.x, fmaf(m.m[9], a.y, m.m[10] * a.z)));
}
__device__ __forceinline__ V3 rot_t(const Xf& m, V3 a) {
  return mk(fmaf(m.m[0], a.x, fmaf(m.m[4], a.y, m.m[8] * a.z)), fmaf(m.m[1], a.x, fmaf(m.m[5], a.y, m.m[9] * a.z)),
            fmaf(m.m[2], a.x, fmaf(m.m[6], a.y, m.m[10] * a.z)));
}
__device__ __forceinline__ V3 apply(const Xf& m, V3 a) {
  V3 r = rot(m, a);
  return mk(r.x + m.m[3], r.y + m.m[7], r.z + m.m[11]);
}
__device__ __forceinline__ V3 trans(const Xf& m) { return mk(m.m[3], m.m[7], m.m[11]); }

// ------------------------------------------------------------------ joint transforms
// Joint type codes are include/ramp_b200.h RampJointType.
__device__ __forceinline__ void rot_aligned(Xf& o, int axis, float c, float s) {
  if (axis == 0) { o.m[5] = c; o.m[6] = -s; o.m[9] = s; o.m[10] = c; }
  else if (axis == 1) { o.m[0] = c; o.m[2] = s; o.m[8] = -s; o.m[10] = c; }
  else { o.m[0] = c; o.m[1] = -s; o.m[4] = s; o.m[5] = c; }
}

__device__ __forceinline__ void rot_rodrigues(Xf& o, float ax, float ay, float az, float c, float s) {
  const float v = 1.f - c;
  o.m[0] = fmaf(v * ax, ax, c);        o.m[1] = fmaf(v * ax, ay, -s * az); o.m[2] = fmaf(v * ax, az, s * ay);
  o.m[4] = fmaf(v * ay, ax, s * az);   o.m[5] = fmaf(v * ay, ay, c);       o.m[6] = fmaf(v * ay, az, -s * ax);
  o.m[8] = fmaf(v * az, ax, -s * ay);  o.m[9] = fmaf(v * az, ay, s * ax);  o.m[10] = fmaf(v * az, az, c);
}

// Eigen toRotationMatrix: quaternion used as given, not normalised.
__device__ __forceinline__ void rot_quat(Xf& o, float x, float y, float z, float w) {
  const float tx = 2.f * x, ty = 2.f * y, tz = 2.f * z;
  const float twx = tx * w, twy = ty * w, twz = tz * w;
  const float txx = tx * x, txy = ty * x, txz = tz * x;
  const float tyy = ty * y, tyz = tz * y, tzz = tz * z;
  o.m[0] = 1.f - (tyy + tzz); o.m[1] = txy - twz;         o.m[2] = txz + twy;
  o.m[4] = txy + twz;         o.m[5] = 1.f - (txx + tzz); o.m[6] = tyz - twx;
  o.m[8] = txz - twy;         o.m[9] = tyz + twx;         o.m[10] = 1.f - (txx + tyy);
}

// M_i(q): q0..q6 are the joint's q slots (unused ones are don't-care).
// sincosf (not __sincosf): pose parity is 1e-5 and joint ranges reach +-2pi.
__device__ __forceinline__ Xf joint_xf(int type, float ax, float ay, float az, const float* q) {
  Xf o = xf_identity();
  float s, c;
  switch (type) {
    case 1: case 2: case 3: sincosf(q[0], &s, &c); rot_aligned(o, type - 1, c, s); break;
    case 4: sincosf(q[0], &s, &c); rot_rodrigues(o, ax, ay, az, c, s); break;
    case 5: case 6: case 7: rot_aligned(o, type - 5, q[0], q[1]); break;
    case 8: rot_rodrigues(o, ax, ay, az, q[0], q[1]); break;
    case 9: o.m[3] = q[0]; break;
    case 10: o.m[7] = q[0]; break;
    case 11: o.m[11] = q[0]; break;
    case 12: o.m[3] = q[0] * ax; o.m[7] = q[0] * ay; o.m[11] = q[0] * az; break;
    case 13: rot_aligned(o, 2, q[2], q[3]); o.m[3] = q[0]; o.m[7] = q[1]; break;
    case 14: rot_quat(o, q[3], q[4], q[5], q[6]); o.m[3] = q[0]; o.m[7] = q[1]; o.m[11] = q[2]; break;
    case 15: rot_quat(o, q[0], q[1], q[2], q[3]); break;
    case 16: {
      Xf rz = xf_identity(), ry = xf_identity(), rx = xf_identity();
      sincosf(q[0], &s, &c); rot_aligned(rz, 2, c, s);
      sincosf(q[1], &s, &c); rot_aligned(ry, 1, c, s);
      sincosf(q[2], &s, &c); rot_aligned(rx, 0, c, s);
      o = xf_mul(xf_mul(rz, ry), rx);
    } break;
    case 17: o.m[3] = q[0]; o.m[7] = q[1]; o.m[11] = q[2]; break;
    default: break;
  }
  return o;
}

__device__ __forceinline__ int joint_nq(int type) {
  switch (type) {
    case 5: case 6: case 7: case 8: return 2;
    case 13: return 4;
    case 14: return 7;
    case 15: return 4;
    case 16: case 17: return 3;
    default: return 1;
  }
}

// ------------------------------------------------------------------------- shapes
// Shape type codes are include/ramp_b200.h RampShapeType.
enum { SH_SPHERE = 0, SH_CAPSULE = 1, SH_BOX = 2, SH_CYLINDER = 3, SH_CONVEX = 4, SH_ELLIPSOID = 5 };

struct Shape {
  int type;
  float p0, p1, p2;  // params
  int voff, vcnt;    // convex vertices (float4 units) in the staged vertex table
};

__device__ __forceinline__ float swept_radius(const Shape& s) {
  return (s.type == SH_SPHERE || s.type == SH_CAPSULE) ? s.p0 : 0.f;
}

// Support point of the shape CORE (sphere -> point, capsule -> segment) along d, local frame.
__device__ __forceinline__ V3 support(const Shape& s, const float4* __restrict__ verts, V3 d) {
  switch (s.type) {
    case SH_SPHERE: return mk(0.f, 0.f, 0.f);
    case SH_CAPSULE: return mk(0.f, 0.f, d.z > 0.f ? s.p1 : (d.z < 0.f ? -s.p1 : 0.f));
    case SH_BOX: return mk(d.x > 0.f ? s.p0 : -s.p0, d.y > 0.f ? s.p1 : -s.p1, d.z > 0.f ? s.p2 : -s.p2);
    case SH_CYLINDER: {
      const float zz = d.z > 0.f ? s.p1 : (d.z < 0.f ? -s.p1 : 0.f);
      const float n2 = fmaf(d.x, d.x, d.y * d.y);
      if (n2 > 0.f) {
        const float k = s.p0 * rsqrtf(n2);
        return mk(d.x * k, d.y * k, zz);
      }
      return mk(0.f, 0.f, zz);
    }
    case SH_ELLIPSOID: {
      const V3 v = mk(s.p0 * s.p0 * d.x, s.p1 * s.p1 * d.y, s.p2 * s.p2 * d.z);
      const float den = dot(v, d);
      return den > 0.f ? v * rsqrtf(den) : mk(0.f, 0.f, 0.f);
    }
    default: {  // SH_CONVEX: brute-force scan of the staged vertices
      const float4* __restrict__ p = verts + s.voff;
      float best = -3.0e38f;
      float4 bv = p[0];
      int i = 0;
      for (; i + 4 <= s.vcnt; i += 4) {
        const float4 a = p[i], b = p[i + 1], c = p[i + 2], e = p[i + 3];
        const float da = fmaf(a.x, d.x, fmaf(a.y, d.y, a.z * d.z));
        const float db = fmaf(b.x, d.x, fmaf(b.y, d.y, b.z * d.z));
        const float dc = fmaf(c.x, d.x, fmaf(c.y, d.y, c.z * d.z));
        const float de = fmaf(e.x, d.x, fmaf(e.y, d.y, e.z * d.z));
        if (da > best) { best = da; bv = a; }
        if (db > best) { best = db; bv = b; }
        if (dc > best) { best = dc; bv = c; }
        if (de > best) { best = de; bv = e; }
      }
      for (; i < s.vcnt; ++i) {
        const float4 a = p[i];
        const float da = fmaf(a.x, d.x, fmaf(a.y, d.y, a.z * d.z));
        if (da > best) { best = da; bv = a; }
      }
      return mk(bv.x, bv.y, bv.z);
    }
  }
}

// ------------------------------------------------------------- closed-form distances
__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

// distance of point p to segment (0,0,-h)..(0,0,h)
__device__ __forceinline__ float point_zseg_dist(V3 p, float h) {
  const float dz = p.z - clampf(p.z, -h, h);
  return sqrtf(fmaf(p.x, p.x, fmaf(p.y, p.y, dz * dz)));
}

__device__ __forceinline__ float sphere_box_dist(V3 c, float hx, float hy, float hz, float r) {
  const float ax = fabsf(c.x) - hx, ay = fabsf(c.y) - hy, az = fabsf(c.z) - hz;
  const float ox = fmaxf(ax, 0.f), oy = fmaxf(ay, 0.f), oz = fmaxf(az, 0.f);
  const float out2 = fmaf(ox, ox, fmaf(oy, oy, oz * oz));
  if (out2 > 0.f) return sqrtf(out2) - r;
  return fmaxf(ax, fmaxf(ay, az)) - r;  // inside: minus depth to the nearest face
}

__device__ __forceinline__ float sphere_cyl_dist(V3 c, float R, float h, float r) {
  const float dr = sqrtf(fmaf(c.x, c.x, c.y * c.y)) - R, dz = fabsf(c.z) - h;
  if (dr <= 0.f && dz <= 0.f) return fmaxf(dr, dz) - r;
  if (dr <= 0.f) return dz - r;
  if (dz <= 0.f) return dr - r;
  return sqrtf(fmaf(dr, dr, dz * dz)) - r;
}

// closest distance between segments p1 + s d1 and p2 + t d2, s,t in [0,1]
__device__ __forceinline__ float seg_seg_dist(V3 p1, V3 d1, V3 p2, V3 d2) {
  const V3 r = p1 - p2;
  const float a = dot(d1, d1), e = dot(d2, d2), f = dot(d2, r);
  const float EPS = 1e-20f;
  float s, t;
  if (a <= EPS && e <= EPS) return sqrtf(dot(r, r));
  if (a <= EPS) { s = 0.f; t = clampf(f / e, 0.f, 1.f); }
  else {
    const float c = dot(d1, r);
    if (e <= EPS) { t = 0.f; s = clampf(-c / a, 0.f, 1.f); }
    else {
      const float b = dot(d1, d2), den = a * e - b * b;
      s = den > EPS ? clampf((b * f - c * e) / den, 0.f, 1.f) : 0.f;
      t = (b * s + f) / e;
      if (t < 0.f) { t = 0.f; s = clampf(-c / a, 0.f, 1.f); }
      else if (t > 1.f) { t = 1.f; s = clampf((b - c) / a, 0.f, 1.f); }
    }
  }
  const V3 w = (p1 + d1 * s) - (p2 + d2 * t);
  return sqrtf(dot(w, w));
}

// Oriented-box overlap (separating axis theorem, 15 axes).  Box A: half extents a, at the
// origin, axis aligned; box B: half extents b, pose m (B in A's frame).  Returns true when
// NO separating axis exists (touching counts as overlap).  `slack` widens every test:
// > 0 makes a conservative cull, 0 the exact box-box decision.
__device__ __forceinline__ bool obb_overlap(V3 a, V3 b, const Xf& m, float slack) {
  const float R[3][3] = {{m.m[0], m.m[1], m.m[2]}, {m.m[4], m.m[5], m.m[6]}, {m.m[8], m.m[9], m.m[10]}};
  float A[3][3];
  const float eps = 1e-6f;  // guards the cross-product axes of (nearly) parallel edges
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) A[i][j] = fabsf(R[i][j]) + eps;
  const float t[3] = {m.m[3], m.m[7], m.m[11]};
  const float ae[3] = {a.x, a.y, a.z}, be[3] = {b.x, b.y, b.z};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const float rb = fmaf(be[0], A[i][0], fmaf(be[1], A[i][1], be[2] * A[i][2]));
    if (fabsf(t[i]) > ae[i] + rb + slack) return false;
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float ra = fmaf(ae[0], A[0][j], fmaf(ae[1], A[1][j], ae[2] * A[2][j]));
    const float tp = fmaf(t[0], R[0][j], fmaf(t[1], R[1][j], t[2] * R[2][j]));
    if (fabsf(tp) > ra + be[j] + slack) return false;
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      const float ra = ae[i1] * A[i2][j] + ae[i2] * A[i1][j];
      const float rb = be[j1] * A[i][j2] + be[j2] * A[i][j1];
      const float tp = t[i2] * R[i1][j] - t[i1] * R[i2][j];
      if (fabsf(tp) > ra + rb + slack) return false;
    }
  }
  return true;
}

// ------------------------------------------------------------------------------ GJK
// Closest point of a simplex to the origin, Voronoi-region form; the simplex is reduced
// in place to its supporting sub-simplex.
struct Simplex {
  V3 p[4];
  int n;
};

__device__ __forceinline__ V3 closest_segment(Simplex& s) {
  const V3 a = s.p[0], b = s.p[1], ab = b - a;
  const float t = -dot(a, ab), den = dot(ab, ab);
  if (t <= 0.f || den <= 0.f) { s.n = 1; return a; }
  if (t >= den) { s.p[0] = b; s.n = 1; return b; }
  return a + ab * (t / den);
}

__device__ __forceinline__ V3 closest_triangle(V3 a, V3 b, V3 c, Simplex& o) {
  const V3 ab = b - a, ac = c - a;
  const float d1 = -dot(ab, a), d2 = -dot(ac, a);
  if (d1 <= 0.f && d2 <= 0.f) { o.p[0] = a; o.n = 1; return a; }
  const float d3 = -dot(ab, b), d4 = -dot(ac, b);
  if (d3 >= 0.f && d4 <= d3) { o.p[0] = b; o.n = 1; return b; }
  const float vc = d1 * d4 - d3 * d2;
  if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) {
    o.p[0] = a; o.p[1] = b; o.n = 2;
    return a + ab * (d1 / (d1 - d3));
  }
  const float d5 = -dot(ab, c), d6 = -dot(ac, c);
  if (d6 >= 0.f && d5 <= d6) { o.p[0] = c; o.n = 1; return c; }
  const float vb = d5 * d2 - d1 * d6;
  if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) {
    o.p[0] = a; o.p[1] = c; o.n = 2;
    return a + ac * (d2 / (d2 - d6));
  }
  const float va = d3 * d6 - d5 * d4;
  if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) {
    o.p[0] = b; o.p[1] = c; o.n = 2;
    return b + (c - b) * ((d4 - d3) / ((d4 - d3) + (d5 - d6)));
  }
  // interior: project the origin on the plane (more accurate in FP32 than barycentric sums)
  const V3 n = cross(ab, ac);
  const float nn = dot(n, n);
  o.p[0] = a; o.p[1] = b; o.p[2] = c; o.n = 3;
  if (nn > 0.f) return n * (dot(a, n) / nn);
  const float den = 1.f / (va + vb + vc);
  return a + ab * (vb * den) + ac * (vc * den);
}

// returns true when the origin is inside (or on) the tetrahedron
__device__ __forceinline__ bool closest_tetra(Simplex& s, V3& out) {
  float best = 3.0e38f;
  Simplex bs;
  bs.n = 0;
  V3 bp = mk(0.f, 0.f, 0.f);
  bool outside_any = false;
#pragma unroll
  for (int f = 0; f < 4; ++f) {
    // faces (0,1,2|3) (0,2,3|1) (0,3,1|2) (1,3,2|0)
    const int ia = f == 3 ? 1 : 0, ib = f == 0 ? 1 : (f == 1 ? 2 : 3), ic = f == 0 ? 2 : (f == 1 ? 3 : (f == 2 ? 1 : 2)),
              id = f == 0 ? 3 : (f == 1 ? 1 : (f == 2 ? 2 : 0));
    const V3 a = s.p[ia], b = s.p[ib], c = s.p[ic], d = s.p[id];
    const V3 n = cross(b - a, c - a);
    const float so = -dot(a, n), sd = dot(d - a, n);
    if (so * sd < 0.f || sd == 0.f) {
      outside_any = true;
      Simplex o;
      const V3 p = closest_triangle(a, b, c, o);
      const float dd = dot(p, p);
      if (dd < best) { best = dd; bs = o; bp = p; }
    }
  }
  if (!outside_any) return true;
  s = bs;
  out = bp;
  return false;
}

struct GjkResult {
  bool hit;     // swept shapes overlap (distance <= 0)
  float dist;   // core distance estimate minus swept radii; meaningful when !hit
  int iters;
};

// Boolean GJK with a distance estimate.  Shape A at the origin of its own frame, shape B at
// pose m (B in A's frame).  `exact`: iterate to convergence even after a separating plane is
// known (distance output); otherwise stop as soon as separation or overlap is proven.
__device__ __forceinline__ GjkResult gjk(const Shape& A, const Shape& B, const Xf& m, const float4* __restrict__ verts,
                                       bool exact) {
  const float rsum = swept_radius(A) + swept_radius(B);
  const float rs2 = rsum * rsum;
  const float EPS_REL = 1e-5f, TINY2 = 1e-12f;
  V3 v = -trans(m);  // centre(A) - centre(B): a point of the Minkowski difference
  float vv = dot(v, v);
  if (vv < TINY2) { v = mk(1.f, 0.f, 0.f); vv = 1.f; }
  Simplex s;
  s.n = 0;
  GjkResult r;
  r.hit = false; r.dist = 0.f; r.iters = 0;
  float lower = 0.f;  // best proven lower bound on the core distance
  for (int it = 0; it < 48; ++it) {
    r.iters = it + 1;
    const V3 dir = -v;
    const V3 sa = support(A, verts, dir);
    const V3 sb = apply(m, support(B, verts, rot_t(m, v)));
    const V3 w = sa - sb;
    const float vw = dot(v, w);
    if (vw > 0.f) {
      const float om = vw * rsqrtf(vv);
      lower = fmaxf(lower, om);
      if (!exact && om > rsum) { r.dist = om - rsum; return r; }  // separating plane
    }
    if (s.n > 0 && vv - vw <= EPS_REL * vv) break;  // converged: |v| is the core distance
    s.p[s.n++] = w;
    V3 nv;
    if (s.n == 1) nv = s.p[0];
    else if (s.n == 2) nv = closest_segment(s);
    else if (s.n == 3) { Simplex o; nv = closest_triangle(s.p[0], s.p[1], s.p[2], o); s = o; }
    else if (closest_tetra(s, nv)) { r.hit = true; r.dist = -rsum; return r; }
    const float nvv = dot(nv, nv);
    if (nvv <= fmaxf(rs2, TINY2) && !exact) { r.hit = true; r.dist = sqrtf(nvv) - rsum; return r; }
    if (nvv <= TINY2) { r.hit = true; r.dist = -rsum; return r; }
    if (it > 0 && nvv >= vv) break;  // no progress at FP32 resolution
    v = nv; vv = nvv;
  }
  const float d = sqrtf(vv);
  r.dist = d - rsum;
  r.hit = d <= rsum;
  // a stalled iteration far from convergence: trust the proven lower bound
  if (r.hit && lower > rsum) { r.hit = false; r.dist = lower - rsum; }
  return r;
}

}  // namespace ramp
