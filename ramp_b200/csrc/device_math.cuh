// device_math.cuh -- FP32 SE(3), joint transforms, closed-form shape tests and GJK
// for the sm_100a validity kernels.  Everything here is register math: no global or
// shared memory traffic except the convex-vertex scans, which read the vertex table
// the caller staged in shared memory.
#pragma once
#include <cuda_fp16.h>
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

// sin/cos with Cody-Waite reduction to [-pi/4, pi/4] and minimax polynomials (the fast path
// of sinf/cosf without their large-argument slow path, which alone is several hundred
// instructions).  |error| ~ 1e-7 for |x| < 1e4 -- joint angles; pose parity is 1e-5.
__device__ __noinline__ void sincos_small(float x, float* sn, float* cs) {
  const float k = rintf(x * 0.636619772367581343f);
  float r = fmaf(-k, 1.5703125f, x);  // pi/2 split in three: k * 1.5703125 is exact
  r = fmaf(-k, 4.837512969970703125e-4f, r);
  r = fmaf(-k, 7.54978995489188216e-8f, r);
  const int qd = static_cast<int>(k) & 3;
  const float r2 = r * r;
  float s = fmaf(fmaf(fmaf(-1.9515295891e-4f, r2, 8.3321608736e-3f), r2, -1.6666654611e-1f), r2 * r, r);
  float c = fmaf(fmaf(fmaf(fmaf(2.443315711809948e-5f, r2, -1.388731625493765e-3f), r2, 4.166664568298827e-2f), r2, -0.5f), r2, 1.f);
  if (qd & 1) { const float t = s; s = c; c = -t; }
  if (qd & 2) { s = -s; c = -c; }
  *sn = s; *cs = c;
}

// Hardware sine / cosine for the validity kernels: the argument is first brought into [-pi, pi]
// (two-term reduction by 2 pi), where MUFU's absolute error is 2^-21.4 ~ 4e-7 -- a few micrometres
// at the end of an arm, a few per cent of the +-1e-4 m margin band.  Kernels that REPORT poses or
// distances use sincos_small.
__device__ __forceinline__ void sincos_fast(float x, float* sn, float* cs) {
  const float k = rintf(x * 0.159154943091895336f);
  float r = fmaf(-k, 6.28318548202514648f, x);
  r = fmaf(k, 1.74845553146e-7f, r);  // 2 pi = 6.28318548202514648 - 1.74845553146e-7
  __sincosf(r, sn, cs);
}

// M_i(q): q0..q6 are the joint's q slots (unused ones are don't-care).
__device__ __forceinline__ Xf joint_xf(int type, float ax, float ay, float az, const float* q) {
  Xf o = xf_identity();
  float s, c;
  switch (type) {
    case 1: case 2: case 3: sincos_small(q[0], &s, &c); rot_aligned(o, type - 1, c, s); break;
    case 4: sincos_small(q[0], &s, &c); rot_rodrigues(o, ax, ay, az, c, s); break;
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
      sincos_small(q[0], &s, &c); rot_aligned(rz, 2, c, s);
      sincos_small(q[1], &s, &c); rot_aligned(ry, 1, c, s);
      sincos_small(q[2], &s, &c); rot_aligned(rx, 0, c, s);
      o = xf_mul(xf_mul(rz, ry), rx);
    } break;
    case 17: o.m[3] = q[0]; o.m[7] = q[1]; o.m[11] = q[2]; break;
    default: break;
  }
  return o;
}

__host__ __device__ __forceinline__ int joint_nq(int type) {
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

// Support point of a shape CORE along d in the shape's local frame.  Every polytope core is a
// vertex list in the staged table -- sphere: 1 vertex (its centre), capsule: the 2 segment
// ends, box: its 8 corners, convex: the hull vertices -- so one scan loop serves them all and
// lanes holding different shape types do not diverge.  Cylinder and ellipsoid have closed
// forms.  (type, p0..p2) are the geometry's shape code and parameters.
__device__ __forceinline__ V3 support(int type, float p0, float p1, float p2, const float4* __restrict__ p, int vcnt,
                                      V3 d) {
  if (type == SH_CYLINDER) {
    const float zz = d.z > 0.f ? p1 : (d.z < 0.f ? -p1 : 0.f);
    const float n2 = fmaf(d.x, d.x, d.y * d.y);
    const float k = n2 > 0.f ? p0 * rsqrtf(n2) : 0.f;
    return mk(d.x * k, d.y * k, zz);
  }
  if (type == SH_ELLIPSOID) {
    const V3 v = mk(p0 * p0 * d.x, p1 * p1 * d.y, p2 * p2 * d.z);
    const float den = dot(v, d);
    return den > 0.f ? v * rsqrtf(den) : mk(0.f, 0.f, 0.f);
  }
  // Four vertices per trip: only the largest of their four dot products is compared with the best
  // so far, and what is remembered is WHERE the winning group starts -- which of its four vertices
  // it was is settled once, after the scan, by looking at that group again.  5.5 instructions per
  // vertex instead of 9 (a compare and four selects per vertex before).
  float best = -3.0e38f;
  float bx = 0.f, by = 0.f, bz = 0.f;
  int i = 0, bg = -1;
#pragma unroll 1
  for (; i + 4 <= vcnt; i += 4) {
    const float4 a = p[i], b = p[i + 1], c = p[i + 2], e = p[i + 3];
    const float da = fmaf(a.x, d.x, fmaf(a.y, d.y, a.z * d.z));
    const float db = fmaf(b.x, d.x, fmaf(b.y, d.y, b.z * d.z));
    const float dc = fmaf(c.x, d.x, fmaf(c.y, d.y, c.z * d.z));
    const float de = fmaf(e.x, d.x, fmaf(e.y, d.y, e.z * d.z));
    const float m = fmaxf(fmaxf(da, db), fmaxf(dc, de));
    if (m > best) { best = m; bg = i; }
  }
  if (bg >= 0) {
    const float4 a = p[bg], b = p[bg + 1], c = p[bg + 2], e = p[bg + 3];
    const float da = fmaf(a.x, d.x, fmaf(a.y, d.y, a.z * d.z));
    const float db = fmaf(b.x, d.x, fmaf(b.y, d.y, b.z * d.z));
    const float dc = fmaf(c.x, d.x, fmaf(c.y, d.y, c.z * d.z));
    float4 w = a;
    float dw = da;
    if (db > dw) { dw = db; w = b; }
    if (dc > dw) { dw = dc; w = c; }
    if (dw < best) w = e;  // none of the first three reaches the group's maximum: it was the fourth
    bx = w.x; by = w.y; bz = w.z;
  }
#pragma unroll 1
  for (; i < vcnt; ++i) {  // the one to three vertices after the last whole group
    const float4 a = p[i];
    const float da = fmaf(a.x, d.x, fmaf(a.y, d.y, a.z * d.z));
    if (da > best) { best = da; bx = a.x; by = a.y; bz = a.z; }
  }
  return mk(bx, by, bz);
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

// closest points between segments p1 + s d1 and p2 + t d2, s,t in [0,1]: returns the witness
// vector c1 - c2 (its length is the distance)
__device__ __forceinline__ V3 seg_seg_witness(V3 p1, V3 d1, V3 p2, V3 d2) {
  const V3 r = p1 - p2;
  const float a = dot(d1, d1), e = dot(d2, d2), f = dot(d2, r), c = dot(d1, r), b = dot(d1, d2);
  const float EPS = 1e-20f;
  // branch-light form of the clamped 2x2 solve (Ericson, Real-Time Collision Detection 5.1.9)
  const float den = fmaf(a, e, -b * b);
  // __fdividef (2 ulp): the parameters only place the closest points; every user of the
  // result compares against a radius sum with at least 1e-5 m of slack or of margin band
  float s = (a > EPS && den > EPS) ? clampf(__fdividef(b * f - c * e, den), 0.f, 1.f) : 0.f;
  float t = e > EPS ? __fdividef(fmaf(b, s, f), e) : 0.f;
  const float tc = clampf(t, 0.f, 1.f);
  if (tc != t || e <= EPS) s = a > EPS ? clampf(__fdividef(b * tc - c, a), 0.f, 1.f) : 0.f;
  return (r + d1 * s) - d2 * tc;
}

__device__ __forceinline__ float seg_seg_dist(V3 p1, V3 d1, V3 p2, V3 d2) {
  const V3 w = seg_seg_witness(p1, d1, p2, d2);
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
// Closest point of a simplex to the origin, Voronoi-region form.  Written so that every
// helper has ONE call site (small instruction footprint: the validity kernel must stay
// inside the instruction cache while its warps sit in different phases).
struct Simplex {
  V3 p[4];
  int n;
};

// closest point of triangle (a,b,c) to the origin; o = supporting sub-simplex
__device__ __forceinline__ V3 closest_triangle(V3 a, V3 b, V3 c, V3& o0, V3& o1, V3& o2, int& on) {
  const V3 ab = b - a, ac = c - a;
  const float d1 = -dot(ab, a), d2 = -dot(ac, a);
  const float d3 = -dot(ab, b), d4 = -dot(ac, b);
  const float d5 = -dot(ab, c), d6 = -dot(ac, c);
  const float vc = d1 * d4 - d3 * d2, vb = d5 * d2 - d1 * d6, va = d3 * d6 - d5 * d4;
  o0 = a; o1 = b; o2 = c;
  if (d1 <= 0.f && d2 <= 0.f) { on = 1; return a; }
  if (d3 >= 0.f && d4 <= d3) { o0 = b; on = 1; return b; }
  if (d6 >= 0.f && d5 <= d6) { o0 = c; on = 1; return c; }
  if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) { on = 2; return a + ab * (d1 / (d1 - d3)); }
  if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) { o1 = c; on = 2; return a + ac * (d2 / (d2 - d6)); }
  if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) {
    o0 = b; o1 = c; on = 2;
    return b + (c - b) * ((d4 - d3) / ((d4 - d3) + (d5 - d6)));
  }
  // interior: barycentric combination.  It must stay a CONVEX combination of the vertices (a
  // point of the Minkowski difference, so |v| never drops below the true distance); projecting
  // the origin on the plane instead breaks that for sliver triangles, whose FP32 normal is noise.
  on = 3;
  const float den = 1.f / (va + vb + vc);
  const float bv = clampf(vb * den, 0.f, 1.f), bw = clampf(vc * den, 0.f, 1.f - bv);
  return a + ab * bv + ac * bw;
}

// Reduce the simplex to the sub-simplex closest to the origin and return that closest point.
// Returns true when the origin is enclosed by a tetrahedron.
__device__ __forceinline__ bool solve_simplex(Simplex& s, V3& out) {
  if (s.n == 1) { out = s.p[0]; return false; }
  if (s.n == 2) {
    const V3 a = s.p[0], b = s.p[1], ab = b - a;
    const float t = -dot(a, ab), den = dot(ab, ab);
    if (t <= 0.f || den <= 0.f) { s.n = 1; out = a; }
    else if (t >= den) { s.p[0] = b; s.n = 1; out = b; }
    else out = a + ab * (t / den);
    return false;
  }
  // triangle: one face; tetrahedron: the four faces, visited by rotating the vertices so the
  // face code below exists once
  const bool tetra = s.n == 4;
  const int nfaces = tetra ? 4 : 1;
  V3 p0 = s.p[0], p1 = s.p[1], p2 = s.p[2], p3 = s.p[3];
  float best = 3.0e38f;
  bool outside_any = false;
  V3 b0 = p0, b1 = p1, b2 = p2, bp = mk(0.f, 0.f, 0.f);
  int bn = 0;
#pragma unroll 1
  for (int f = 0; f < nfaces; ++f) {
    bool outside = true;
    if (tetra) {
      // The origin is outside this face when it lies on the other side than the opposite
      // vertex.  A sliver tetrahedron (tiny height over the face) or an origin within rounding
      // of the face plane makes those signs noise in FP32: treat both as "outside", i.e. fall
      // back to the projection on the face, so a false "origin enclosed" cannot happen.
      const V3 n = cross(p1 - p0, p2 - p0), e = p3 - p0;
      const float so = -dot(p0, n), sd = dot(e, n), nn = dot(n, n);
      outside = so * sd < 0.f || sd * sd <= 1e-10f * dot(e, e) * nn || so * so <= 1e-10f * dot(p0, p0) * nn;
    }
    if (outside) {
      outside_any = true;
      V3 o0, o1, o2;
      int on;
      const V3 q = closest_triangle(p0, p1, p2, o0, o1, o2, on);
      const float dd = dot(q, q);
      if (dd < best) { best = dd; b0 = o0; b1 = o1; b2 = o2; bn = on; bp = q; }
    }
    const V3 t = p0; p0 = p1; p1 = p2; p2 = p3; p3 = t;  // next face
  }
  if (!outside_any) return true;
  s.p[0] = b0; s.p[1] = b1; s.p[2] = b2; s.n = bn;
  out = bp;
  return false;
}

// Per-lane GJK state: a lane keeps one (configuration, pair) work item and advances it one
// iteration per trip of the warp loop; finished lanes fetch the next item, so lanes stay
// busy even though iteration counts differ per item.
struct GjkLane {
  Xf rel;      // pose of shape B in shape A's frame
  V3 v;        // current closest point of the simplex (a point of A (-) B)
  float vv;
  Simplex s;
  float rsum;  // swept radii of the two cores
  float lower; // best proven lower bound on the core distance
  int it;
};

struct GjkShape {
  int type, vcnt;
  float p0, p1, p2;
  const float4* verts;
};

// v0: a point of the Minkowski difference A (-) B to start from (the closer to the origin the
// fewer iterations): the witness vector of the two inner capsules, or centre(A) - centre(B).
__device__ __forceinline__ void gjk_init(GjkLane& g, const Xf& rel, float rsum, V3 v0) {
  g.rel = rel;
  g.v = v0;
  g.vv = dot(g.v, g.v);
  if (g.vv < 1e-12f) { g.v = mk(1.f, 0.f, 0.f); g.vv = 1.f; }
  g.s.n = 0;
  g.rsum = rsum;
  g.lower = 0.f;
  g.it = 0;
}

enum { GJK_CONTINUE = 0, GJK_HIT = 1, GJK_MISS = 2, GJK_STALL = 3 };  // STALL: exact mode only (see gjk_advance)

// The simplex half of one GJK iteration: sa, sb are the support points of A along -v and of B
// along +v (each in its own frame).  `exact`: run to convergence (distance output) instead of
// stopping at the first proof of separation / overlap.  On GJK_MISS / GJK_HIT with exact, dist
// holds the core distance minus the swept radii.
__device__ __forceinline__ int gjk_advance(GjkLane& g, V3 sa, V3 sb, bool exact, float& dist) {
  // |v| below 2e-5 m counts as touching: FP32 cannot resolve a simplex face through the origin
  // any better, and such a pair is inside the +-1e-4 m margin band either way
  const float EPS_REL = 1e-5f, TINY2 = 4e-10f;
  const float rs2 = g.rsum * g.rsum;
  const V3 w = sa - apply(g.rel, sb);
  const float vw = dot(g.v, w);
  if (vw > 0.f) {
    const float om = vw * rsqrtf(g.vv);
    g.lower = fmaxf(g.lower, om);
    if (!exact && om > g.rsum) { dist = om - g.rsum; return GJK_MISS; }  // separating plane
  }
  const bool converged = g.s.n > 0 && g.vv - vw <= EPS_REL * g.vv;  // |v| is the core distance
  bool done = converged, dup = false;
  // A support point that is already a vertex of the simplex (polytope corners come back bit for
  // bit) means v cannot be improved; added again it would make a degenerate triangle whose
  // closest-point solve is 0 / 0 (seen as "origin enclosed": box / box distances 2 - 14 % short
  // when the relative gap at the last step was between 1e-5 and 2e-5).
  if (exact) {  // (the boolean verdicts are validated as they are: 37M configurations, profiles/r02_parity_sweep.jsonl)
    const V3 p0 = g.s.p[0], p1 = g.s.p[1], p2 = g.s.p[2];
    const bool d0 = g.s.n > 0 && w.x == p0.x && w.y == p0.y && w.z == p0.z;
    const bool d1 = g.s.n > 1 && w.x == p1.x && w.y == p1.y && w.z == p1.z;
    const bool d2 = g.s.n > 2 && w.x == p2.x && w.y == p2.y && w.z == p2.z;
    dup = d0 || d1 || d2;
    done = done || dup;
  }
  if (!done) {
    // no dynamic register indexing (it would spill the simplex to local memory)
    if (g.s.n == 0) g.s.p[0] = w; else if (g.s.n == 1) g.s.p[1] = w; else if (g.s.n == 2) g.s.p[2] = w; else g.s.p[3] = w;
    ++g.s.n;
    V3 nv;
    if (solve_simplex(g.s, nv)) {  // origin enclosed by the simplex
      // ... which cannot be when a separating plane beyond the swept radii was already proven
      if (g.lower > g.rsum) { dist = sqrtf(g.vv) - g.rsum; return GJK_MISS; }
      g.vv = 0.f; dist = -g.rsum; return GJK_HIT;
    }
    const float nvv = dot(nv, nv);
    if (nvv <= fmaxf(rs2, TINY2) && !exact) { dist = sqrtf(nvv) - g.rsum; return GJK_HIT; }
    if (nvv <= TINY2) { g.vv = 0.f; dist = -g.rsum; return GJK_HIT; }
    if (g.it > 0 && nvv >= g.vv) done = true;  // no progress at FP32 resolution
    else { g.v = nv; g.vv = nvv; }
  }
  if (!done && ++g.it < 48) return GJK_CONTINUE;
  const float d = sqrtf(g.vv);
  dist = d - g.rsum;
  // a stalled iteration far from convergence: trust the proven lower bound
  if (d <= g.rsum && g.lower > g.rsum) { dist = g.lower - g.rsum; return GJK_MISS; }
  // The boolean verdict of an iteration that ended WITHOUT a proof (no separating plane beyond the
  // swept radii, upper and lower bound not met: a stall at FP32 resolution, or the iteration cap):
  // |v| is a point of A (-) B, so the true distance is at most |v| - rsum.  If that is inside the
  // 1e-4 m margin band, "colliding" is right or in the band, whereas "free" is wrong whenever the
  // shapes overlap -- FP32 GJK stalls about 1e-4 m short of enclosing the origin of shallow
  // overlaps (1 in 5e6 configurations, depths 1.1 - 2.6e-4 m: profiles/r02_parity_sweep.jsonl).
  // Not wider than the band: separated pairs 1 - 2e-4 m apart end the same way by the hundred per
  // 5e6 configurations, and for them `d <= rsum` below is the right answer.
  if (!exact && !converged && g.lower <= g.rsum && d <= g.rsum + 1e-4f) return GJK_HIT;
  // exact mode: the same unproven ending within a millimetre of contact is handed back as such --
  // gjk_signed_distance then lets EPA look at the simplex (shallow overlaps came back as small
  // positive distances: 7 sign changes in 4.5e7 pair distances of the UR5 stand-in)
  if (exact && !converged && g.lower <= g.rsum && d <= 1e-3f) return GJK_STALL;  // (d: the CORE distance; a repeated vertex is no proof here)
  return d <= g.rsum ? GJK_HIT : GJK_MISS;
}

// One GJK iteration: both support scans, then the simplex update.
__device__ __forceinline__ int gjk_step(GjkLane& g, const GjkShape& A, const GjkShape& B, bool exact, float& dist) {
  V3 sa = mk(0.f, 0.f, 0.f), sb = sa;
#pragma unroll 1
  for (int side = 0; side < 2; ++side) {
    const GjkShape& S = side ? B : A;
    const V3 d = side ? rot_t(g.rel, g.v) : -g.v;
    const V3 r = support(S.type, S.p0, S.p1, S.p2, S.verts, S.vcnt, d);
    if (side) sb = r; else sa = r;
  }
  return gjk_advance(g, sa, sb, exact, dist);
}

// ------------------------------------------------------------------------------ EPA
// Penetration depth of two overlapping cores (expanding polytope algorithm).  Only the signed
// distance output (ramp_distance_batch) uses it; the boolean paths never do.  The polytope lives
// in per-thread local memory.
constexpr int kEpaMaxV = 68, kEpaMaxF = 160, kEpaIters = 64;
constexpr float kEpaTol = 1e-5f;  // relative gap at which the expansion stops

struct EpaFace {
  unsigned char a, b, c, alive;
  V3 n;
  float d;
};

__device__ __forceinline__ V3 mink_support(const GjkShape& A, const GjkShape& B, const Xf& rel, V3 d) {
  const V3 sa = support(A.type, A.p0, A.p1, A.p2, A.verts, A.vcnt, d);
  const V3 sb = support(B.type, B.p0, B.p1, B.p2, B.verts, B.vcnt, rot_t(rel, -d));
  return sa - apply(rel, sb);
}

// Face (a, b, c), its normal turned AWAY from `inside` (a point strictly inside the polytope: the
// centroid of the start tetrahedron).  Orienting by that test rather than by the winding of the
// horizon edge keeps normals outward when the new vertex lies in the plane of a face next to the
// horizon (flat faces of box / cylinder Minkowski sums give many coplanar support points).
__device__ __noinline__ bool epa_face(const V3* vt, int a, int b, int c, V3 inside, EpaFace& f) {
  const V3 e1 = vt[b] - vt[a], e2 = vt[c] - vt[a];
  V3 n = cross(e1, e2);
  const float l2 = dot(n, n);
  f.a = a; f.b = b; f.c = c; f.alive = 0;
  if (l2 < 1e-30f || l2 <= 1e-10f * dot(e1, e1) * dot(e2, e2)) return false;  // a sliver: no plane to speak of
  n = n * rsqrtf(l2);
  if (dot(n, vt[a] - inside) < 0.f) { n = -n; f.b = c; f.c = b; }
  f.n = n; f.d = dot(n, vt[a]); f.alive = 1;
  return true;
}

// Grow a degenerate terminal simplex of GJK into a tetrahedron around the origin.
__device__ __noinline__ bool epa_enclose(const GjkShape& A, const GjkShape& B, const Xf& rel, Simplex& s) {
  const V3 AX[6] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
  if (s.n == 0) { s.p[0] = mink_support(A, B, rel, AX[0]); s.n = 1; }
  if (s.n == 1) {
    for (int k = 0; k < 6 && s.n == 1; ++k) {
      const V3 w = mink_support(A, B, rel, AX[k]), e = w - s.p[0];
      if (dot(e, e) > 1e-12f) { s.p[1] = w; s.n = 2; }
    }
    if (s.n == 1) return false;
  }
  if (s.n == 2) {
    const V3 e = s.p[1] - s.p[0];
    for (int k = 0; k < 6 && s.n == 2; k += 2) {
      const V3 dir = cross(e, AX[k]);
      if (dot(dir, dir) < 1e-12f) continue;
      for (int sg = 0; sg < 2 && s.n == 2; ++sg) {
        const V3 w = mink_support(A, B, rel, sg ? -dir : dir), n = cross(e, w - s.p[0]);
        if (dot(n, n) > 1e-14f) { s.p[2] = w; s.n = 3; }
      }
    }
    if (s.n == 2) return false;
  }
  if (s.n == 3) {
    const V3 n = cross(s.p[1] - s.p[0], s.p[2] - s.p[0]);
    for (int sg = 0; sg < 2 && s.n == 3; ++sg) {
      const V3 w = mink_support(A, B, rel, sg ? -n : n);
      if (fabsf(dot(w - s.p[0], n)) > 1e-7f * (1.f + dot(n, n))) { s.p[3] = w; s.n = 4; }
    }
    if (s.n == 3) return false;
  }
  return true;
}

// `restart`: when given and the starting tetrahedron does NOT hold the origin (GJK's exact mode
// can report a false "enclosed" on a sliver tetrahedron of two shapes a few mm apart), nothing
// is expanded: *restart receives the point of the offending face plane nearest to the origin --
// a good place to start GJK again from -- and the result is -1.
__device__ __noinline__ float epa_depth(const GjkShape& A, const GjkShape& B, const Xf& rel, Simplex s, V3* restart) {
  V3 vt[kEpaMaxV];
  EpaFace fc[kEpaMaxF];
  int nv = 0, nf = 0;
#ifdef RAMP_TRACE_DIST
  const bool tr = (threadIdx.x & 31) == 0 && blockIdx.x == 0;
  if (tr) printf("trace epa: start n %d\n", s.n);
#endif
  if (!epa_enclose(A, B, rel, s)) return 0.f;
  for (int i = 0; i < 4; ++i) vt[nv++] = s.p[i];
  const V3 inside = (vt[0] + vt[1] + vt[2] + vt[3]) * 0.25f;
  for (int i = 0; i < 4; ++i) {  // faces (0,1,2) (0,1,3) (0,2,3) (1,2,3)
    const int a = i == 3 ? 1 : 0, b = i < 2 ? 1 : 2, c = i == 0 ? 2 : 3;
    if (!epa_face(vt, a, b, c, inside, fc[nf++])) return 0.f;
  }
  if (restart) {
    int worst = 0;
    for (int i = 1; i < 4; ++i)
      if (fc[i].d < fc[worst].d) worst = i;
    if (fc[worst].d < -1e-6f) { *restart = fc[worst].n * fc[worst].d; return -1.f; }
  }
  float best = 0.f;
  for (int it = 0; it < kEpaIters; ++it) {
    int bi = -1;
    float bd = 3.0e38f;
    for (int i = 0; i < nf; ++i)
      if (fc[i].alive && fc[i].d < bd) { bd = fc[i].d; bi = i; }
    if (bi < 0) break;
    best = bd;
    const V3 w = mink_support(A, B, rel, fc[bi].n);
#ifdef RAMP_TRACE_DIST
    if (tr) printf("trace epa: it %d nf %d nv %d bd %.9g gap %.9g n %.4g %.4g %.4g\n", it, nf, nv, bd, dot(fc[bi].n, w) - bd, fc[bi].n.x, fc[bi].n.y, fc[bi].n.z);
#endif
    // FP32: face normals and plane offsets carry ~1e-6 of noise, a tighter test never fires
    if (dot(fc[bi].n, w) - bd <= kEpaTol * (1.f + bd) || nv >= kEpaMaxV) break;
    bool dup = false;  // a support point the polytope already has: nothing left to expand
    for (int k = 0; k < nv; ++k) dup = dup || (vt[k].x == w.x && vt[k].y == w.y && vt[k].z == w.z);
    if (dup) break;
    // drop the faces that see w; the horizon is the set of edges with exactly ONE dropped
    // neighbour, matched without regard to direction (so it does not depend on windings)
    unsigned char eh[96][2];
    int ne = 0;
    for (int i = 0; i < nf; ++i) {
      if (!fc[i].alive || dot(fc[i].n, w) - fc[i].d <= 1e-7f) continue;
      fc[i].alive = 0;
      const unsigned char ed[3][2] = {{fc[i].a, fc[i].b}, {fc[i].b, fc[i].c}, {fc[i].c, fc[i].a}};
      for (int k = 0; k < 3; ++k) {
        const unsigned char lo = ed[k][0] < ed[k][1] ? ed[k][0] : ed[k][1], hi = ed[k][0] < ed[k][1] ? ed[k][1] : ed[k][0];
        int found = -1;
        for (int e = 0; e < ne; ++e)
          if (eh[e][0] == lo && eh[e][1] == hi) { found = e; break; }
        if (found >= 0) { eh[found][0] = eh[ne - 1][0]; eh[found][1] = eh[ne - 1][1]; --ne; }
        else if (ne < 96) { eh[ne][0] = lo; eh[ne][1] = hi; ++ne; }
      }
    }
    if (ne == 0) break;
    vt[nv] = w;
    bool ok = true;
    for (int e = 0; e < ne && ok; ++e) {
      int slot = -1;
      for (int i = 0; i < nf; ++i) if (!fc[i].alive) { slot = i; break; }
      if (slot < 0) { if (nf >= kEpaMaxF) { ok = false; break; } slot = nf++; }
      // (a sliver stays dead: it has no plane to test.)  Expanding the polytope can only move its
      // faces outwards: a new face closer to the origin than the one it replaces is rounding noise
      // (w almost in the plane of the old face) -- stop with the depth reached, which by then is
      // within the noise of the converged value.  (bd < 0: the start tetrahedron does not hold the
      // origin and the expansion walks to the face the origin lies beyond; leave that alone.)
      if (epa_face(vt, eh[e][0], eh[e][1], nv, inside, fc[slot]) && bd > 0.f && fc[slot].d < bd - kEpaTol * (1.f + bd)) ok = false;
    }
    ++nv;
    if (!ok) break;
  }
  return best;
}

// Signed distance of two shapes through GJK (+ EPA when the cores overlap).
__device__ __forceinline__ float gjk_signed_distance(const GjkShape& A, const GjkShape& B, const Xf& rel, float rsum) {
  GjkLane g;
  V3 v0 = -trans(rel);
#pragma unroll 1
  for (int attempt = 0;; ++attempt) {
    gjk_init(g, rel, rsum, v0);
    float dist = 0.f;
    int st = GJK_CONTINUE;
    while (st == GJK_CONTINUE) {
      st = gjk_step(g, A, B, true, dist);
#ifdef RAMP_TRACE_DIST
      if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) printf("trace dist: attempt %d it %d n %d |v| %.9g lower %.9g status %d dist %.9g\n", attempt, g.it, g.s.n, sqrtf(g.vv), g.lower, st, dist);
#endif
    }
    if (st == GJK_MISS || (st != GJK_STALL && g.vv > 0.f)) return dist;  // cores apart: |v| - radii (negative if the sweeps overlap)
    // "enclosed": believed only if the simplex really holds the origin; else start over from
    // the face the origin lies beyond (twice at most, then take EPA's walk to that face)
    V3 from;
    const bool last = attempt >= 2;
    const float depth = epa_depth(A, B, rel, g.s, last ? nullptr : &from);
    // (on the last attempt a negative "depth" is the offset of the face EPA walked to)
    if (last || depth != -1.f) return -depth - rsum;
    v0 = from;
  }
}

}  // namespace ramp
