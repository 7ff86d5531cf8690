// ik.cuh -- ik_kernel: batched differential inverse kinematics over many seeds (damped least squares).
// Part of the single translation unit ramp_b200.cu (included there, in this order).
#pragma once

namespace ramp {

// ------------------------------------------------------------------ kernel: batched IK seeds
// The step before the hot path: goal states for the planner come from inverse kinematics (the
// reference runs one pink / quadprog solve at a time, src/ramp/robot_state.py:237-304, inside a
// rejection loop, examples/so_arm100_motion_planning.py:50-54).  Here every lane owns one seed and
// iterates   q <- clamp(q + step * J^T (J J^T + lambda^2 I)^-1 e)   on the actuated values of a
// group, where e is the pose error of the tool frame in that frame (translation, rotation
// vector) and J its 6 x n frame Jacobian in the LOCAL frame -- columns built from the joint
// transforms the FK walk leaves in the lane's staging rows.  J is never stored: one pass over the
// columns accumulates J J^T (21 numbers), a 6 x 6 Cholesky solve gives y, a second pass applies
// dq_k = col_k . y.  One-degree-of-freedom joints only (revolute, continuous, prismatic, aligned
// or not); mimic joints add their column to the joint they follow.
constexpr int kIkMaxCols = 16, kIkMaxMimic = 8;

struct IkArgs {
  const unsigned char* img;
  const float* seed;      // [n, ld] actuated values of the group
  int64_t n;
  int32_t ld;
  const float* target;    // [n or 1][12] 3x4 row-major world pose of the tool frame
  int32_t target_stride;  // 12: one target per seed, 0: one for all
  float* q_out;           // [n, n_cols]
  float* err_out;         // [n] norm of the pose error at the returned q (nullable)
  uint8_t* ok;            // [n] error below the tolerance (nullable)
  int32_t* iters_out;     // [n] iterations used (nullable)
  int32_t frame_joint;    // parent joint of the tool frame (-1: universe)
  float frame_pl[12];     // placement of the frame in that joint
  int32_t max_iter;
  float damping2, step, tol, max_dq;
  int32_t n_cols;
  int32_t col_joint[kIkMaxCols];  // actuated column -> joint it drives
  float lo[kIkMaxCols], hi[kIkMaxCols];
  int32_t n_mimic;
  int32_t mim_col[kIkMaxMimic], mim_joint[kIkMaxMimic];
  float mim_mult[kIkMaxMimic];
  uint64_t chain;  // bit j: joint j is the frame's parent joint or one of its ancestors
  SlabLayout L;
};

__device__ __forceinline__ Xf read_staged_xf(const Warp& w, int slot) {
  const float4* p = reinterpret_cast<const float4*>(w.stage) + w.lane * w.stage_stride + slot * 3;
  const float4 r0 = p[0], r1 = p[1], r2 = p[2];
  Xf o;
  o.m[0] = r0.x; o.m[1] = r0.y; o.m[2] = r0.z; o.m[3] = r0.w;
  o.m[4] = r1.x; o.m[5] = r1.y; o.m[6] = r1.z; o.m[7] = r1.w;
  o.m[8] = r2.x; o.m[9] = r2.y; o.m[10] = r2.z; o.m[11] = r2.w;
  return o;
}

// rotation vector of a rotation matrix given by its rows (log map of SO(3))
__device__ __forceinline__ V3 log_so3(const Xf& r) {
  const V3 vee = mk(0.5f * (r.m[9] - r.m[6]), 0.5f * (r.m[2] - r.m[8]), 0.5f * (r.m[4] - r.m[1]));
  const float c = clampf(0.5f * (r.m[0] + r.m[5] + r.m[10] - 1.f), -1.f, 1.f);
  const float s = sqrtf(dot(vee, vee));
  const float th = atan2f(s, c);
  if (s > 1e-4f) return vee * (th / s);
  if (c > 0.f) return vee;  // small angle: sin(th) ~ th
  // close to pi: the axis from the diagonal, signs from the symmetric part
  const float k = 1.f / (1.f - c);
  V3 a = mk(sqrtf(fmaxf((r.m[0] - c) * k, 0.f)), sqrtf(fmaxf((r.m[5] - c) * k, 0.f)), sqrtf(fmaxf((r.m[10] - c) * k, 0.f)));
  if (a.x >= a.y && a.x >= a.z) { a.y = copysignf(a.y, r.m[1] + r.m[4]); a.z = copysignf(a.z, r.m[2] + r.m[8]); }
  else if (a.y >= a.z) { a.x = copysignf(a.x, r.m[1] + r.m[4]); a.z = copysignf(a.z, r.m[6] + r.m[9]); }
  else { a.x = copysignf(a.x, r.m[2] + r.m[8]); a.y = copysignf(a.y, r.m[6] + r.m[9]); }
  // (the largest component is positive: within 1e-4 rad of a half turn +axis and -axis describe the
  // same rotation to 2e-4 rad, and sin(th) * axis is rounding noise -- a fixed choice keeps FP32 and
  // FP64 iterations on the same branch, e.g. from a stretched-out arm to a flipped tool)
  return a * th;
}

// column of the frame Jacobian (LOCAL frame: linear part, then angular part) for joint j
__device__ __forceinline__ void ik_column(const Warp& w, const IkArgs& a, int j, const Xf& tcp, V3& lin, V3& ang) {
  lin = ang = mk(0.f, 0.f, 0.f);
  if (j < 0 || !((a.chain >> j) & 1ull)) return;
  const DJoint* dj = &w.t.joints[j];
  const int type = dj->type;
  const Xf oj = read_staged_xf(w, j);
  V3 ax;
  const int al = type <= 3 ? type - 1 : (type >= 5 && type <= 7 ? type - 5 : (type >= 9 && type <= 11 ? type - 9 : -1));
  if (al >= 0) ax = mk(al == 0 ? 1.f : 0.f, al == 1 ? 1.f : 0.f, al == 2 ? 1.f : 0.f);
  else ax = mk(dj->ax, dj->ay, dj->az);
  const V3 aw = rot(oj, ax);
  if (type >= 9) {  // prismatic
    lin = rot_t(tcp, aw);
  } else {
    lin = rot_t(tcp, cross(aw, trans(tcp) - trans(oj)));
    ang = rot_t(tcp, aw);
  }
}

__device__ __forceinline__ void ik_full_column(const Warp& w, const IkArgs& a, int k, const Xf& tcp, float col[6]) {
  V3 lin, ang;
  ik_column(w, a, a.col_joint[k], tcp, lin, ang);
  for (int mi = 0; mi < a.n_mimic; ++mi)
    if (a.mim_col[mi] == k) {
      V3 l2, a2;
      ik_column(w, a, a.mim_joint[mi], tcp, l2, a2);
      lin = lin + l2 * a.mim_mult[mi];
      ang = ang + a2 * a.mim_mult[mi];
    }
  col[0] = lin.x; col[1] = lin.y; col[2] = lin.z; col[3] = ang.x; col[4] = ang.y; col[5] = ang.z;
}

// Persistent lanes: seeds need anything from 3 to max_iter iterations (a seed that runs into a
// joint limit or a singularity never converges), so a lane that is finished writes its result and
// takes the next seed -- warps draw 32 seeds at a time from a cursor -- instead of the whole tile
// waiting for its slowest seed (one tile in four has a seed that runs to 500 iterations when 1 %
// of the seeds do).  `cursor` is zeroed by the host.
__global__ void __launch_bounds__(512, 1) ik_kernel(const __grid_constant__ IkArgs a, unsigned long long* __restrict__ cursor) {
  uint64_t* mbar = reinterpret_cast<uint64_t*>(g_smem);
  unsigned char* tbl = g_smem + 16;
  stage_tables(tbl, a.img, a.L.table_bytes, mbar);
  const Tables t = make_tables(tbl);
  const int wid = threadIdx.x >> 5;
  float* slab = reinterpret_cast<float*>(tbl + a.L.table_bytes) + static_cast<size_t>(wid) * a.L.words;
  const Warp w = make_warp(t, slab, a.L);
  const int l = w.lane;
  const unsigned FULL = 0xffffffffu;
  Xf fpl;
#pragma unroll
  for (int k = 0; k < 12; ++k) fpl.m[k] = a.frame_pl[k];
  int64_t row = -1, lo = 0, hi = 0;  // this lane's seed (-1: idle); [lo, hi): seeds the warp holds but has not started
  bool exhausted = false;
  Xf target = xf_identity();
  int it = 0;
  for (int k = 0; k < a.n_cols; ++k) w.act[k * 32 + l] = 0.f;  // idle lanes run the kinematics too: on defined values
  while (true) {
    // ---- idle lanes take the next seeds
    const unsigned need = __ballot_sync(FULL, row < 0);
    if (need && !(exhausted && lo == hi)) {
      const int rank = __popc(need & ((1u << l) - 1)), nneed = __popc(need);
      if (lo == hi && !exhausted) {
        unsigned long long base = 0;
        if (l == 0) base = atomicAdd(cursor, 32ull);
        base = __shfl_sync(FULL, base, 0);
        if (static_cast<int64_t>(base) >= a.n) exhausted = true;
        else { lo = static_cast<int64_t>(base); hi = imin64(lo + 32, a.n); }
      }
      if (row < 0 && lo + rank < hi) {
        row = lo + rank;
        const float* sd = a.seed + row * a.ld;
        for (int k = 0; k < a.n_cols; ++k) w.act[k * 32 + l] = __ldg(sd + k);
        const float* tg = a.target + row * a.target_stride;
#pragma unroll
        for (int k = 0; k < 12; ++k) target.m[k] = __ldg(tg + k);
        it = 0;
      }
      lo = imin64(lo + nneed, hi);
    }
    if (__all_sync(FULL, row < 0)) {
      if (exhausted && lo == hi) break;
      continue;
    }
    __syncwarp();
    // ---- one iteration for every lane that holds a seed
    expand_q(w);
    fk_lane<1>(w);
    __syncwarp();
    const Xf tcp = a.frame_joint < 0 ? fpl : xf_mul(read_staged_xf(w, a.frame_joint), fpl);
    const Xf d = xf_inv_mul(tcp, target);  // the target seen from the frame
    const V3 ep = trans(d), er = log_so3(d);
    const float e[6] = {ep.x, ep.y, ep.z, er.x, er.y, er.z};
    const float err = sqrtf(dot(ep, ep) + dot(er, er));
    const bool done = err < a.tol;
    if (row >= 0 && (done || it >= a.max_iter)) {
      for (int k = 0; k < a.n_cols; ++k) a.q_out[row * a.n_cols + k] = w.act[k * 32 + l];
      if (a.err_out) a.err_out[row] = err;
      if (a.ok) a.ok[row] = done;
      if (a.iters_out) a.iters_out[row] = it;
      row = -1;
    }
    if (row >= 0) {
      // A = J J^T + lambda^2 I (lower triangle, row-major packed)
      float A[21];
#pragma unroll
      for (int i = 0; i < 21; ++i) A[i] = 0.f;
#pragma unroll
      for (int i = 0; i < 6; ++i) A[i * (i + 1) / 2 + i] = a.damping2;
#pragma unroll 1
      for (int k = 0; k < a.n_cols; ++k) {
        float c[6];
        ik_full_column(w, a, k, tcp, c);
#pragma unroll
        for (int i = 0; i < 6; ++i)
#pragma unroll
          for (int j = 0; j <= i; ++j) A[i * (i + 1) / 2 + j] = fmaf(c[i], c[j], A[i * (i + 1) / 2 + j]);
      }
      // Cholesky A = L L^T in place, then L z = e, L^T y = z
#pragma unroll
      for (int i = 0; i < 6; ++i) {
#pragma unroll
        for (int j = 0; j <= i; ++j) {
          float s = A[i * (i + 1) / 2 + j];
#pragma unroll
          for (int k = 0; k < j; ++k) s = fmaf(-A[i * (i + 1) / 2 + k], A[j * (j + 1) / 2 + k], s);
          A[i * (i + 1) / 2 + j] = i == j ? sqrtf(fmaxf(s, 1e-20f)) : s / A[j * (j + 1) / 2 + j];
        }
      }
      float y[6];
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        float s = e[i];
#pragma unroll
        for (int k = 0; k < i; ++k) s = fmaf(-A[i * (i + 1) / 2 + k], y[k], s);
        y[i] = s / A[i * (i + 1) / 2 + i];
      }
#pragma unroll
      for (int i = 5; i >= 0; --i) {
        float s = y[i];
#pragma unroll
        for (int k = i + 1; k < 6; ++k) s = fmaf(-A[k * (k + 1) / 2 + i], y[k], s);
        y[i] = s / A[i * (i + 1) / 2 + i];
      }
#pragma unroll 1
      for (int k = 0; k < a.n_cols; ++k) {
        float c[6];
        ik_full_column(w, a, k, tcp, c);
        float dq = 0.f;
#pragma unroll
        for (int i = 0; i < 6; ++i) dq = fmaf(c[i], y[i], dq);
        dq = clampf(a.step * dq, -a.max_dq, a.max_dq);
        w.act[k * 32 + l] = clampf(w.act[k * 32 + l] + dq, a.lo[k], a.hi[k]);
      }
      ++it;
    }
    __syncwarp();
  }
}

}  // namespace ramp
