// Helpers shared by the tile kernels (one CTA per group, one lane per K sorted observations):
// camera-constant loads, the projection, asynchronous staging of a chunk's tables.
#pragma once
#include "ba_internal.h"
#include "sfmba_device.cuh"

namespace sfmba {

// R (9, row-major, camera-to-world) and C (3) of one camera: six 16-byte loads of its table row
__device__ __forceinline__ void load_cam_rc(const double* __restrict__ camtab, int cam, double R[9], double C[3]) {
  const double2* row = reinterpret_cast<const double2*>(camtab + (size_t)CAMTAB_STRIDE * cam);
  const double2 a = row[0], b = row[1], c = row[2], e = row[3], f = row[4], g = row[5];
  R[0] = a.x; R[1] = a.y; R[2] = b.x; R[3] = b.y; R[4] = c.x; R[5] = c.y; R[6] = e.x; R[7] = e.y; R[8] = f.x;
  C[0] = f.y; C[1] = g.x; C[2] = g.y;
}

__device__ __forceinline__ void load_cam_q(const double* __restrict__ camtab, int cam, double q[4]) {
  const double2* row = reinterpret_cast<const double2*>(camtab + (size_t)CAMTAB_STRIDE * cam);
  const double2 a = row[6], b = row[7];
  q[0] = a.x; q[1] = a.y; q[2] = b.x; q[3] = b.y;
}

// 1 / x: MUFU.RCP64H seed (rcp.approx.ftz.f64, ~20 bits) + two Newton steps = full f64 accuracy to
// ~1 ulp in 5 instructions, no slow path (x == 0 gives inf and is flagged by the caller).
__device__ __forceinline__ double fast_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}

// projection of one point: d = X - C, p = R^T d, iw = 1 / p_w, (u, v) = p_uv * iw
struct Proj {
  double d0, d1, d2, iw, u, v;
  bool ok;
};

__device__ __forceinline__ Proj project(const double R[9], const double C[3], double x0, double x1, double x2) {
  Proj p;
  p.d0 = x0 - C[0];
  p.d1 = x1 - C[1];
  p.d2 = x2 - C[2];
  const double pu = R[0] * p.d0 + R[3] * p.d1 + R[6] * p.d2;
  const double pv = R[1] * p.d0 + R[4] * p.d1 + R[7] * p.d2;
  const double pw = R[2] * p.d0 + R[5] * p.d1 + R[8] * p.d2;
  p.iw = fast_rcp(pw);
  p.u = pu * p.iw;
  p.v = pv * p.iw;
  p.ok = pw != 0.0;
  return p;
}

}  // namespace sfmba

// ---- asynchronous global -> shared staging (cp.async; LDGSTS in SASS) ------------------------------
namespace sfmba {

__device__ __forceinline__ void cp_async4(void* smem, const void* g) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(g));
}
__device__ __forceinline__ void cp_async8(void* smem, const void* g) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(g));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(g));
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// chunk-specific staging: record words, segment cameras, run tables
struct ChunkStage {
  uint32_t* ixs;     // [k_cap * RUN_LANES]
  int32_t* segc;     // [seg_cap rounded to 4]
  int32_t* rpos;     // [seg_cap rounded to 4] row of each segment in the camera-sorted row arrays, or nullptr
  uint16_t* srun;    // [SEG_RUN_STRIDE] first lane of each segment (+ end), or nullptr
  double2* bs;       // [k_cap * RUN_LANES] normalised observations, or nullptr
  double* ws = nullptr;   // [k_cap * RUN_LANES] Huber weights (robust matvec), or nullptr
};

__device__ __forceinline__ void stage_chunk_async(const BaDev& d, int c, const ChunkInfo& ci, const ChunkStage& cs) {
  const int K = ci.K;
  const uint32_t* ix = d.rec_ix + (size_t)c * CHUNK_CAP;
  for (int i = threadIdx.x; i < K * (RUN_LANES / 4); i += blockDim.x) cp_async16(cs.ixs + 4 * i, ix + 4 * i);
  const int32_t* sc = d.seg_cam + (size_t)c * SEG_CAP;
  for (int i = threadIdx.x; i < (ci.nseg + 3) / 4; i += blockDim.x) cp_async16(cs.segc + 4 * i, sc + 4 * i);
  if (cs.rpos) {
    const int32_t* rp = d.rowpos + (size_t)c * SEG_CAP;
    for (int i = threadIdx.x; i < (ci.nseg + 3) / 4; i += blockDim.x) cp_async16(cs.rpos + 4 * i, rp + 4 * i);
  }
  if (cs.srun) {
    const uint16_t* sr = d.seg_lane0 + (size_t)c * SEG_RUN_STRIDE;
    for (int i = threadIdx.x; i < (ci.nseg + 2) / 2; i += blockDim.x) cp_async4(cs.srun + 2 * i, sr + 2 * i);
  }
  if (cs.bs) {
    const double2* b = d.rec_b + (size_t)c * CHUNK_CAP;
    for (int i = threadIdx.x; i < K * RUN_LANES; i += blockDim.x) cp_async16(cs.bs + i, b + i);
  }
  if (cs.ws) {
    const double* w = d.rec_w + (size_t)c * CHUNK_CAP;
    for (int i = threadIdx.x; i < K * (RUN_LANES / 2); i += blockDim.x) cp_async16(cs.ws + 2 * i, w + 2 * i);
  }
}

// camtab row of camera c from its parameters (C, q): R(q) row-major, C, qj = R -> q (the Jacobian's quaternion,
// campose_processor.py:464)
__device__ __forceinline__ void cam_table_row(const BaDev& d, int c, const double cam[7]) {
  double q[4] = {cam[3], cam[4], cam[5], cam[6]};
  double R[9], qj[4];
  quat_to_rot(q, R);                              // ba_processor.py:323 (q as stored, not renormalised)
  if (!rot_to_quat(R, qj)) atomicOr(&d.sc->flags, FLAG_QW_NEAR_ZERO);
  double* t = d.camtab + (size_t)CAMTAB_STRIDE * c;
#pragma unroll
  for (int i = 0; i < 9; ++i) t[i] = R[i];
  t[9] = cam[0];
  t[10] = cam[1];
  t[11] = cam[2];
#pragma unroll
  for (int i = 0; i < 4; ++i) t[12 + i] = qj[i];
}

// ---- per-camera rows of a camera-space vector (shared by the PCG kernels and the fused camera finish) -------------
// Q[j][3 c + k] = d R[k][c] / d q_j for q2 = 2 (w,x,y,z)        (campose_processor.py:636-702)
__device__ __forceinline__ void quat_dRt(const double q2[4], double Q[4][9]) {
  const double w2 = q2[0], x2 = q2[1], y2 = q2[2], z2 = q2[3];
  Q[0][0] = 0.0;       Q[0][1] = z2;        Q[0][2] = -y2;
  Q[0][3] = -z2;       Q[0][4] = 0.0;       Q[0][5] = x2;
  Q[0][6] = y2;        Q[0][7] = -x2;       Q[0][8] = 0.0;
  Q[1][0] = 0.0;       Q[1][1] = y2;        Q[1][2] = z2;
  Q[1][3] = y2;        Q[1][4] = -2.0 * x2; Q[1][5] = w2;
  Q[1][6] = z2;        Q[1][7] = -w2;       Q[1][8] = -2.0 * x2;
  Q[2][0] = -2.0 * y2; Q[2][1] = x2;        Q[2][2] = -w2;
  Q[2][3] = x2;        Q[2][4] = 0.0;       Q[2][5] = z2;
  Q[2][6] = w2;        Q[2][7] = z2;        Q[2][8] = -2.0 * y2;
  Q[3][0] = -2.0 * z2; Q[3][1] = w2;        Q[3][2] = x2;
  Q[3][3] = -w2;       Q[3][4] = -2.0 * z2; Q[3][5] = y2;
  Q[3][6] = x2;        Q[3][7] = y2;        Q[3][8] = 0.0;
}

// camit row of one camera from its 7 entries dl of a camera-space vector: E = sum_j dl_q[j] Q_j, w = R^T dl_C
__device__ __forceinline__ void camit_row(const double* __restrict__ t, const double dl[7], int dvdc_quirk,
                                          double o[14]) {
  const double q2[4] = {2.0 * t[12], 2.0 * t[13], 2.0 * t[14], 2.0 * t[15]};
  double Q[4][9];
  quat_dRt(q2, Q);
#pragma unroll
  for (int e = 0; e < 9; ++e) o[e] = dl[3] * Q[0][e] + dl[4] * Q[1][e] + dl[5] * Q[2][e] + dl[6] * Q[3][e];
#pragma unroll
  for (int c = 0; c < 3; ++c) o[9 + c] = t[c] * dl[0] + t[3 + c] * dl[1] + t[6 + c] * dl[2];
  o[12] = dvdc_quirk ? 2.0 * o[11] : 0.0;
  o[13] = dvdc_quirk ? 1.0 : -1.0;
}

}  // namespace sfmba
