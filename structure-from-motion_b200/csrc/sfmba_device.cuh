// Device-side maths shared by every kernel of libsfmba: camera table, per-observation residual and
// Jacobian evaluation, small symmetric solves, block reductions.
//
// Reference maths (cited per function): ba_processor.py:317-349, campose_processor.py:462-482 and
// 636-808, triangulation_processor.py:237-271, utils.py:28-97.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "ba_internal.h"

namespace sfmba {

// f(integral_constant<int, I>) for I = BEGIN .. END - 1: loop bodies whose array indices MUST be compile-time constants
// (small register arrays; `#pragma unroll` is only a request, and a loop nest it declines puts the array in local memory)
template <int BEGIN, int END, typename F>
__device__ __forceinline__ void static_for(F&& f) {
  if constexpr (BEGIN < END) {
    f(std::integral_constant<int, BEGIN>{});
    static_for<BEGIN + 1, END>(f);
  }
}

constexpr uint32_t FLAG_QW_NEAR_ZERO = 1u;
constexpr uint32_t FLAG_DEPTH_ZERO = 2u;
constexpr uint32_t FLAG_NONFINITE = 4u;
constexpr uint32_t FLAG_PCG_MAXITER = 8u;
constexpr uint32_t FLAG_SINGULAR = 16u;
constexpr uint32_t FLAG_PEER_TIMEOUT = 32u;
constexpr uint32_t FLAG_BAD_INDEX = 64u;      // a caller-supplied index outside its array (RANSAC sample tables)

// q (w,x,y,z) -> R row-major, utils.py:64-97
__device__ __forceinline__ void quat_to_rot(const double q[4], double R[9]) {
  const double w = q[0], x = q[1], y = q[2], z = q[3];
  R[0] = 1.0 - 2.0 * z * z - 2.0 * y * y;
  R[1] = -2.0 * z * w + 2.0 * y * x;
  R[2] = 2.0 * y * w + 2.0 * z * x;
  R[3] = 2.0 * x * y + 2.0 * w * z;
  R[4] = 1.0 - 2.0 * z * z - 2.0 * x * x;
  R[5] = 2.0 * z * y - 2.0 * x * w;
  R[6] = 2.0 * x * z - 2.0 * w * y;
  R[7] = 2.0 * y * z + 2.0 * w * x;
  R[8] = 1.0 - 2.0 * y * y - 2.0 * x * x;
}

// R -> q, qw > 0 branch only, utils.py:47-58.  Returns false when |qw| < 1e-6 (the reference raises).
__device__ __forceinline__ bool rot_to_quat(const double R[9], double q[4]) {
  const double qw = sqrt(1.0 + R[0] + R[4] + R[8]) / 2.0;
  q[0] = qw;
  const double s = 4.0 * qw;
  q[1] = (R[7] - R[5]) / s;
  q[2] = (R[2] - R[6]) / s;
  q[3] = (R[3] - R[1]) / s;
  return !(fabs(qw) < 1e-6) && (qw == qw);
}

// Per-observation evaluation.  ct: camera-table row (R, C, qj).  X: the point.  b: normalised
// observation.  Outputs the residual r = b - f (ba_processor.py:346-349,376), the point block Jx
// (2x3, triangulation_processor.py:259-271 with the K-free projection of ba_processor.py:328) and
// the pose block Jp (2x7 = [d/dC | d/dq], campose_processor.py:462-482).
//   Jx[r][k]   = (R[k][r] - f_r R[k][2]) / p_w
//   Jpc[0][k]  = -Jx[0][k]
//   Jpc[1][k]  = -(R[k][1] + v R[k][2]) / p_w      <- reference sign quirk (campose_processor.py:802-804)
//              = -Jx[1][k]                          when the quirk is off
//   Jpq[r][:]  = (G_r - f_r G_2) / p_w,  G_c = sum_k d_k dR[k][c]/dq  (campose_processor.py:636-767)
template <bool WANT_JAC>
__device__ __forceinline__ bool eval_obs(const double* __restrict__ ct, const double X[3], const double b[2],
                                         bool dvdc_quirk, double r[2], double Jp[14], double Jx[6]) {
  const double d0 = X[0] - ct[9], d1 = X[1] - ct[10], d2 = X[2] - ct[11];
  const double R00 = ct[0], R01 = ct[1], R02 = ct[2];
  const double R10 = ct[3], R11 = ct[4], R12 = ct[5];
  const double R20 = ct[6], R21 = ct[7], R22 = ct[8];
  const double pu = R00 * d0 + R10 * d1 + R20 * d2;
  const double pv = R01 * d0 + R11 * d1 + R21 * d2;
  const double pw = R02 * d0 + R12 * d1 + R22 * d2;
  const double iw = 1.0 / pw;
  const double u = pu * iw, v = pv * iw;
  r[0] = b[0] - u;
  r[1] = b[1] - v;
  if (WANT_JAC) {
    Jx[0] = (R00 - u * R02) * iw;
    Jx[1] = (R10 - u * R12) * iw;
    Jx[2] = (R20 - u * R22) * iw;
    Jx[3] = (R01 - v * R02) * iw;
    Jx[4] = (R11 - v * R12) * iw;
    Jx[5] = (R21 - v * R22) * iw;
    Jp[0] = -Jx[0];
    Jp[1] = -Jx[1];
    Jp[2] = -Jx[2];
    if (dvdc_quirk) {
      Jp[7] = -(R01 + v * R02) * iw;
      Jp[8] = -(R11 + v * R12) * iw;
      Jp[9] = -(R21 + v * R22) * iw;
    } else {
      Jp[7] = -Jx[3];
      Jp[8] = -Jx[4];
      Jp[9] = -Jx[5];
    }
    const double w2 = 2.0 * ct[12], x2 = 2.0 * ct[13], y2 = 2.0 * ct[14], z2 = 2.0 * ct[15];
    // G_c[j] = d0 * dR[0][c]/dq_j + d1 * dR[1][c]/dq_j + d2 * dR[2][c]/dq_j
    const double G00 = d1 * z2 - d2 * y2;
    const double G01 = d1 * y2 + d2 * z2;
    const double G02 = -2.0 * d0 * y2 + d1 * x2 - d2 * w2;
    const double G03 = -2.0 * d0 * z2 + d1 * w2 + d2 * x2;
    const double G10 = -d0 * z2 + d2 * x2;
    const double G11 = d0 * y2 - 2.0 * d1 * x2 + d2 * w2;
    const double G12 = d0 * x2 + d2 * z2;
    const double G13 = -d0 * w2 - 2.0 * d1 * z2 + d2 * y2;
    const double G20 = d0 * y2 - d1 * x2;
    const double G21 = d0 * z2 - d1 * w2 - 2.0 * d2 * x2;
    const double G22 = d0 * w2 + d1 * z2 - 2.0 * d2 * y2;
    const double G23 = d0 * x2 + d1 * y2;
    Jp[3] = (G00 - u * G20) * iw;
    Jp[4] = (G01 - u * G21) * iw;
    Jp[5] = (G02 - u * G22) * iw;
    Jp[6] = (G03 - u * G23) * iw;
    Jp[10] = (G10 - v * G20) * iw;
    Jp[11] = (G11 - v * G21) * iw;
    Jp[12] = (G12 - v * G22) * iw;
    Jp[13] = (G13 - v * G23) * iw;
  }
  return pw != 0.0;
}

// Huber loss of the squared residual norm s: returns the IRLS weight rho'(s) and rho(s) itself.
__device__ __forceinline__ double huber_weight(double s, double delta, double& rho) {
  if (s <= delta * delta) {
    rho = s;
    return 1.0;
  }
  const double rs = sqrt(s);
  rho = 2.0 * delta * rs - delta * delta;
  return delta / rs;
}

// b = (K^-1 [u,v,1]^T) / z, ba_processor.py:339-342.  kinv NULL: (u,v) already normalised.
__device__ __forceinline__ void normalise_obs(const double* __restrict__ kinv, float u, float v, double b[2]) {
  if (kinv == nullptr) {
    b[0] = (double)u;
    b[1] = (double)v;
  } else {
    const double x = (double)u, y = (double)v;
    const double cx = kinv[0] * x + kinv[1] * y + kinv[2];
    const double cy = kinv[3] * x + kinv[4] * y + kinv[5];
    const double cz = kinv[6] * x + kinv[7] * y + kinv[8];
    b[0] = cx / cz;
    b[1] = cy / cz;
  }
}

// Inverse of a symmetric 3x3 given as (a00,a01,a02,a11,a12,a22); output in the same packing.
__device__ __forceinline__ bool sym3_inverse(const double a[6], double inv[6]) {
  const double c00 = a[3] * a[5] - a[4] * a[4];
  const double c01 = a[2] * a[4] - a[1] * a[5];
  const double c02 = a[1] * a[4] - a[2] * a[3];
  const double det = a[0] * c00 + a[1] * c01 + a[2] * c02;
  const double id = 1.0 / det;
  inv[0] = c00 * id;
  inv[1] = c01 * id;
  inv[2] = c02 * id;
  inv[3] = (a[0] * a[5] - a[2] * a[2]) * id;
  inv[4] = (a[1] * a[2] - a[0] * a[4]) * id;
  inv[5] = (a[0] * a[3] - a[1] * a[1]) * id;
  return det > 0.0;
}

// y = Sym3(a) * x
__device__ __forceinline__ void sym3_mul(const double a[6], const double x[3], double y[3]) {
  y[0] = a[0] * x[0] + a[1] * x[1] + a[2] * x[2];
  y[1] = a[1] * x[0] + a[3] * x[1] + a[4] * x[2];
  y[2] = a[2] * x[0] + a[4] * x[1] + a[5] * x[2];
}

// In-place Cholesky inverse of a dense SPD n x n (n <= 7) stored row-major in a[]; returns false if a
// pivot is not positive.  Used for the 7x7 block-Jacobi blocks and the PnP normal equations.
template <int N>
__device__ __forceinline__ bool spd_inverse(double* a) {
  double L[N * N];
  bool ok = true;
#pragma unroll
  for (int i = 0; i < N; ++i) {
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      double s = a[i * N + j];
#pragma unroll
      for (int k = 0; k < j; ++k) s -= L[i * N + k] * L[j * N + k];
      if (i == j) {
        if (!(s > 0.0)) ok = false;
        L[i * N + i] = sqrt(s);
      } else {
        L[i * N + j] = s / L[j * N + j];
      }
    }
  }
  // invert L (lower triangular) into Li
  double Li[N * N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
#pragma unroll
    for (int j = 0; j < N; ++j) Li[i * N + j] = 0.0;
  }
#pragma unroll
  for (int j = 0; j < N; ++j) {
    Li[j * N + j] = 1.0 / L[j * N + j];
#pragma unroll
    for (int i = j + 1; i < N; ++i) {
      double s = 0.0;
#pragma unroll
      for (int k = j; k < i; ++k) s -= L[i * N + k] * Li[k * N + j];
      Li[i * N + j] = s / L[i * N + i];
    }
  }
  // A^-1 = Li^T Li
#pragma unroll
  for (int i = 0; i < N; ++i) {
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      double s = 0.0;
#pragma unroll
      for (int k = i; k < N; ++k) s += Li[k * N + i] * Li[k * N + j];
      a[i * N + j] = s;
      a[j * N + i] = s;
    }
  }
  return ok;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Deterministic block-wide sum; `scratch` holds >= 32 doubles of shared memory.  Result valid in all
// threads.  Contains __syncthreads(): every thread of the block must call it.
__device__ __forceinline__ double block_sum(double v, double* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  double t = 0.0;
  for (int w = 0; w < nwarps; ++w) t += scratch[w];
  return t;
}

}  // namespace sfmba
