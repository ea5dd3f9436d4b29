// Batched local refiners: nonlinear triangulation (one thread per point) and nonlinear PnP (one warp
// per view).  Both are independent small damped Gauss-Newton problems with a fixed damping factor and
// a fixed iteration count, exactly as the reference runs them; everything is f64 (pixel-unit
// residuals of a few hundred leave no room for f32, and the kernels are latency- not HBM-bound).
//
// Reference: triangulation_processor.py:160-309, campose_processor.py:308-482.
#include "ba_internal.h"
#include "sfmba_device.cuh"

namespace sfmba {

// rowptr must be a CSR over [0, num_obs] and idx (when given) must lie in [0, bound): the refiners index their
// observation and projection arrays through them.  They return at once when this kernel has raised the flag.
__global__ void k_validate_csr(int n_items, int num_obs, const int32_t* __restrict__ rowptr,
                               const int32_t* __restrict__ idx, int bound, uint32_t* flags) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  bool bad = false;
  if (j <= n_items) {
    const int a = rowptr[j];
    bad = a < 0 || a > num_obs;
    if (j == 0 && a != 0) bad = true;
    if (j == n_items && a != num_obs) bad = true;
    if (j < n_items && a > rowptr[j + 1]) bad = true;
  }
  if (idx) {
    const int stride = gridDim.x * blockDim.x;
    for (int o = j; o < num_obs; o += stride)
      if ((uint32_t)idx[o] >= (uint32_t)bound) bad = true;
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flags, FLAG_BAD_INDEX);
}

void launch_validate_csr(int n_items, int num_obs, const int32_t* rowptr, const int32_t* idx, int bound,
                         uint32_t* flags, cudaStream_t st, LaunchCounter& lc) {
  k_validate_csr<<<(n_items + 1 + 255) / 256, 256, 0, st>>>(n_items, num_obs, rowptr, idx, bound, flags);
  lc.launches++;
}

// X -= inv(J^T J + lambda I3) J^T (proj - obs)     triangulation_processor.py:220-228
__global__ void __launch_bounds__(128) k_triangulate(int num_pts, const double* __restrict__ projs,
                                                     const int32_t* __restrict__ rowptr,
                                                     const int32_t* __restrict__ obs_view,
                                                     const double* __restrict__ obs_uv, double* __restrict__ pts,
                                                     double lambda, int iterations, uint32_t* flags) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= num_pts || (*flags & FLAG_BAD_INDEX)) return;
  const int a = rowptr[j], e = rowptr[j + 1];
  double X[3] = {pts[3 * (size_t)j], pts[3 * (size_t)j + 1], pts[3 * (size_t)j + 2]};
  uint32_t f = 0;
  for (int it = 0; it < iterations; ++it) {
    double H[6] = {lambda, 0, 0, lambda, 0, lambda}, g[3] = {0, 0, 0};
    for (int o = a; o < e; ++o) {
      const double* P = projs + 12 * (size_t)obs_view[o];
      const double pu = P[0] * X[0] + P[1] * X[1] + P[2] * X[2] + P[3];
      const double pv = P[4] * X[0] + P[5] * X[1] + P[6] * X[2] + P[7];
      const double pw = P[8] * X[0] + P[9] * X[1] + P[10] * X[2] + P[11];
      if (pw == 0.0) f |= FLAG_DEPTH_ZERO;
      const double pw2 = pw * pw;
      const double eu = pu / pw - obs_uv[2 * (size_t)o];
      const double ev = pv / pw - obs_uv[2 * (size_t)o + 1];
      // triangulation_processor.py:264-267
      const double ju0 = (pw * P[0] - pu * P[8]) / pw2, ju1 = (pw * P[1] - pu * P[9]) / pw2,
                   ju2 = (pw * P[2] - pu * P[10]) / pw2;
      const double jv0 = (pw * P[4] - pv * P[8]) / pw2, jv1 = (pw * P[5] - pv * P[9]) / pw2,
                   jv2 = (pw * P[6] - pv * P[10]) / pw2;
      H[0] += ju0 * ju0 + jv0 * jv0;
      H[1] += ju0 * ju1 + jv0 * jv1;
      H[2] += ju0 * ju2 + jv0 * jv2;
      H[3] += ju1 * ju1 + jv1 * jv1;
      H[4] += ju1 * ju2 + jv1 * jv2;
      H[5] += ju2 * ju2 + jv2 * jv2;
      g[0] += ju0 * eu + jv0 * ev;
      g[1] += ju1 * eu + jv1 * ev;
      g[2] += ju2 * eu + jv2 * ev;
    }
    double Hi[6], dl[3];
    if (!sym3_inverse(H, Hi)) f |= FLAG_SINGULAR;
    sym3_mul(Hi, g, dl);
    X[0] -= dl[0];
    X[1] -= dl[1];
    X[2] -= dl[2];
  }
  if (!(X[0] == X[0]) || !(X[1] == X[1]) || !(X[2] == X[2])) f |= FLAG_NONFINITE;
  pts[3 * (size_t)j] = X[0];
  pts[3 * (size_t)j + 1] = X[1];
  pts[3 * (size_t)j + 2] = X[2];
  if (f) atomicOr(flags, f);
}

// One warp per view.  params = [C, q]; each iteration: err = b - f, J (2x7) per point,
// delta = inv(J^T J + lambda I7) J^T err, params += delta, q normalised, R = q->R.
// Row-packing quirk (campose_processor.py:404-405): only the u-rows of all points plus the v-row of
// the LAST point enter J and err.  Iteration 0 uses the given rotation as is (campose_processor.py:367).
constexpr int PNP_WARPS = 4;

__global__ void __launch_bounds__(32 * PNP_WARPS) k_pnp(int num_views, const int32_t* __restrict__ rowptr,
                                                        const double* __restrict__ obs_xy,
                                                        const double* __restrict__ obs_X, double* __restrict__ rot,
                                                        double* __restrict__ loc, double lambda, int iterations,
                                                        int row_quirk, int dvdc_quirk, uint32_t* flags) {
  __shared__ double s_ct[PNP_WARPS][CAMTAB_STRIDE];
  __shared__ double s_par[PNP_WARPS][8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int view = blockIdx.x * PNP_WARPS + w;
  if (view >= num_views || (*flags & FLAG_BAD_INDEX)) return;   // whole warp exits together; no block-level barrier below
  const int a = rowptr[view], e = rowptr[view + 1];
  uint32_t f = 0;
  if (lane == 0) {
    double R[9], q[4];
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = rot[9 * (size_t)view + i];
    if (!rot_to_quat(R, q)) f |= FLAG_QW_NEAR_ZERO;
    const double nrm = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
#pragma unroll
    for (int i = 0; i < 9; ++i) s_ct[w][i] = R[i];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      s_ct[w][9 + i] = loc[3 * (size_t)view + i];
      s_par[w][i] = loc[3 * (size_t)view + i];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      s_ct[w][12 + i] = q[i];           // Jacobian quaternion = R->q of the rotation in use
      s_par[w][3 + i] = q[i] / nrm;     // campose_processor.py:361-363
    }
  }
  __syncwarp();
  for (int it = 0; it < iterations; ++it) {
    double ct[CAMTAB_STRIDE];
#pragma unroll
    for (int i = 0; i < CAMTAB_STRIDE; ++i) ct[i] = s_ct[w][i];
    double acc[35];
#pragma unroll
    for (int i = 0; i < 35; ++i) acc[i] = 0.0;
    for (int o = a + lane; o < e; o += 32) {
      const double X[3] = {obs_X[3 * (size_t)o], obs_X[3 * (size_t)o + 1], obs_X[3 * (size_t)o + 2]};
      const double b[2] = {obs_xy[2 * (size_t)o], obs_xy[2 * (size_t)o + 1]};
      double r[2], Jp[14], Jx[6];
      if (!eval_obs<true>(ct, X, b, dvdc_quirk != 0, r, Jp, Jx)) f |= FLAG_DEPTH_ZERO;
      const bool use_v = !row_quirk || (o == e - 1);
      int k = 0;
#pragma unroll
      for (int i = 0; i < 7; ++i) {
#pragma unroll
        for (int jn = i; jn < 7; ++jn, ++k) {
          double v = Jp[i] * Jp[jn];
          if (use_v) v += Jp[7 + i] * Jp[7 + jn];
          acc[k] += v;
        }
      }
#pragma unroll
      for (int i = 0; i < 7; ++i) {
        double v = Jp[i] * r[0];
        if (use_v) v += Jp[7 + i] * r[1];
        acc[28 + i] += v;
      }
    }
#pragma unroll
    for (int i = 0; i < 35; ++i) acc[i] = warp_sum(acc[i]);
    if (lane == 0) {
      double H[49], g[7];
      int k = 0;
      for (int i = 0; i < 7; ++i) {
        for (int jn = i; jn < 7; ++jn, ++k) {
          H[i * 7 + jn] = acc[k];
          H[jn * 7 + i] = acc[k];
        }
        H[i * 7 + i] += lambda;
        g[i] = acc[28 + i];
      }
      if (!spd_inverse<7>(H)) f |= FLAG_SINGULAR;
      double par[7];
      for (int i = 0; i < 7; ++i) {
        double dl = 0.0;
        for (int jn = 0; jn < 7; ++jn) dl += H[i * 7 + jn] * g[jn];
        par[i] = s_par[w][i] + dl;
      }
      const double nrm = sqrt(par[3] * par[3] + par[4] * par[4] + par[5] * par[5] + par[6] * par[6]);
      for (int i = 3; i < 7; ++i) par[i] /= nrm;
      double R[9], qj[4];
      quat_to_rot(par + 3, R);
      if (!rot_to_quat(R, qj)) f |= FLAG_QW_NEAR_ZERO;
      for (int i = 0; i < 7; ++i) {
        s_par[w][i] = par[i];
        if (!(par[i] == par[i])) f |= FLAG_NONFINITE;
      }
      for (int i = 0; i < 9; ++i) s_ct[w][i] = R[i];
      for (int i = 0; i < 3; ++i) s_ct[w][9 + i] = par[i];
      for (int i = 0; i < 4; ++i) s_ct[w][12 + i] = qj[i];
    }
    __syncwarp();
  }
  if (lane == 0) {
    double R[9];
    double q[4] = {s_par[w][3], s_par[w][4], s_par[w][5], s_par[w][6]};
    quat_to_rot(q, R);   // campose_processor.py:458
    for (int i = 0; i < 9; ++i) rot[9 * (size_t)view + i] = R[i];
    for (int i = 0; i < 3; ++i) loc[3 * (size_t)view + i] = s_par[w][i];
  }
  f = __reduce_or_sync(0xffffffffu, f);
  if (lane == 0 && f) atomicOr(flags, f);
}

// ---- f2: linear (DLT) triangulation + cheirality, one thread per point --------
// triangulation_processor.py:136-155: rows u P[2] - P[0] and v P[2] - P[1] of every view of the point,
// the right singular vector of the smallest singular value, divided by its last entry.  The 2n x 4
// system is never stored: each row is rotated into a 4 x 4 upper-triangular factor (Givens QR, which has
// the same right singular vectors and no squaring of the condition number), and the factor's SVD comes
// from one-sided Jacobi rotations of its columns.  front[j] = 1 when the point has positive depth in
// every one of its views (campose_processor.py:181-189 for the two-view case).
__device__ __forceinline__ void givens_row_into_r(double R[4][4], double r[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (r[i] != 0.0) {
      const double h = hypot(R[i][i], r[i]);
      const double c = R[i][i] / h, s = r[i] / h;
      R[i][i] = h;
#pragma unroll
      for (int k = i + 1; k < 4; ++k) {
        const double t = c * R[i][k] + s * r[k];
        r[k] = c * r[k] - s * R[i][k];
        R[i][k] = t;
      }
    }
  }
}

__global__ void __launch_bounds__(128) k_triangulate_linear(int num_pts, const double* __restrict__ projs,
                                                            const int32_t* __restrict__ rowptr,
                                                            const int32_t* __restrict__ obs_view,
                                                            const double* __restrict__ obs_uv,
                                                            double* __restrict__ pts, int32_t* __restrict__ front,
                                                            uint32_t* flags) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= num_pts || (*flags & FLAG_BAD_INDEX)) return;
  const int a = rowptr[j], e = rowptr[j + 1];
  double U[4][4], V[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      U[i][k] = 0.0;
      V[i][k] = i == k ? 1.0 : 0.0;
    }
  }
  for (int o = a; o < e; ++o) {
    const double* P = projs + 12 * (size_t)obs_view[o];
    const double u = obs_uv[2 * (size_t)o], v = obs_uv[2 * (size_t)o + 1];
    double ru[4], rv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      ru[k] = u * P[8 + k] - P[k];
      rv[k] = v * P[8 + k] - P[4 + k];
    }
    givens_row_into_r(U, ru);
    givens_row_into_r(U, rv);
  }
  // one-sided Jacobi: rotate column pairs of U (and of V) until all columns are mutually orthogonal
  for (int sweep = 0; sweep < 12; ++sweep) {
    double off = 0.0;
#pragma unroll
    for (int p = 0; p < 3; ++p) {
#pragma unroll
      for (int q = p + 1; q < 4; ++q) {
        double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          al += U[k][p] * U[k][p];
          be += U[k][q] * U[k][q];
          ga += U[k][p] * U[k][q];
        }
        if (ga != 0.0 && fabs(ga) > 1e-17 * sqrt(al * be)) {
          off = fmax(off, fabs(ga) / sqrt(al * be));
          const double zeta = (be - al) / (2.0 * ga);
          const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
          const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const double up = U[k][p], uq = U[k][q];
            U[k][p] = c * up - sn * uq;
            U[k][q] = sn * up + c * uq;
            const double vp = V[k][p], vq = V[k][q];
            V[k][p] = c * vp - sn * vq;
            V[k][q] = sn * vp + c * vq;
          }
        }
      }
    }
    if (off < 1e-15) break;
  }
  int best = 0;
  double smin = 0.0;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    double n2 = 0.0;
#pragma unroll
    for (int k = 0; k < 4; ++k) n2 += U[k][q] * U[k][q];
    if (q == 0 || n2 < smin) {
      smin = n2;
      best = q;
    }
  }
  double x[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    x[k] = V[k][0];
#pragma unroll
    for (int q = 1; q < 4; ++q)
      if (q == best) x[k] = V[k][q];
  }
  uint32_t f = 0;
  if (x[3] == 0.0) f |= FLAG_DEPTH_ZERO;
  const double X0 = x[0] / x[3], X1 = x[1] / x[3], X2 = x[2] / x[3];
  if (!(X0 == X0) || !(X1 == X1) || !(X2 == X2) || isinf(X0) || isinf(X1) || isinf(X2)) f |= FLAG_NONFINITE;
  pts[3 * (size_t)j] = X0;
  pts[3 * (size_t)j + 1] = X1;
  pts[3 * (size_t)j + 2] = X2;
  if (front) {
    int ok = 1;
    for (int o = a; o < e; ++o) {
      const double* P = projs + 12 * (size_t)obs_view[o];
      if (!(P[8] * X0 + P[9] * X1 + P[10] * X2 + P[11] > 0.0)) ok = 0;
    }
    front[j] = ok;
  }
  if (f) atomicOr(flags, f);
}

void launch_triangulate_linear(int num_pts, const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                               const double* obs_uv, double* pts, int32_t* front, uint32_t* flags, cudaStream_t st,
                               LaunchCounter& lc) {
  if (num_pts <= 0) return;
  k_triangulate_linear<<<(num_pts + 127) / 128, 128, 0, st>>>(num_pts, projs, rowptr, obs_view, obs_uv, pts, front,
                                                              flags);
  lc.launches++;
}

void launch_triangulate(int num_pts, const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                        const double* obs_uv, double* pts, double lambda, int iterations, uint32_t* flags,
                        cudaStream_t st, LaunchCounter& lc) {
  if (num_pts <= 0) return;
  k_triangulate<<<(num_pts + 127) / 128, 128, 0, st>>>(num_pts, projs, rowptr, obs_view, obs_uv, pts, lambda,
                                                       iterations, flags);
  lc.launches++;
}

void launch_pnp(int num_views, const int32_t* rowptr, const double* obs_xy, const double* obs_X, double* rot,
                double* loc, double lambda, int iterations, bool row_packing_quirk, bool dvdc_quirk,
                uint32_t* flags, cudaStream_t st, LaunchCounter& lc) {
  if (num_views <= 0) return;
  k_pnp<<<(num_views + PNP_WARPS - 1) / PNP_WARPS, 32 * PNP_WARPS, 0, st>>>(
      num_views, rowptr, obs_xy, obs_X, rot, loc, lambda, iterations, row_packing_quirk ? 1 : 0,
      dvdc_quirk ? 1 : 0, flags);
  lc.launches++;
}

}  // namespace sfmba
