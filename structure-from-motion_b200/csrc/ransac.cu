// f3: RANSAC hypotheses evaluated in parallel.
//
// Reference loops (one hypothesis after another, a Python loop over every point inside):
//   CamposeProcessor.__linear_determine_cam_pos   campose_processor.py:485-563  (6-point DLT PnP)
//   CamposeProcessor.__estimate_six_pts           campose_processor.py:565-633  (12x12 SVD, 3x3 SVD)
//   EpipolarProcessor.__estimate_ransac           epipolar_processor.py:194-247 (8-point F)
//   EpipolarProcessor.__estimate_eight_pts        epipolar_processor.py:140-191 (8x9 SVD, rank-2 projection)
// The sample indices are drawn on the host with Python's random.sample, in the reference's order, and
// handed over as an [H][6] / [H][8] table, so the draws are the reference's.  Here
//   k_*_models   one thread per hypothesis: the minimal-sample model.  Null vectors come from a one-sided
//                Jacobi SVD (columns of W are rotated until mutually orthogonal; the right singular vector of
//                the smallest singular value is the column of V that belongs to the shortest column), which
//                has LAPACK's accuracy without forming W^T W
//   k_*_score    one CTA per hypothesis: inlier count over all points (integer, fixed order)
//   k_ransac_pick   first hypothesis with the strictly largest count (the reference's `>` update rule)
//   k_*_mask     inlier mask of the winner
// LAPACK-sign quirk of the reference (PnP only): the sign of the last right singular vector decides whether
// campose_processor.py:620-631 returns (R, C) or (R, -C); the model kernel computes the sign-free pair and
// multiplies C by loc_sign[h] (+1 / -1, NULL = +1, the geometrically consistent choice).
#include "ba_internal.h"
#include "sfmba_device.cuh"

namespace sfmba {

// One-sided (Hestenes) Jacobi on the columns of A (M x N, row-major), V (N x N) accumulates the rotations.
// On return the columns of A are mutually orthogonal: A_in V = A_out, column norms = singular values.
template <int M, int N>
__device__ void jacobi_columns(double* A, double* V) {
  for (int i = 0; i < N; ++i)
    for (int k = 0; k < N; ++k) V[i * N + k] = i == k ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 40; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < N - 1; ++p) {
      for (int q = p + 1; q < N; ++q) {
        double al = 0.0, be = 0.0, ga = 0.0;
        for (int k = 0; k < M; ++k) {
          const double ap = A[k * N + p], aq = A[k * N + q];
          al = fma(ap, ap, al);
          be = fma(aq, aq, be);
          ga = fma(ap, aq, ga);
        }
        const double lim = sqrt(al * be);
        if (ga != 0.0 && fabs(ga) > 1e-17 * lim) {
          off = fmax(off, fabs(ga) / lim);
          const double zeta = (be - al) / (2.0 * ga);
          const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
          const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
          for (int k = 0; k < M; ++k) {
            const double ap = A[k * N + p], aq = A[k * N + q];
            A[k * N + p] = c * ap - sn * aq;
            A[k * N + q] = sn * ap + c * aq;
          }
          for (int k = 0; k < N; ++k) {
            const double vp = V[k * N + p], vq = V[k * N + q];
            V[k * N + p] = c * vp - sn * vq;
            V[k * N + q] = sn * vp + c * vq;
          }
        }
      }
    }
    if (off < 1e-15) break;
  }
}

// right singular vector of the smallest singular value of W (M x N); W is destroyed
template <int M, int N>
__device__ void null_vector(double* W, double* V, double* x) {
  jacobi_columns<M, N>(W, V);
  int best = 0;
  double smin = 0.0;
  for (int q = 0; q < N; ++q) {
    double n2 = 0.0;
    for (int k = 0; k < M; ++k) n2 = fma(W[k * N + q], W[k * N + q], n2);
    if (q == 0 || n2 < smin) {
      smin = n2;
      best = q;
    }
  }
  for (int k = 0; k < N; ++k) x[k] = V[k * N + best];
}

// 3x3 SVD by the same rotations: B = A V has orthogonal columns of norm s[i]
__device__ __forceinline__ void svd3(const double A[9], double B[9], double V[9], double s[3]) {
  for (int i = 0; i < 9; ++i) B[i] = A[i];
  jacobi_columns<3, 3>(B, V);
  for (int q = 0; q < 3; ++q) s[q] = sqrt(B[q] * B[q] + B[3 + q] * B[3 + q] + B[6 + q] * B[6 + q]);
}

__device__ __forceinline__ double det3(const double a[9]) {
  return a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
}

// ---- 6-point DLT pose (campose_processor.py:565-633) -----------------------------------------------
// models[h] = rot[9] (row-major, camera-to-world) | loc[3];  projs[h] = K [rot^T | -rot^T loc] (3x4)
__global__ void __launch_bounds__(32) k_pnp_models(int H, int n, const double* __restrict__ key, const double* __restrict__ X,
                                                   const double* __restrict__ Kmat, const double* __restrict__ Kinv,
                                                   const int32_t* __restrict__ samples,
                                                   const int32_t* __restrict__ loc_sign, double* __restrict__ models,
                                                   double* __restrict__ projs, uint32_t* flags) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= H) return;
  double W[144], V[144], v[12];
  for (int i = 0; i < 144; ++i) W[i] = 0.0;
  for (int i = 0; i < 6; ++i) {
    int s = samples[6 * (size_t)h + i];
    if (s < 0 || s >= n) {                       // reported to the caller as SFMBA_ERR_INVALID_ARG
      atomicOr(flags, FLAG_BAD_INDEX);
      s = 0;
    }
    const double kx = key[3 * (size_t)s], ky = key[3 * (size_t)s + 1], kz = key[3 * (size_t)s + 2];
    // campose_processor.py:537: key points of the sample in camera coordinates
    const double px = Kinv[0] * kx + Kinv[1] * ky + Kinv[2] * kz;
    const double py = Kinv[3] * kx + Kinv[4] * ky + Kinv[5] * kz;
    const double pz = Kinv[6] * kx + Kinv[7] * ky + Kinv[8] * kz;
    const double x0 = X[4 * (size_t)s], x1 = X[4 * (size_t)s + 1], x2 = X[4 * (size_t)s + 2];
    double* r0 = W + 24 * i;
    double* r1 = r0 + 12;
    r0[0] = pz * x0;  r0[1] = pz * x1;  r0[2] = pz * x2;  r0[3] = pz;       // :589-600 (4th column is p2d_z)
    r0[8] = -px * x0; r0[9] = -px * x1; r0[10] = -px * x2; r0[11] = -px;
    r1[4] = pz * x0;  r1[5] = pz * x1;  r1[6] = pz * x2;  r1[7] = pz;       // :602-613
    r1[8] = -py * x0; r1[9] = -py * x1; r1[10] = -py * x2; r1[11] = -py;
  }
  null_vector<12, 12>(W, V, v);
  double Mm[9] = {v[0], v[1], v[2], v[4], v[5], v[6], v[8], v[9], v[10]};
  double t[3] = {v[3], v[7], v[11]};
  if (det3(Mm) < 0.0) {               // sign-free form: det(rot_t_noise) > 0
    for (int i = 0; i < 9; ++i) Mm[i] = -Mm[i];
    for (int i = 0; i < 3; ++i) t[i] = -t[i];
  }
  double B[9], V3[9], s[3];
  svd3(Mm, B, V3, s);
  // Q = uu vvh = sum_i (b_i / s_i) v_i^T  (:622);  rot = Q^T
  double Q[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c)
      Q[3 * r + c] = B[3 * r] / s[0] * V3[3 * c] + B[3 * r + 1] / s[1] * V3[3 * c + 1] + B[3 * r + 2] / s[2] * V3[3 * c + 2];
  const double smax = fmax(s[0], fmax(s[1], s[2]));
  const double sg = loc_sign ? (double)loc_sign[h] : 1.0;
  double loc[3];
  for (int r = 0; r < 3; ++r) loc[r] = sg * (-(Q[r] * t[0] + Q[3 + r] * t[1] + Q[6 + r] * t[2])) / smax;   // :625
  double* mo = models + 12 * (size_t)h;
  bool bad = false;
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      mo[3 * r + c] = Q[3 * c + r];
      bad |= !(fabs(Q[3 * c + r]) <= 2.0);
    }
  for (int r = 0; r < 3; ++r) {
    mo[9 + r] = loc[r];
    bad |= !(fabs(loc[r]) < 1e300);
  }
  if (bad) atomicOr(flags, FLAG_NONFINITE);
  // proj = K [rot^T | rot^T (-loc)] = K [Q | -Q loc]   (:540)
  double E[12];
  for (int r = 0; r < 3; ++r) {
    E[4 * r] = Q[3 * r];
    E[4 * r + 1] = Q[3 * r + 1];
    E[4 * r + 2] = Q[3 * r + 2];
    E[4 * r + 3] = -(Q[3 * r] * loc[0] + Q[3 * r + 1] * loc[1] + Q[3 * r + 2] * loc[2]);
  }
  double* po = projs + 12 * (size_t)h;
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c)
      po[4 * r + c] = Kmat[3 * r] * E[c] + Kmat[3 * r + 1] * E[4 + c] + Kmat[3 * r + 2] * E[8 + c];
}

// reprojection test of campose_processor.py:546-556: err = || pt_2d - proj X / (proj X)_z || < threshold
__device__ __forceinline__ bool pnp_inlier(const double* __restrict__ P, const double* __restrict__ key,
                                           const double* __restrict__ X, int i, double thr) {
  const double x0 = X[4 * (size_t)i], x1 = X[4 * (size_t)i + 1], x2 = X[4 * (size_t)i + 2], x3 = X[4 * (size_t)i + 3];
  const double a = P[0] * x0 + P[1] * x1 + P[2] * x2 + P[3] * x3;
  const double b = P[4] * x0 + P[5] * x1 + P[6] * x2 + P[7] * x3;
  const double c = P[8] * x0 + P[9] * x1 + P[10] * x2 + P[11] * x3;
  const double e0 = key[3 * (size_t)i] - a / c, e1 = key[3 * (size_t)i + 1] - b / c, e2 = key[3 * (size_t)i + 2] - c / c;
  return sqrt(e0 * e0 + e1 * e1 + e2 * e2) < thr;
}

// epipolar test of epipolar_processor.py:231-236: | x_r^T F x_l | < threshold, pairs = [xl, yl, xr, yr]
__device__ __forceinline__ bool fund_inlier(const double* __restrict__ F, const double* __restrict__ pairs, int i,
                                            double thr) {
  const double xl = pairs[4 * (size_t)i], yl = pairs[4 * (size_t)i + 1];
  const double xr = pairs[4 * (size_t)i + 2], yr = pairs[4 * (size_t)i + 3];
  const double a = F[0] * xl + F[1] * yl + F[2];
  const double b = F[3] * xl + F[4] * yl + F[5];
  const double c = F[6] * xl + F[7] * yl + F[8];
  return fabs(xr * a + yr * b + c) < thr;
}

template <bool PNP>
__global__ void __launch_bounds__(256) k_ransac_score(int n, const double* __restrict__ model, int model_stride,
                                                      const double* __restrict__ a, const double* __restrict__ b,
                                                      double thr, int32_t* __restrict__ counts) {
  __shared__ double P[12];
  __shared__ int wsum[8];
  const int h = blockIdx.x;
  if (threadIdx.x < model_stride) P[threadIdx.x] = model[(size_t)model_stride * h + threadIdx.x];
  __syncthreads();
  int cnt = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) cnt += PNP ? pnp_inlier(P, a, b, i, thr) : fund_inlier(P, a, i, thr);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < 8; ++w) t += wsum[w];
    counts[h] = t;
  }
}

// winner = first hypothesis whose count is strictly larger than every earlier one (campose_processor.py:558,
// epipolar_processor.py:239); best = -1 when no hypothesis has an inlier
__global__ void __launch_bounds__(1024) k_ransac_pick(int H, const int32_t* __restrict__ counts, int32_t* __restrict__ best) {
  __shared__ int sc[1024], sh[1024];
  int c = 0, hh = -1;
  for (int h = threadIdx.x; h < H; h += blockDim.x) {
    if (counts[h] > c) {
      c = counts[h];
      hh = h;
    }
  }
  sc[threadIdx.x] = c;
  sh[threadIdx.x] = hh;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      const int c2 = sc[threadIdx.x + o], h2 = sh[threadIdx.x + o];
      const int c1 = sc[threadIdx.x], h1 = sh[threadIdx.x];
      if (c2 > c1 || (c2 == c1 && h2 >= 0 && (h1 < 0 || h2 < h1))) {
        sc[threadIdx.x] = c2;
        sh[threadIdx.x] = h2;
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    best[0] = sh[0];
    best[1] = sc[0];
  }
}

template <bool PNP>
__global__ void __launch_bounds__(256) k_ransac_mask(int n, const double* __restrict__ model, int model_stride,
                                                     const int32_t* __restrict__ best, const double* __restrict__ a,
                                                     const double* __restrict__ b, double thr, uint8_t* __restrict__ mask) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int h = best[0];
  if (h < 0) {
    mask[i] = 0;
    return;
  }
  const double* P = model + (size_t)model_stride * h;
  mask[i] = (PNP ? pnp_inlier(P, a, b, i, thr) : fund_inlier(P, a, i, thr)) ? 1 : 0;
}

// ---- 8-point fundamental matrix (epipolar_processor.py:140-191) ------------------------------------
__global__ void __launch_bounds__(32) k_fund_models(int H, int n, const double* __restrict__ pairs,
                                                    const int32_t* __restrict__ samples, double* __restrict__ models,
                                                    uint32_t* flags) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= H) return;
  double W[72], V[81], f[9];
  for (int i = 0; i < 8; ++i) {
    int s = samples[8 * (size_t)h + i];
    if (s < 0 || s >= n) {
      atomicOr(flags, FLAG_BAD_INDEX);
      s = 0;
    }
    const double x1 = pairs[4 * (size_t)s], y1 = pairs[4 * (size_t)s + 1];
    const double x1_ = pairs[4 * (size_t)s + 2], y1_ = pairs[4 * (size_t)s + 3];
    double* r = W + 9 * i;
    r[0] = x1 * x1_; r[1] = y1 * x1_; r[2] = x1_;
    r[3] = x1 * y1_; r[4] = y1 * y1_; r[5] = y1_;
    r[6] = x1;       r[7] = y1;       r[8] = 1.0;
  }
  null_vector<8, 9>(W, V, f);
  // rank-2 projection (:176-177): drop the smallest singular value, f__ = sum over the two largest of b_i v_i^T
  double B[9], V3[9], s[3];
  svd3(f, B, V3, s);
  int drop = 0;
  if (s[1] < s[drop]) drop = 1;
  if (s[2] < s[drop]) drop = 2;
  const double smax = fmax(s[0], fmax(s[1], s[2]));
  const double smid = s[0] + s[1] + s[2] - smax - s[drop];   // the middle singular value
  double F2[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      double a = 0.0;
      for (int q = 0; q < 3; ++q)
        if (q != drop) a += B[3 * r + q] * V3[3 * c + q];
      F2[3 * r + c] = a;
    }
  // :179-183 raises when rank(f__) != 2 (numpy.linalg.matrix_rank: s > s_max * 3 * eps)
  uint32_t fl = 0;
  if (!(smid > smax * 3.0 * 2.220446049250313e-16)) fl |= FLAG_SINGULAR;
  double* mo = models + 9 * (size_t)h;
  for (int i = 0; i < 9; ++i) {
    mo[i] = F2[i] / F2[8];                             // :185
  }
  if (fl) atomicOr(flags, fl);
}

void launch_pnp_ransac(int n, int H, const double* key, const double* X, const double* Kmat, const double* Kinv,
                       const int32_t* samples, const int32_t* loc_sign, double thr, double* models, double* projs,
                       int32_t* counts, int32_t* best, uint8_t* mask, uint32_t* flags, cudaStream_t st,
                       LaunchCounter& lc) {
  if (H <= 0) return;
  k_pnp_models<<<(H + 31) / 32, 32, 0, st>>>(H, n, key, X, Kmat, Kinv, samples, loc_sign, models, projs, flags);
  k_ransac_score<true><<<H, 256, 0, st>>>(n, projs, 12, key, X, thr, counts);
  k_ransac_pick<<<1, 1024, 0, st>>>(H, counts, best);
  lc.launches += 3;
  if (n > 0) {
    k_ransac_mask<true><<<(n + 255) / 256, 256, 0, st>>>(n, projs, 12, best, key, X, thr, mask);
    lc.launches++;
  }
}

void launch_fund_ransac(int n, int H, const double* pairs, const int32_t* samples, double thr, double* models,
                        int32_t* counts, int32_t* best, uint8_t* mask, uint32_t* flags, cudaStream_t st,
                        LaunchCounter& lc) {
  if (H <= 0) return;
  k_fund_models<<<(H + 31) / 32, 32, 0, st>>>(H, n, pairs, samples, models, flags);
  k_ransac_score<false><<<H, 256, 0, st>>>(n, models, 9, pairs, nullptr, thr, counts);
  k_ransac_pick<<<1, 1024, 0, st>>>(H, counts, best);
  lc.launches += 3;
  if (n > 0) {
    k_ransac_mask<false><<<(n + 255) / 256, 256, 0, st>>>(n, models, 9, best, pairs, nullptr, thr, mask);
    lc.launches++;
  }
}

}  // namespace sfmba
