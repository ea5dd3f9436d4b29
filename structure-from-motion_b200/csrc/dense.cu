// Dense direct solve of the reduced camera system (sfmba_options.solver = SFMBA_SOLVER_DENSE_CHOLESKY).
//
// Reference: ba_processor.py:376-382 forms A = J_p^T J_p + lambda I, B = J_p^T J_x and e_p densely and
// computes delta_p = inv(A - B D^-1 B^T)(e_p - B D^-1 e_x).  Here the reduced matrix
//     S = A + lambda I - sum_j W_j D_j^-1 W_j^T,   W_ij = J_p,ij^T J_x,ij  (7x3)
// is assembled explicitly (lower triangle, column-major, padded to a multiple of 128 with an identity
// block), factored S = L L^T by a blocked right-looking Cholesky whose O(n^3) part runs on the FP64
// tensor cores (mma.sync m8n8k4 f64: DMMA), and solved by blocked substitution.  The right-hand side
// g = e_p - B D^-1 e_x and the diagonal blocks S_ii come from the linearisation kernel (cam_acc), so
// only the off-diagonal blocks are accumulated here.
//
//   k_dense_offdiag   one warp per point: W_i, Z_i = W_i D^-1 of its observations in shared memory,
//                     then every camera pair's 7x7 block Z_a W_b^T goes to S with red.global.add.f64
//                     (the only data atomics of the library: this mode is not bit-reproducible)
//   k_dense_diag      S_ii += A_i + lambda I - Ssub_i from cam_acc; identity padding; rhs = g
//   per 128-column panel k:
//     k_dense_potrf   one CTA: L_kk (pivots first, scaling afterwards: one barrier per column) and
//                     L_kk^-1 by a row sweep with the inverse held in registers
//     k_dense_gemm<1> L_ik = S_ik L_kk^-T for the tiles below the diagonal (a GEMM with the inverse)
//     k_dense_gemm<0> S_ij -= L_ik L_jk^T for the trailing lower tiles (i >= j > k)
//   k_dense_fwd / k_dense_bwd   blocked substitution, one launch per panel
#include "ba_internal.h"
#include "sfmba_device.cuh"
#include "tile.cuh"

#include <algorithm>
#include <vector>

namespace sfmba {

constexpr int DT = 128;            // tile / panel width
constexpr int DA_WARPS = 4;        // points per assembly CTA
constexpr int DA_TILE = 32;        // observations of a point held in shared memory at a time

// W = J_p^T J_x (7x3, row-major) of one observation of point X
__device__ __forceinline__ bool obs_w(const BaDev& d, int o, const double X[3], bool quirk, int& cam, double W[21]) {
  cam = d.obs_cam[o];
  double b[2], r[2], Jp[14], Jx[6];
  normalise_obs(d.kinv ? d.kinv + 9 * (size_t)cam : nullptr, d.obs_uv[o], d.obs_uv[d.m + o], b);
  const bool ok = eval_obs<true>(d.camtab + (size_t)CAMTAB_STRIDE * cam, X, b, quirk, r, Jp, Jx);
#pragma unroll
  for (int a = 0; a < 7; ++a) {
#pragma unroll
    for (int t = 0; t < 3; ++t) W[3 * a + t] = Jp[a] * Jx[t] + Jp[7 + a] * Jx[3 + t];
  }
  return ok;
}

__global__ void __launch_bounds__(32 * DA_WARPS) k_dense_offdiag(BaDev d, double* __restrict__ S, size_t ld,
                                                                 int dvdc_quirk) {
  __shared__ double sZ[DA_WARPS][DA_TILE][21];
  __shared__ double sW[DA_WARPS][DA_TILE][21];
  __shared__ int sCa[DA_WARPS][DA_TILE], sCb[DA_WARPS][DA_TILE];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * DA_WARPS + warp;
  if (j >= d.N) return;                                  // (no block-wide barrier below)
  const int o0 = d.rowptr[j], o1 = d.rowptr[j + 1];
  const int k = o1 - o0;
  if (k < 2) return;
  const size_t N = (size_t)d.N;
  double X[3], Dinv[6];
#pragma unroll
  for (int e = 0; e < 3; ++e) X[e] = d.pts[3 * (size_t)j + e];
#pragma unroll
  for (int e = 0; e < 6; ++e) Dinv[e] = d.dinv[e * N + j];
  const bool quirk = dvdc_quirk != 0;
  for (int a0 = 0; a0 < k; a0 += DA_TILE) {
    const int na = min(DA_TILE, k - a0);
    double W[21];
    int cam = -1;
    __syncwarp();
    if (lane < na) {
      obs_w(d, o0 + a0 + lane, X, quirk, cam, W);
#pragma unroll
      for (int a = 0; a < 7; ++a) {
        double z[3];
        sym3_mul(Dinv, W + 3 * a, z);                    // Z[a][:] = W[a][:] D^-1 (D^-1 symmetric)
        sZ[warp][lane][3 * a] = z[0];
        sZ[warp][lane][3 * a + 1] = z[1];
        sZ[warp][lane][3 * a + 2] = z[2];
      }
      sCa[warp][lane] = cam;
    }
    for (int b0 = 0; b0 <= a0; b0 += DA_TILE) {
      const int nb = min(DA_TILE, k - b0);
      __syncwarp();
      if (b0 == a0) {
        if (lane < na) {
#pragma unroll
          for (int e = 0; e < 21; ++e) sW[warp][lane][e] = W[e];
          sCb[warp][lane] = cam;
        }
      } else if (lane < nb) {
        double Wb[21];
        int camb;
        obs_w(d, o0 + b0 + lane, X, quirk, camb, Wb);
#pragma unroll
        for (int e = 0; e < 21; ++e) sW[warp][lane][e] = Wb[e];
        sCb[warp][lane] = camb;
      }
      __syncwarp();
      for (int a = 0; a < na; ++a) {
        const int nbb = (b0 == a0) ? a : nb;              // b < a inside the diagonal tile
        const int cam_a = sCa[warp][a];
        const double* Za = sZ[warp][a];
        for (int e = lane; e < nbb * 49; e += 32) {
          const int b = e / 49, el = e - 49 * b;
          const int c = el / 7, r = el - 7 * c;           // consecutive lanes: consecutive rows of one column
          const int cam_b = sCb[warp][b];
          const double* Wb = sW[warp][b];
          const double v_rc = Za[3 * r] * Wb[3 * c] + Za[3 * r + 1] * Wb[3 * c + 1] + Za[3 * r + 2] * Wb[3 * c + 2];
          const double v_cr = Za[3 * c] * Wb[3 * r] + Za[3 * c + 1] * Wb[3 * r + 1] + Za[3 * c + 2] * Wb[3 * r + 2];
          double v;
          size_t row, col;
          if (cam_a > cam_b) {                            // block (cam_a, cam_b) = W_a D^-1 W_b^T
            v = v_rc;
            row = 7 * (size_t)cam_a + r;
            col = 7 * (size_t)cam_b + c;
          } else if (cam_a < cam_b) {                     // block (cam_b, cam_a) = its transpose
            v = v_cr;
            row = 7 * (size_t)cam_b + r;
            col = 7 * (size_t)cam_a + c;
          } else {                                        // one camera twice: both orders land on the diagonal block
            v = v_rc + v_cr;
            row = 7 * (size_t)cam_a + r;
            col = 7 * (size_t)cam_a + c;
          }
          atomicAdd(S + col * ld + row, -v);
        }
      }
    }
  }
}

// diagonal blocks from the linearisation's per-camera sums, identity padding, right-hand side
__global__ void __launch_bounds__(256) k_dense_diag(BaDev d, double* __restrict__ S, size_t ld, int n, int npad,
                                                    double* __restrict__ rhs) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < d.V * 49) {
    const int cam = idx / 49, el = idx - 49 * cam;
    const int c = el / 7, r = el - 7 * c;
    const int a = min(r, c), b = max(r, c);
    const int k = 7 * a - a * (a - 1) / 2 + (b - a);      // packed upper triangle of k_linearize
    const double* acc = d.cam_acc + (size_t)cam * CAM_ACC;
    double v = acc[k] - acc[28 + k];
    if (r == c) v += d.sc->lambda;
    S[(7 * (size_t)cam + c) * ld + 7 * (size_t)cam + r] += v;
  }
  if (idx < npad) {
    rhs[idx] = idx < n ? d.gvec[idx] : 0.0;
    if (idx >= n) S[(size_t)idx * ld + idx] = 1.0;
  }
}

// ---- diagonal tile: Cholesky factor and its inverse -------------------------------------------------
// Thread (r, cg): row r = tid & 127, column class cg = tid >> 7 (columns c = cg, cg + 8, ...).
// Elimination keeps the columns unscaled (a_rc -= a_rj a_cj / a_jj touches only columns > j), so a
// step needs one barrier; L_rc = a_rc / sqrt(a_cc) afterwards.  The inverse solves L X = I row by row:
// thread (r, cg) keeps X[r][c] of its 16 columns in registers, the finished row t is broadcast
// through a double-buffered shared row.
constexpr size_t POTRF_SMEM = ((size_t)DT * DT + 3 * DT) * sizeof(double);

__global__ void __launch_bounds__(1024) k_dense_potrf(double* __restrict__ S, size_t ld, int k,
                                                      double* __restrict__ linv_all, Scalars* sc) {
  extern __shared__ __align__(16) double sm[];
  double* A = sm;                       // [c][r] column-major 128 x 128
  double* xrow = A + DT * DT;           // [2][128]
  double* dsq = xrow + 2 * DT;          // [128] sqrt of the pivots
  const int tid = threadIdx.x, r = tid & (DT - 1), cg = tid >> 7;
  double* G = S + (size_t)k * DT * ld + (size_t)k * DT;
  for (int c = cg; c < DT; c += 8) A[c * DT + r] = G[(size_t)c * ld + r];
  __syncthreads();
  for (int j = 0; j < DT - 1; ++j) {
    if (r > j) {
      const double lrj = A[j * DT + r] / A[j * DT + j];
      for (int c = j + 1 + ((cg - (j + 1)) & 7); c <= r; c += 8) A[c * DT + r] -= lrj * A[j * DT + c];
    }
    __syncthreads();
  }
  if (tid < DT) {
    const double p = A[tid * DT + tid];
    if (!(p > 0.0)) atomicOr(&sc->flags, FLAG_SINGULAR);
    dsq[tid] = sqrt(p);
  }
  __syncthreads();
  for (int c = cg; c < DT; c += 8) {
    const double s = dsq[c];
    const double v = r > c ? A[c * DT + r] / s : (r == c ? s : 0.0);
    A[c * DT + r] = v;
    G[(size_t)c * ld + r] = v;                           // (upper part of the diagonal tile: zeros)
  }
  __syncthreads();
  // inverse
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = (cg + 8 * i == r) ? 1.0 : 0.0;
  for (int t = 0; t < DT; ++t) {
    double* xr = xrow + (t & 1) * DT;
    if (r == t) {
      const double invd = 1.0 / A[t * DT + t];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int c = cg + 8 * i;
        if (c <= t) {
          acc[i] *= invd;
          xr[c] = acc[i];
        }
      }
    }
    __syncthreads();
    if (r > t) {
      const double l = A[t * DT + r];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int c = cg + 8 * i;
        if (c <= t) acc[i] -= l * xr[c];
      }
    }
  }
  double* Li = linv_all + (size_t)k * DT * DT;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int c = cg + 8 * i;
    Li[c * DT + r] = c <= r ? acc[i] : 0.0;
  }
}

// ---- tile GEMM on the FP64 tensor cores ---------------------------------------------------------------
// D (128 x 128) = X Y^T with X, Y 128 x 128 (column-major, K along the columns).  16 warps, warp tile
// 32 x 32 = 4 x 4 DMMA tiles (m8n8k4), operands staged in K chunks of 32 columns with cp.async, double
// buffered; shared rows padded to 132 doubles so that the fragment loads (lane -> [k0 + lane % 4][r0 +
// lane / 4]) hit 16 distinct 8-byte banks per half warp.
constexpr int GK = 32;
constexpr int GS = DT + 4;
constexpr size_t GEMM_SMEM = (size_t)2 * 2 * GK * GS * sizeof(double);

__device__ __forceinline__ void dmma_m8n8k4(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// MODE 0 (trailing update): tile (ti, tj), i >= j > k:  S_ij -= S_ik S_jk^T
//        part 0: every such tile;  part 1: only the next panel's column j = k + 1 (what the next diagonal
//        factor and panel solve wait for);  part 2: the columns j >= k + 2 (runs beside them: look-ahead)
// MODE 1 (panel solve):     tile (ti, k),  i > k:       S_ik  = S_ik Linv_k^T
// CTA b takes tiles b, b + gridDim.x, ... (the K chunks of consecutive tiles form one stream through the two
// shared-memory stages); the launches below use one tile per CTA.
template <int MODE>
__global__ void __launch_bounds__(512) k_dense_gemm(double* __restrict__ S, size_t ld, int k,
                                                    const double* __restrict__ linv_all, int part, int ntiles) {
  extern __shared__ __align__(16) double sm[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  auto decode = [&](int b, int& ti, int& tj) {
    if (MODE == 0 && part == 1) {
      ti = k + 1 + b;
      tj = k + 1;
    } else if (MODE == 0) {
      int i = (int)((sqrt(8.0 * (double)b + 1.0) - 1.0) * 0.5);
      while ((long long)(i + 1) * (i + 2) / 2 <= b) ++i;
      while ((long long)i * (i + 1) / 2 > b) --i;
      const int base = part == 2 ? k + 2 : k + 1;      // part 2: the triangle without its first column
      ti = base + i;
      tj = base + (b - (int)((long long)i * (i + 1) / 2));
    } else {
      ti = k + 1 + b;
      tj = k;
    }
  };
  const size_t ldy = MODE == 0 ? ld : (size_t)DT;
  auto load = [&](int stage, int ti, int tj, int kc) {
    const double* X = S + (size_t)k * DT * ld + (size_t)ti * DT;
    const double* Y = MODE == 0 ? S + (size_t)k * DT * ld + (size_t)tj * DT : linv_all + (size_t)k * DT * DT;
    double* xs = sm + (size_t)stage * 2 * GK * GS;
    double* ys = xs + GK * GS;
    for (int p = tid; p < GK * (DT / 2); p += 512) {
      const int col = p >> 6, piece = p & 63;
      cp_async16(xs + col * GS + 2 * piece, X + (size_t)(kc * GK + col) * ld + 2 * piece);
      cp_async16(ys + col * GS + 2 * piece, Y + (size_t)(kc * GK + col) * ldy + 2 * piece);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const int r0 = (warp & 3) * 32 + (lane >> 2), c0 = (warp >> 2) * 32 + (lane >> 2), kl = lane & 3;
  constexpr int NCH = DT / GK;
  int t = blockIdx.x;
  if (t >= ntiles) return;
  int ti, tj;
  decode(t, ti, tj);
  int seq = 0;                                           // chunks consumed so far: stage = seq & 1
  load(0, ti, tj, 0);
  for (; t < ntiles; t += gridDim.x) {
    const int tn = t + gridDim.x;
    int tin = 0, tjn = 0;
    if (tn < ntiles) decode(tn, tin, tjn);
    double acc[4][4][2];
#pragma unroll
    for (int mi = 0; mi < 4; ++mi)
#pragma unroll
      for (int ni = 0; ni < 4; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
    for (int kc = 0; kc < NCH; ++kc, ++seq) {
      bool more = true;
      if (kc + 1 < NCH) {
        load((seq + 1) & 1, ti, tj, kc + 1);
      } else if (tn < ntiles) {
        load((seq + 1) & 1, tin, tjn, 0);
      } else {
        more = false;
      }
      if (more) {
        asm volatile("cp.async.wait_group 1;" ::: "memory");
      } else {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
      }
      __syncthreads();
      const double* xs = sm + (size_t)(seq & 1) * 2 * GK * GS;
      const double* ys = xs + GK * GS;
#pragma unroll
      for (int k0 = 0; k0 < GK; k0 += 4) {
        double a[4], b[4];
#pragma unroll
        for (int mi = 0; mi < 4; ++mi) a[mi] = xs[(k0 + kl) * GS + r0 + 8 * mi];
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) b[ni] = ys[(k0 + kl) * GS + c0 + 8 * ni];
#pragma unroll
        for (int mi = 0; mi < 4; ++mi)
#pragma unroll
          for (int ni = 0; ni < 4; ++ni) dmma_m8n8k4(acc[mi][ni][0], acc[mi][ni][1], a[mi], b[ni]);
      }
      __syncthreads();                                   // the stage is refilled by the next iteration's load
    }
    // epilogue: lane holds D[row][col], D[row][col + 1], row = lane / 4, col = 2 (lane % 4) of each 8 x 8 tile
    double* C = MODE == 0 ? S + (size_t)tj * DT * ld + (size_t)ti * DT : S + (size_t)k * DT * ld + (size_t)ti * DT;
#pragma unroll
    for (int mi = 0; mi < 4; ++mi) {
#pragma unroll
      for (int ni = 0; ni < 4; ++ni) {
        const int row = (warp & 3) * 32 + 8 * mi + (lane >> 2);
        const int col = (warp >> 2) * 32 + 8 * ni + 2 * (lane & 3);
        double* p0 = C + (size_t)col * ld + row;
        double* p1 = p0 + ld;
        if (MODE == 0) {
          *p0 -= acc[mi][ni][0];
          *p1 -= acc[mi][ni][1];
        } else {
          *p0 = acc[mi][ni][0];
          *p1 = acc[mi][ni][1];
        }
      }
    }
    ti = tin;
    tj = tjn;
  }
}

// ---- blocked substitution ---------------------------------------------------------------------------
// forward step k (grid T - k): every CTA forms y_k = Linv_k rhs_k itself (128 x 128, from L2); CTA 0
// stores it, CTA b > 0 updates rhs_{k+b} -= L_{k+b,k} y_k.
__global__ void __launch_bounds__(DT) k_dense_fwd(const double* __restrict__ S, size_t ld, int k,
                                                  const double* __restrict__ linv_all, double* __restrict__ rhs,
                                                  double* __restrict__ y) {
  __shared__ double sr[DT], sy[DT];
  const int tid = threadIdx.x;
  sr[tid] = rhs[(size_t)k * DT + tid];
  __syncthreads();
  const double* Li = linv_all + (size_t)k * DT * DT;
  double a = 0.0;
#pragma unroll 8
  for (int c = 0; c < DT; ++c) a += Li[c * DT + tid] * sr[c];
  if (blockIdx.x == 0) {
    y[(size_t)k * DT + tid] = a;
    return;
  }
  sy[tid] = a;
  __syncthreads();
  const int ti = k + blockIdx.x;
  const double* L = S + (size_t)k * DT * ld + (size_t)ti * DT;
  double u = 0.0;
#pragma unroll 8
  for (int c = 0; c < DT; ++c) u += L[(size_t)c * ld + tid] * sy[c];
  rhs[(size_t)ti * DT + tid] -= u;
}

// backward step k (grid k + 1): x_k = Linv_k^T y_k in every CTA; CTA 0 stores it, CTA b > 0 updates
// y_{b-1} -= L_{k,b-1}^T x_k.  A warp owns 32 columns; lanes run down the rows (coalesced) and the
// column sums are reduced with shuffles.
__global__ void __launch_bounds__(DT) k_dense_bwd(const double* __restrict__ S, size_t ld, int k,
                                                  const double* __restrict__ linv_all, double* __restrict__ y,
                                                  double* __restrict__ x) {
  __shared__ double sy[DT], sx[DT];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  sy[tid] = y[(size_t)k * DT + tid];
  __syncthreads();
  const double* Li = linv_all + (size_t)k * DT * DT;
  for (int c = warp * 32; c < warp * 32 + 32; ++c) {
    double a = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) a += Li[c * DT + lane + 32 * i] * sy[lane + 32 * i];
    a = warp_sum(a);
    if (lane == 0) sx[c] = a;
  }
  __syncthreads();
  if (blockIdx.x == 0) {
    x[(size_t)k * DT + tid] = sx[tid];
    return;
  }
  const int tj = blockIdx.x - 1;
  const double* L = S + (size_t)tj * DT * ld + (size_t)k * DT;
  for (int c = warp * 32; c < warp * 32 + 32; ++c) {
    double a = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) a += L[(size_t)c * ld + lane + 32 * i] * sx[lane + 32 * i];
    a = warp_sum(a);
    if (lane == 0) y[(size_t)tj * DT + c] -= a;
  }
}

__global__ void k_dense_take(const double* __restrict__ x, int n, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = x[i];
}

// ---- host side ----------------------------------------------------------------------------------------
size_t dense_padded_dim(int V) { return ((size_t)7 * V + DT - 1) / DT * DT; }

void launch_dense_assemble(const BaDev& d, const DenseWork& w, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc) {
  cudaMemsetAsync(w.S, 0, w.ld * (size_t)w.npad * sizeof(double), st);
  if (d.N > 0) {
    k_dense_offdiag<<<(d.N + DA_WARPS - 1) / DA_WARPS, 32 * DA_WARPS, 0, st>>>(d, w.S, w.ld, dvdc_quirk ? 1 : 0);
    lc.launches++;
  }
}

void launch_dense_factor_solve(const BaDev& d, const DenseWork& w, DenseStreams& ds, cudaStream_t st,
                               LaunchCounter& lc) {
  // (the opt-in shared memory size is per function and per device; setting it again is cheap next to a factorisation)
  cudaFuncSetAttribute(k_dense_potrf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)POTRF_SMEM);
  cudaFuncSetAttribute(k_dense_gemm<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GEMM_SMEM);
  cudaFuncSetAttribute(k_dense_gemm<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GEMM_SMEM);
  const int T = w.npad / DT;
  double* rhs = w.vec;
  double* y = w.vec + w.npad;
  double* x = w.vec + 2 * (size_t)w.npad;
  const int cover = std::max(d.V * 49, w.npad);
  k_dense_diag<<<(cover + 255) / 256, 256, 0, st>>>(d, w.S, w.ld, w.n, w.npad, rhs);
  lc.launches++;
  // Right-looking factorisation with one panel of look-ahead: the diagonal factor and the panel solve of
  // panel k + 1 only need column k + 1 of update k, so the rest of update k (columns >= k + 2: the bulk of the
  // flops) stays on the caller's stream while the critical path (column update, diagonal factor, panel solve)
  // runs on an internal high-priority stream -- its few CTAs must not queue behind the thousands of the bulk
  // update.  Without the look-ahead the one-CTA diagonal factor (0.2 ms) sits between any two updates.
  if (!ds.crit) {
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    cudaStreamCreateWithPriority(&ds.crit, cudaStreamNonBlocking, hi);
  }
  while ((int)ds.ev_panel.size() < T + 1) {
    cudaEvent_t e1, e2;
    cudaEventCreateWithFlags(&e1, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&e2, cudaEventDisableTiming);
    ds.ev_panel.push_back(e1);
    ds.ev_upd.push_back(e2);
  }
  const std::vector<cudaEvent_t>& ev_panel_ = ds.ev_panel;
  const std::vector<cudaEvent_t>& ev_upd_ = ds.ev_upd;
  cudaStream_t sc = ds.crit;
  // (the critical stream starts after everything queued on st so far: the assembly and the diagonal blocks)
  cudaEventRecord(ev_panel_[T], st);
  cudaStreamWaitEvent(sc, ev_panel_[T], 0);
  for (int k = 0; k < T; ++k) {
    const int R = T - k - 1;
    if (k > 0) {
      // column k of update k - 1, after the bulk stream has finished update k - 2 on it
      if (k > 1) cudaStreamWaitEvent(sc, ev_upd_[k - 2], 0);
      k_dense_gemm<0><<<(unsigned)(R + 1), 512, GEMM_SMEM, sc>>>(w.S, w.ld, k - 1, w.linv, 1, R + 1);
      lc.launches++;
    }
    k_dense_potrf<<<1, 1024, POTRF_SMEM, sc>>>(w.S, w.ld, k, w.linv, d.sc);
    lc.launches++;
    if (R > 0) {
      k_dense_gemm<1><<<(unsigned)R, 512, GEMM_SMEM, sc>>>(w.S, w.ld, k, w.linv, 0, R);
      lc.launches++;
      cudaEventRecord(ev_panel_[k], sc);
      cudaStreamWaitEvent(st, ev_panel_[k], 0);
      if (R > 1) {
        // columns >= k + 2 of update k (ordered after update k - 1 on the same stream)
        const int nt = (int)((long long)(R - 1) * R / 2);
        // (one tile per CTA: CTAs that hold their SM for the whole update -- grid = SM count -- leave the critical
        // stream nothing to run on and were not faster, 56.2 ms against 54.9 ms per factorisation at Venice)
        k_dense_gemm<0><<<(unsigned)nt, 512, GEMM_SMEM, st>>>(w.S, w.ld, k, w.linv, 2, nt);
        lc.launches++;
      }
      cudaEventRecord(ev_upd_[k], st);
    } else {
      cudaEventRecord(ev_panel_[k], sc);
      cudaStreamWaitEvent(st, ev_panel_[k], 0);      // the last diagonal factor joins the caller's stream
    }
  }
  for (int k = 0; k < T; ++k) k_dense_fwd<<<T - k, DT, 0, st>>>(w.S, w.ld, k, w.linv, rhs, y);
  for (int k = T - 1; k >= 0; --k) k_dense_bwd<<<k + 1, DT, 0, st>>>(w.S, w.ld, k, w.linv, y, x);
  k_dense_take<<<(w.n + 255) / 256, 256, 0, st>>>(x, w.n, d.x);
  lc.launches += 2 * T + 1;
}

}  // namespace sfmba
