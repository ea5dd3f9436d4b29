// Reduced-camera-system solve: implicit-Schur matvec (K3), block-Jacobi PCG, back-substitution and
// parameter update (K4).
//
// Reference: ba_processor.py:382 delta_p = inv(A - B D^-1 B^T)(e_p - B D^-1 e_x); 383-392 p += delta_p,
// quaternion renormalised; 405-406 delta_x = D^-1 (e_x - B^T delta_p), x += delta_x.
//
// The matvec q = S p = A p - sum_j W_j D_j^-1 W_j^T p never forms S or B:
//   pass A (point-major tiles): a_o = J_p,o p_cam(o);  t_j = sum_o J_x,o^T a_o;  u_j = D_j^-1 t_j;
//                               b_o = J_x,o u_j  -> scattered (16 B) to the camera-major slot
//   pass B (camera-major chunks, one camera each): y_chunk = sum_o J_p,o^T b_o   (f64, fixed order)
//   finalize: svec_i = sum of the camera's chunks;   step kernel: q = A p - svec
// J values are the stored f32 ones; every product and sum is f64, so the operator is exactly linear
// and symmetric and PCG converges to f64 round-off.
#include "ba_internal.h"
#include "sfmba_device.cuh"

namespace sfmba {

struct MatvecSmem {
  float jx[6][TILE_CAP];
  double a[2][TILE_CAP];
  double red[32];
  double pt[4];
};

template <bool BACKSUB>
__global__ void __launch_bounds__(TILE_THREADS) k_matvec_a(BaDev d, const double* __restrict__ xvec, int force_slow) {
  __shared__ MatvecSmem s;
  if (!BACKSUB && d.sc->pcg_done) return;
  const int t = blockIdx.x;
  const int j0 = d.tile_pt[t], j1 = d.tile_pt[t + 1];
  const int o0 = d.rowptr[j0], o1 = d.rowptr[j1];
  const int L = o1 - o0;
  const int tid = threadIdx.x;
  const size_t mp = (size_t)d.mp;

  if (L <= TILE_CAP && !force_slow) {
    for (int sidx = tid; sidx < L; sidx += TILE_THREADS) {
      const int o = o0 + sidx;
      const double* xv = xvec + 7 * (size_t)d.obs_cam[o];
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int c = 0; c < 7; ++c) {
        const double xc = xv[c];
        a0 += (double)d.jp_pm[c * mp + o] * xc;
        a1 += (double)d.jp_pm[(7 + c) * mp + o] * xc;
      }
      s.a[0][sidx] = a0;
      s.a[1][sidx] = a1;
#pragma unroll
      for (int e = 0; e < 6; ++e) s.jx[e][sidx] = d.jx_pm[e * mp + o];
    }
    __syncthreads();
    for (int j = j0 + tid; j < j1; j += TILE_THREADS) {
      const int a = d.rowptr[j] - o0, e = d.rowptr[j + 1] - o0;
      double tv[3] = {0, 0, 0};
      for (int k = a; k < e; ++k) {
        const double a0 = s.a[0][k], a1 = s.a[1][k];
        tv[0] += (double)s.jx[0][k] * a0 + (double)s.jx[3][k] * a1;
        tv[1] += (double)s.jx[1][k] * a0 + (double)s.jx[4][k] * a1;
        tv[2] += (double)s.jx[2][k] * a0 + (double)s.jx[5][k] * a1;
      }
      double Dinv[6], u[3];
#pragma unroll
      for (int q = 0; q < 6; ++q) Dinv[q] = (double)d.dinv[(size_t)q * d.N + j];
      sym3_mul(Dinv, tv, u);
      if (BACKSUB) {
        double* p = d.pts + 3 * (size_t)j;
#pragma unroll
        for (int q = 0; q < 3; ++q) p[q] += (double)d.vj[(size_t)q * d.N + j] - u[q];
      } else {
        for (int k = a; k < e; ++k) {
          const double b0 = (double)s.jx[0][k] * u[0] + (double)s.jx[1][k] * u[1] + (double)s.jx[2][k] * u[2];
          const double b1 = (double)s.jx[3][k] * u[0] + (double)s.jx[4][k] * u[1] + (double)s.jx[5][k] * u[2];
          s.a[0][k] = b0;
          s.a[1][k] = b1;
        }
      }
    }
    if (!BACKSUB) {
      __syncthreads();
      for (int sidx = tid; sidx < L; sidx += TILE_THREADS)
        d.b_cm[d.pm2cm[o0 + sidx]] = make_double2(s.a[0][sidx], s.a[1][sidx]);
    }
  } else {
    for (int j = j0; j < j1; ++j) {
      const int a = d.rowptr[j], e = d.rowptr[j + 1];
      double tv[3] = {0, 0, 0};
      for (int o = a + tid; o < e; o += TILE_THREADS) {
        const double* xv = xvec + 7 * (size_t)d.obs_cam[o];
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int c = 0; c < 7; ++c) {
          const double xc = xv[c];
          a0 += (double)d.jp_pm[c * mp + o] * xc;
          a1 += (double)d.jp_pm[(7 + c) * mp + o] * xc;
        }
        tv[0] += (double)d.jx_pm[0 * mp + o] * a0 + (double)d.jx_pm[3 * mp + o] * a1;
        tv[1] += (double)d.jx_pm[1 * mp + o] * a0 + (double)d.jx_pm[4 * mp + o] * a1;
        tv[2] += (double)d.jx_pm[2 * mp + o] * a0 + (double)d.jx_pm[5 * mp + o] * a1;
      }
#pragma unroll
      for (int q = 0; q < 3; ++q) tv[q] = block_sum(tv[q], s.red);
      double Dinv[6], u[3];
#pragma unroll
      for (int q = 0; q < 6; ++q) Dinv[q] = (double)d.dinv[(size_t)q * d.N + j];
      sym3_mul(Dinv, tv, u);
      if (BACKSUB) {
        if (tid == 0) {
          double* p = d.pts + 3 * (size_t)j;
#pragma unroll
          for (int q = 0; q < 3; ++q) p[q] += (double)d.vj[(size_t)q * d.N + j] - u[q];
        }
      } else {
        for (int o = a + tid; o < e; o += TILE_THREADS) {
          const double b0 = (double)d.jx_pm[0 * mp + o] * u[0] + (double)d.jx_pm[1 * mp + o] * u[1] +
                            (double)d.jx_pm[2 * mp + o] * u[2];
          const double b1 = (double)d.jx_pm[3 * mp + o] * u[0] + (double)d.jx_pm[4 * mp + o] * u[1] +
                            (double)d.jx_pm[5 * mp + o] * u[2];
          d.b_cm[d.pm2cm[o]] = make_double2(b0, b1);
        }
      }
    }
  }
}

// pass B: one chunk (one camera) per CTA, one lane per observation
__global__ void __launch_bounds__(256) k_matvec_b(BaDev d) {
  __shared__ double s_w[8][8];
  if (d.sc->pcg_done) return;
  const int c = blockIdx.x;
  if (c >= d.sc->num_chunks) return;
  const int cam = d.chunk_cam[c];
  const int beg = d.chunk_beg[c];
  int end = beg + CHUNK_OBS;
  const int cam_end = d.cam_rowptr[cam + 1];
  if (end > cam_end) end = cam_end;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double y[7] = {0, 0, 0, 0, 0, 0, 0};
  const int slot = beg + tid;
  if (slot < end) {
    const float4* rec = d.rec_cm + 4 * (size_t)slot;
    const float4 a = rec[0], b = rec[1], cc = rec[2], dd = rec[3];
    const double2 bb = d.b_cm[slot];
    y[0] = (double)a.x * bb.x + (double)b.w * bb.y;
    y[1] = (double)a.y * bb.x + (double)cc.x * bb.y;
    y[2] = (double)a.z * bb.x + (double)cc.y * bb.y;
    y[3] = (double)a.w * bb.x + (double)cc.z * bb.y;
    y[4] = (double)b.x * bb.x + (double)cc.w * bb.y;
    y[5] = (double)b.y * bb.x + (double)dd.x * bb.y;
    y[6] = (double)b.z * bb.x + (double)dd.y * bb.y;
  }
#pragma unroll
  for (int q = 0; q < 7; ++q) y[q] = warp_sum(y[q]);
  if (lane == 0) {
#pragma unroll
    for (int q = 0; q < 7; ++q) s_w[warp][q] = y[q];
  }
  __syncthreads();
  if (tid < 7) {
    double a = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += s_w[w][tid];
    d.chunk_y[(size_t)c * 8 + tid] = a;
  }
}

__global__ void k_matvec_finalize(BaDev d) {
  if (d.sc->pcg_done) return;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 7 * d.V) return;
  const int cam = idx / 7, e = idx % 7;
  double a = 0.0;
  const int c0 = d.cam_chunk_ptr[cam], c1 = d.cam_chunk_ptr[cam + 1];
  for (int c = c0; c < c1; ++c) a += d.chunk_y[(size_t)c * 8 + e];
  d.svec[idx] = a;
}

// z = Minv r for element k (block product over the camera's 7 entries)
__device__ __forceinline__ double blk7_row(const double* __restrict__ blk, const double* __restrict__ v, int k) {
  const int cam = k / 7, a = k % 7;
  const double* row = blk + (size_t)cam * 49 + a * 7;
  const double* vv = v + 7 * (size_t)cam;
  double sacc = 0.0;
#pragma unroll
  for (int b = 0; b < 7; ++b) sacc += row[b] * vv[b];
  return sacc;
}

// x = 0, r = g, z = Minv g, p = z, rz = r.z, gg = g.g          (single CTA)
__global__ void __launch_bounds__(1024) k_pcg_init(BaDev d, unsigned epoch) {
  __shared__ double red[32];
  const int n = 7 * d.V;
  double rz = 0.0, gg = 0.0;
  for (int k = threadIdx.x; k < n; k += blockDim.x) {
    const double g = d.gvec[k];
    const double z = blk7_row(d.minv, d.gvec, k);
    d.x[k] = 0.0;
    d.r[k] = g;
    d.z[k] = z;
    d.p[k] = z;
    rz += g * z;
    gg += g * g;
  }
  rz = block_sum(rz, red);
  gg = block_sum(gg, red);
  if (threadIdx.x == 0) {
    Scalars* sc = d.sc;
    sc->rz = rz;
    sc->gg = gg;
    sc->rr = gg;
    sc->pcg_iters = 0;
    const int done = (gg == 0.0 || !(gg == gg)) ? 1 : 0;
    sc->pcg_done = done;
    if (!(gg == gg)) atomicOr(&sc->flags, FLAG_NONFINITE);
    if (d.host_sig) *(volatile unsigned long long*)d.host_sig = ((unsigned long long)epoch << 32) | (unsigned)done;
  }
}

// one PCG iteration after the matvec: q = A p - svec; alpha; x, r; z = Minv r; beta; p   (single CTA)
__global__ void __launch_bounds__(1024) k_pcg_step(BaDev d, double tol, int max_iter, unsigned epoch) {
  __shared__ double red[32];
  Scalars* sc = d.sc;
  if (sc->pcg_done) return;
  const int n = 7 * d.V;
  const double rz = sc->rz;
  double pq = 0.0;
  for (int k = threadIdx.x; k < n; k += blockDim.x) {
    const double q = blk7_row(d.ablk, d.p, k) - d.svec[k];
    d.z[k] = q;                  // z is free here: reuse it as q
    pq += d.p[k] * q;
  }
  pq = block_sum(pq, red);
  const double alpha = rz / pq;
  double rr = 0.0;
  for (int k = threadIdx.x; k < n; k += blockDim.x) {
    d.x[k] += alpha * d.p[k];
    const double r = d.r[k] - alpha * d.z[k];
    d.r[k] = r;
    rr += r * r;
  }
  rr = block_sum(rr, red);      // also orders the r writes before the block products below
  double rz_new = 0.0;
  for (int k = threadIdx.x; k < n; k += blockDim.x) {
    const double z = blk7_row(d.minv, d.r, k);
    d.z[k] = z;
    rz_new += d.r[k] * z;
  }
  rz_new = block_sum(rz_new, red);
  const double beta = rz_new / rz;
  for (int k = threadIdx.x; k < n; k += blockDim.x) d.p[k] = d.z[k] + beta * d.p[k];
  if (threadIdx.x == 0) {
    sc->rz = rz_new;
    sc->rr = rr;
    sc->pq = pq;
    sc->alpha = alpha;
    sc->beta = beta;
    const int it = sc->pcg_iters + 1;
    sc->pcg_iters = it;
    sc->pcg_iters_total += 1;
    bool done = false;
    if (!(pq > 0.0) || !(rr == rr)) {
      atomicOr(&sc->flags, FLAG_NONFINITE);
      done = true;
    }
    if (rr <= tol * tol * sc->gg) done = true;
    if (!done && it >= max_iter) {
      atomicOr(&sc->flags, FLAG_PCG_MAXITER);
      done = true;
    }
    if (done) sc->pcg_done = 1;
    if (d.host_sig)
      *(volatile unsigned long long*)d.host_sig =
          ((unsigned long long)epoch << 32) | ((unsigned long long)(unsigned)it << 1) | (done ? 1ull : 0ull);
  }
}

// cams += delta_p; q /= |q|                      (ba_processor.py:383-392)
__global__ void k_update_cams(BaDev d) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d.V) return;
  double* cam = d.cams + 7 * (size_t)c;
  const double* dx = d.x + 7 * (size_t)c;
  double v[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) v[k] = cam[k] + dx[k];
  const double nrm = sqrt(v[3] * v[3] + v[4] * v[4] + v[5] * v[5] + v[6] * v[6]);
#pragma unroll
  for (int k = 3; k < 7; ++k) v[k] /= nrm;
  bool bad = false;
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    cam[k] = v[k];
    bad |= !(v[k] == v[k]);
  }
  if (bad) atomicOr(&d.sc->flags, FLAG_NONFINITE);
}

// ---- adaptive LM (extension; the reference has fixed lambda and no accept/reject) -----------------
__global__ void k_lm_backup(BaDev d) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (size_t)7 * d.V) d.cams_backup[i] = d.cams[i];
  if (i < (size_t)3 * d.N) d.pts_backup[i] = d.pts[i];
}

// predicted decrease of the damped Gauss-Newton model, camera part: delta_p . (lambda delta_p + g)
// (the point part is folded in through the Schur complement: for the exact solve, the model
// decrease is  delta^T (lambda delta + J^T r)  and J^T r restricted to points is consumed by g)
__global__ void __launch_bounds__(1024) k_lm_model_decrease(BaDev d) {
  __shared__ double red[32];
  const int n = 7 * d.V;
  const double lam = d.sc->lambda;
  double a = 0.0;
  for (int k = threadIdx.x; k < n; k += blockDim.x) a += d.x[k] * (lam * d.x[k] + d.gvec[k]);
  a = block_sum(a, red);
  if (threadIdx.x == 0) d.sc->model_decrease = a;
}

// accept / reject on device: gain ratio from the actual decrease; Nielsen's lambda update
__global__ void k_lm_decide(BaDev d, double lam_min, double lam_max) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  Scalars* sc = d.sc;
  const double cur = sc->cost_cur, nw = sc->cost_new;
  const bool accept = (nw == nw) && (nw < cur);
  if (!accept) {
    if (i < (size_t)7 * d.V) d.cams[i] = d.cams_backup[i];
    if (i < (size_t)3 * d.N) d.pts[i] = d.pts_backup[i];
  }
  if (i == 0) {
    double lam = sc->lambda;
    if (accept) {
      const double md = sc->model_decrease;
      double rho = md > 0.0 ? (cur - nw) / md : 1.0;
      double f = 1.0 - (2.0 * rho - 1.0) * (2.0 * rho - 1.0) * (2.0 * rho - 1.0);
      if (f < 1.0 / 3.0) f = 1.0 / 3.0;
      lam *= f;
      sc->nu = 2.0;
      sc->lm_accepted += 1;
      sc->lm_last_accepted = 1;
    } else {
      lam *= sc->nu;
      sc->nu *= 2.0;
      sc->lm_last_accepted = 0;
    }
    if (lam < lam_min) lam = lam_min;
    if (lam > lam_max) lam = lam_max;
    sc->lambda = lam;
  }
}

void launch_pcg_init(const BaDev& d, unsigned epoch, cudaStream_t st, LaunchCounter& lc) {
  k_pcg_init<<<1, 1024, 0, st>>>(d, epoch);
  lc.launches++;
}

void launch_matvec_a(const BaDev& d, const double* xvec, bool backsub, bool force_slow, cudaStream_t st,
                     LaunchCounter& lc) {
  if (backsub)
    k_matvec_a<true><<<d.ntiles, TILE_THREADS, 0, st>>>(d, xvec, force_slow ? 1 : 0);
  else
    k_matvec_a<false><<<d.ntiles, TILE_THREADS, 0, st>>>(d, xvec, force_slow ? 1 : 0);
  lc.launches++;
}

void launch_matvec_b(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  if (d.nchunk_bound > 0) {
    k_matvec_b<<<d.nchunk_bound, 256, 0, st>>>(d);
    lc.launches++;
  }
}

void launch_matvec_finalize(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  const int n = 7 * d.V;
  k_matvec_finalize<<<(n + 255) / 256, 256, 0, st>>>(d);
  lc.launches++;
}

void launch_pcg_step(const BaDev& d, double tol, int max_iter, unsigned epoch, cudaStream_t st, LaunchCounter& lc) {
  k_pcg_step<<<1, 1024, 0, st>>>(d, tol, max_iter, epoch);
  lc.launches++;
}

void launch_update_cams(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  k_update_cams<<<(d.V + 127) / 128, 128, 0, st>>>(d);
  lc.launches++;
}

void launch_lm_backup(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  size_t n = (size_t)3 * d.N;
  if ((size_t)7 * d.V > n) n = (size_t)7 * d.V;
  k_lm_backup<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d);
  lc.launches++;
}

void launch_lm_model_decrease(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  k_lm_model_decrease<<<1, 1024, 0, st>>>(d);
  lc.launches++;
}

void launch_lm_decide(const BaDev& d, double lam_min, double lam_max, cudaStream_t st, LaunchCounter& lc) {
  size_t n = (size_t)3 * d.N;
  if ((size_t)7 * d.V > n) n = (size_t)7 * d.V;
  k_lm_decide<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d, lam_min, lam_max);
  lc.launches++;
}

}  // namespace sfmba
