// Linearisation: camera table, fused residual/Jacobian + per-point Schur elimination (K1+K2a), the
// stand-alone residual/Jacobian kernel (K1) and the cost pass.
//
// Reference: ba_processor.py:304-363 (per point: gather visible views, J_p, J_x, b, f, D_j += J_x^T J_x,
// D_j += lambda I, D_j^-1).  Here one CTA owns a point-major tile of ~256 observations:
//   P1  one lane per observation: r (f64), J_p, J_x; J written point-major (SoA, coalesced)
//   P2  one thread per point: D_j, e_x,j in f64 from shared memory, closed-form 3x3 inverse,
//       v_j = D^-1 e_x; per observation rho = r - J_x v_j and Y = J_x D^-1 J_x^T
//   P3  one lane per observation: the 64-byte camera-major record [J_p | rho] and Y are scattered
//       to their camera-major slot (full-sector writes)
// Tiles whose observation range does not fit the shared-memory slots (a track longer than ~256)
// take a block-cooperative path that handles one point at a time.
#include "ba_internal.h"
#include "sfmba_device.cuh"

namespace sfmba {

// camtab row = R(q) row-major, C, qj = R->q (the Jacobian's quaternion, campose_processor.py:464)
__global__ void k_cam_prepare(BaDev d) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d.V) return;
  const double* cam = d.cams + 7 * (size_t)c;
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

struct LinSmem {
  float jx[6][TILE_CAP];
  double r[2][TILE_CAP];
  float out[5][TILE_CAP];    // rho0 rho1 yuu yuv yvv
  float tail[2][TILE_CAP];   // Jp[12], Jp[13] (last float4 of the camera-major record)
  double red[32];
  double pt[16];             // slow path broadcast
};

__device__ __forceinline__ void load_obs(const BaDev& d, int o, int& cam, double X[3], double b[2]) {
  cam = d.obs_cam[o];
  const int j = d.obs_pt[o];
  const double* p = d.pts + 3 * (size_t)j;
  X[0] = p[0];
  X[1] = p[1];
  X[2] = p[2];
  normalise_obs(d.kinv ? d.kinv + 9 * (size_t)cam : nullptr, d.obs_uv[o], d.obs_uv[d.m + o], b);
}

__device__ __forceinline__ void store_pm(const BaDev& d, int o, const double Jp[14], const double Jx[6]) {
#pragma unroll
  for (int e = 0; e < 14; ++e) d.jp_pm[(size_t)e * d.mp + o] = (float)Jp[e];
#pragma unroll
  for (int e = 0; e < 6; ++e) d.jx_pm[(size_t)e * d.mp + o] = (float)Jx[e];
}

// per-point epilogue shared by both paths: D (with lambda) -> Dinv, v; writes dinv/vj
__device__ __forceinline__ bool point_solve(const BaDev& d, int j, double D[6], const double ex[3], double lam,
                                            double Dinv[6], double v[3]) {
  D[0] += lam;
  D[3] += lam;
  D[5] += lam;
  const bool ok = sym3_inverse(D, Dinv);
  // the stored (f32) inverse is the one every later kernel uses, so v and rho use it as well
#pragma unroll
  for (int e = 0; e < 6; ++e) {
    const float f = (float)Dinv[e];
    d.dinv[(size_t)e * d.N + j] = f;
    Dinv[e] = (double)f;
  }
  sym3_mul(Dinv, ex, v);
#pragma unroll
  for (int e = 0; e < 3; ++e) {
    const float f = (float)v[e];
    d.vj[(size_t)e * d.N + j] = f;
    v[e] = (double)f;
  }
  return ok;
}

// rho = r - Jx v ; Y = Jx Dinv Jx^T (Jx taken as the stored f32 values)
__device__ __forceinline__ void obs_schur_terms(const double jx[6], const double r[2], const double Dinv[6],
                                                const double v[3], float out[5]) {
  out[0] = (float)(r[0] - (jx[0] * v[0] + jx[1] * v[1] + jx[2] * v[2]));
  out[1] = (float)(r[1] - (jx[3] * v[0] + jx[4] * v[1] + jx[5] * v[2]));
  double t0[3], t1[3];
  sym3_mul(Dinv, jx, t0);
  sym3_mul(Dinv, jx + 3, t1);
  out[2] = (float)(jx[0] * t0[0] + jx[1] * t0[1] + jx[2] * t0[2]);
  out[3] = (float)(jx[0] * t1[0] + jx[1] * t1[1] + jx[2] * t1[2]);
  out[4] = (float)(jx[3] * t1[0] + jx[4] * t1[1] + jx[5] * t1[2]);
}

__device__ __forceinline__ void store_cm(const BaDev& d, int o, const float jp[14], const float out[5]) {
  const int slot = d.pm2cm[o];
  float4* rec = d.rec_cm + 4 * (size_t)slot;
  rec[0] = make_float4(jp[0], jp[1], jp[2], jp[3]);
  rec[1] = make_float4(jp[4], jp[5], jp[6], jp[7]);
  rec[2] = make_float4(jp[8], jp[9], jp[10], jp[11]);
  rec[3] = make_float4(jp[12], jp[13], out[0], out[1]);
  d.y_cm[slot] = make_float4(out[2], out[3], out[4], 0.f);
}

__global__ void __launch_bounds__(TILE_THREADS) k_linearize(BaDev d, int dvdc_quirk, int force_slow) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  LinSmem& s = *reinterpret_cast<LinSmem*>(smem_raw);
  const int t = blockIdx.x;
  const int j0 = d.tile_pt[t], j1 = d.tile_pt[t + 1];
  const int o0 = d.rowptr[j0], o1 = d.rowptr[j1];
  const int L = o1 - o0;
  const double lam = d.sc->lambda;
  const int tid = threadIdx.x;
  double cost = 0.0;
  uint32_t flags = 0;

  if (L <= TILE_CAP && !force_slow) {
    // ---- P1: one lane per observation; first three float4 of the camera-major record go out now
    for (int sidx = tid; sidx < L; sidx += TILE_THREADS) {
      const int o = o0 + sidx;
      int cam;
      double X[3], b[2], r[2], Jp[14], Jx[6];
      load_obs(d, o, cam, X, b);
      if (!eval_obs<true>(d.camtab + (size_t)CAMTAB_STRIDE * cam, X, b, dvdc_quirk != 0, r, Jp, Jx))
        flags |= FLAG_DEPTH_ZERO;
      store_pm(d, o, Jp, Jx);
      cost += r[0] * r[0] + r[1] * r[1];
#pragma unroll
      for (int e = 0; e < 6; ++e) s.jx[e][sidx] = (float)Jx[e];
      s.r[0][sidx] = r[0];
      s.r[1][sidx] = r[1];
      s.tail[0][sidx] = (float)Jp[12];
      s.tail[1][sidx] = (float)Jp[13];
      float4* rec = d.rec_cm + 4 * (size_t)d.pm2cm[o];
      rec[0] = make_float4((float)Jp[0], (float)Jp[1], (float)Jp[2], (float)Jp[3]);
      rec[1] = make_float4((float)Jp[4], (float)Jp[5], (float)Jp[6], (float)Jp[7]);
      rec[2] = make_float4((float)Jp[8], (float)Jp[9], (float)Jp[10], (float)Jp[11]);
    }
    __syncthreads();
    // ---- P2: one thread per point
    for (int j = j0 + tid; j < j1; j += TILE_THREADS) {
      const int a = d.rowptr[j] - o0, e = d.rowptr[j + 1] - o0;
      double D[6] = {0, 0, 0, 0, 0, 0}, ex[3] = {0, 0, 0};
      for (int k = a; k < e; ++k) {
        const double x0 = s.jx[0][k], x1 = s.jx[1][k], x2 = s.jx[2][k];
        const double y0 = s.jx[3][k], y1 = s.jx[4][k], y2 = s.jx[5][k];
        const double r0 = s.r[0][k], r1 = s.r[1][k];
        D[0] += x0 * x0 + y0 * y0;
        D[1] += x0 * x1 + y0 * y1;
        D[2] += x0 * x2 + y0 * y2;
        D[3] += x1 * x1 + y1 * y1;
        D[4] += x1 * x2 + y1 * y2;
        D[5] += x2 * x2 + y2 * y2;
        ex[0] += x0 * r0 + y0 * r1;
        ex[1] += x1 * r0 + y1 * r1;
        ex[2] += x2 * r0 + y2 * r1;
      }
      double Dinv[6], v[3];
      if (!point_solve(d, j, D, ex, lam, Dinv, v)) flags |= FLAG_SINGULAR;
      for (int k = a; k < e; ++k) {
        const double jx[6] = {s.jx[0][k], s.jx[1][k], s.jx[2][k], s.jx[3][k], s.jx[4][k], s.jx[5][k]};
        const double r[2] = {s.r[0][k], s.r[1][k]};
        float out[5];
        obs_schur_terms(jx, r, Dinv, v, out);
#pragma unroll
        for (int q = 0; q < 5; ++q) s.out[q][k] = out[q];
      }
    }
    __syncthreads();
    // ---- P3: finish the camera-major record
    for (int sidx = tid; sidx < L; sidx += TILE_THREADS) {
      const int slot = d.pm2cm[o0 + sidx];
      d.rec_cm[4 * (size_t)slot + 3] = make_float4(s.tail[0][sidx], s.tail[1][sidx], s.out[0][sidx], s.out[1][sidx]);
      d.y_cm[slot] = make_float4(s.out[2][sidx], s.out[3][sidx], s.out[4][sidx], 0.f);
    }
  } else {
    // ---- long-track path: the block handles one point at a time
    for (int j = j0; j < j1; ++j) {
      const int a = d.rowptr[j], e = d.rowptr[j + 1];
      double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int o = a + tid; o < e; o += TILE_THREADS) {
        int cam;
        double X[3], b[2], r[2], Jp[14], Jx[6];
        load_obs(d, o, cam, X, b);
        if (!eval_obs<true>(d.camtab + (size_t)CAMTAB_STRIDE * cam, X, b, dvdc_quirk != 0, r, Jp, Jx))
          flags |= FLAG_DEPTH_ZERO;
        cost += r[0] * r[0] + r[1] * r[1];
        double jx[6];
#pragma unroll
        for (int q = 0; q < 6; ++q) jx[q] = (double)(float)Jx[q];
        acc[0] += jx[0] * jx[0] + jx[3] * jx[3];
        acc[1] += jx[0] * jx[1] + jx[3] * jx[4];
        acc[2] += jx[0] * jx[2] + jx[3] * jx[5];
        acc[3] += jx[1] * jx[1] + jx[4] * jx[4];
        acc[4] += jx[1] * jx[2] + jx[4] * jx[5];
        acc[5] += jx[2] * jx[2] + jx[5] * jx[5];
        acc[6] += jx[0] * r[0] + jx[3] * r[1];
        acc[7] += jx[1] * r[0] + jx[4] * r[1];
        acc[8] += jx[2] * r[0] + jx[5] * r[1];
      }
#pragma unroll
      for (int q = 0; q < 9; ++q) acc[q] = block_sum(acc[q], s.red);
      if (tid == 0) {
        double D[6] = {acc[0], acc[1], acc[2], acc[3], acc[4], acc[5]};
        double ex[3] = {acc[6], acc[7], acc[8]};
        double Dinv[6], v[3];
        if (!point_solve(d, j, D, ex, lam, Dinv, v)) flags |= FLAG_SINGULAR;
#pragma unroll
        for (int q = 0; q < 6; ++q) s.pt[q] = Dinv[q];
#pragma unroll
        for (int q = 0; q < 3; ++q) s.pt[6 + q] = v[q];
      }
      __syncthreads();
      double Dinv[6], v[3];
#pragma unroll
      for (int q = 0; q < 6; ++q) Dinv[q] = s.pt[q];
#pragma unroll
      for (int q = 0; q < 3; ++q) v[q] = s.pt[6 + q];
      for (int o = a + tid; o < e; o += TILE_THREADS) {
        int cam;
        double X[3], b[2], r[2], Jp[14], Jx[6];
        load_obs(d, o, cam, X, b);
        eval_obs<true>(d.camtab + (size_t)CAMTAB_STRIDE * cam, X, b, dvdc_quirk != 0, r, Jp, Jx);
        store_pm(d, o, Jp, Jx);
        double jx[6];
        float jpf[14], out[5];
#pragma unroll
        for (int q = 0; q < 6; ++q) jx[q] = (double)(float)Jx[q];
#pragma unroll
        for (int q = 0; q < 14; ++q) jpf[q] = (float)Jp[q];
        obs_schur_terms(jx, r, Dinv, v, out);
        store_cm(d, o, jpf, out);
      }
      __syncthreads();
    }
  }
  cost = block_sum(cost, s.red);
  if (tid == 0) d.tile_cost[t] = cost;
  if (flags) atomicOr(&d.sc->flags, flags);
}

// deterministic sum of per-tile costs into sc->costs[slot] and sc->cost_cur (slot < 0: cost_new only)
__global__ void k_sum_cost(const double* __restrict__ part, int n, Scalars* sc, int slot, double* out) {
  __shared__ double red[32];
  double a = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) a += part[i];
  a = block_sum(a, red);
  if (threadIdx.x == 0) {
    if (out) *out = a;
    if (sc) {
      if (slot >= 0) {
        sc->cost_cur = a;
        if (slot < SFMBA_MAX_COST_HISTORY) sc->costs[slot] = a;
      } else {
        sc->cost_new = a;
      }
      if (!(a == a) || isinf(a)) atomicOr(&sc->flags, FLAG_NONFINITE);
    }
  }
}

// cost only: one lane per observation, per-block partials (tile_cost is reused as scratch: the
// grid is capped at ntiles blocks)
__global__ void __launch_bounds__(256) k_cost(BaDev d) {
  __shared__ double red[32];
  double cost = 0.0;
  for (int o = blockIdx.x * blockDim.x + threadIdx.x; o < d.m; o += gridDim.x * blockDim.x) {
    int cam;
    double X[3], b[2], r[2];
    load_obs(d, o, cam, X, b);
    eval_obs<false>(d.camtab + (size_t)CAMTAB_STRIDE * cam, X, b, false, r, nullptr, nullptr);
    cost += r[0] * r[0] + r[1] * r[1];
  }
  cost = block_sum(cost, red);
  if (threadIdx.x == 0) d.tile_cost[blockIdx.x] = cost;
}

// K1 alone (sfmba_residual_jacobian): SoA outputs, f32 (and optional f64 residuals)
__global__ void __launch_bounds__(256) k_residual_jacobian(BaDev d, int dvdc_quirk, float* __restrict__ out_r,
                                                           float* __restrict__ out_jp, float* __restrict__ out_jx,
                                                           double* __restrict__ out_r64) {
  __shared__ double red[32];
  double cost = 0.0;
  uint32_t flags = 0;
  const size_t m = (size_t)d.m;
  for (int o = blockIdx.x * blockDim.x + threadIdx.x; o < d.m; o += gridDim.x * blockDim.x) {
    int cam;
    double X[3], b[2], r[2], Jp[14], Jx[6];
    load_obs(d, o, cam, X, b);
    if (!eval_obs<true>(d.camtab + (size_t)CAMTAB_STRIDE * cam, X, b, dvdc_quirk != 0, r, Jp, Jx))
      flags |= FLAG_DEPTH_ZERO;
    cost += r[0] * r[0] + r[1] * r[1];
    if (out_r) {
      out_r[o] = (float)r[0];
      out_r[m + o] = (float)r[1];
    }
    if (out_r64) {
      out_r64[o] = r[0];
      out_r64[m + o] = r[1];
    }
    if (out_jp) {
#pragma unroll
      for (int e = 0; e < 14; ++e) out_jp[e * m + o] = (float)Jp[e];
    }
    if (out_jx) {
#pragma unroll
      for (int e = 0; e < 6; ++e) out_jx[e * m + o] = (float)Jx[e];
    }
  }
  cost = block_sum(cost, red);
  if (threadIdx.x == 0) d.tile_cost[blockIdx.x] = cost;
  if (flags) atomicOr(&d.sc->flags, flags);
}

void launch_cam_prepare(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  k_cam_prepare<<<(d.V + 127) / 128, 128, 0, st>>>(d);
  lc.launches++;
}

void launch_linearize(const BaDev& d, bool dvdc_quirk, bool force_slow, int cost_slot, cudaStream_t st,
                      LaunchCounter& lc) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(k_linearize, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LinSmem));
    attr_set = true;
  }
  k_linearize<<<d.ntiles, TILE_THREADS, sizeof(LinSmem), st>>>(d, dvdc_quirk ? 1 : 0, force_slow ? 1 : 0);
  k_sum_cost<<<1, 1024, 0, st>>>(d.tile_cost, d.ntiles, d.sc, cost_slot, nullptr);
  lc.launches += 2;
}

static int cost_grid(const BaDev& d) {
  int g = (d.m + 255) / 256;
  if (g > d.ntiles) g = d.ntiles;
  if (g < 1) g = 1;
  return g;
}

void launch_cost(const BaDev& d, double* out_cost_dev, cudaStream_t st, LaunchCounter& lc) {
  const int g = cost_grid(d);
  k_cost<<<g, 256, 0, st>>>(d);
  k_sum_cost<<<1, 1024, 0, st>>>(d.tile_cost, g, d.sc, -1, out_cost_dev);
  lc.launches += 2;
}

void launch_k1(const BaDev& d, bool dvdc_quirk, float* out_r, float* out_jp, float* out_jx, double* out_r64,
               double* out_cost, cudaStream_t st, LaunchCounter& lc) {
  const int g = cost_grid(d);
  k_residual_jacobian<<<g, 256, 0, st>>>(d, dvdc_quirk ? 1 : 0, out_r, out_jp, out_jx, out_r64);
  k_sum_cost<<<1, 1024, 0, st>>>(d.tile_cost, g, nullptr, -1, out_cost);
  lc.launches += 2;
}

}  // namespace sfmba
