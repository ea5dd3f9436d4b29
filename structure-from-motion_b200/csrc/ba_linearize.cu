// Linearisation, point side: camera table, per-point Schur elimination (K2a: D_j, D_j^-1, v_j and the
// cost), the index-free cost pass and the stand-alone residual/Jacobian kernel (K1).
//
// Reference: ba_processor.py:304-363 (per point: gather visible views, J_x, b, f, D_j += J_x^T J_x,
// D_j += lambda I, D_j^-1).  k_point walks a group's camera-sorted records (ba_internal.h, tile.cuh):
// each lane evaluates r and J_x of its observations in f64 with the camera constants in registers,
// scatters the 9 per-observation terms of (D_j, e_x,j) to shared-memory slots at their point-major
// position, and one thread per point then sums its slots in order and solves the 3x3 block in
// closed form.  Nothing per-observation is written to HBM.
#include "ba_internal.h"
#include "sfmba_device.cuh"
#include "tile.cuh"

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

// ---- K2a: per-point Schur elimination ------------------------------------------------------------
// dynamic shared memory: Xs[3][pt_cap] | pacc[9][pt_cap] | slots[9][slot_cap] | red[32]
template <bool COST_ONLY>
__global__ void __launch_bounds__(RUN_LANES) k_point(BaDev d, int slot_cap, int pt_cap) {
  extern __shared__ __align__(16) double smem[];
  double* Xs = smem;
  double* pacc = Xs + 3 * pt_cap;
  double* slots = COST_ONLY ? pacc : pacc + 9 * pt_cap;
  double* red = COST_ONLY ? pacc : slots + 9 * slot_cap;
  const GroupRange gr = group_range(d, blockIdx.x);
  const int tid = threadIdx.x;
  stage_points(d, gr, Xs, pt_cap);
  if (!COST_ONLY) {
    for (int i = tid; i < 9 * pt_cap; i += RUN_LANES) pacc[i] = 0.0;
  }
  __syncthreads();
  double cost = 0.0;
  uint32_t flags = 0;
  for (int c = gr.c0; c < gr.c1; ++c) {
    const ChunkInfo ci = d.chunk_info[c];
    const int K = ci.K;
    const uint32_t* ixp = d.rec_ix + (size_t)c * CHUNK_CAP + tid;
    const double2* bp = d.rec_b + (size_t)c * CHUNK_CAP + tid;
    double R[9], C[3];
    uint32_t ix = ixp[0];
    double2 b = bp[0];
    if (ix != REC_INVALID) load_cam_rc(d.camtab, d.seg_cam[(size_t)c * SEG_CAP + rec_seg(ix)], R, C);   // one camera per lane
    for (int k = 0; k < K && ix != REC_INVALID; ++k) {
      const uint32_t ix_cur = ix;
      const double2 b_cur = b;
      if (k + 1 < K) {                                   // prefetch the next record
        ix = ixp[(k + 1) * RUN_LANES];
        b = bp[(k + 1) * RUN_LANES];
      } else {
        ix = REC_INVALID;
      }
      const int ptl = rec_ptl(ix_cur);
      const Proj pr = project(R, C, Xs[ptl], Xs[pt_cap + ptl], Xs[2 * pt_cap + ptl]);
      if (!pr.ok) flags |= FLAG_DEPTH_ZERO;
      const double r0 = b_cur.x - pr.u, r1 = b_cur.y - pr.v;
      cost += r0 * r0 + r1 * r1;
      if (!COST_ONLY) {
        // J_x rows (triangulation_processor.py:259-271 with the K-free projection)
        const double x0 = (R[0] - pr.u * R[2]) * pr.iw, x1 = (R[3] - pr.u * R[5]) * pr.iw,
                     x2 = (R[6] - pr.u * R[8]) * pr.iw;
        const double y0 = (R[1] - pr.v * R[2]) * pr.iw, y1 = (R[4] - pr.v * R[5]) * pr.iw,
                     y2 = (R[7] - pr.v * R[8]) * pr.iw;
        double* s = slots + rec_pos(ix_cur);
        s[0 * slot_cap] = x0 * x0 + y0 * y0;
        s[1 * slot_cap] = x0 * x1 + y0 * y1;
        s[2 * slot_cap] = x0 * x2 + y0 * y2;
        s[3 * slot_cap] = x1 * x1 + y1 * y1;
        s[4 * slot_cap] = x1 * x2 + y1 * y2;
        s[5 * slot_cap] = x2 * x2 + y2 * y2;
        s[6 * slot_cap] = x0 * r0 + y0 * r1;
        s[7 * slot_cap] = x1 * r0 + y1 * r1;
        s[8 * slot_cap] = x2 * r0 + y2 * r1;
      }
    }
    if (!COST_ONLY) {
      __syncthreads();
      for (int j = tid; j < gr.np; j += RUN_LANES) {
        int lo, hi;
        point_slots(d, gr.pt0 + j, ci, lo, hi);
        if (lo < hi) {
          double a[9];
#pragma unroll
          for (int e = 0; e < 9; ++e) a[e] = pacc[e * pt_cap + j];
          for (int s = lo; s < hi; ++s) {
#pragma unroll
            for (int e = 0; e < 9; ++e) a[e] += slots[e * slot_cap + s];
          }
#pragma unroll
          for (int e = 0; e < 9; ++e) pacc[e * pt_cap + j] = a[e];
        }
      }
      __syncthreads();
    }
  }
  if (!COST_ONLY) {
    const double lam = d.sc->lambda;
    const size_t N = (size_t)d.N;
    for (int j = tid; j < gr.np; j += RUN_LANES) {
      double D[6], ex[3], Dinv[6], v[3];
#pragma unroll
      for (int e = 0; e < 6; ++e) D[e] = pacc[e * pt_cap + j];
#pragma unroll
      for (int e = 0; e < 3; ++e) ex[e] = pacc[(6 + e) * pt_cap + j];
      D[0] += lam;                                     // ba_processor.py:359 (every point, observed or not)
      D[3] += lam;
      D[5] += lam;
      if (!sym3_inverse(D, Dinv)) flags |= FLAG_SINGULAR;
      sym3_mul(Dinv, ex, v);
      const size_t pt = (size_t)gr.pt0 + j;
#pragma unroll
      for (int e = 0; e < 6; ++e) d.dinv[e * N + pt] = Dinv[e];
#pragma unroll
      for (int e = 0; e < 3; ++e) d.vj[e * N + pt] = v[e];
    }
  }
  cost = block_sum(cost, red);
  if (tid == 0) d.grp_cost[blockIdx.x] = cost;
  if (flags) atomicOr(&d.sc->flags, flags);
}

// deterministic sum of per-block costs into sc->costs[slot] and sc->cost_cur (slot < 0: cost_new only)
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

// ---- index-free kernels behind sfmba_ba_cost / sfmba_residual_jacobian ---------------------------
__device__ __forceinline__ void load_obs(const BaDev& d, int o, int& cam, double X[3], double b[2]) {
  cam = d.obs_cam[o];
  const int j = d.obs_pt[o];
  const double* p = d.pts + 3 * (size_t)j;
  X[0] = p[0];
  X[1] = p[1];
  X[2] = p[2];
  normalise_obs(d.kinv ? d.kinv + 9 * (size_t)cam : nullptr, d.obs_uv[o], d.obs_uv[d.m + o], b);
}

// cost only: one lane per observation in point-major order, per-block partials in grp_cost
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
  if (threadIdx.x == 0) d.grp_cost[blockIdx.x] = cost;
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
  if (threadIdx.x == 0) d.grp_cost[blockIdx.x] = cost;
  if (flags) atomicOr(&d.sc->flags, flags);
}

void launch_cam_prepare(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  k_cam_prepare<<<(d.V + 127) / 128, 128, 0, st>>>(d);
  lc.launches++;
}

static size_t point_smem(const TileShape& ts, bool cost_only) {
  const size_t pt = (size_t)ts.max_group_pts, sl = (size_t)ts.max_chunk_len;
  return cost_only ? (3 * pt + 32) * sizeof(double) : (3 * pt + 9 * pt + 9 * sl + 32) * sizeof(double);
}

void launch_point(const BaDev& d, const TileShape& ts, int cost_slot, cudaStream_t st, LaunchCounter& lc) {
  const size_t smem = point_smem(ts, false);
  cudaFuncSetAttribute(k_point<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (d.G > 0) k_point<false><<<d.G, RUN_LANES, smem, st>>>(d, ts.max_chunk_len, ts.max_group_pts);
  k_sum_cost<<<1, 1024, 0, st>>>(d.grp_cost, d.G, d.sc, cost_slot, nullptr);
  lc.launches += 2;
}

void launch_cost_tiles(const BaDev& d, const TileShape& ts, double* out_cost_dev, cudaStream_t st, LaunchCounter& lc) {
  const size_t smem = point_smem(ts, true);
  cudaFuncSetAttribute(k_point<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (d.G > 0) k_point<true><<<d.G, RUN_LANES, smem, st>>>(d, ts.max_chunk_len, ts.max_group_pts);
  k_sum_cost<<<1, 1024, 0, st>>>(d.grp_cost, d.G, d.sc, -1, out_cost_dev);
  lc.launches += 2;
}

static int cost_grid(const BaDev& d) {
  int g = (d.m + 255) / 256;
  if (g > 148 * 8) g = 148 * 8;
  if (g < 1) g = 1;
  return g;
}

void launch_cost(const BaDev& d, double* out_cost_dev, cudaStream_t st, LaunchCounter& lc) {
  const int g = cost_grid(d);
  k_cost<<<g, 256, 0, st>>>(d);
  k_sum_cost<<<1, 1024, 0, st>>>(d.grp_cost, g, d.sc, -1, out_cost_dev);
  lc.launches += 2;
}

void launch_k1(const BaDev& d, bool dvdc_quirk, float* out_r, float* out_jp, float* out_jx, double* out_r64,
               double* out_cost, cudaStream_t st, LaunchCounter& lc) {
  const int g = cost_grid(d);
  k_residual_jacobian<<<g, 256, 0, st>>>(d, dvdc_quirk ? 1 : 0, out_r, out_jp, out_jx, out_r64);
  k_sum_cost<<<1, 1024, 0, st>>>(d.grp_cost, g, nullptr, -1, out_cost);
  lc.launches += 2;
}

}  // namespace sfmba
