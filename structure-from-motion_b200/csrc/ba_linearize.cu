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
// The group's points, row pointers and records are staged with cp.async; the passes read shared
// memory only.  dynamic shared memory (doubles, then 16-byte, then 4-byte items):
//   Xa[3 pt_cap] | pacc[9][pt_cap] | slots[9][slot_cap] | red[32] | bs[k_cap 128] | ixs | segc | rowp
struct PtShape {
  int slot_cap, pt_cap, seg_cap, k_cap;
};

__host__ __device__ inline size_t pt_smem_bytes(const PtShape& s, bool cost_only) {
  size_t dbl = (size_t)3 * s.pt_cap + 32 + (cost_only ? 0 : (size_t)9 * s.pt_cap + (size_t)9 * s.slot_cap);
  size_t b = dbl * 8;
  b += (size_t)16 * s.k_cap * RUN_LANES;           // bs
  b += (size_t)4 * s.k_cap * RUN_LANES;            // ixs
  b += (size_t)4 * ((s.seg_cap + 3) / 4 * 4);      // segc
  b += (size_t)4 * (s.pt_cap + 4);                 // rowp
  return b;
}

template <bool COST_ONLY>
__global__ void __launch_bounds__(RUN_LANES) k_point(BaDev d, PtShape sh) {
  extern __shared__ __align__(16) double smem[];
  const int pt_cap = sh.pt_cap, slot_cap = sh.slot_cap;
  double* Xa = smem;
  double* pacc = Xa + 3 * pt_cap;
  double* slots = COST_ONLY ? pacc : pacc + 9 * pt_cap;
  double* red = COST_ONLY ? pacc : slots + 9 * slot_cap;
  ChunkStage cs;
  cs.bs = reinterpret_cast<double2*>(red + 32);
  cs.ixs = reinterpret_cast<uint32_t*>(cs.bs + (size_t)sh.k_cap * RUN_LANES);
  cs.segc = reinterpret_cast<int32_t*>(cs.ixs + (size_t)sh.k_cap * RUN_LANES);
  cs.srun = nullptr;
  int32_t* rowp = cs.segc + (sh.seg_cap + 3) / 4 * 4;
  const int tid = threadIdx.x;
  const GroupMeta gm = d.grp_meta[blockIdx.x];
  const int np = gm.np, pt0 = gm.pt0, c0 = gm.c0, c1 = gm.c0 + gm.nc;
  {
    const double* src = d.pts + 3 * (size_t)pt0;
    for (int i = tid; i < 3 * np; i += RUN_LANES) cp_async8(Xa + i, src + i);
    if (!COST_ONLY) {
      for (int i = tid; i <= np; i += RUN_LANES) cp_async4(rowp + i, d.rowptr + pt0 + i);
      for (int i = tid; i < 9 * pt_cap; i += RUN_LANES) pacc[i] = 0.0;
    }
  }
  ChunkInfo ci;
  ci.obs0 = gm.obs0;
  ci.len = gm.len0;
  ci.nseg = gm.nseg0;
  ci.K = gm.K0;
  double cost = 0.0;
  uint32_t flags = 0;
  for (int c = c0; c < c1; ++c) {
    if (c > c0) ci = d.chunk_info[c];
    stage_chunk_async(d, c, ci, cs);
    cp_async_wait_all();
    __syncthreads();
    const int K = ci.K;
    double R[9], C[3];
    const uint32_t ix_first = cs.ixs[tid];
    if (ix_first != REC_INVALID) load_cam_rc(d.camtab, cs.segc[rec_seg(ix_first)], R, C);   // one camera per lane
    for (int k = 0; k < K; ++k) {
      const uint32_t ix = cs.ixs[k * RUN_LANES + tid];
      if (ix == REC_INVALID) break;
      const double2 b = cs.bs[k * RUN_LANES + tid];
      const int ptl = rec_ptl(ix);
      const Proj pr = project(R, C, Xa[3 * ptl], Xa[3 * ptl + 1], Xa[3 * ptl + 2]);
      if (!pr.ok) flags |= FLAG_DEPTH_ZERO;
      const double r0 = b.x - pr.u, r1 = b.y - pr.v;
      cost += r0 * r0 + r1 * r1;
      if (!COST_ONLY) {
        // J_x rows (triangulation_processor.py:259-271 with the K-free projection)
        const double x0 = (R[0] - pr.u * R[2]) * pr.iw, x1 = (R[3] - pr.u * R[5]) * pr.iw,
                     x2 = (R[6] - pr.u * R[8]) * pr.iw;
        const double y0 = (R[1] - pr.v * R[2]) * pr.iw, y1 = (R[4] - pr.v * R[5]) * pr.iw,
                     y2 = (R[7] - pr.v * R[8]) * pr.iw;
        double* s = slots + rec_pos(ix);
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
    __syncthreads();
    if (!COST_ONLY) {
      for (int j = tid; j < np; j += RUN_LANES) {
        const int lo = max(rowp[j], ci.obs0) - ci.obs0;
        const int hi = min(rowp[j + 1], ci.obs0 + ci.len) - ci.obs0;
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
    for (int j = tid; j < np; j += RUN_LANES) {
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
      const size_t pt = (size_t)pt0 + j;
#pragma unroll
      for (int e = 0; e < 6; ++e) d.dinv[e * N + pt] = Dinv[e];
#pragma unroll
      for (int e = 0; e < 3; ++e) d.vj[e * N + pt] = v[e];
    }
  }
  cp_async_wait_all();
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

void launch_sum_cost(const BaDev& d, int cost_slot, cudaStream_t st, LaunchCounter& lc) {
  k_sum_cost<<<1, 1024, 0, st>>>(d.grp_cost, d.G, d.sc, cost_slot, nullptr);
  lc.launches++;
}

static PtShape point_shape(const TileShape& ts) {
  PtShape sh;
  sh.slot_cap = (ts.max_chunk_len + 1) & ~1;
  sh.pt_cap = (ts.max_group_pts + 1) & ~1;
  sh.seg_cap = ts.max_nseg;
  sh.k_cap = ts.max_k;
  return sh;
}

void launch_point(const BaDev& d, const TileShape& ts, int cost_slot, cudaStream_t st, LaunchCounter& lc) {
  const PtShape sh = point_shape(ts);
  const size_t smem = pt_smem_bytes(sh, false);
  cudaFuncSetAttribute(k_point<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (d.G > 0) k_point<false><<<d.G, RUN_LANES, smem, st>>>(d, sh);
  k_sum_cost<<<1, 1024, 0, st>>>(d.grp_cost, d.G, d.sc, cost_slot, nullptr);
  lc.launches += 2;
}

void launch_cost_tiles(const BaDev& d, const TileShape& ts, double* out_cost_dev, cudaStream_t st, LaunchCounter& lc) {
  const PtShape sh = point_shape(ts);
  const size_t smem = pt_smem_bytes(sh, true);
  cudaFuncSetAttribute(k_point<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (d.G > 0) k_point<true><<<d.G, RUN_LANES, smem, st>>>(d, sh);
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
