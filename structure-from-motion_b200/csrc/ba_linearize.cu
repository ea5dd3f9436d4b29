// Camera table (k_cam_prepare), the cost pass over the solve's index (k_cost_tiles), the index-free
// cost pass and the stand-alone residual/Jacobian kernel (K1) behind sfmba_ba_cost /
// sfmba_residual_jacobian.  The linearisation itself (point blocks and camera Gram rows) is k_linearize
// in ba_schur.cu.
//
// Reference: ba_processor.py:323 (R from q per observation -> once per camera here), 339-349 (b, f,
// r = b - f), campose_processor.py:462-482 and triangulation_processor.py:237-271 (Jacobian blocks).
#include "ba_internal.h"
#include "sfmba_device.cuh"
#include "tile.cuh"

namespace sfmba {

// camtab row = R(q) row-major, C, qj = R->q (cam_table_row, tile.cuh)
__global__ void k_cam_prepare(BaDev d) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d.V) return;
  const double* cam = d.cams + 7 * (size_t)c;
  const double v[7] = {cam[0], cam[1], cam[2], cam[3], cam[4], cam[5], cam[6]};
  cam_table_row(d, c, v);
}

// ---- cost over the solve's index: sum ||b - f||^2 per group (ba_processor.py:346-349,376) ------------
// The group's points and records are staged with cp.async; one lane walks its K observations of one
// camera.  dynamic shared memory: Xa[3 pt_cap] | red[32] | bs[k_cap 128] | ixs | segc
struct PtShape {
  int pt_cap, seg_cap, k_cap;
};

__host__ __device__ inline size_t pt_smem_bytes(const PtShape& s) {
  size_t b = ((size_t)3 * s.pt_cap + 32) * 8;
  b += (size_t)16 * s.k_cap * RUN_LANES;           // bs
  b += (size_t)4 * s.k_cap * RUN_LANES;            // ixs
  b += (size_t)4 * ((s.seg_cap + 3) / 4 * 4);      // segc
  return b;
}

template <bool ROBUST>
__global__ void __launch_bounds__(RUN_LANES) k_cost_tiles(BaDev d, PtShape sh) {
  extern __shared__ __align__(16) double smem[];
  double* Xa = smem;
  double* red = Xa + 3 * sh.pt_cap;
  ChunkStage cs;
  cs.bs = reinterpret_cast<double2*>(red + 32);
  cs.ixs = reinterpret_cast<uint32_t*>(cs.bs + (size_t)sh.k_cap * RUN_LANES);
  cs.segc = reinterpret_cast<int32_t*>(cs.ixs + (size_t)sh.k_cap * RUN_LANES);
  cs.srun = nullptr;
  cs.rpos = nullptr;
  const int tid = threadIdx.x;
  const GroupMeta gm = d.grp_meta[blockIdx.x];
  const int np = gm.np, c0 = gm.c0, c1 = gm.c0 + gm.nc;
  {
    const double* src = d.pts + 3 * (size_t)gm.pt0;
    for (int i = tid; i < 3 * np; i += RUN_LANES) cp_async8(Xa + i, src + i);
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
      if (ROBUST) {
        double rho;
        huber_weight(r0 * r0 + r1 * r1, d.huber_delta, rho);
        cost += rho;
      } else {
        cost += r0 * r0 + r1 * r1;
      }
    }
    __syncthreads();
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
// false (nothing loaded) when the observation's camera or point index lies outside the arrays
__device__ __forceinline__ bool load_obs(const BaDev& d, int o, int& cam, double X[3], double b[2]) {
  cam = d.obs_cam[o];
  const int j = d.obs_pt[o];
  if ((uint32_t)cam >= (uint32_t)d.V || (uint32_t)j >= (uint32_t)d.N) return false;
  const double* p = d.pts + 3 * (size_t)j;
  X[0] = p[0];
  X[1] = p[1];
  X[2] = p[2];
  normalise_obs(d.kinv ? d.kinv + 9 * (size_t)cam : nullptr, d.obs_uv[o], d.obs_uv[d.m + o], b);
  return true;
}

// cost only: one lane per observation in point-major order, per-block partials in grp_cost
__global__ void __launch_bounds__(256) k_cost(BaDev d) {
  __shared__ double red[32];
  double cost = 0.0;
  bool bad = false;
  for (int o = blockIdx.x * blockDim.x + threadIdx.x; o < d.m; o += gridDim.x * blockDim.x) {
    int cam;
    double X[3], b[2], r[2];
    if (!load_obs(d, o, cam, X, b)) {
      bad = true;
      continue;
    }
    eval_obs<false>(d.camtab + (size_t)CAMTAB_STRIDE * cam, X, b, false, r, nullptr, nullptr);
    cost += r[0] * r[0] + r[1] * r[1];
  }
  cost = block_sum(cost, red);
  if (threadIdx.x == 0) d.grp_cost[blockIdx.x] = cost;
  if (bad) atomicOr(&d.sc->flags, FLAG_BAD_INDEX);
}

// K1 alone (sfmba_residual_jacobian): SoA outputs, f32 (and optional f64 residuals)
__global__ void __launch_bounds__(256) k_residual_jacobian(BaDev d, int dvdc_quirk, float* __restrict__ out_r,
                                                           float* __restrict__ out_jp, float* __restrict__ out_jx,
                                                           double* __restrict__ out_r64, double* __restrict__ out_jp64,
                                                           double* __restrict__ out_jx64) {
  __shared__ double red[32];
  double cost = 0.0;
  uint32_t flags = 0;
  const size_t m = (size_t)d.m;
  for (int o = blockIdx.x * blockDim.x + threadIdx.x; o < d.m; o += gridDim.x * blockDim.x) {
    int cam;
    double X[3], b[2], r[2], Jp[14], Jx[6];
    if (!load_obs(d, o, cam, X, b)) {               // outputs of this observation stay as they were
      flags |= FLAG_BAD_INDEX;
      continue;
    }
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
    if (out_jp64) {
#pragma unroll
      for (int e = 0; e < 14; ++e) out_jp64[e * m + o] = Jp[e];
    }
    if (out_jx64) {
#pragma unroll
      for (int e = 0; e < 6; ++e) out_jx64[e * m + o] = Jx[e];
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

void launch_cost_tiles(const BaDev& d, const TileShape& ts, double* out_cost_dev, cudaStream_t st, LaunchCounter& lc) {
  PtShape sh;
  sh.pt_cap = (ts.max_group_pts + 1) & ~1;
  sh.seg_cap = ts.max_nseg;
  sh.k_cap = ts.max_k;
  const size_t smem = pt_smem_bytes(sh);
  if (d.huber_delta > 0.0) {
    cudaFuncSetAttribute(k_cost_tiles<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (d.G > 0) k_cost_tiles<true><<<d.G, RUN_LANES, smem, st>>>(d, sh);
  } else {
    cudaFuncSetAttribute(k_cost_tiles<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (d.G > 0) k_cost_tiles<false><<<d.G, RUN_LANES, smem, st>>>(d, sh);
  }
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
               double* out_jp64, double* out_jx64, double* out_cost, cudaStream_t st, LaunchCounter& lc) {
  const int g = cost_grid(d);
  k_residual_jacobian<<<g, 256, 0, st>>>(d, dvdc_quirk ? 1 : 0, out_r, out_jp, out_jx, out_r64, out_jp64, out_jx64);
  k_sum_cost<<<1, 1024, 0, st>>>(d.grp_cost, g, nullptr, -1, out_cost);
  lc.launches += 2;
}

}  // namespace sfmba
