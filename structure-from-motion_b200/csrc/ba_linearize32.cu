// Linearisation with FP32 Jacobian arithmetic (BASELINE.json north_star: "FP32 with FP64 accumulation of the normal
// equations").  Same tile structure, index and outputs as k_linearize<false, false> (ba_schur.cu); what changes is the
// arithmetic type of everything that only has to be good to the accuracy of a linearisation:
//   FP64  the projection of every observation (d = X - C, p = R^T d, 1 / p_w, u, v), its residual r = b - f and the cost
//         -- the gradient J^T r is only as good as r, and r is a difference of two numbers a thousand times larger;
//   FP32  the Jacobian blocks J_x (2x3) and J_p (2x7) built from (d, u, v, 1 / p_w) and R, q rounded to FP32, every
//         product that enters D_j, e_x, A_i, g_i, and the sums of those products inside one lane's run or one point's
//         track (a dozen terms);
//   FP64  everything above: D_j + lambda I and its inverse, v_j, the segment rows (sums over a camera's lanes), the
//         per-camera sums (k_cam_finish), the reduced system, the PCG.
// The normal equations it produces differ from the FP64 kernel's by ~1e-7 relative, the same class of perturbation as the
// FP32 matvec's; sfmba_options.matvec_precision gates both.  One projection per phase, nine FP32 slot planes in one
// round, 35 FP32 accumulators per lane: about half the registers and half the shared memory of the FP64 kernel.
#include "ba_internal.h"
#include "sfmba_device.cuh"
#include "tile.cuh"

namespace sfmba {

#ifndef LIN32_MINB
#define LIN32_MINB 5     // resident CTAs per SM the register budget is compiled for
#endif

struct Lin32Shape {
  int pt_cap, seg_cap, k_cap, slot_cap, un_floats;
};
constexpr int GRAM32 = 35;          // A(28) g(7): floats per lane / doubles per segment row (ROW_COMPACT with its pad)

// dynamic shared memory: Xa f64[3 pt_cap] | red f64[32] | Vf f32[3][pt_cap] | un f32[max(9 slot_cap, RUN_LANES 35)] | ixs |
// segc | rpos | srun | rowp
__host__ __device__ inline size_t lin32_smem_bytes(const Lin32Shape& s) {
  size_t b = ((size_t)3 * s.pt_cap + 32) * 8;
  b += (size_t)4 * 3 * s.pt_cap;
  b += (size_t)4 * s.un_floats;
  b += (size_t)4 * s.k_cap * RUN_LANES;            // ixs
  b += (size_t)8 * ((s.seg_cap + 3) / 4 * 4);      // segc, rpos
  b += (size_t)2 * (SEG_RUN_STRIDE + 6);           // srun
  b += (size_t)4 * (s.pt_cap + 4);                 // rowp
  return b;
}

// FP32 Jacobian blocks of one observation from its FP64 projection
struct Obs32 {
  float u, v, iw, d0, d1, d2, r0, r1;
};
__device__ __forceinline__ Obs32 to_f32(const Proj& pr, double r0, double r1) {
  Obs32 o;
  o.u = (float)pr.u;
  o.v = (float)pr.v;
  o.iw = (float)pr.iw;
  o.d0 = (float)pr.d0;
  o.d1 = (float)pr.d1;
  o.d2 = (float)pr.d2;
  o.r0 = (float)r0;
  o.r1 = (float)r1;
  return o;
}

__global__ void __launch_bounds__(RUN_LANES, LIN32_MINB) k_linearize32(BaDev d, Lin32Shape sh, int dvdc_quirk, int g0) {
  extern __shared__ __align__(16) double smem[];
  const int pt_cap = sh.pt_cap, slot_cap = sh.slot_cap;
  double* Xa = smem;
  double* red = Xa + 3 * pt_cap;
  float* Vf = reinterpret_cast<float*>(red + 32);
  float* un = Vf + 3 * pt_cap;
  ChunkStage cs;
  cs.bs = nullptr;
  cs.ixs = reinterpret_cast<uint32_t*>(un + sh.un_floats);
  cs.segc = reinterpret_cast<int32_t*>(cs.ixs + (size_t)sh.k_cap * RUN_LANES);
  cs.rpos = cs.segc + (sh.seg_cap + 3) / 4 * 4;
  cs.srun = reinterpret_cast<uint16_t*>(cs.rpos + (sh.seg_cap + 3) / 4 * 4);
  int32_t* rowp = reinterpret_cast<int32_t*>(cs.srun + SEG_RUN_STRIDE + 6);
  const int tid = threadIdx.x;
  const int grp = g0 + blockIdx.x;
  const GroupMeta gm = d.grp_meta[grp];
  const int np = gm.np, pt0 = gm.pt0, c0 = gm.c0, c1 = gm.c0 + gm.nc;
  {
    const double* src = d.pts + 3 * (size_t)pt0;
    for (int i = tid; i < 3 * np; i += RUN_LANES) cp_async8(Xa + i, src + i);
    for (int i = tid; i <= np; i += RUN_LANES) cp_async4(rowp + i, d.rowptr + pt0 + i);
  }
  ChunkInfo ci;
  ci.obs0 = gm.obs0;
  ci.len = gm.len0;
  ci.nseg = gm.nseg0;
  ci.K = gm.K0;
  const bool single = gm.nc == 1;
  // multi-chunk groups (very long tracks, more than SEG_CAP cameras) accumulate D_j, e_x over their chunks in global
  // scratch: dinv (6 planes) and vj (3 planes) of the group's points, which the point solve below then overwrites
  const size_t N = (size_t)d.N;
  if (!single) {
    for (int j = tid; j < np; j += RUN_LANES) {
#pragma unroll
      for (int e = 0; e < 6; ++e) d.dinv[e * N + pt0 + j] = 0.0;
#pragma unroll
      for (int e = 0; e < 3; ++e) d.vj[e * N + pt0 + j] = 0.0;
    }
  }
  double cost = 0.0, exv = 0.0;
  uint32_t flags = 0;
  const double lam = d.sc->lambda;
  // the point solve of one point from its sums (FP64): D^-1, v = D^-1 e_x, outputs
  auto solve_point = [&](int j, const double a[9]) {
    double D[6] = {a[0] + lam, a[1], a[2], a[3] + lam, a[4], a[5] + lam};    // ba_processor.py:359
    const double ex[3] = {a[6], a[7], a[8]};
    double Dinv[6], v[3];
    if (!sym3_inverse(D, Dinv)) flags |= FLAG_SINGULAR;
    sym3_mul(Dinv, ex, v);
    exv += ex[0] * v[0] + ex[1] * v[1] + ex[2] * v[2];
    const size_t pt = (size_t)pt0 + j;
    if (d.xrel) {
      d.xrel[pt] = make_float4((float)(Xa[3 * j] - Xa[0]), (float)(Xa[3 * j + 1] - Xa[1]),
                               (float)(Xa[3 * j + 2] - Xa[2]), 0.f);
      float4* dg = reinterpret_cast<float4*>(d.dinv32 + 8 * pt);
      dg[0] = make_float4((float)Dinv[0], (float)Dinv[1], (float)Dinv[2], (float)Dinv[3]);
      dg[1] = make_float4((float)Dinv[4], (float)Dinv[5], 0.f, 0.f);
    }
#pragma unroll
    for (int e = 0; e < 6; ++e) d.dinv[e * N + pt] = Dinv[e];
#pragma unroll
    for (int e = 0; e < 3; ++e) {
      d.vj[e * N + pt] = v[e];
      Vf[e * pt_cap + j] = (float)v[e];
    }
  };
  // ---- phase A: point blocks.  One round: nine FP32 slot planes (D00 D01 D02 D11 D12 D22 ex0 ex1 ex2)
  for (int c = c0; c < c1; ++c) {
    if (c > c0) ci = d.chunk_info[c];
    stage_chunk_async(d, c, ci, cs);
    cp_async_wait_all();
    __syncthreads();
    const int K = ci.K;
    const double2* gb = d.rec_b + (size_t)c * CHUNK_CAP + tid;
    double R[9], C[3];
    float Rf[9];
    const uint32_t ix_first = cs.ixs[tid];
    double2 b_next = make_double2(0.0, 0.0);
    if (ix_first != REC_INVALID) {
      load_cam_rc(d.camtab, cs.segc[rec_seg(ix_first)], R, C);
      b_next = gb[0];
    }
#pragma unroll
    for (int e = 0; e < 9; ++e) Rf[e] = (float)R[e];
    for (int k = 0; k < K; ++k) {
      const uint32_t ix = cs.ixs[k * RUN_LANES + tid];
      if (ix == REC_INVALID) break;
      const double2 b = b_next;
      if (k + 1 < K) b_next = gb[(k + 1) * RUN_LANES];     // one trip ahead
      const int ptl = rec_ptl(ix);
      const Proj pr = project(R, C, Xa[3 * ptl], Xa[3 * ptl + 1], Xa[3 * ptl + 2]);
      if (!pr.ok) flags |= FLAG_DEPTH_ZERO;
      const double r0 = b.x - pr.u, r1 = b.y - pr.v;
      cost += r0 * r0 + r1 * r1;
      const Obs32 o = to_f32(pr, r0, r1);
      const float x0 = (Rf[0] - o.u * Rf[2]) * o.iw, x1 = (Rf[3] - o.u * Rf[5]) * o.iw, x2 = (Rf[6] - o.u * Rf[8]) * o.iw;
      const float y0 = (Rf[1] - o.v * Rf[2]) * o.iw, y1 = (Rf[4] - o.v * Rf[5]) * o.iw, y2 = (Rf[7] - o.v * Rf[8]) * o.iw;
      float* sl = un + rec_pos(ix);
      sl[0] = x0 * x0 + y0 * y0;
      sl[slot_cap] = x0 * x1 + y0 * y1;
      sl[2 * slot_cap] = x0 * x2 + y0 * y2;
      sl[3 * slot_cap] = x1 * x1 + y1 * y1;
      sl[4 * slot_cap] = x1 * x2 + y1 * y2;
      sl[5 * slot_cap] = x2 * x2 + y2 * y2;
      sl[6 * slot_cap] = x0 * o.r0 + y0 * o.r1;
      sl[7 * slot_cap] = x1 * o.r0 + y1 * o.r1;
      sl[8 * slot_cap] = x2 * o.r0 + y2 * o.r1;
    }
    __syncthreads();
    // one work item per (point, plane triple): the slots of a point are summed in observation order (FP32, a track's
    // worth of terms), then join the FP64 sums
    if (single) {
      // the whole track is here: one thread per point sums all nine planes and solves the point right away
      // (three work items per point that leave their sums in the point's first slot, then one solve per point, was
      // measured: 220 against 215 us -- the extra barrier costs more than the idle lanes)
      for (int j = tid; j < np; j += RUN_LANES) {
        const int lo = rowp[j] - ci.obs0, hi = rowp[j + 1] - ci.obs0;
        float a[9];
#pragma unroll
        for (int e = 0; e < 9; ++e) a[e] = 0.f;
        for (int s = lo; s < hi; ++s) {
#pragma unroll
          for (int e = 0; e < 9; ++e) a[e] += un[e * slot_cap + s];
        }
        double ad[9];
#pragma unroll
        for (int e = 0; e < 9; ++e) ad[e] = (double)a[e];
        solve_point(j, ad);
      }
    } else {
      for (int idx = tid; idx < 3 * np; idx += RUN_LANES) {
        const int j = idx / 3, e = 3 * (idx - 3 * j);
        const int lo = max(rowp[j], ci.obs0) - ci.obs0;
        const int hi = min(rowp[j + 1], ci.obs0 + ci.len) - ci.obs0;
        if (lo < hi) {
          float a0 = 0.f, a1 = 0.f, a2 = 0.f;
          for (int s = lo; s < hi; ++s) {
            a0 += un[e * slot_cap + s];
            a1 += un[(e + 1) * slot_cap + s];
            a2 += un[(e + 2) * slot_cap + s];
          }
          double* g = e < 6 ? d.dinv + e * N + pt0 + j : d.vj + (e - 6) * N + pt0 + j;
          g[0] += (double)a0;                              // (this CTA owns the group's points: no other writer)
          g[N] += (double)a1;
          g[2 * N] += (double)a2;
        }
      }
    }
    __syncthreads();
  }
  cp_async_wait_all();
  __syncthreads();
  if (!single) {
    // (also groups without any chunk: unobserved points get D = lambda I, ba_processor.py:359)
    for (int j = tid; j < np; j += RUN_LANES) {
      double ad[9];
#pragma unroll
      for (int e = 0; e < 6; ++e) ad[e] = gm.nc > 0 ? d.dinv[e * N + pt0 + j] : 0.0;
#pragma unroll
      for (int e = 0; e < 3; ++e) ad[6 + e] = gm.nc > 0 ? d.vj[e * N + pt0 + j] : 0.0;
      solve_point(j, ad);
    }
  }
  cost = block_sum(cost, red);     // (contains the barrier that publishes Vf)
  if (tid == 0) d.grp_cost[grp] = cost;
  if (d.grp_exv) {                 // adaptive LM only (uniform branch)
    exv = block_sum(exv, red);
    if (tid == 0) d.grp_exv[grp] = exv;
  }
  if (flags) atomicOr(&d.sc->flags, flags);
  // ---- phase B: camera-side Gram sums, 35 FP32 accumulators per lane
  float* runpart = un;
  for (int c = c0; c < c1; ++c) {
    if (gm.nc > 1) {                       // single-chunk groups still hold their chunk from phase A
      ci = d.chunk_info[c];
      stage_chunk_async(d, c, ci, cs);
      cp_async_wait_all();
      __syncthreads();
    }
    const int K = ci.K;
    const double2* gb = d.rec_b + (size_t)c * CHUNK_CAP + tid;
    double R[9], C[3];
    float Rf[9], q2[4];
    float acc[GRAM32];
#pragma unroll
    for (int e = 0; e < GRAM32; ++e) acc[e] = 0.f;
#pragma unroll
    for (int e = 0; e < 9; ++e) R[e] = 0.0;
    C[0] = C[1] = C[2] = 0.0;
    q2[0] = q2[1] = q2[2] = q2[3] = 0.f;
    const uint32_t ix_first = cs.ixs[tid];
    const bool live = ix_first != REC_INVALID;
    double2 b_next = make_double2(0.0, 0.0);
    if (live) {                                            // one camera per lane
      const int cam = cs.segc[rec_seg(ix_first)];
      b_next = gb[0];
      load_cam_rc(d.camtab, cam, R, C);
      double qd[4];
      load_cam_q(d.camtab, cam, qd);
#pragma unroll
      for (int e = 0; e < 4; ++e) q2[e] = (float)(2.0 * qd[e]);
      if (d.segst && cs.srun[rec_seg(ix_first)] == tid) {  // first lane of the segment: its FP32 row for the matvec
        float4* row = reinterpret_cast<float4*>(d.segst + ((size_t)c * SEG_CAP + rec_seg(ix_first)) * 12);
        row[0] = make_float4((float)R[0], (float)R[1], (float)R[2], (float)R[3]);
        row[1] = make_float4((float)R[4], (float)R[5], (float)R[6], (float)R[7]);
        row[2] = make_float4((float)R[8], (float)(C[0] - Xa[0]), (float)(C[1] - Xa[1]), (float)(C[2] - Xa[2]));
      }
    }
#pragma unroll
    for (int e = 0; e < 9; ++e) Rf[e] = (float)R[e];
    for (int k = 0; k < K; ++k) {
      const uint32_t ix_cur = cs.ixs[k * RUN_LANES + tid];
      if (ix_cur == REC_INVALID) break;
      const double2 b_cur = b_next;
      if (k + 1 < K) b_next = gb[(k + 1) * RUN_LANES];     // one trip ahead
      const int ptl = rec_ptl(ix_cur);
      const Proj pr = project(R, C, Xa[3 * ptl], Xa[3 * ptl + 1], Xa[3 * ptl + 2]);
      const Obs32 o = to_f32(pr, b_cur.x - pr.u, b_cur.y - pr.v);
      const float iw = o.iw, u = o.u, v = o.v;
      // J_x (2x3)
      float jx[6];
      jx[0] = (Rf[0] - u * Rf[2]) * iw;
      jx[1] = (Rf[3] - u * Rf[5]) * iw;
      jx[2] = (Rf[6] - u * Rf[8]) * iw;
      jx[3] = (Rf[1] - v * Rf[2]) * iw;
      jx[4] = (Rf[4] - v * Rf[5]) * iw;
      jx[5] = (Rf[7] - v * Rf[8]) * iw;
      // J_p (2x7): [d/dC | d/dq], u-row in jp[0..6], v-row in jp[7..13]
      float jp[14];
      jp[0] = -jx[0];
      jp[1] = -jx[1];
      jp[2] = -jx[2];
      if (dvdc_quirk) {                                   // campose_processor.py:802-804
        jp[7] = -(Rf[1] + v * Rf[2]) * iw;
        jp[8] = -(Rf[4] + v * Rf[5]) * iw;
        jp[9] = -(Rf[7] + v * Rf[8]) * iw;
      } else {
        jp[7] = -jx[3];
        jp[8] = -jx[4];
        jp[9] = -jx[5];
      }
      {
        const float d0 = o.d0, d1 = o.d1, d2 = o.d2;
        const float w2 = q2[0], x2 = q2[1], y2 = q2[2], z2 = q2[3];
        // G_c[j] = sum_k d_k dR[k][c]/dq_j   (campose_processor.py:636-767)
        const float G00 = d1 * z2 - d2 * y2;
        const float G01 = d1 * y2 + d2 * z2;
        const float G02 = -2.f * d0 * y2 + d1 * x2 - d2 * w2;
        const float G03 = -2.f * d0 * z2 + d1 * w2 + d2 * x2;
        const float G10 = -d0 * z2 + d2 * x2;
        const float G11 = d0 * y2 - 2.f * d1 * x2 + d2 * w2;
        const float G12 = d0 * x2 + d2 * z2;
        const float G13 = -d0 * w2 - 2.f * d1 * z2 + d2 * y2;
        const float G20 = d0 * y2 - d1 * x2;
        const float G21 = d0 * z2 - d1 * w2 - 2.f * d2 * x2;
        const float G22 = d0 * w2 + d1 * z2 - 2.f * d2 * y2;
        const float G23 = d0 * x2 + d1 * y2;
        jp[3] = (G00 - u * G20) * iw;
        jp[4] = (G01 - u * G21) * iw;
        jp[5] = (G02 - u * G22) * iw;
        jp[6] = (G03 - u * G23) * iw;
        jp[10] = (G10 - v * G20) * iw;
        jp[11] = (G11 - v * G21) * iw;
        jp[12] = (G12 - v * G22) * iw;
        jp[13] = (G13 - v * G23) * iw;
      }
      // rho = r - J_x v_j
      const float v0 = Vf[ptl], v1 = Vf[pt_cap + ptl], v2 = Vf[2 * pt_cap + ptl];
      const float rho0 = o.r0 - (jx[0] * v0 + jx[1] * v1 + jx[2] * v2);
      const float rho1 = o.r1 - (jx[3] * v0 + jx[4] * v1 + jx[5] * v2);
      {
        int e = 0;
#pragma unroll
        for (int a = 0; a < 7; ++a) {
#pragma unroll
          for (int bcol = a; bcol < 7; ++bcol, ++e) acc[e] = fmaf(jp[a], jp[bcol], acc[e]);
        }
#pragma unroll
        for (int a = 0; a < 7; ++a) acc[28 + a] = fmaf(jp[a], rho0, acc[28 + a]);
        e = 0;
#pragma unroll
        for (int a = 0; a < 7; ++a) {
#pragma unroll
          for (int bcol = a; bcol < 7; ++bcol, ++e) acc[e] = fmaf(jp[7 + a], jp[7 + bcol], acc[e]);
        }
#pragma unroll
        for (int a = 0; a < 7; ++a) acc[28 + a] = fmaf(jp[7 + a], rho1, acc[28 + a]);
      }
    }
    if (gm.nc > 1) __syncthreads();        // (multi-chunk: every lane is done with the previous chunk's lane sums)
    if (live) {
      float* dst = runpart + tid * GRAM32;
#pragma unroll
      for (int e = 0; e < GRAM32; ++e) dst[e] = acc[e];
    }
    __syncthreads();
    // per segment: sum its lanes in lane order -> one row of ROW_F32 floats (the sums over a camera's rows are FP64,
    // k_cam_finish)
    {
      const int lane = tid & 31;
      for (int s = tid >> 5; s < ci.nseg; s += RUN_LANES / 32) {
        const int ra = cs.srun[s], rb = cs.srun[s + 1];
        float a0 = 0.f, a1 = 0.f;
        for (int rr = ra; rr < rb; ++rr) {
          a0 += runpart[rr * GRAM32 + lane];
          if (lane + 32 < GRAM32) a1 += runpart[rr * GRAM32 + 32 + lane];
        }
        float* row = d.part64f + (size_t)cs.rpos[s] * ROW_F32;
        row[lane] = a0;
        if (lane + 32 < ROW_F32) row[32 + lane] = a1;          // (entries 35..39: zero padding)
      }
    }
    __syncthreads();
  }
}

void launch_linearize32(const BaDev& d, const TileShape& ts, bool dvdc_quirk, int cost_slot, cudaStream_t st,
                        LaunchCounter& lc, int g0, int g1, bool sum_cost) {
  Lin32Shape sh;
  sh.pt_cap = (ts.max_group_pts + 3) & ~3;
  sh.seg_cap = ts.max_nseg;
  sh.k_cap = ts.max_k;
  sh.slot_cap = ((ts.max_chunk_len + 31) & ~31) + 1;     // = 1 mod 32: the nine slot planes start in different banks
  const int a = 9 * sh.slot_cap, bb = RUN_LANES * GRAM32;
  sh.un_floats = ((a > bb ? a : bb) + 3) & ~3;
  const size_t smem = lin32_smem_bytes(sh);
  if (g1 < 0) {
    g0 = 0;
    g1 = d.G;
  }
  if (g1 > g0) {
    cudaFuncSetAttribute(k_linearize32, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_linearize32<<<g1 - g0, RUN_LANES, smem, st>>>(d, sh, dvdc_quirk ? 1 : 0, g0);
    lc.launches++;
  }
  if (sum_cost) launch_sum_cost(d, cost_slot, st, lc);
}

}  // namespace sfmba
