// Linearisation, camera side (K1 + K2b fused): the 63 numbers per camera the reduced system needs,
//   A_i   = sum J_p^T J_p                          (28, upper triangle)
//   Ssub  = sum J_p^T Y J_p,  Y = J_x D^-1 J_x^T    (28)  -> S_ii = A_i + lambda I - Ssub (block-Jacobi)
//   g_i   = sum J_p^T rho,    rho = r - J_x D^-1 e_x (7)  -> right-hand side e_p - B D^-1 e_x
// Reference: ba_processor.py:376-382 forms A = J_p^T J_p + lambda I, e_p, B = J_p^T J_x densely and
// inverts A - B D^-1 B^T; campose_processor.py:462-482, 636-808 build J_p.
//
// k_camgram walks a group's camera-sorted records: a lane keeps the camera constants AND the 63
// running sums of its current run in registers, so an observation costs ~340 FP64 FMAs and no
// memory traffic beyond its 20-byte record and shared-memory reads of its point (X, D^-1, v).
// Lanes flush once to shared memory, are summed per segment (camera) in lane order and written as one
// 64-double row per (chunk, camera); k_camsum64 then adds each camera's rows in row order.
#include "ba_internal.h"
#include "sfmba_device.cuh"
#include "peer.cuh"
#include "tile.cuh"

namespace sfmba {

constexpr int SLOT_PLANES = 5;     // slot planes live at once in the point phase (two rounds: 5 + 4 of D(6) e_x(3))

// Gram sums per lane / per segment row.  SCHUR: A(28) S(28) g(7) = 63, rows of CAM_ACC doubles; otherwise the 28 S
// entries do not exist: A(28) g(7) = 35, rows of ROW_COMPACT doubles.  Both counts are odd: a lane's flush into
// shared memory with that stride is conflict-free.
template <bool SCHUR> struct Gram {
  static constexpr int N = SCHUR ? 63 : 35;
  static constexpr int G0 = SCHUR ? 56 : 28;          // first g entry
  static constexpr int ROW = SCHUR ? CAM_ACC : ROW_COMPACT;
};

// dynamic shared memory (doubles): Xa[3 pt_cap] | Dv[9][pt_cap] | red[32] | un[max(5 slot_cap, RUN_LANES Gram::N)]
// then ixs | segc | rpos | srun | rowp.   `un` holds the point-pass slots, later the lane sums.  The observations
// (rec_b) are read from global memory, lane-transposed and coalesced, one trip ahead.
struct GramShape {
  int pt_cap, seg_cap, k_cap, slot_cap, un_doubles;
};

__host__ __device__ inline size_t gram_smem_bytes(const GramShape& s) {
  size_t b = ((size_t)12 * s.pt_cap + 32 + (size_t)s.un_doubles) * 8;
  b += (size_t)4 * s.k_cap * RUN_LANES;            // ixs
  b += (size_t)8 * ((s.seg_cap + 3) / 4 * 4);      // segc, rpos
  b += (size_t)2 * (SEG_RUN_STRIDE + 6);           // srun
  b += (size_t)4 * (s.pt_cap + 4);                 // rowp
  return b;
}

// Phase A (K2a, ba_processor.py:355-363): per point D_j = sum J_x^T J_x + lambda I, D_j^-1, v_j = D_j^-1 e_x
// and the cost; phase B: the camera-side Gram sums.  One launch linearises the whole problem.
// ROBUST: every observation enters with its Huber weight w = rho'(|r|^2) (J^T W J, J^T W r); the cost is sum rho
// SCHUR: also the 28 sums of J_p^T Y J_p (diagonal blocks of B D^-1 B^T)
// Three CTAs per SM without SCHUR (<= 168 registers, <= 75 KB of shared memory), two with it.
template <bool ROBUST, bool SCHUR>
__global__ void __launch_bounds__(RUN_LANES, SCHUR ? 2 : 3) k_linearize(BaDev d, GramShape sh, int dvdc_quirk, int g0) {
  extern __shared__ __align__(16) double smem[];
  constexpr int GRAM = Gram<SCHUR>::N, RP_STRIDE = Gram<SCHUR>::N, G0 = Gram<SCHUR>::G0, ROW = Gram<SCHUR>::ROW;
  const int pt_cap = sh.pt_cap, slot_cap = sh.slot_cap;
  double* Xa = smem;
  double* Dv = Xa + 3 * pt_cap;
  double* red = Dv + 9 * pt_cap;
  double* runpart = red + 32;
  double* slots = runpart;
  ChunkStage cs;
  cs.bs = nullptr;
  cs.ixs = reinterpret_cast<uint32_t*>(runpart + sh.un_doubles);
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
    for (int i = tid; i < 9 * pt_cap; i += RUN_LANES) Dv[i] = 0.0;
  }
  ChunkInfo ci;
  ci.obs0 = gm.obs0;
  ci.len = gm.len0;
  ci.nseg = gm.nseg0;
  ci.K = gm.K0;
  double cost = 0.0;
  uint32_t flags = 0;
  // ---- phase A: point blocks, in two rounds over the chunk (planes 0..4, then 5..8: D00 D01 D02 D11 D12 | D22 e_x)
  for (int c = c0; c < c1; ++c) {
    if (c > c0) ci = d.chunk_info[c];
    stage_chunk_async(d, c, ci, cs);
    cp_async_wait_all();
    __syncthreads();
    const int K = ci.K;
    const double2* gb = d.rec_b + (size_t)c * CHUNK_CAP + tid;
    double R[9], C[3];
    const uint32_t ix_first = cs.ixs[tid];
    if (ix_first != REC_INVALID) load_cam_rc(d.camtab, cs.segc[rec_seg(ix_first)], R, C);
    // J_x of one observation and, in the second round, its residual
    auto jx_terms = [&](uint32_t ix, double x[3], double y[3], Proj& pr) {
      const int ptl = rec_ptl(ix);
      pr = project(R, C, Xa[3 * ptl], Xa[3 * ptl + 1], Xa[3 * ptl + 2]);
      x[0] = (R[0] - pr.u * R[2]) * pr.iw;
      x[1] = (R[3] - pr.u * R[5]) * pr.iw;
      x[2] = (R[6] - pr.u * R[8]) * pr.iw;
      y[0] = (R[1] - pr.v * R[2]) * pr.iw;
      y[1] = (R[4] - pr.v * R[5]) * pr.iw;
      y[2] = (R[7] - pr.v * R[8]) * pr.iw;
    };
#pragma unroll
    for (int round = 0; round < 2; ++round) {
      // two observations of the lane per trip: their dependent chains (projection, reciprocal, products) interleave
      double2 nb_a = make_double2(0.0, 0.0), nb_b = nb_a;
      if (round == 1 || ROBUST) {
        nb_a = gb[0];
        if (K > 1) nb_b = gb[RUN_LANES];
      }
      for (int k = 0; k < K; k += 2) {
        const uint32_t ix_a = cs.ixs[k * RUN_LANES + tid];
        if (ix_a == REC_INVALID) break;
        uint32_t ix_b = k + 1 < K ? cs.ixs[(k + 1) * RUN_LANES + tid] : REC_INVALID;
        const bool two = ix_b != REC_INVALID;
        if (!two) ix_b = ix_a;
        const double2 b_a = nb_a, b_b = two ? nb_b : nb_a;
        if (round == 1 || ROBUST) {                          // the next trip's observations are on their way
          if (k + 2 < K) nb_a = gb[(k + 2) * RUN_LANES];
          if (k + 3 < K) nb_b = gb[(k + 3) * RUN_LANES];
        }
        double xa[3], ya[3], xb[3], yb[3];
        Proj pa, pb;
        jx_terms(ix_a, xa, ya, pa);
        jx_terms(ix_b, xb, yb, pb);
        double wa = 1.0, wb = 1.0;
        const double ra0 = b_a.x - pa.u, ra1 = b_a.y - pa.v, rb0 = b_b.x - pb.u, rb1 = b_b.y - pb.v;
        if (ROBUST) {
          double rho;
          wa = huber_weight(ra0 * ra0 + ra1 * ra1, d.huber_delta, rho);
          if (round == 1) cost += rho;
          wb = huber_weight(rb0 * rb0 + rb1 * rb1, d.huber_delta, rho);
          if (round == 1 && two) cost += rho;
        } else if (round == 1) {
          cost += ra0 * ra0 + ra1 * ra1;
          if (two) cost += rb0 * rb0 + rb1 * rb1;
        }
        auto emit = [&](uint32_t ix, const double x[3], const double y[3], double w, double r0, double r1) {
          double* sl = slots + rec_pos(ix);
          const double wx0 = ROBUST ? w * x[0] : x[0], wx1 = ROBUST ? w * x[1] : x[1], wx2 = ROBUST ? w * x[2] : x[2];
          const double wy0 = ROBUST ? w * y[0] : y[0], wy1 = ROBUST ? w * y[1] : y[1], wy2 = ROBUST ? w * y[2] : y[2];
          if (round == 0) {
            sl[0] = wx0 * x[0] + wy0 * y[0];
            sl[slot_cap] = wx0 * x[1] + wy0 * y[1];
            sl[2 * slot_cap] = wx0 * x[2] + wy0 * y[2];
            sl[3 * slot_cap] = wx1 * x[1] + wy1 * y[1];
            sl[4 * slot_cap] = wx1 * x[2] + wy1 * y[2];
          } else {
            sl[0] = wx2 * x[2] + wy2 * y[2];
            sl[slot_cap] = wx0 * r0 + wy0 * r1;
            sl[2 * slot_cap] = wx1 * r0 + wy1 * r1;
            sl[3 * slot_cap] = wx2 * r0 + wy2 * r1;
          }
        };
        if (round == 1 && (!pa.ok || !pb.ok)) flags |= FLAG_DEPTH_ZERO;
        emit(ix_a, xa, ya, wa, ra0, ra1);
        if (two) emit(ix_b, xb, yb, wb, rb0, rb1);
      }
      __syncthreads();
      // one work item per (point, plane group): the slots of a point are summed in observation order, up to three
      // planes side by side (the row-pointer lookups are paid once per item).  Round 0: planes {0,1,2} {3,4};
      // round 1: {5,6} {7,8}
      for (int idx = tid; idx < 2 * np; idx += RUN_LANES) {
        const int j = idx >> 1, half = idx & 1;
        const int lo = max(rowp[j], ci.obs0) - ci.obs0;
        const int hi = min(rowp[j + 1], ci.obs0 + ci.len) - ci.obs0;
        if (lo < hi) {
          const int sp = round == 0 ? 3 * half : 2 * half;                 // first slot plane of the item
          const int dp = round == 0 ? 3 * half : 5 + 2 * half;             // first Dv plane
          const bool three = round == 0 && half == 0;
          const double* sl = slots + sp * slot_cap;
          double a0 = Dv[dp * pt_cap + j], a1 = Dv[(dp + 1) * pt_cap + j], a2 = three ? Dv[(dp + 2) * pt_cap + j] : 0.0;
          for (int sidx = lo; sidx < hi; ++sidx) {
            a0 += sl[sidx];
            a1 += sl[slot_cap + sidx];
            if (three) a2 += sl[2 * slot_cap + sidx];
          }
          Dv[dp * pt_cap + j] = a0;
          Dv[(dp + 1) * pt_cap + j] = a1;
          if (three) Dv[(dp + 2) * pt_cap + j] = a2;
        }
      }
      __syncthreads();
    }
  }
  cp_async_wait_all();
  __syncthreads();
  double exv = 0.0;                                    // sum_j e_x^T D^-1 e_x (adaptive LM: point part of the model decrease)
  {
    const double lam = d.sc->lambda;
    const size_t N = (size_t)d.N;
    for (int j = tid; j < np; j += RUN_LANES) {
      double D[6], ex[3], Dinv[6], v[3];
#pragma unroll
      for (int e = 0; e < 6; ++e) D[e] = Dv[e * pt_cap + j];
#pragma unroll
      for (int e = 0; e < 3; ++e) ex[e] = Dv[(6 + e) * pt_cap + j];
      D[0] += lam;                                     // ba_processor.py:359 (every point, observed or not)
      D[3] += lam;
      D[5] += lam;
      if (!sym3_inverse(D, Dinv)) flags |= FLAG_SINGULAR;
      sym3_mul(Dinv, ex, v);
      exv += ex[0] * v[0] + ex[1] * v[1] + ex[2] * v[2];
      const size_t pt = (size_t)pt0 + j;
      if (d.xrel) {                                    // FP32 matvec: the point relative to the group's first point, D^-1
        d.xrel[pt] = make_float4((float)(Xa[3 * j] - Xa[0]), (float)(Xa[3 * j + 1] - Xa[1]),
                                 (float)(Xa[3 * j + 2] - Xa[2]), 0.f);
        float4* dg = reinterpret_cast<float4*>(d.dinv32 + 8 * pt);
        dg[0] = make_float4((float)Dinv[0], (float)Dinv[1], (float)Dinv[2], (float)Dinv[3]);
        dg[1] = make_float4((float)Dinv[4], (float)Dinv[5], 0.f, 0.f);
      }
#pragma unroll
      for (int e = 0; e < 6; ++e) {
        if (SCHUR) Dv[e * pt_cap + j] = Dinv[e];
        d.dinv[e * N + pt] = Dinv[e];
      }
#pragma unroll
      for (int e = 0; e < 3; ++e) {
        Dv[(6 + e) * pt_cap + j] = v[e];
        d.vj[e * N + pt] = v[e];
      }
    }
  }
  cost = block_sum(cost, red);     // (contains the barrier that publishes Dv)
  if (tid == 0) d.grp_cost[grp] = cost;
  if (d.grp_exv) {                 // adaptive LM only (uniform branch)
    exv = block_sum(exv, red);
    if (tid == 0) d.grp_exv[grp] = exv;
  }
  if (flags) atomicOr(&d.sc->flags, flags);
  // ---- phase B: camera-side Gram sums
  for (int c = c0; c < c1; ++c) {
    if (gm.nc > 1) {                       // single-chunk groups still hold their chunk from phase A
      ci = d.chunk_info[c];
      stage_chunk_async(d, c, ci, cs);
      cp_async_wait_all();
      __syncthreads();
    }
    const int K = ci.K;
    const double2* gb = d.rec_b + (size_t)c * CHUNK_CAP + tid;
    double R[9], C[3], q2[4];
    double acc[GRAM];
#pragma unroll
    for (int e = 0; e < GRAM; ++e) acc[e] = 0.0;
    const uint32_t ix_first = cs.ixs[tid];
    const bool live = ix_first != REC_INVALID;
    double2 b_next = make_double2(0.0, 0.0);
    if (live) {                                            // one camera per lane
      const int cam = cs.segc[rec_seg(ix_first)];
      b_next = gb[0];
      load_cam_rc(d.camtab, cam, R, C);
      load_cam_q(d.camtab, cam, q2);
#pragma unroll
      for (int e = 0; e < 4; ++e) q2[e] *= 2.0;
      if (d.segst && cs.srun[rec_seg(ix_first)] == tid) {  // first lane of the segment: its FP32 row for the matvec
        float4* row = reinterpret_cast<float4*>(d.segst + ((size_t)c * SEG_CAP + rec_seg(ix_first)) * 12);
        row[0] = make_float4((float)R[0], (float)R[1], (float)R[2], (float)R[3]);
        row[1] = make_float4((float)R[4], (float)R[5], (float)R[6], (float)R[7]);
        row[2] = make_float4((float)R[8], (float)(C[0] - Xa[0]), (float)(C[1] - Xa[1]), (float)(C[2] - Xa[2]));
      }
    }
    for (int k = 0; k < K; ++k) {
      const uint32_t ix_cur = cs.ixs[k * RUN_LANES + tid];
      if (ix_cur == REC_INVALID) break;
      const double2 b_cur = b_next;
      if (k + 1 < K) b_next = gb[(k + 1) * RUN_LANES];     // one trip ahead
      const int ptl = rec_ptl(ix_cur);
      const Proj pr = project(R, C, Xa[3 * ptl], Xa[3 * ptl + 1], Xa[3 * ptl + 2]);
      const double iw = pr.iw, u = pr.u, v = pr.v;
      const double r0 = b_cur.x - u, r1 = b_cur.y - v;
      // J_x (2x3)
      double jx[6];
      jx[0] = (R[0] - u * R[2]) * iw;
      jx[1] = (R[3] - u * R[5]) * iw;
      jx[2] = (R[6] - u * R[8]) * iw;
      jx[3] = (R[1] - v * R[2]) * iw;
      jx[4] = (R[4] - v * R[5]) * iw;
      jx[5] = (R[7] - v * R[8]) * iw;
      // J_p (2x7): [d/dC | d/dq], u-row in jp[0..6], v-row in jp[7..13]
      double jp[14];
      jp[0] = -jx[0];
      jp[1] = -jx[1];
      jp[2] = -jx[2];
      if (dvdc_quirk) {                                   // campose_processor.py:802-804
        jp[7] = -(R[1] + v * R[2]) * iw;
        jp[8] = -(R[4] + v * R[5]) * iw;
        jp[9] = -(R[7] + v * R[8]) * iw;
      } else {
        jp[7] = -jx[3];
        jp[8] = -jx[4];
        jp[9] = -jx[5];
      }
      {
        const double d0 = pr.d0, d1 = pr.d1, d2 = pr.d2;
        const double w2 = q2[0], x2 = q2[1], y2 = q2[2], z2 = q2[3];
        // G_c[j] = sum_k d_k dR[k][c]/dq_j   (campose_processor.py:636-767)
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
        jp[3] = (G00 - u * G20) * iw;
        jp[4] = (G01 - u * G21) * iw;
        jp[5] = (G02 - u * G22) * iw;
        jp[6] = (G03 - u * G23) * iw;
        jp[10] = (G10 - v * G20) * iw;
        jp[11] = (G11 - v * G21) * iw;
        jp[12] = (G12 - v * G22) * iw;
        jp[13] = (G13 - v * G23) * iw;
      }
      // point terms: Y = J_x D^-1 J_x^T, rho = r - J_x v_j
      double vv[3];
#pragma unroll
      for (int e = 0; e < 3; ++e) vv[e] = Dv[(6 + e) * pt_cap + ptl];
      double yuu = 0.0, yuv = 0.0, yvv = 0.0;
      if (SCHUR) {
        double Dinv[6];
#pragma unroll
        for (int e = 0; e < 6; ++e) Dinv[e] = Dv[e * pt_cap + ptl];
        double t0[3], t1[3];
        sym3_mul(Dinv, jx, t0);
        sym3_mul(Dinv, jx + 3, t1);
        yuu = jx[0] * t0[0] + jx[1] * t0[1] + jx[2] * t0[2];
        yuv = jx[0] * t1[0] + jx[1] * t1[1] + jx[2] * t1[2];
        yvv = jx[3] * t1[0] + jx[4] * t1[1] + jx[5] * t1[2];
      }
      const double rho0 = r0 - (jx[0] * vv[0] + jx[1] * vv[1] + jx[2] * vv[2]);
      const double rho1 = r1 - (jx[3] * vv[0] + jx[4] * vv[1] + jx[5] * vv[2]);
      // weighted sums: the left factor of every product below is w J_p, Y carries the second w of J_p^T (w Y w) J_p
      double jw[14];
#pragma unroll
      for (int a = 0; a < 14; ++a) jw[a] = jp[a];
      if (ROBUST) {
        double rho_unused;
        const double w = huber_weight(r0 * r0 + r1 * r1, d.huber_delta, rho_unused);
        d.rec_w[(size_t)c * CHUNK_CAP + k * RUN_LANES + tid] = w;
#pragma unroll
        for (int a = 0; a < 14; ++a) jw[a] *= w;
        yuu *= w;
        yuv *= w;
        yvv *= w;
      }
      double tu[7], tv[7];
#pragma unroll
      for (int a = 0; a < 7; ++a) {
        tu[a] = SCHUR ? yuu * jp[a] + yuv * jp[7 + a] : 0.0;
        tv[a] = SCHUR ? yuv * jp[a] + yvv * jp[7 + a] : 0.0;
      }
      // two FMA sweeps (u-row products, then v-row products): 2 instructions per sum, all independent
      {
        int e = 0;
#pragma unroll
        for (int a = 0; a < 7; ++a) {
#pragma unroll
          for (int bcol = a; bcol < 7; ++bcol, ++e) {
            acc[e] = fma(jw[a], jp[bcol], acc[e]);
            if (SCHUR) acc[28 + e] = fma(jw[a], tu[bcol], acc[28 + e]);
          }
        }
#pragma unroll
        for (int a = 0; a < 7; ++a) acc[G0 + a] = fma(jw[a], rho0, acc[G0 + a]);
        e = 0;
#pragma unroll
        for (int a = 0; a < 7; ++a) {
#pragma unroll
          for (int bcol = a; bcol < 7; ++bcol, ++e) {
            acc[e] = fma(jw[7 + a], jp[7 + bcol], acc[e]);
            if (SCHUR) acc[28 + e] = fma(jw[7 + a], tv[bcol], acc[28 + e]);
          }
        }
#pragma unroll
        for (int a = 0; a < 7; ++a) acc[G0 + a] = fma(jw[7 + a], rho1, acc[G0 + a]);
      }
    }
    if (live) {
      double* dst = runpart + tid * RP_STRIDE;
#pragma unroll
      for (int e = 0; e < GRAM; ++e) dst[e] = acc[e];
    }
    __syncthreads();
    // per segment: sum its lanes in lane order -> one row of ROW doubles
    const uint16_t* run0 = cs.srun;
    // one warp per segment, two components per lane: no index arithmetic per element, coalesced rows
    {
      const int lane = tid & 31;
      for (int s = tid >> 5; s < ci.nseg; s += RUN_LANES / 32) {
        const int ra = run0[s], rb = run0[s + 1];
        double a0 = 0.0, a1 = 0.0;
        for (int rr = ra; rr < rb; ++rr) {
          a0 += runpart[rr * RP_STRIDE + lane];
          if (lane + 32 < GRAM) a1 += runpart[rr * RP_STRIDE + 32 + lane];
        }
        double* row = d.part64 + (size_t)cs.rpos[s] * ROW;
        row[lane] = a0;
        if (lane + 32 < ROW) row[32 + lane] = a1;
      }
    }
    __syncthreads();
  }
}

// Sum of one camera's (contiguous) segment rows, element e: four row slots (slot = tid >> 6) take rows
// r0+slot, r0+slot+4, ... with eight loads in flight each and are combined in slot order through shared
// memory.  One CTA of 256 threads per camera; the result is valid in the threads of slot 0 with e < row length.
// Rows are CAM_ACC doubles with the Schur-diagonal sums and ROW_COMPACT doubles (A | g) without them.
__device__ __forceinline__ double cam_rows_sum64(const BaDev& d, int cam, double (*sh)[CAM_ACC]) {
  const int e = threadIdx.x & 63, slot = threadIdx.x >> 6;
  const int r0 = d.camrow_ptr[cam], r1 = d.camrow_ptr[cam + 1];
  const size_t row = d.schur_diag ? CAM_ACC : ROW_COMPACT;
  const double* p = d.part64 + e;
  double a = 0.0;
  int k = r0 + slot;
  if (e < (int)row) {
    for (; k + 28 < r1; k += 32) {
      double v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = p[(size_t)(k + 4 * u) * row];
#pragma unroll
      for (int u = 0; u < 8; ++u) a += v[u];
    }
    if (k < r1) {                                        // the last (up to seven) rows: loads first, same order of sums
      double v[7];
#pragma unroll
      for (int u = 0; u < 7; ++u) v[u] = k + 4 * u < r1 ? p[(size_t)(k + 4 * u) * row] : 0.0;
#pragma unroll
      for (int u = 0; u < 7; ++u)
        if (k + 4 * u < r1) a += v[u];
    }
  }
  if (slot) sh[slot - 1][e] = a;
  __syncthreads();
  if (slot == 0) a = ((a + sh[0][e]) + sh[1][e]) + sh[2][e];
  return a;
}

// position in a cam_acc row (A(28) S(28) g(7) pad) of element e of a segment row
__device__ __forceinline__ int acc_index(const BaDev& d, int e) { return (d.schur_diag || e < 28) ? e : e + 28; }

// cam_acc[cam][:] = sum over the camera's segment rows
__global__ void __launch_bounds__(256) k_camsum64(BaDev d) {
  __shared__ double sh[3][CAM_ACC];
  const int cam = blockIdx.x;
  const double a = cam_rows_sum64(d, cam, sh);
  const int row = d.schur_diag ? CAM_ACC : ROW_COMPACT;
  if ((int)threadIdx.x < row) d.cam_acc[(size_t)cam * CAM_ACC + acc_index(d, threadIdx.x)] = a;
}

// per camera: A_i + lambda I (full 7x7), block-Jacobi inverse of S_ii, right-hand side g_i
__global__ void k_cam_blocks(BaDev d) {
  const int cam = blockIdx.x * blockDim.x + threadIdx.x;
  if (cam >= d.V) return;
  const double lam = d.sc->lambda;
  const double* acc = d.cam_acc + (size_t)cam * CAM_ACC;
  double A[49], S[49];
  int k = 0;
  for (int a = 0; a < 7; ++a) {
    for (int b = a; b < 7; ++b, ++k) {
      double va = acc[k], vs = d.schur_diag ? acc[k] - acc[28 + k] : acc[k];
      if (a == b) {
        va += lam;
        vs += lam;
      }
      A[a * 7 + b] = va;
      A[b * 7 + a] = va;
      S[a * 7 + b] = vs;
      S[b * 7 + a] = vs;
    }
  }
  if (!spd_inverse<7>(S)) atomicOr(&d.sc->flags, FLAG_SINGULAR);
  double* ao = d.ablk + (size_t)cam * 49;
  double* mo = d.minv + (size_t)cam * 49;
  for (int i = 0; i < 49; ++i) {
    ao[i] = A[i];
    mo[i] = S[i];
  }
  for (int a = 0; a < 7; ++a) d.gvec[7 * cam + a] = acc[56 + a];
}

__global__ void k_camsum64_peer(BaDev d);

// ---- camera finish, fused (PCG solver): segment rows (or the ranks' window slots) -> cam_acc -> A_i + lambda I, the
// block-Jacobi inverse, the right-hand side -> the start of the PCG solve for this camera (x = 0, r = g, z = Minv g and
// its camit rows, p = s = 0) -> |g|^2 by the CTA that finishes last: what k_camsum64 (or k_cam_gather_peers),
// k_cam_blocks and k_pcg_init do in three launches.  One WARP per camera: lane e holds elements e and e + 32 of the
// camera's sums; every lane factors the 7x7 block (Cholesky, from values fetched by shuffles) and lane c < 7 solves for
// column c of the inverse, so the 7x7 work is seven-fold parallel and lives in registers.
constexpr int FIN_WARPS = 8;

#ifdef SFMBA_STEP_TRACE
// profiling build only (profiles/step_trace.py): warp 0 of every CTA stamps the phases of the last k_cam_finish launch
__device__ unsigned long long g_fin_trace[2][512][8];
#define FIN_STAMP(i)                                                     \
  do {                                                                   \
    if (threadIdx.x == 0 && blockIdx.x < 512) {                          \
      g_fin_trace[0][blockIdx.x][i] = global_timer_ns();                 \
      g_fin_trace[1][blockIdx.x][i] = (unsigned long long)clock64();     \
    }                                                                    \
  } while (0)
#else
#define FIN_STAMP(i) do {} while (0)
#endif

template <bool PEER>
__global__ void __launch_bounds__(32 * FIN_WARPS) k_cam_finish(BaDev d, int cost_slot, unsigned epoch, int dvdc_quirk,
                                                              int rows_f32, int cost_partials) {
  __shared__ double red[32];
  __shared__ float4 fin_buf[FIN_WARPS][5 * 32];      // lane sums of the FP32 row gather
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  Scalars* sc = d.sc;
  FIN_STAMP(0);
  // (the CTA with the fewest cameras does the extra work)
  const bool cost_cta = blockIdx.x == gridDim.x - 1;
  if (cost_partials > 0 && cost_cta) {
    // the cost of this linearisation point: the per-group partials of the linearisation, summed in a fixed order (what
    // k_sum_cost did in a launch of its own); four loads in flight per thread
    double a = 0.0;
    constexpr int T = 32 * FIN_WARPS;
    int i = tid;
    for (; i + 3 * T < cost_partials; i += 4 * T) {
      const double v0 = __ldcg(d.grp_cost + i), v1 = __ldcg(d.grp_cost + i + T), v2 = __ldcg(d.grp_cost + i + 2 * T);
      const double v3 = __ldcg(d.grp_cost + i + 3 * T);
      a += v0;
      a += v1;
      a += v2;
      a += v3;
    }
    for (; i < cost_partials; i += T) a += __ldcg(d.grp_cost + i);
    a = block_sum(a, red);
    if (tid == 0) {
      sc->cost_cur = a;
      if (cost_slot < SFMBA_MAX_COST_HISTORY) sc->costs[cost_slot] = a;
      if (!(a == a) || isinf(a)) atomicOr(&sc->flags, FLAG_NONFINITE);
    }
    __syncthreads();
  }
  const bool schur = d.schur_diag != 0;
  const int row = schur ? CAM_ACC : ROW_COMPACT;
  const bool hi = lane + 32 < row;
  int par = 0;
  unsigned seq = 0;
  if (PEER) {
    // several GPUs: this rank's sums go out as flag-in-data words and every rank's words of the same camera come back in
    // the same warp (peer.cuh); the cost travels as one more entry behind the cameras
    const unsigned long long seq64 = *d.xw.counter() + 1ull;  // advanced by the CTA that finishes last
    seq = (unsigned)seq64;
    par = (int)(seq64 & 1ull);
    if (cost_cta && tid == 0) {
      const PeerWindow& xw = d.xw;
      const size_t k2 = 2 * (size_t)d.V * row;
      for (int w = 0; w < xw.world; ++w) ll_store(ll_slot(xw, w, par, xw.rank, xw.ll_cam) + k2, sc->cost_cur, seq);
      const double c = ll_sum_ranks(ll_slot(xw, xw.rank, par, 0, xw.ll_cam) + k2, xw.cap, xw.world, seq, xw.timeout_ns, sc);
      sc->cost_cur = c;
      if (cost_slot < SFMBA_MAX_COST_HISTORY) sc->costs[cost_slot] = c;
    }
  }
  const double lam = sc->lambda;
  double rr = 0.0;
  FIN_STAMP(1);
  for (int cam = blockIdx.x * FIN_WARPS + warp; cam < d.V; cam += gridDim.x * FIN_WARPS) {
    double a0 = 0.0, a1 = 0.0;                       // elements lane, lane + 32 of the camera's row sum
    {
      const int r0 = d.camrow_ptr[cam], r1 = d.camrow_ptr[cam + 1];
      const double* p = d.part64 + lane;
      const float* pf = d.part64f + lane;            // rows of the FP32 linearisation (ROW_F32 floats each)
      if (rows_f32) {
        // The camera's rows are contiguous: the warp streams them as 16-byte words, ten loads per lane in flight.  Word
        // i = 32 t + lane of the stream holds floats 4 (i mod 10) .. + 3 of some row, and i mod 10 = (lane + 2 t) mod 10
        // repeats with t mod 5: five FP32 accumulators per lane (a handful of terms each), combined across lanes in FP64
        // in a fixed order through shared memory.  (The earlier form -- scalar loads, eight rows at a time -- was a chain
        // of fifteen L2 round trips for a typical camera: 11 of the kernel's 21 us.)
        const float4* src = reinterpret_cast<const float4*>(d.part64f + (size_t)r0 * ROW_F32);
        const int n4 = (r1 - r0) * (ROW_F32 / 4);
        const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
        float4 acc[5] = {zero, zero, zero, zero, zero};
        for (int base = 0; base < n4; base += 320) {
          float4 v[10];
#pragma unroll
          for (int u = 0; u < 10; ++u) {
            const int i = base + 32 * u + lane;
            v[u] = i < n4 ? __ldcg(src + i) : zero;
          }
#pragma unroll
          for (int u = 0; u < 10; ++u) {
            acc[u % 5].x += v[u].x;
            acc[u % 5].y += v[u].y;
            acc[u % 5].z += v[u].z;
            acc[u % 5].w += v[u].w;
          }
        }
        float4* fb = fin_buf[warp];
        __syncwarp();
#pragma unroll
        for (int ph = 0; ph < 5; ++ph) fb[ph * 32 + lane] = acc[ph];
        __syncwarp();
        const float* ff = reinterpret_cast<const float*>(fb);
        auto element = [&](int e) {                    // lanes (cg - 2 ph) mod 10, + 10, ... of phase ph hold column group cg
          const int cg = e >> 2, j = e & 3;
          double sum = 0.0;
#pragma unroll
          for (int ph = 0; ph < 5; ++ph) {
            int l0 = (cg - 2 * ph) % 10;
            if (l0 < 0) l0 += 10;
            for (int l = l0; l < 32; l += 10) sum += (double)ff[(ph * 32 + l) * 4 + j];
          }
          return sum;
        };
        a0 = element(lane);
        a1 = lane < 8 ? element(32 + lane) : 0.0;
        __syncwarp();
      } else {
        for (int k = r0; k < r1; k += 8) {             // eight rows in flight, summed in row order
          double v[8], w[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const bool ok = k + u < r1;
            v[u] = ok ? p[(size_t)(k + u) * row] : 0.0;
            w[u] = (ok && hi) ? p[(size_t)(k + u) * row + 32] : 0.0;
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            a0 += v[u];
            a1 += w[u];
          }
        }
      }
    }
    if (a0 != 12345.0) FIN_STAMP(2);
    if (PEER) {
      const PeerWindow& xw = d.xw;
      const size_t k2 = 2 * ((size_t)cam * row + lane);
      for (int w = 0; w < xw.world; ++w) {
        double* dst = ll_slot(xw, w, par, xw.rank, xw.ll_cam) + k2;
        ll_store(dst, a0, seq);
        if (hi) ll_store(dst + 64, a1, seq);
      }
      ll_flush();
      const double* src = ll_slot(xw, xw.rank, par, 0, xw.ll_cam) + k2;
      a0 = ll_sum_ranks(src, xw.cap, xw.world, seq, xw.timeout_ns, sc);
      a1 = hi ? ll_sum_ranks(src + 64, xw.cap, xw.world, seq, xw.timeout_ns, sc) : 0.0;
    }
    d.cam_acc[(size_t)cam * CAM_ACC + acc_index(d, lane)] = a0;
    if (hi) d.cam_acc[(size_t)cam * CAM_ACC + acc_index(d, lane + 32)] = a1;
    // element e of the row (compile-time e) in every lane
    auto get = [&](int e) { return e < 32 ? __shfl_sync(0xffffffffu, a0, e) : __shfl_sync(0xffffffffu, a1, e - 32); };
    // S_ii = A + lambda I - (Schur-diagonal sums), packed upper triangle k(a, b) = a (15 - a) / 2 + b - a, a <= b
    // (static_for: every index below is a compile-time constant, so S and L live in registers)
    double S[28];
    static_for<0, 28>([&](auto K) {
      constexpr int k = K;
      S[k] = get(k);
      if (schur) S[k] -= get(28 + k);
    });
    static_for<0, 7>([&](auto A) {
      constexpr int a = A;
      S[a * (15 - a) / 2] += lam;
    });
    // Cholesky S = L L^T, L[i][j] (j <= i) stored at S-index k(j, i): L overwrites the packed triangle column by column
    bool ok = true;
    double L[28];
    static_for<0, 7>([&](auto J) {
      constexpr int j = J;
      double dj = S[j * (15 - j) / 2];
      static_for<0, j>([&](auto K) {
        constexpr int k = K;
        dj -= L[k * (15 - k) / 2 + j - k] * L[k * (15 - k) / 2 + j - k];
      });
      if (!(dj > 0.0)) ok = false;
      const double ljj = sqrt(dj);
      const double inv = 1.0 / ljj;
      L[j * (15 - j) / 2] = ljj;
      static_for<j + 1, 7>([&](auto I) {
        constexpr int i = I;
        double v = S[j * (15 - j) / 2 + i - j];
        static_for<0, j>([&](auto K) {
          constexpr int k = K;
          v -= L[k * (15 - k) / 2 + i - k] * L[k * (15 - k) / 2 + j - k];
        });
        L[j * (15 - j) / 2 + i - j] = v * inv;          // L[i][j]
      });
    });
    if (L[27] != 12345.0) FIN_STAMP(3);
    if (!ok && lane == 0) atomicOr(&sc->flags, FLAG_SINGULAR);
    // lane c < 7: column c of S^-1 (L y = e_c, L^T x = y)
    const int c = lane < 7 ? lane : 0;
    double y[7], x[7];
    static_for<0, 7>([&](auto I) {
      constexpr int i = I;
      double v = (i == c) ? 1.0 : 0.0;
      static_for<0, i>([&](auto J) {
        constexpr int j = J;
        v -= L[j * (15 - j) / 2 + i - j] * y[j];
      });
      y[i] = v / L[i * (15 - i) / 2];
    });
    static_for<0, 7>([&](auto R) {
      constexpr int i = 6 - R;
      double v = y[i];
      static_for<i + 1, 7>([&](auto J) {
        constexpr int j = J;
        v -= L[i * (15 - i) / 2 + j - i] * x[j];
      });
      x[i] = v / L[i * (15 - i) / 2];
    });
    if (x[0] != 12345.0) FIN_STAMP(4);
    // right-hand side g and z = Minv g (row c of the symmetric inverse is x)
    const int g0 = schur ? 56 : 28;
    double g[7], zc = 0.0;
#pragma unroll
    for (int bb = 0; bb < 7; ++bb) {
      g[bb] = schur ? get(56 + bb) : get(28 + bb);
      zc += x[bb] * g[bb];
    }
    (void)g0;
    double gc = 0.0;
#pragma unroll
    for (int bb = 0; bb < 7; ++bb)
      if (bb == c) gc = g[bb];
    // A_i + lambda I, row c: entry (c, b) sits at packed index k(min, max) of a0
    double arow[7];
#pragma unroll
    for (int bb = 0; bb < 7; ++bb) {
      const int lo = c < bb ? c : bb, hh = c < bb ? bb : c;
      arow[bb] = __shfl_sync(0xffffffffu, a0, lo * (15 - lo) / 2 + hh - lo) + (bb == c ? lam : 0.0);
    }
    // camit row of z: every lane needs all of z
    double zl[7], o[16];
#pragma unroll
    for (int bb = 0; bb < 7; ++bb) zl[bb] = __shfl_sync(0xffffffffu, zc, bb);
    camit_row(d.camtab + (size_t)CAMTAB_STRIDE * cam, zl, dvdc_quirk, o);
    o[14] = 0.0;
    o[15] = 0.0;
    double mine = 0.0;
#pragma unroll
    for (int e = 0; e < 16; ++e)
      if (lane == e) mine = o[e];
    if (lane < 16) {
      d.camit[(size_t)CAMIT_STRIDE * cam + lane] = mine;
      if (d.camit32) d.camit32[(size_t)CAMIT_STRIDE * cam + lane] = (float)mine;
    }
    if (lane < 7) {
      double* ao = d.ablk + (size_t)cam * 49 + 7 * lane;
      double* mo = d.minv + (size_t)cam * 49 + 7 * lane;
#pragma unroll
      for (int bb = 0; bb < 7; ++bb) {
        ao[bb] = arow[bb];
        mo[bb] = x[bb];
      }
      const size_t k7 = 7 * (size_t)cam + lane;
      d.gvec[k7] = gc;
      d.x[k7] = 0.0;
      d.p[k7] = 0.0;
      d.s[k7] = 0.0;
      d.r[k7] = gc;
      d.z[k7] = zc;
      rr += gc * gc;
    }
  }
  FIN_STAMP(5);
  rr = block_sum(rr, red);
  if (tid == 0) d.red_b[blockIdx.x] = rr;
  const bool fin_last = last_cta_arrives(&sc->ticket_b);
  FIN_STAMP(6);
  if (!fin_last) return;
  // ---- the last CTA: |g|^2 in CTA order, the iteration state
  double a = 0.0;
  for (int k = tid; k < (int)gridDim.x; k += blockDim.x) a += __ldcg(d.red_b + k);
  a = block_sum(a, red);
  if (tid != 0) return;
  if (PEER) *d.xw.counter() += 1ull;                     // every CTA has read its slots
  const int done = (a == 0.0 || !(a == a)) ? 1 : 0;
  if (!(a == a)) atomicOr(&sc->flags, FLAG_NONFINITE);
  sc->gg = a;
  sc->rr = a;
  sc->rz = 0.0;
  sc->alpha = 0.0;
  sc->beta = 0.0;
  sc->pcg_iters = 0;
  sc->bar_a = 0u;
  sc->pcg_done = done;
  if (d.host_sig) *(volatile unsigned long long*)d.host_sig = ((unsigned long long)epoch << 32) | (done ? 1ull : 0ull);
  FIN_STAMP(7);
}

static int fin_grid(const BaDev& d) {
  int g = (d.V + FIN_WARPS - 1) / FIN_WARPS;
  return g < 1 ? 1 : (g > 2048 ? 2048 : g);           // red_b holds one partial per CTA
}

void launch_cam_finish(const BaDev& d, bool peer, bool rows_f32, bool sum_cost, int cost_slot, unsigned epoch,
                       bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc) {
  const int rf = rows_f32 ? 1 : 0;
  const int cp = sum_cost ? d.G : 0;                   // the linearisation left one cost partial per group
  if (peer) {
    k_cam_finish<true><<<fin_grid(d), 32 * FIN_WARPS, 0, st>>>(d, cost_slot, epoch, dvdc_quirk ? 1 : 0, rf, cp);
    lc.launches++;
  } else {
    k_cam_finish<false><<<fin_grid(d), 32 * FIN_WARPS, 0, st>>>(d, cost_slot, epoch, dvdc_quirk ? 1 : 0, rf, cp);
    lc.launches++;
  }
}

// Linearises the groups [g0, g1) (g1 < 0: all).  The host-buffer solve calls it once per upload range, so the
// first linearisation overlaps the arrival of the later points; sum_cost closes the last range.
void launch_linearize(const BaDev& d, const TileShape& ts, bool dvdc_quirk, int cost_slot, cudaStream_t st,
                      LaunchCounter& lc, int g0, int g1, bool sum_cost) {
  GramShape sh;
  sh.pt_cap = (ts.max_group_pts + 1) & ~1;
  sh.seg_cap = ts.max_nseg;
  sh.k_cap = ts.max_k;
  sh.slot_cap = ((ts.max_chunk_len + 15) & ~15) + 1;   // = 1 mod 16: the slot planes start in different banks
  const int a = SLOT_PLANES * sh.slot_cap;
  const int bb = RUN_LANES * (d.schur_diag ? Gram<true>::N : Gram<false>::N);
  sh.un_doubles = ((a > bb ? a : bb) + 1) & ~1;
  const size_t smem = gram_smem_bytes(sh);
  if (g1 < 0) {
    g0 = 0;
    g1 = d.G;
  }
  if (g1 > g0) {
    auto go = [&](auto kernel) {
      cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      kernel<<<g1 - g0, RUN_LANES, smem, st>>>(d, sh, dvdc_quirk ? 1 : 0, g0);
    };
    if (d.huber_delta > 0.0) {
      if (d.schur_diag) go(k_linearize<true, true>); else go(k_linearize<true, false>);
    } else {
      if (d.schur_diag) go(k_linearize<false, true>); else go(k_linearize<false, false>);
    }
    lc.launches++;
  }
  if (sum_cost) launch_sum_cost(d, cost_slot, st, lc);
}

// Several GPUs with peer windows: this rank's sums (and its cost, in the trailing entry) go straight into
// every rank's window over NVLink; k_cam_gather_peers then adds the ranks' slots in rank order.
__global__ void __launch_bounds__(256) k_camsum64_peer(BaDev d) {
  const PeerWindow& xw = d.xw;
  const unsigned long long seq = *xw.counter() + 1ull;      // advanced by k_cam_gather_peers
  const int par = (int)(seq & 1ull);
  __shared__ double sh[3][CAM_ACC];
  const int cam = blockIdx.x;
  const double a = cam_rows_sum64(d, cam, sh);
  const int row = d.schur_diag ? CAM_ACC : ROW_COMPACT;     // only the live entries travel
  if ((int)threadIdx.x < row)
    for (int w = 0; w < xw.world; ++w) xw.slot(w, par, xw.rank)[(size_t)cam * row + threadIdx.x] = a;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const double c = d.sc->cost_cur;
    for (int w = 0; w < xw.world; ++w) xw.slot(w, par, xw.rank)[(size_t)d.V * row] = c;
  }
  peer_publish(xw, seq, &d.sc->ticket_a);
}

__global__ void __launch_bounds__(256) k_cam_gather_peers(BaDev d, int cost_slot) {
  const PeerWindow& xw = d.xw;
  Scalars* sc = d.sc;
  const unsigned long long seq = *xw.counter() + 1ull;
  const int par = (int)(seq & 1ull);
  peer_wait(xw, seq, sc);
  const int row = d.schur_diag ? CAM_ACC : ROW_COMPACT;
  const size_t n = (size_t)d.V * row + 1;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    double a = 0.0;
    for (int w = 0; w < xw.world; ++w) a += __ldcg(xw.slot(xw.rank, par, w) + i);
    if (i + 1 < n) {
      const size_t cam = i / row;
      d.cam_acc[cam * CAM_ACC + acc_index(d, (int)(i - cam * row))] = a;
    } else {
      sc->cost_cur = a;
      if (cost_slot < SFMBA_MAX_COST_HISTORY) sc->costs[cost_slot] = a;
    }
  }
  if (last_cta_arrives(&sc->ticket_b) && threadIdx.x == 0) *xw.counter() = seq;   // every CTA has read it
}

void launch_camsum64_peer(const BaDev& d, int cost_slot, cudaStream_t st, LaunchCounter& lc) {
  k_camsum64_peer<<<d.V, 256, 0, st>>>(d);
  int g = (int)(((size_t)d.V * (d.schur_diag ? CAM_ACC : ROW_COMPACT) + 256) / 256);
  if (g > 592) g = 592;
  k_cam_gather_peers<<<g, 256, 0, st>>>(d, cost_slot);
  lc.launches += 2;
}

void launch_camsum64(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  k_camsum64<<<d.V, 256, 0, st>>>(d);
  lc.launches++;
}

void launch_cam_blocks(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  k_cam_blocks<<<(d.V + 63) / 64, 64, 0, st>>>(d);
  lc.launches++;
}

}  // namespace sfmba

#ifdef SFMBA_STEP_TRACE
extern "C" __attribute__((visibility("default"))) int sfmba_debug_fin_trace(unsigned long long* out_host) {
  return (int)cudaMemcpyFromSymbol(out_host, sfmba::g_fin_trace, sizeof(sfmba::g_fin_trace));
}
#endif
