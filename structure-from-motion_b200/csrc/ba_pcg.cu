// Reduced-camera-system solve: matrix-free implicit-Schur matvec (K3), block-Jacobi PCG,
// back-substitution and parameter update (K4).
//
// Reference: ba_processor.py:382 delta_p = inv(A - B D^-1 B^T)(e_p - B D^-1 e_x); 383-392 p += delta_p,
// quaternion renormalised; 405-406 delta_x = D^-1 (e_x - B^T delta_p), x += delta_x.
//
// q = S p = A p - sum_j W_j D_j^-1 W_j^T p never forms S, B or even J.  With d = X - C, p_c = R^T d,
// (u, v) = p_uv / p_w, iw = 1 / p_w and M = [[1,0,-u],[0,1,-v]] the blocks factor as
//     J_x = iw M R^T,   J_pC = -iw M' R^T,   J_pq = iw M [Q_0 d .. Q_3 d],   Q_j = d(R^T)/dq_j
// (M' = M with the reference's sign quirk in its last column, campose_processor.py:802-804), so
//   pass 1  a   = J_p p_cam = iw (M (E d - w) - quirk term),  E = sum_j p_q[j] Q_j,  w = R^T p_C
//           t_j = sum_o J_x^T a = sum_o R (iw M^T a)                      (slots -> per-point sum)
//   solve   u_j = D_j^-1 t_j
//   pass 2  b   = J_x u_j = iw M (R^T u_j);   c~ = iw M^T b,  c' = iw M'^T b
//           y_C = -R sum_o c',   y_q[j] = < Q_j , sum_o c~ d^T >          (12 sums per camera)
// E and w are per-camera, per-iteration constants (k_cam_iter); the 12 sums of a run stay in
// registers.  Everything is f64, so the operator is linear and symmetric to round-off and PCG
// converges to f64 accuracy.  All reductions have a fixed order.
#include "ba_internal.h"
#include "sfmba_device.cuh"

#include <stdio.h>
#include <stdlib.h>

#include <mutex>

#include "launch.cuh"
#include "peer.cuh"
#include "tile.cuh"

namespace sfmba {

constexpr int MV_STRIDE = 13;   // doubles per run row in shared memory (odd: conflict-free flushes)

// camit rows of an arbitrary 7V-vector (the direction of the tile back-substitution), one thread per camera
__global__ void __launch_bounds__(128) k_cam_rows(BaDev d, const double* __restrict__ vec, int dvdc_quirk) {
  const int cam = blockIdx.x * blockDim.x + threadIdx.x;
  if (cam >= d.V) return;
  double dl[7], o[14];
#pragma unroll
  for (int k = 0; k < 7; ++k) dl[k] = vec[7 * (size_t)cam + k];
  camit_row(d.camtab + (size_t)CAMTAB_STRIDE * cam, dl, dvdc_quirk, o);
  double* dst = d.camit + (size_t)CAMIT_STRIDE * cam;
#pragma unroll
  for (int e = 0; e < 14; ++e) dst[e] = o[e];
  dst[14] = 0.0;
  dst[15] = 0.0;
}

// ---- K3: matrix-free implicit-Schur matvec / back-substitution -----------------------------------
// Everything a group needs is staged into shared memory up front with cp.async (points, D^-1, row
// pointers, record words, segment cameras, run tables), then the camera constants of the chunk's
// segments are gathered; the two passes touch global memory only to write their results.
// BACKSUB: pass 1 with delta_p, then pts_j += v_j - D_j^-1 t_j        (ba_processor.py:405-406)
struct MvShape {
  int slot_cap, pt_cap, seg_cap, k_cap, un_doubles, vn_doubles;
  int ws_doubles;            // Huber weights of the chunk's records (robust solves), else 0
};
constexpr int CAMC = 26;   // doubles per staged camera row: R[9] C[3] E[9] w[3] wq pad

__host__ __device__ inline size_t mv_smem_bytes(const MvShape& s) {
  size_t dbl = (size_t)3 * s.pt_cap + 3 * s.pt_cap + 6 * s.pt_cap + s.vn_doubles + s.un_doubles + (size_t)CAMC * s.seg_cap +
               s.ws_doubles;
  size_t b = dbl * 8;
  b += (size_t)4 * (s.pt_cap + 4);                 // rowp
  b += (size_t)4 * s.k_cap * RUN_LANES;            // ixs
  b += (size_t)8 * ((s.seg_cap + 3) / 4 * 4);      // segc, rpos
  b += (size_t)2 * (SEG_RUN_STRIDE + 6);           // srun (padded to 16 B)
  return b;
}

// One group (blockIdx-independent, so a persistent kernel can walk over groups).  Ends with a
// block-wide barrier unless BACKSUB, so the shared memory can be restaged right away.
template <bool BACKSUB, bool ROBUST>
__device__ __forceinline__ void matvec_group(const BaDev& d, const MvShape& sh, int dvdc_quirk, int grp, double* smem) {
  // The shape words and the lane id pass through an empty asm: when this body sits inside a persistent
  // kernel's group loop, the compiler would otherwise hoist every address derived from them out of the
  // loop and keep them live across it, which costs the inner loops their registers.
  int pt_cap = sh.pt_cap, slot_cap = sh.slot_cap, seg_cap = sh.seg_cap, k_cap = sh.k_cap;
  int un_doubles = sh.un_doubles, vn_doubles = sh.vn_doubles, ws_doubles = sh.ws_doubles;
  int tid = threadIdx.x;
  asm volatile("" : "+r"(pt_cap), "+r"(slot_cap), "+r"(seg_cap), "+r"(k_cap), "+r"(un_doubles), "+r"(vn_doubles),
               "+r"(tid), "+r"(ws_doubles));
  double* Xa = smem;                               // AoS [np][3]
  double* tacc = Xa + 3 * pt_cap;                  // SoA [3][pt_cap]
  double* Dn = tacc + 3 * pt_cap;                  // SoA [6][pt_cap]
  double* Vn = Dn + 6 * pt_cap;                    // SoA [3][pt_cap] v_j (back-substitution only)
  double* un = Vn + vn_doubles;
  double* camc = un + un_doubles;
  ChunkStage cs;
  cs.ws = ROBUST ? camc + (size_t)CAMC * seg_cap : nullptr;
  cs.ixs = reinterpret_cast<uint32_t*>(camc + (size_t)CAMC * seg_cap + ws_doubles);
  cs.segc = reinterpret_cast<int32_t*>(cs.ixs + (size_t)k_cap * RUN_LANES);
  cs.rpos = cs.segc + (seg_cap + 3) / 4 * 4;
  cs.srun = reinterpret_cast<uint16_t*>(cs.rpos + (seg_cap + 3) / 4 * 4);
  cs.bs = nullptr;
  int32_t* rowp = reinterpret_cast<int32_t*>(cs.srun + SEG_RUN_STRIDE + 6);
  const GroupMeta gm = d.grp_meta[grp];
  const int np = gm.np, pt0 = gm.pt0, c0 = gm.c0, c1 = gm.c0 + gm.nc;
  {
    const size_t N = (size_t)d.N;
    const double* src = d.pts + 3 * (size_t)pt0;
    for (int i = tid; i < 3 * np; i += RUN_LANES) cp_async8(Xa + i, src + i);
    for (int i = tid; i <= np; i += RUN_LANES) cp_async4(rowp + i, d.rowptr + pt0 + i);
    for (int j = tid; j < np; j += RUN_LANES) {             // one trip stages all planes of a point
      const double* gd = d.dinv + pt0 + j;
      double* sd = Dn + j;
#pragma unroll
      for (int e = 0; e < 6; ++e) cp_async8(sd + e * pt_cap, gd + e * N);
      if (BACKSUB) {
        const double* gv = d.vj + pt0 + j;
        double* sv = Vn + j;
#pragma unroll
        for (int e = 0; e < 3; ++e) cp_async8(sv + e * pt_cap, gv + e * N);
      }
    }
    for (int i = tid; i < 3 * pt_cap; i += RUN_LANES) tacc[i] = 0.0;
  }
  ChunkInfo ci;
  ci.obs0 = gm.obs0;
  ci.len = gm.len0;
  ci.nseg = gm.nseg0;
  ci.K = gm.K0;
  // Everything staged so far is constant during the PCG solve, so under a programmatic dependent launch
  // it overlaps the previous kernel's tail; camit and the done flag are that kernel's results.
  pdl_wait();
  if ((!BACKSUB && d.sc->pcg_done) || (BACKSUB && (d.sc->flags & FLAG_PEER_TIMEOUT))) {
    cp_async_wait_all();
    return;
  }
  // PCG matvec number pcg_iters of this solve (0-based): its u_j = D_j^-1 W_j^T p are kept, written the moment they
  // are formed.  x = sum alpha_i p_i, so D^-1 B^T delta_p = sum alpha_i u^(i) and the back-substitution of a short
  // solve is one pass over the points (k_backsub_points)
  double* ukeep = nullptr;
  if (!BACKSUB) {
    const int pcg_it = d.sc->pcg_iters;
    if (pcg_it < PCG_UBUF) ukeep = d.ubuf + (size_t)pcg_it * 3 * (size_t)d.N + pt0;
  }
  // stage one chunk: async tables, then gather the camera constants of its segments
  auto stage = [&](int c, bool with_iter) {
    stage_chunk_async(d, c, ci, cs);
    cp_async_wait_all();
    __syncthreads();
    const int parts = with_iter ? 13 : 6;
    const int total = ci.nseg * parts;
    auto fetch = [&](int i, int& s, int& part) -> double2 {
      s = i / parts;
      part = i - s * parts;
      const int cam = cs.segc[s];
      return part < 6 ? reinterpret_cast<const double2*>(d.camtab + (size_t)CAMTAB_STRIDE * cam)[part]
                      : reinterpret_cast<const double2*>(d.camit + (size_t)CAMIT_STRIDE * cam)[part - 6];
    };
    // four rounds of loads in flight before the first store: a round-by-round loop would pay the L2 latency
    // once per round (a typical chunk has ~30 cameras = 3 to 4 rounds)
    for (int i0 = tid; i0 < total; i0 += 4 * RUN_LANES) {
      double2 v[4];
      int sg[4], pa[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * RUN_LANES;
        sg[u] = -1;
        if (i < total) v[u] = fetch(i, sg[u], pa[u]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (sg[u] >= 0) reinterpret_cast<double2*>(camc + (size_t)CAMC * sg[u])[pa[u]] = v[u];
      }
    }
    __syncthreads();
  };
  const bool single = gm.nc == 1;
  // R and C of the lane's camera stay in registers from pass 1 to pass 2 when the group is a single chunk
  // (the lane owns the same run in both passes): the middle phases need few registers
  double R[9], C[3];
#pragma unroll
  for (int e = 0; e < 9; ++e) R[e] = 0.0;
  C[0] = C[1] = C[2] = 0.0;
  // ---- pass 1: t_j = sum_o J_x^T (J_p p_cam)
  for (int c = c0; c < c1; ++c) {
    if (c > c0) ci = d.chunk_info[c];
    stage(c, true);
    const int K = ci.K;
    double E[9], w[3], wq = 0.0;
    const uint32_t ix_first = cs.ixs[tid];
    if (ix_first != REC_INVALID) {                         // one camera per lane
      const double2* row = reinterpret_cast<const double2*>(camc + (size_t)CAMC * rec_seg(ix_first));
      const double2 a = row[0], b = row[1], cc = row[2], e = row[3], f = row[4], g = row[5];
      R[0] = a.x; R[1] = a.y; R[2] = b.x; R[3] = b.y; R[4] = cc.x; R[5] = cc.y; R[6] = e.x; R[7] = e.y; R[8] = f.x;
      C[0] = f.y; C[1] = g.x; C[2] = g.y;
      const double2 h = row[6], i2 = row[7], j2 = row[8], k2 = row[9], l2 = row[10], m2 = row[11], n2 = row[12];
      E[0] = h.x; E[1] = h.y; E[2] = i2.x; E[3] = i2.y; E[4] = j2.x; E[5] = j2.y; E[6] = k2.x; E[7] = k2.y; E[8] = l2.x;
      w[0] = l2.y; w[1] = m2.x; w[2] = m2.y;
      wq = n2.x;
    }
    // two observations per trip: their dependency chains are independent and interleave
    auto pass1 = [&](uint32_t ix, double wgt, double t[3]) {
      const int ptl = rec_ptl(ix);
      const Proj pr = project(R, C, Xa[3 * ptl], Xa[3 * ptl + 1], Xa[3 * ptl + 2]);
      const double g0 = E[0] * pr.d0 + E[1] * pr.d1 + E[2] * pr.d2 - w[0];
      const double g1 = E[3] * pr.d0 + E[4] * pr.d1 + E[5] * pr.d2 - w[1];
      const double g2 = E[6] * pr.d0 + E[7] * pr.d1 + E[8] * pr.d2 - w[2];
      const double a0 = pr.iw * (g0 - pr.u * g2);
      const double a1 = pr.iw * (g1 - pr.v * (g2 + wq));
      const double iww = ROBUST ? pr.iw * wgt : pr.iw;        // J_x^T W (J_p p)
      const double q0 = iww * a0, q1 = iww * a1;
      const double q2 = -(pr.u * q0 + pr.v * q1);
      t[0] = R[0] * q0 + R[1] * q1 + R[2] * q2;
      t[1] = R[3] * q0 + R[4] * q1 + R[5] * q2;
      t[2] = R[6] * q0 + R[7] * q1 + R[8] * q2;
    };
    for (int k = 0; k < K; k += 2) {
      const uint32_t ixa = cs.ixs[k * RUN_LANES + tid];
      if (ixa == REC_INVALID) break;
      uint32_t ixb = (k + 1 < K) ? cs.ixs[(k + 1) * RUN_LANES + tid] : REC_INVALID;
      const bool vb = ixb != REC_INVALID;
      if (!vb) ixb = ixa;
      double ta[3], tb[3];
      const double wa = ROBUST ? cs.ws[k * RUN_LANES + tid] : 1.0;
      const double wb = ROBUST && vb ? cs.ws[(k + 1) * RUN_LANES + tid] : 1.0;
      pass1(ixa, wa, ta);
      pass1(ixb, wb, tb);
      double* sa = un + rec_pos(ixa);
      sa[0] = ta[0];
      sa[slot_cap] = ta[1];
      sa[2 * slot_cap] = ta[2];
      if (vb) {
        double* sb = un + rec_pos(ixb);
        sb[0] = tb[0];
        sb[slot_cap] = tb[1];
        sb[2 * slot_cap] = tb[2];
      }
    }
    __syncthreads();
    // one thread per point: its slots are summed in observation order, three components side by side.  A
    // single-chunk group has the whole t_j here, so the same thread goes on to u_j = D_j^-1 t_j (and, in the
    // back-substitution, to the point update): no second pass over the points, no barrier in between.
    for (int j = tid; j < np; j += RUN_LANES) {
      const int lo = max(rowp[j], ci.obs0) - ci.obs0;
      const int hi = min(rowp[j + 1], ci.obs0 + ci.len) - ci.obs0;
      double a0 = tacc[j], a1 = tacc[pt_cap + j], a2 = tacc[2 * pt_cap + j];
      for (int s = lo; s < hi; s += 3) {                 // up to nine loads in flight, sums in observation order
        const int n = hi - s;
        const double* q = un + s;
        const double v00 = q[0], v10 = q[slot_cap], v20 = q[2 * slot_cap];
        const double v01 = n > 1 ? q[1] : 0.0, v11 = n > 1 ? q[slot_cap + 1] : 0.0, v21 = n > 1 ? q[2 * slot_cap + 1] : 0.0;
        const double v02 = n > 2 ? q[2] : 0.0, v12 = n > 2 ? q[slot_cap + 2] : 0.0, v22 = n > 2 ? q[2 * slot_cap + 2] : 0.0;
        a0 += v00;
        a1 += v10;
        a2 += v20;
        if (n > 1) {
          a0 += v01;
          a1 += v11;
          a2 += v21;
        }
        if (n > 2) {
          a0 += v02;
          a1 += v12;
          a2 += v22;
        }
      }
      if (single) {
        double Dinv[6], u[3];
        const double t[3] = {a0, a1, a2};
#pragma unroll
        for (int e = 0; e < 6; ++e) Dinv[e] = Dn[e * pt_cap + j];
        sym3_mul(Dinv, t, u);
        if (BACKSUB) {
          double* p = d.pts + 3 * ((size_t)pt0 + j);
#pragma unroll
          for (int e = 0; e < 3; ++e) p[e] = Xa[3 * j + e] + (Vn[e * pt_cap + j] - u[e]);
        } else {
          a0 = u[0];
          a1 = u[1];
          a2 = u[2];
          if (ukeep) {
#pragma unroll
            for (int e = 0; e < 3; ++e) ukeep[e * (size_t)d.N + j] = u[e];
          }
        }
      }
      if (!(BACKSUB && single)) {
        tacc[j] = a0;
        tacc[pt_cap + j] = a1;
        tacc[2 * pt_cap + j] = a2;
      }
    }
    __syncthreads();
  }
  if (!single) {
    // ---- u_j = D_j^-1 t_j once every chunk has added its part (a group of unobserved points has no chunk at
    // all: nothing has waited for the staged points yet)
    cp_async_wait_all();
    __syncthreads();
    for (int j = tid; j < np; j += RUN_LANES) {
      double Dinv[6], t[3], u[3];
#pragma unroll
      for (int e = 0; e < 6; ++e) Dinv[e] = Dn[e * pt_cap + j];
      t[0] = tacc[j];
      t[1] = tacc[pt_cap + j];
      t[2] = tacc[2 * pt_cap + j];
      sym3_mul(Dinv, t, u);
      if (BACKSUB) {
        double* p = d.pts + 3 * ((size_t)pt0 + j);
#pragma unroll
        for (int e = 0; e < 3; ++e) p[e] = Xa[3 * j + e] + (Vn[e * pt_cap + j] - u[e]);
      } else {
        tacc[j] = u[0];
        tacc[pt_cap + j] = u[1];
        tacc[2 * pt_cap + j] = u[2];
        if (ukeep) {
#pragma unroll
          for (int e = 0; e < 3; ++e) ukeep[e * (size_t)d.N + j] = u[e];
        }
      }
    }
    if (!BACKSUB) __syncthreads();
  }
  if (BACKSUB) return;
  // ---- pass 2: per camera sums of c~ d^T and c'
  const double sgn = dvdc_quirk ? 1.0 : -1.0;
  for (int c = c0; c < c1; ++c) {
    if (gm.nc > 1) {                       // single-chunk groups still hold their chunk from pass 1
      ci = d.chunk_info[c];
      stage(c, false);
    }
    const int K = ci.K;
    double acc[MV_ACC];
#pragma unroll
    for (int e = 0; e < MV_ACC; ++e) acc[e] = 0.0;
    const uint32_t ix_first = cs.ixs[tid];
    const bool any = ix_first != REC_INVALID;
    if (any && gm.nc > 1) {
      const double2* row = reinterpret_cast<const double2*>(camc + (size_t)CAMC * rec_seg(ix_first));
      const double2 a = row[0], b = row[1], cc = row[2], e = row[3], f = row[4], g = row[5];
      R[0] = a.x; R[1] = a.y; R[2] = b.x; R[3] = b.y; R[4] = cc.x; R[5] = cc.y; R[6] = e.x; R[7] = e.y; R[8] = f.x;
      C[0] = f.y; C[1] = g.x; C[2] = g.y;
    }
    auto pass2 = [&](uint32_t ix, double scale) {
      const int ptl = rec_ptl(ix);
      const Proj pr = project(R, C, Xa[3 * ptl], Xa[3 * ptl + 1], Xa[3 * ptl + 2]);
      const double u0 = tacc[ptl], u1 = tacc[pt_cap + ptl], u2 = tacc[2 * pt_cap + ptl];
      const double e0 = R[0] * u0 + R[3] * u1 + R[6] * u2;
      const double e1 = R[1] * u0 + R[4] * u1 + R[7] * u2;
      const double e2 = R[2] * u0 + R[5] * u1 + R[8] * u2;
      const double iws = pr.iw * scale;
      const double b0 = pr.iw * (e0 - pr.u * e2);
      const double b1 = pr.iw * (e1 - pr.v * e2);
      const double ct0 = iws * b0, ct1 = iws * b1;
      const double ct2 = -(pr.u * ct0 + pr.v * ct1);
      acc[0] += ct0 * pr.d0; acc[1] += ct0 * pr.d1; acc[2] += ct0 * pr.d2;
      acc[3] += ct1 * pr.d0; acc[4] += ct1 * pr.d1; acc[5] += ct1 * pr.d2;
      acc[6] += ct2 * pr.d0; acc[7] += ct2 * pr.d1; acc[8] += ct2 * pr.d2;
      acc[9] += ct0;
      acc[10] += ct1;
      acc[11] += sgn * pr.v * ct1 - pr.u * ct0;
    };
    for (int k = 0; k < K; k += 2) {
      const uint32_t ixa = cs.ixs[k * RUN_LANES + tid];
      if (ixa == REC_INVALID) break;
      uint32_t ixb = (k + 1 < K) ? cs.ixs[(k + 1) * RUN_LANES + tid] : REC_INVALID;
      const bool vb = ixb != REC_INVALID;
      if (!vb) ixb = ixa;
      const double wa = ROBUST ? cs.ws[k * RUN_LANES + tid] : 1.0;
      const double wb = ROBUST && vb ? cs.ws[(k + 1) * RUN_LANES + tid] : 1.0;
      pass2(ixa, wa);
      pass2(ixb, vb ? wb : 0.0);
    }
    if (any) {
      double* dst = un + tid * MV_STRIDE;
#pragma unroll
      for (int e = 0; e < MV_ACC; ++e) dst[e] = acc[e];
    }
    __syncthreads();
    // one thread per (segment, component): its runs are added in lane order; consecutive threads write
    // consecutive doubles of a row
    for (int i = tid; i < ci.nseg * MV_ACC; i += RUN_LANES) {
      const int s = i / MV_ACC, e = i - s * MV_ACC;
      const int ra = cs.srun[s], rb = cs.srun[s + 1];
      const double* src = un + ra * MV_STRIDE + e;
      double a = 0.0;
      for (int rr = ra; rr < rb; rr += 4, src += 4 * MV_STRIDE) {   // up to four loads in flight, sums in run order
        const int n = rb - rr;
        const double v0 = src[0];
        const double v1 = n > 1 ? src[MV_STRIDE] : 0.0;
        const double v2 = n > 2 ? src[2 * MV_STRIDE] : 0.0;
        const double v3 = n > 3 ? src[3 * MV_STRIDE] : 0.0;
        a += v0;
        if (n > 1) a += v1;
        if (n > 2) a += v2;
        if (n > 3) a += v3;
      }
      d.part12[(size_t)cs.rpos[s] * MV_ACC + e] = a;
    }
    __syncthreads();
  }
}

template <bool BACKSUB, bool ROBUST>
__global__ void __launch_bounds__(RUN_LANES, 4) k_matvec(BaDev d, MvShape sh, int dvdc_quirk) {
  extern __shared__ __align__(16) double smem[];
  pdl_trigger();                                   // the next kernel may queue up behind this one's last wave
  matvec_group<BACKSUB, ROBUST>(d, sh, dvdc_quirk, blockIdx.x, smem);
}

// y_i of a camera from the 12 sums over its segment rows (component c in lane c): y_C = -R s, y_q[j] = <Q_j, T>.
// Returns y in lanes 0..6 (0 elsewhere).
__device__ __forceinline__ double cam_y_from_sums(const double* t, int lane, double tsum) {
  double T[MV_ACC];
#pragma unroll
  for (int k = 0; k < MV_ACC; ++k) T[k] = __shfl_sync(0xffffffffu, tsum, k);
  double y = 0.0;
  if (lane < 7) {
    if (lane < 3) {
      y = -(t[3 * lane] * T[9] + t[3 * lane + 1] * T[10] + t[3 * lane + 2] * T[11]);
    } else {
      const double q2[4] = {2.0 * t[12], 2.0 * t[13], 2.0 * t[14], 2.0 * t[15]};
      double Q[4][9];
      quat_dRt(q2, Q);
      // all four products with constant indices and a select at the end: indexing Q by the lane would put the table in
      // local memory (288 bytes per thread, stored and re-read through L1 / L2 -- it was a fifth of k_pcg_step)
      double yq[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        double a = 0.0;
#pragma unroll
        for (int k = 0; k < 9; ++k) a += Q[j][k] * T[k];
        yq[j] = a;
      }
      y = lane == 3 ? yq[0] : (lane == 4 ? yq[1] : (lane == 5 ? yq[2] : yq[3]));
    }
  }
  return y;
}
__device__ __forceinline__ double cam_y_from_sums(const BaDev& d, int cam, int lane, double tsum) {
  return cam_y_from_sums(d.camtab + (size_t)CAMTAB_STRIDE * cam, lane, tsum);
}

// y_i from the camera's segment rows.  The rows of one camera are contiguous in part12 (camrow_ptr), so the whole warp
// streams them with 16-byte loads: lane l's t-th load covers the doubles 2 (32 t + l), +1 of the block, i.e. component
// pair (2 t + l) mod 6 of some row -- three accumulators per lane (t mod 3), six loads in flight, then one fixed-order
// combination of the 96 lane sums through shared memory (`buf`: 96 double2 per warp).
// y_C = -R s, y_q[j] = <Q_j, T>.  Returns y in lanes 0..6 (0 elsewhere).
__device__ __forceinline__ double cam_y_gather(const BaDev& d, int r0, int r1, const double* t, int lane, double2* buf) {
  const double2* src = reinterpret_cast<const double2*>(d.part12 + (size_t)r0 * MV_ACC);
  const int n2 = (r1 - r0) * (MV_ACC / 2);
  double2 a0 = make_double2(0.0, 0.0), a1 = a0, a2 = a0;
  const double2 zero = a0;
  int e = lane;
  for (; e + 5 * 32 < n2; e += 6 * 32) {
    const double2 v0 = __ldcg(src + e), v1 = __ldcg(src + e + 32), v2 = __ldcg(src + e + 64);
    const double2 v3 = __ldcg(src + e + 96), v4 = __ldcg(src + e + 128), v5 = __ldcg(src + e + 160);
    a0.x += v0.x; a0.y += v0.y; a1.x += v1.x; a1.y += v1.y; a2.x += v2.x; a2.y += v2.y;
    a0.x += v3.x; a0.y += v3.y; a1.x += v4.x; a1.y += v4.y; a2.x += v5.x; a2.y += v5.y;
  }
  {                                                        // the last (up to five) loads of the lane, same order
    double2 v[5];
#pragma unroll
    for (int u = 0; u < 5; ++u) v[u] = e + 32 * u < n2 ? __ldcg(src + e + 32 * u) : zero;
    a0.x += v[0].x; a0.y += v[0].y; a1.x += v[1].x; a1.y += v[1].y; a2.x += v[2].x; a2.y += v[2].y;
    a0.x += v[3].x; a0.y += v[3].y; a1.x += v[4].x; a1.y += v[4].y;
  }
  __syncwarp();
  buf[lane] = a0;
  buf[32 + lane] = a1;
  buf[64 + lane] = a2;
  __syncwarp();
  double tsum = 0.0;
  if (lane < MV_ACC) {                                     // lane = component: its 16 contributions in (t, l) order
    const int c = lane >> 1;
#pragma unroll
    for (int ph = 0; ph < 3; ++ph) {
      int l0 = (c - 2 * ph) % 6;
      if (l0 < 0) l0 += 6;
      for (int l = l0; l < 32; l += 6) {
        const double2 v = buf[32 * ph + l];
        tsum += (lane & 1) ? v.y : v.x;
      }
    }
  }
  return cam_y_from_sums(t, lane, tsum);
}
__device__ __forceinline__ double cam_y_gather(const BaDev& d, int cam, int lane, double2* buf) {
  return cam_y_gather(d, d.camrow_ptr[cam], d.camrow_ptr[cam + 1], d.camtab + (size_t)CAMTAB_STRIDE * cam, lane, buf);
}

// The same for the FP32 segment rows the FP32 matvec writes (MV_ROW32 = 16 floats = four 16-byte words per row, the
// last one padding): lane l always reads word l mod 4 of some row, eight rows per warp-load, six loads in flight; FP32
// sums inside a lane (about a dozen terms), FP64 across the lanes in a fixed order.
__device__ __forceinline__ double cam_y_gather32(const BaDev& d, int cam, int lane, double2* buf) {
  const int r0 = d.camrow_ptr[cam], r1 = d.camrow_ptr[cam + 1];
  const float4* src = reinterpret_cast<const float4*>(d.part12f + (size_t)r0 * MV_ROW32);
  const int n4 = (r1 - r0) * (MV_ROW32 / 4);
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 a0 = zero;
  auto add = [](float4& a, const float4 v) { a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; };
  int e = lane;
  for (; e + 5 * 32 < n4; e += 6 * 32) {
    const float4 v0 = __ldcg(src + e), v1 = __ldcg(src + e + 32), v2 = __ldcg(src + e + 64);
    const float4 v3 = __ldcg(src + e + 96), v4 = __ldcg(src + e + 128), v5 = __ldcg(src + e + 160);
    add(a0, v0); add(a0, v1); add(a0, v2); add(a0, v3); add(a0, v4); add(a0, v5);
  }
  {
    float4 v[5];
#pragma unroll
    for (int u = 0; u < 5; ++u) v[u] = e + 32 * u < n4 ? __ldcg(src + e + 32 * u) : zero;
    add(a0, v[0]); add(a0, v[1]); add(a0, v[2]); add(a0, v[3]); add(a0, v[4]);
  }
  float4* bf = reinterpret_cast<float4*>(buf);               // [32] float4
  __syncwarp();
  bf[lane] = a0;
  __syncwarp();
  double tsum = 0.0;
  if (lane < MV_ACC) {                                     // lane = component: word lane / 4 lives in lanes = it mod 4
    const int q = lane >> 2, j = lane & 3;
    const float* f = reinterpret_cast<const float*>(bf);
#pragma unroll
    for (int l = 0; l < 8; ++l) tsum += (double)f[(4 * l + q) * 4 + j];
  }
  return cam_y_from_sums(d, cam, lane, tsum);
}

#ifdef SFMBA_STEP_TRACE
// profiling build only (profiles/step_trace.py): thread 0 of every CTA stamps the phases of the last k_pcg_step launch
__device__ unsigned long long g_step_trace[2][256][16];
#define STEP_STAMP(i)                                                    \
  do {                                                                   \
    if (threadIdx.x == 0 && blockIdx.x < 256) {                          \
      g_step_trace[0][blockIdx.x][i] = global_timer_ns();                \
      g_step_trace[1][blockIdx.x][i] = (unsigned long long)clock64();    \
    }                                                                    \
  } while (0)
#else
#define STEP_STAMP(i) do {} while (0)
#endif

// FP32 rows again, for a warp that owns ONE camera and has registers to spare (k_pcg_step1): three lanes per row (the
// padding word is not read), ten rows per warp-load, GATHER_DEPTH loads of a lane in flight -- the gather is a chain of
// L2 round trips, and this makes it a quarter as long.  FP32 sums inside a lane (two accumulators), FP64 across lanes in
// a fixed order.  Row range and camera-table row come from the caller (fetched before the wait on the matvec).
constexpr int GATHER_DEPTH = 16;
__device__ __forceinline__ double cam_y_gather32_deep(const BaDev& d, int r0, int r1, const double* t, int lane,
                                                      double2* buf) {
  const int nrow = r1 - r0;
  const int sub = lane / 3, word = lane - 3 * sub;
  const float4* mine = reinterpret_cast<const float4*>(d.part12f + (size_t)r0 * MV_ROW32) + 4 * sub + word;
  const int lim = lane < 30 ? nrow - sub : 0;                // this lane reads rows sub + 10 j while 10 j < lim
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 a0 = zero, a1 = zero;
  auto add = [](float4& a, const float4 v) { a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; };
  for (int base = 0; base < nrow; base += 10 * GATHER_DEPTH) {
    float4 v[GATHER_DEPTH];
#pragma unroll
    for (int u = 0; u < GATHER_DEPTH; ++u) v[u] = base + 10 * u < lim ? __ldcg(mine + 4 * (base + 10 * u)) : zero;
#pragma unroll
    for (int u = 0; u < GATHER_DEPTH; u += 2) {
      add(a0, v[u]);
      add(a1, v[u + 1]);
    }
#ifdef SFMBA_STEP_TRACE
    if (base == 0 && threadIdx.x == 0 && blockIdx.x < 256) {
      g_step_trace[1][blockIdx.x][13] = a0.x + a1.x != 12345.f ? (unsigned long long)clock64() : 0ull;
      g_step_trace[1][blockIdx.x][12] = (unsigned long long)nrow;
    }
#endif
  }
  add(a0, a1);
  float4* bf = reinterpret_cast<float4*>(buf);               // [32] float4
  __syncwarp();
  bf[lane] = a0;
  __syncwarp();
  double tsum = 0.0;
  if (lane < MV_ACC) {                                     // lane = component: word lane / 4 lives in lanes = it mod 3
    const int q = lane >> 2, j = lane & 3;
    const float* f = reinterpret_cast<const float*>(bf);
#pragma unroll
    for (int g = 0; g < 10; ++g) tsum += (double)f[(3 * g + q) * 4 + j];
  }
#ifdef SFMBA_STEP_TRACE
  if (threadIdx.x == 0 && blockIdx.x < 256) g_step_trace[1][blockIdx.x][14] = tsum != 12345.0 ? (unsigned long long)clock64() : 0ull;
#endif
  return cam_y_from_sums(t, lane, tsum);
}

constexpr int VEC_THREADS = 512;                 // vector kernels: one warp per camera
constexpr int VEC_WARPS = VEC_THREADS / 32;

// svec = y only (several GPUs without peer windows: all-reduced by NCCL before k_pcg_step<2>)
__global__ void __launch_bounds__(VEC_THREADS) k_camsum12(BaDev d) {
  __shared__ double2 gat[VEC_WARPS][96];
  if (d.sc->pcg_done) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int cam = blockIdx.x * VEC_WARPS + warp; cam < d.V; cam += gridDim.x * VEC_WARPS) {
    const double y = cam_y_gather(d, cam, lane, gat[warp]);
    if (lane < 7) d.svec[7 * (size_t)cam + lane] = y;
  }
}

// Every block sums the per-block partials itself, in block order, so all blocks hold identical scalars.
__device__ __forceinline__ double all_blocks_sum(const double* __restrict__ partials, int stride, int which,
                                                 double* red) {
  double a = 0.0;
  for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) a += __ldcg(partials + (size_t)b * stride + which);
  return block_sum(a, red);
}

// a[lane][:] . v for the camera's 7x7 block (row-major), v held one entry per lane
__device__ __forceinline__ double row7_dot(const double* __restrict__ M49, int lane, double v_lane) {
  double a = 0.0;
#pragma unroll
  for (int b = 0; b < 7; ++b) {
    const double vb = __shfl_sync(0xffffffffu, v_lane, b);
    if (lane < 7) a += M49[lane * 7 + b] * vb;
  }
  return a;
}

// the same with the lane's row of the block already in registers
__device__ __forceinline__ double row7_dot_reg(const double row[7], double v_lane) {
  double a = 0.0;
#pragma unroll
  for (int b = 0; b < 7; ++b) a += row[b] * __shfl_sync(0xffffffffu, v_lane, b);
  return a;
}

// two block-wide sums in one exchange (scratch: 64 doubles); results valid in all threads
__device__ __forceinline__ void block_sum2(double& a, double& b, double* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = (blockDim.x + 31) >> 5;
  a = warp_sum(a);
  b = warp_sum(b);
  __syncthreads();
  if (lane == 0) {
    scratch[warp] = a;
    scratch[32 + warp] = b;
  }
  __syncthreads();
  double ta = 0.0, tb = 0.0;
  for (int w = 0; w < nwarps; ++w) {
    ta += scratch[w];
    tb += scratch[32 + w];
  }
  a = ta;
  b = tb;
}

// NS block-wide sums in one exchange (scratch: NS * 32 + NS doubles); results valid in all threads.  Same order as
// block_sum2 for each of them (xor tree inside a warp, then the warps in order).
template <int NS>
__device__ __forceinline__ void block_sum_n(double (&v)[NS], double* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int i = 0; i < NS; ++i) v[i] = warp_sum(v[i]);
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NS; ++i) scratch[i * 32 + warp] = v[i];
  }
  __syncthreads();
  if (threadIdx.x < NS) {                          // one thread per sum: the warps in order
    double t = 0.0;
    for (int w = 0; w < nwarps; ++w) t += scratch[threadIdx.x * 32 + w];
    scratch[NS * 32 + threadIdx.x] = t;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < NS; ++i) v[i] = scratch[NS * 32 + i];
}

// camit row of a camera whose vector entries sit in lanes 0..6 (every lane computes the row, lanes 0..15 store it)
__device__ __forceinline__ void warp_camit_row(const BaDev& d, int cam, int lane, double v_lane, int dvdc_quirk,
                                               const double* t_pre = nullptr) {
  double dl[7], o[16];
#pragma unroll
  for (int k = 0; k < 7; ++k) dl[k] = __shfl_sync(0xffffffffu, v_lane, k);
  if (t_pre) camit_row(t_pre, dl, dvdc_quirk, o); else camit_row(d.camtab + (size_t)CAMTAB_STRIDE * cam, dl, dvdc_quirk, o);
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
}

// All CTAs of the grid are resident (the launcher sizes the grid to the SM count): arrive, then spin until everyone
// has.  The counter is put back to zero by the CTA that finishes the kernel last.  Bounded like the peer wait.
__device__ __forceinline__ void grid_barrier(uint32_t* counter, Scalars* sc) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    const unsigned long long t0 = global_timer_ns();
    unsigned spins = 0;
    while (*(volatile uint32_t*)counter < gridDim.x) {
      if ((++spins & 0xfffu) == 0 && global_timer_ns() - t0 > PEER_TIMEOUT_NS) {
        atomicOr(&sc->flags, FLAG_PEER_TIMEOUT);
        break;
      }
    }
    __threadfence();
  }
  __syncthreads();
}

// The same with a running arrival count: wait until `target` CTAs have arrived since the counter was last zero.
__device__ __forceinline__ void grid_barrier_gen(uint32_t* counter, uint32_t target, Scalars* sc) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    const unsigned long long t0 = global_timer_ns();
    unsigned spins = 0;
    while (*(volatile uint32_t*)counter < target) {
      if ((++spins & 0xfffu) == 0 && global_timer_ns() - t0 > PEER_TIMEOUT_NS) {
        atomicOr(&sc->flags, FLAG_PEER_TIMEOUT);
        break;
      }
    }
    __threadfence();
  }
  __syncthreads();
}

// y of one PCG matvec, by mode.
//   MODE 1  from this GPU's segment rows
//   MODE 2  svec, all-reduced by NCCL
//   MODE 3  peer windows over NVLink, flag-in-data ("LL") protocol: the warp that owns a camera gathers this rank's partial
//           y, stores each double as one 16-byte word {lo, seq, hi, seq} into its slot of EVERY rank's window, then reads
//           the same camera's words of all ranks from its own window, spinning until both sequence tags match, and adds
//           them in rank order -- every rank gets bit-identical sums.  No fence, no ticket, no separate flag, no grid-wide
//           wait: a warp waits only for the seven numbers it needs.  Two parities: a rank can reach exchange s + 2 only
//           after every peer has finished reading exchange s (they send s + 1 after that).
// MODE 3, first half: the warp gathers this rank's partial y of ALL its cameras and sends them off (the words of the later
// cameras are under way while the first ones are being waited for)
template <int MODE>
__device__ __forceinline__ void pcg_push_y(const BaDev& d, double2* gat_warp, int rows_f32) {
  if (MODE != 3) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const PeerWindow& xw = d.xw;
  const unsigned long long seq64 = *xw.counter() + 1ull;      // advanced by the CTA that finishes the kernel last
  const unsigned seq = (unsigned)seq64;
  const int par = (int)(seq64 & 1ull);
  for (int cam = blockIdx.x * VEC_WARPS + warp; cam < d.V; cam += gridDim.x * VEC_WARPS) {
    const double y = rows_f32 ? cam_y_gather32(d, cam, lane, gat_warp) : cam_y_gather(d, cam, lane, gat_warp);
    if (lane < 7) {
      const size_t k2 = 2 * (7 * (size_t)cam + lane);
      for (int w = 0; w < xw.world; ++w) ll_store(ll_slot(xw, w, par, xw.rank, xw.ll_y) + k2, y, seq);
    }
  }
}

// y of one camera (lanes 0..6).  MODE 1: from this GPU's rows; 2: svec; 3: the words of all ranks, in rank order
template <int MODE>
__device__ __forceinline__ double pcg_cam_y(const BaDev& d, int cam, int lane, double2* gat_warp, int rows_f32) {
  if (MODE == 2) return lane < 7 ? d.svec[7 * (size_t)cam + lane] : 0.0;
  if (MODE == 1) return rows_f32 ? cam_y_gather32(d, cam, lane, gat_warp) : cam_y_gather(d, cam, lane, gat_warp);
  double y = 0.0;
  const PeerWindow& xw = d.xw;
  const unsigned long long seq64 = *xw.counter() + 1ull;
  const unsigned seq = (unsigned)seq64;
  const int par = (int)(seq64 & 1ull);
  if (lane < 7) {
    const size_t k2 = 2 * (7 * (size_t)cam + lane);
    y = ll_sum_ranks(ll_slot(xw, xw.rank, par, 0, xw.ll_y) + k2, xw.cap, xw.world, seq, xw.timeout_ns, d.sc);
  }
  return y;
}

// What the host reads from pinned memory after every iteration: (epoch, slow, iteration, done) in one 8-byte store.
__device__ __forceinline__ unsigned long long pcg_signal(unsigned epoch, int it, int done, bool slow) {
  return ((unsigned long long)epoch << 32) | (slow ? 0x80000000ull : 0ull) | ((unsigned long long)(unsigned)it << 1) |
         (done ? 1ull : 0ull);
}
// A solve being tried on an FP32 linearisation (abort_slow) is given up as soon as its convergence rate says it will need
// more than PCG_FAST_ITERS iterations -- |r|^2 has to fall from |g|^2 to tol^2 |g|^2 within them, i.e. by 1e-2 per
// iteration on average at the default tolerance: the host then redoes the LM iteration from an FP64 linearisation.  The decision is taken
// on the device (every rank takes the same one from the same numbers) and stops the solve like convergence does, so ranks
// need not agree on how many iterations they have queued.
constexpr int PCG_FAST_ITERS = 8;
__device__ __forceinline__ bool pcg_slow(int abort_slow, int done, int it, double rr, double gg, double tol) {
  return abort_slow && !done && it >= 2 && it <= PCG_FAST_ITERS &&
         rr > gg * exp10(2.0 * log10(tol) * (double)it / (double)PCG_FAST_ITERS);
}

// ---- the vector part of the solve ------------------------------------------------------------------------
// Preconditioned conjugate gradients in the Chronopoulos-Gear form: the matvec is applied to z = Minv r (k_matvec reads
// the camit rows of z), and ONE grid-wide reduction per iteration gives both scalars,
//     w = S z = A z - y,   gamma = r.z,   delta = z.w
//     beta = gamma / gamma_old,   alpha = gamma / (delta - beta gamma / alpha_old)          (first iteration: 0, gamma / delta)
//     p = z + beta p;  s = w + beta s (= S p);  x += alpha p;  r -= alpha s;  z = Minv r;  camit rows of the new z
// -- the iterates of textbook PCG in exact arithmetic.  |r|^2 of the new residual is summed by the CTA that finishes
// last (ticket), which also decides convergence and signals the host: nothing later in the kernel waits for it.
// One kernel per iteration, one warp per camera (lane a < 7 owns entry a of the camera's 7-vectors):
//   phase 0  (MODE 3) this rank's partial y from its rows -> its slot in EVERY rank's peer window over NVLink; flags
//   phase 1  y (MODE 1: from this GPU's rows; 2: svec, all-reduced by NCCL; 3: the slots of all ranks, in rank order),
//            w, per-CTA partials of gamma and delta;  grid barrier;  every CTA sums the partials in CTA order
//   phase 2  the updates above, partial |r|^2, ticket
// WIDE: more cameras than one grid of 1-CTA-per-SM warps (V > 16 x SMs, e.g. the Final shape): compiled for two CTAs per SM
// (64 registers: the per-warp constants are then read where they are used instead of being fetched ahead)
template <int MODE, bool WIDE>
__global__ void __launch_bounds__(VEC_THREADS, WIDE ? 2 : 1) k_pcg_step(BaDev d, double tol, int max_iter, unsigned epoch,
                                                                        int dvdc_quirk, int rows_f32, int abort_slow) {
  __shared__ double red[64];
  __shared__ double2 gat[VEC_WARPS][96];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gwarp = blockIdx.x * VEC_WARPS + warp, nwarps = gridDim.x * VEC_WARPS;
  const bool l7 = lane < 7;
  STEP_STAMP(0);
  // What does not change during the solve is fetched before the wait on the matvec: the lane's rows of A and Minv and the
  // camera-table row of the warp's first camera (its only one unless V exceeds the grid's warps)
  double arow[7], mrow[7], trow[16];
#pragma unroll
  for (int b = 0; b < 7; ++b) arow[b] = mrow[b] = 0.0;
#pragma unroll
  for (int b = 0; b < 16; ++b) trow[b] = 0.0;
  if (!WIDE && gwarp < d.V) {
    if (l7) {
#pragma unroll
      for (int b = 0; b < 7; ++b) {
        arow[b] = d.ablk[49 * (size_t)gwarp + lane * 7 + b];
        mrow[b] = d.minv[49 * (size_t)gwarp + lane * 7 + b];
      }
    }
#pragma unroll
    for (int b = 0; b < 16; ++b) trow[b] = d.camtab[(size_t)CAMTAB_STRIDE * gwarp + b];
  }
  STEP_STAMP(1);
  pdl_wait();
  STEP_STAMP(2);
  Scalars* sc = d.sc;
  if (sc->pcg_done) return;
  // scalars of the previous iteration: rewritten only by the CTA that finishes this kernel last
  const double gam_old = sc->rz, alpha_old = sc->alpha, gg = sc->gg;
  const int it_old = sc->pcg_iters;
  pcg_push_y<MODE>(d, gat[warp], rows_f32);
  STEP_STAMP(3);
  double gam = 0.0, del = 0.0;
  for (int cam = gwarp; cam < d.V; cam += nwarps) {
    const size_t k = 7 * (size_t)cam + lane;
    const double y = pcg_cam_y<MODE>(d, cam, lane, gat[warp], rows_f32);
    const double zv = l7 ? d.z[k] : 0.0;
    const double az = (!WIDE && cam == gwarp) ? row7_dot_reg(arow, zv) : row7_dot(d.ablk + 49 * (size_t)cam, lane, zv);
    if (l7) {
      const double wv = az - y;
      d.q[k] = wv;
      gam += d.r[k] * zv;
      del += zv * wv;
    }
  }
  STEP_STAMP(4);
  block_sum2(gam, del, red);
  STEP_STAMP(5);
  if (tid == 0) {
    d.red_a[2 * blockIdx.x] = gam;
    d.red_a[2 * blockIdx.x + 1] = del;
  }
  grid_barrier(&sc->bar_a, sc);
  STEP_STAMP(6);
  pdl_trigger();                                   // every CTA of this grid is resident: the next kernel may queue up
  gam = del = 0.0;
  for (int b = tid; b < (int)gridDim.x; b += blockDim.x) {
    gam += __ldcg(d.red_a + 2 * b);
    del += __ldcg(d.red_a + 2 * b + 1);
  }
  block_sum2(gam, del, red);                       // every CTA sums the partials in CTA order: identical scalars
  const double beta = it_old == 0 ? 0.0 : gam / gam_old;
  const double den = it_old == 0 ? del : del - beta * gam / alpha_old;
  const double alpha = gam / den;
  STEP_STAMP(7);
  double rr = 0.0;
  for (int cam = gwarp; cam < d.V; cam += nwarps) {
    const size_t k = 7 * (size_t)cam + lane;
    double rn = 0.0;
    if (l7) {
      const double pv = d.z[k] + beta * d.p[k];
      const double sv = d.q[k] + beta * d.s[k];
      d.p[k] = pv;
      d.s[k] = sv;
      d.x[k] += alpha * pv;
      rn = d.r[k] - alpha * sv;
      d.r[k] = rn;
      rr += rn * rn;
    }
    const double zn = (!WIDE && cam == gwarp) ? row7_dot_reg(mrow, rn) : row7_dot(d.minv + 49 * (size_t)cam, lane, rn);
    if (l7) d.z[k] = zn;
    if (!WIDE && cam == gwarp) warp_camit_row(d, cam, lane, zn, dvdc_quirk, trow); else warp_camit_row(d, cam, lane, zn, dvdc_quirk);
  }
  STEP_STAMP(8);
  rr = block_sum(rr, red);
  if (tid == 0) d.red_b[blockIdx.x] = rr;
  STEP_STAMP(9);
  const bool last_cta = last_cta_arrives(&sc->ticket_b);
  STEP_STAMP(10);
  if (!last_cta) return;
  // ---- the last CTA: scalars of the iteration
  rr = all_blocks_sum(d.red_b, 1, 0, red);
  if (tid != 0) return;
  const int it = it_old + 1;
  int done = 0;
  uint32_t newflags = 0;
  if (!(den > 0.0) || !(gam > 0.0) || !(rr == rr)) {      // breakdown (S is not positive definite) or NaN
    newflags |= FLAG_NONFINITE;
    done = 1;
  }
  if (rr <= tol * tol * gg) done = 1;
  if (*(volatile uint32_t*)&sc->flags & FLAG_PEER_TIMEOUT) done = 1;
  if (!done && it >= max_iter) {
    newflags |= FLAG_PCG_MAXITER;
    done = 1;
  }
  const bool slow = pcg_slow(abort_slow, done, it, rr, gg, tol);
  if (slow) done = 1;
  if (MODE == 3) *d.xw.counter() += 1ull;                   // this exchange is finished on this rank
  sc->bar_a = 0u;
  sc->pcg_iters_total += 1;
  sc->pcg_iters = it;
  if (newflags) atomicOr(&sc->flags, newflags);
  sc->pq = del;
  sc->alpha = alpha;
  sc->beta = beta;
  if (it_old < PCG_UBUF) {
    sc->alpha_hist[it_old] = alpha;
    sc->beta_hist[it_old] = beta;
  }
  sc->rz = gam;
  sc->rr = rr;
  sc->pcg_done = done;
  if (d.host_sig) *(volatile unsigned long long*)d.host_sig = pcg_signal(epoch, it, done, slow);
  STEP_STAMP(11);
}

// The same iteration for the usual case of one camera per warp (V <= 16 x the grid; MODE 1 or 3), arranged so that as
// little as possible sits on the chain of dependent memory round trips that this latency-bound kernel is:
//   * before the wait on the matvec: the camera's row range, its blocks A and Minv and its camera-table row (shared
//     memory, so no registers are held through the gather);
//   * right after it: the camera's entries of z, r, p, s, x and the scalars -- in flight during the gather;
//   * the gather of the FP32 rows with GATHER_DEPTH loads per lane in flight (cam_y_gather32_deep);
//   * the updates use the registers loaded above: no loads after the grid-wide sum.
// (Tried and dropped: the grid-wide sum as a flag-in-data exchange, every CTA polling every CTA's tagged partials -- 12 500
// polling threads on a few L2 lines delayed the stores they were waiting for: 5 us from the last partial to alpha instead
// of 2.6 us with the arrival counter.  Prefetches before the wait to warm the address translations: no gain.)
// The arithmetic (and the order of every sum) is that of k_pcg_step, except inside cam_y_gather32_deep.
constexpr int STEP_SUMS = 8;
template <int MODE>
__global__ void __launch_bounds__(VEC_THREADS, 1) k_pcg_step1(BaDev d, double tol, int max_iter, unsigned epoch,
                                                              int dvdc_quirk, int rows_f32, int abort_slow) {
  __shared__ double red[STEP_SUMS * 32 + STEP_SUMS];
  __shared__ double2 gat[VEC_WARPS][96];
  __shared__ double s_a[VEC_WARPS][49], s_m[VEC_WARPS][49], s_t[VEC_WARPS][16];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cam = blockIdx.x * VEC_WARPS + warp;
  const bool has = cam < d.V, l7 = has && lane < 7;
  const size_t k = 7 * (size_t)(has ? cam : 0) + (lane < 7 ? lane : 0);
  STEP_STAMP(0);
  int r0 = 0, r1 = 0;
  if (has) {
    r0 = d.camrow_ptr[cam];
    r1 = d.camrow_ptr[cam + 1];
    const double* A = d.ablk + 49 * (size_t)cam;
    const double* M = d.minv + 49 * (size_t)cam;
    s_a[warp][lane] = A[lane];
    s_m[warp][lane] = M[lane];
    if (lane < 17) {
      s_a[warp][32 + lane] = A[32 + lane];
      s_m[warp][32 + lane] = M[32 + lane];
    }
    if (lane < 16) s_t[warp][lane] = d.camtab[(size_t)CAMTAB_STRIDE * cam + lane];
  }
  __syncwarp();
  STEP_STAMP(1);
  pdl_wait();
  STEP_STAMP(2);
  Scalars* sc = d.sc;
  // the previous iteration's state of this camera and the scalars (rewritten only by the CTA that finishes this kernel
  // last): issued here, needed after the gather
  double zv = 0.0, rv = 0.0, pv = 0.0, sv = 0.0, xv = 0.0;
  if (l7) {
    zv = __ldcg(d.z + k);
    rv = __ldcg(d.r + k);
    pv = __ldcg(d.p + k);
    sv = __ldcg(d.s + k);
    xv = __ldcg(d.x + k);
  }
  const int done_before = *(volatile int32_t*)&sc->pcg_done;
  const double gam_old = sc->rz, alpha_old = sc->alpha, gg = sc->gg;
  const int it_old = sc->pcg_iters;
  unsigned long long xseq = 0ull;
  if (MODE == 3) xseq = *d.xw.counter() + 1ull;               // advanced by the CTA that finishes the kernel last
  double y = 0.0;
  if (has)
    y = rows_f32 ? cam_y_gather32_deep(d, r0, r1, s_t[warp], lane, gat[warp])
                 : cam_y_gather(d, r0, r1, s_t[warp], lane, gat[warp]);
  if (done_before) return;                                   // (the same in every thread of the grid)
  STEP_STAMP(3);
  if (MODE == 3 && l7) {
    // this rank's partial y into the camera's words of EVERY rank's window, then the words of all ranks, in rank order
    const PeerWindow& xw = d.xw;
    const unsigned seq = (unsigned)xseq;
    const int par = (int)(xseq & 1ull);
    const size_t k2 = 2 * k;
    for (int w = 0; w < xw.world; ++w) ll_store(ll_slot(xw, w, par, xw.rank, xw.ll_y) + k2, y, seq);
    ll_flush();
    y = ll_sum_ranks(ll_slot(xw, xw.rank, par, 0, xw.ll_y) + k2, xw.cap, xw.world, seq, xw.timeout_ns, sc);
  }
  // the sums of the iteration, all through ONE grid-wide reduction: gamma = r.z, delta = z.w, and what gives |r_new|^2
  // without a second one (below): r.r, r.w, r.s, w.w, w.s, s.s with s the previous direction image
  double wv = 0.0;
  double sums[STEP_SUMS];
#pragma unroll
  for (int i = 0; i < STEP_SUMS; ++i) sums[i] = 0.0;
  {
    double az = 0.0;
#pragma unroll
    for (int b = 0; b < 7; ++b) {
      const double zb = __shfl_sync(0xffffffffu, zv, b);
      if (lane < 7) az += s_a[warp][lane * 7 + b] * zb;
    }
    if (l7) {
      wv = az - y;
      sums[0] = rv * zv;
      sums[1] = zv * wv;
      sums[2] = rv * rv;
      sums[3] = rv * wv;
      sums[4] = rv * sv;
      sums[5] = wv * wv;
      sums[6] = wv * sv;
      sums[7] = sv * sv;
    }
  }
  STEP_STAMP(4);
  block_sum_n<STEP_SUMS>(sums, red);
  STEP_STAMP(5);
  if (tid < STEP_SUMS) d.red_a[STEP_SUMS * blockIdx.x + tid] = red[STEP_SUMS * 32 + tid];     // (block_sum_n leaves the totals there)
  // grid-wide barrier by generation: the arrival counter is zero when a PCG solve starts (k_cam_finish, k_pcg_init,
  // k_pcg_refresh) and every iteration of the solve adds the grid to it, so nobody has to put it back
  grid_barrier_gen(&sc->bar_a, (uint32_t)gridDim.x * (uint32_t)(it_old + 1), sc);
  STEP_STAMP(6);
  pdl_trigger();                                   // every CTA of this grid is resident: the next kernel may queue up
#pragma unroll
  for (int i = 0; i < STEP_SUMS; ++i) sums[i] = tid < (int)gridDim.x ? __ldcg(d.red_a + STEP_SUMS * tid + i) : 0.0;
  block_sum_n<STEP_SUMS>(sums, red);               // every CTA sums the partials in CTA order: identical scalars
  const double gam = sums[0], del = sums[1];
  const double beta = it_old == 0 ? 0.0 : gam / gam_old;
  const double den = it_old == 0 ? del : del - beta * gam / alpha_old;
  const double alpha = gam / den;
  STEP_STAMP(7);
  double rn = 0.0;
  if (l7) {
    const double pn = zv + beta * pv;
    const double sn = wv + beta * sv;
    d.p[k] = pn;
    d.s[k] = sn;
    d.x[k] = xv + alpha * pn;
    rn = rv - alpha * sn;
    d.r[k] = rn;
  }
  if (has) {
    double zn = 0.0;
#pragma unroll
    for (int b = 0; b < 7; ++b) {
      const double rb = __shfl_sync(0xffffffffu, rn, b);
      if (lane < 7) zn += s_m[warp][lane * 7 + b] * rb;
    }
    if (l7) d.z[k] = zn;
    warp_camit_row(d, cam, lane, zn, dvdc_quirk, s_t[warp]);
  }
  STEP_STAMP(8);
  if (blockIdx.x != 0 || tid != 0) return;
  // ---- one thread: scalars of the iteration.  |r_new|^2 from the sums above (r_new = r - alpha s_new, s_new = w + beta s):
  // no second grid-wide reduction, nobody waits for a last CTA.  The r.r it starts from is summed afresh every iteration,
  // so the rounding of this formula (a few ulp of r.r) does not accumulate; a value that small against r.r is converged.
  const double rs_new = sums[3] + beta * sums[4];
  const double ss_new = sums[5] + 2.0 * beta * sums[6] + beta * beta * sums[7];
  const double rr_new = sums[2] - 2.0 * alpha * rs_new + alpha * alpha * ss_new;
  const int it = it_old + 1;
  int done = 0;
  uint32_t newflags = 0;
  if (!(den > 0.0) || !(gam > 0.0) || !(rr_new == rr_new)) {      // breakdown (S is not positive definite) or NaN
    newflags |= FLAG_NONFINITE;
    done = 1;
  }
  if (rr_new <= tol * tol * gg) done = 1;
  if (*(volatile uint32_t*)&sc->flags & FLAG_PEER_TIMEOUT) done = 1;
  if (!done && it >= max_iter) {
    newflags |= FLAG_PCG_MAXITER;
    done = 1;
  }
  const bool slow = pcg_slow(abort_slow, done, it, rr_new, gg, tol);
  if (slow) done = 1;
  if (MODE == 3) *d.xw.counter() += 1ull;                   // every CTA has read its words of this exchange (barrier)
  sc->pcg_iters_total += 1;
  sc->pcg_iters = it;
  if (newflags) atomicOr(&sc->flags, newflags);
  sc->pq = del;
  sc->alpha = alpha;
  sc->beta = beta;
  if (it_old < PCG_UBUF) {
    sc->alpha_hist[it_old] = alpha;
    sc->beta_hist[it_old] = beta;
  }
  sc->rz = gam;
  sc->rr = rr_new > 0.0 ? rr_new : 0.0;
  sc->pcg_done = done;
  if (d.host_sig) *(volatile unsigned long long*)d.host_sig = pcg_signal(epoch, it, done, slow);
  STEP_STAMP(11);
}

// x = 0, p = s = 0, r = g, z = Minv g and its camit rows; the CTA that finishes last sets |g|^2 and the iteration state
__global__ void __launch_bounds__(VEC_THREADS) k_pcg_init(BaDev d, unsigned epoch, int dvdc_quirk) {
  __shared__ double red[32];
  pdl_wait();
  pdl_trigger();
  Scalars* sc = d.sc;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool l7 = lane < 7;
  double rr = 0.0;
  for (int cam = blockIdx.x * VEC_WARPS + warp; cam < d.V; cam += gridDim.x * VEC_WARPS) {
    const size_t k = 7 * (size_t)cam + lane;
    double rn = 0.0;
    if (l7) {
      rn = d.gvec[k];
      d.x[k] = 0.0;
      d.p[k] = 0.0;
      d.s[k] = 0.0;
      d.r[k] = rn;
      rr += rn * rn;
    }
    const double zn = row7_dot(d.minv + 49 * (size_t)cam, lane, rn);
    if (l7) d.z[k] = zn;
    warp_camit_row(d, cam, lane, zn, dvdc_quirk);
  }
  rr = block_sum(rr, red);
  if (tid == 0) d.red_b[blockIdx.x] = rr;
  if (!last_cta_arrives(&sc->ticket_b)) return;
  rr = all_blocks_sum(d.red_b, 1, 0, red);
  if (tid != 0) return;
  const int done = (rr == 0.0 || !(rr == rr)) ? 1 : 0;
  if (!(rr == rr)) atomicOr(&sc->flags, FLAG_NONFINITE);
  sc->gg = rr;
  sc->rr = rr;
  sc->rz = 0.0;
  sc->alpha = 0.0;
  sc->beta = 0.0;
  sc->pcg_iters = 0;
  sc->bar_a = 0u;
  sc->pcg_done = done;
  if (d.host_sig) *(volatile unsigned long long*)d.host_sig = ((unsigned long long)epoch << 32) | (done ? 1ull : 0ull);
}

// Restart of a PCG solve from its current x with a freshly computed residual (the matvec before this kernel was applied
// to x): r = g - (A x - y), z = Minv r and its camit rows, p = s = 0, iteration count 0.  Used to finish in FP64 a
// solve whose first iterations ran the FP32 matvec: from here on the residual is a true FP64 one.
template <int MODE>
__global__ void __launch_bounds__(VEC_THREADS) k_pcg_refresh(BaDev d, double tol, unsigned epoch, int dvdc_quirk) {
  const int rows_f32 = 0;                            // (the matvec before a refresh is the FP64 one)
  __shared__ double red[32];
  __shared__ double2 gat[VEC_WARPS][96];
  pdl_wait();
  pdl_trigger();
  Scalars* sc = d.sc;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool l7 = lane < 7;
  pcg_push_y<MODE>(d, gat[warp], rows_f32);
  double rr = 0.0;
  for (int cam = blockIdx.x * VEC_WARPS + warp; cam < d.V; cam += gridDim.x * VEC_WARPS) {
    const size_t k = 7 * (size_t)cam + lane;
    const double y = pcg_cam_y<MODE>(d, cam, lane, gat[warp], rows_f32);
    const double xv = l7 ? d.x[k] : 0.0;
    const double ax = row7_dot(d.ablk + 49 * (size_t)cam, lane, xv);
    double rn = 0.0;
    if (l7) {
      rn = d.gvec[k] - (ax - y);
      d.r[k] = rn;
      d.p[k] = 0.0;
      d.s[k] = 0.0;
      rr += rn * rn;
    }
    const double zn = row7_dot(d.minv + 49 * (size_t)cam, lane, rn);
    if (l7) d.z[k] = zn;
    warp_camit_row(d, cam, lane, zn, dvdc_quirk);
  }
  rr = block_sum(rr, red);
  if (tid == 0) d.red_b[blockIdx.x] = rr;
  if (!last_cta_arrives(&sc->ticket_b)) return;
  rr = all_blocks_sum(d.red_b, 1, 0, red);
  if (tid != 0) return;
  int done = (rr <= tol * tol * sc->gg) ? 1 : 0;
  if (!(rr == rr)) {
    atomicOr(&sc->flags, FLAG_NONFINITE);
    done = 1;
  }
  if (*(volatile uint32_t*)&sc->flags & FLAG_PEER_TIMEOUT) done = 1;
  if (MODE == 3) *d.xw.counter() += 1ull;
  sc->rr = rr;
  sc->rz = 0.0;
  sc->alpha = 0.0;
  sc->beta = 0.0;
  sc->pcg_iters = 0;
  sc->bar_a = 0u;
  sc->pcg_done = done;
  if (d.host_sig) *(volatile unsigned long long*)d.host_sig = ((unsigned long long)epoch << 32) | (done ? 1ull : 0ull);
}

// the solve goes on: the matvec kernels of a finished solve exit at their first instruction
__global__ void k_pcg_reopen(BaDev d) {
  d.sc->pcg_done = 0;
  d.sc->pcg_iters = PCG_UBUF;          // (no u_j is kept from the matvec on x)
}

// Back-substitution of a PCG solve of at most PCG_UBUF iterations, pts_j += v_j - D_j^-1 W_j^T delta_p
// (ba_processor.py:405-406), from what the matvecs left.  Matvec i was applied to z_i and kept u^(i) = D^-1 W^T z_i;
// the directions obey p_i = z_i + beta_i p_(i-1) and delta_p = sum alpha_i p_i, so with P_i = u^(i) + beta_i P_(i-1)
// the wanted D^-1 W^T delta_p is sum_i alpha_i P_i, formed in iteration order.  One thread per point.
// NIT > 0: the number of PCG iterations, known to the host when it launches (it has read the solve's last signal) -- the
// loads of all iterations are then issued before the first is used; NIT = 0 reads the count from the device scalars.
template <typename UT, int NIT>
__global__ void __launch_bounds__(256) k_backsub_points(BaDev d, const UT* __restrict__ ubuf) {
  const size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t N = (size_t)d.N;
  if (j >= N) return;
  double u[3] = {0.0, 0.0, 0.0}, P[3] = {0.0, 0.0, 0.0};
  double* p = d.pts + 3 * j;
  if (NIT > 0) {
    UT ub[NIT > 0 ? NIT : 1][3];
#pragma unroll
    for (int i = 0; i < NIT; ++i)
#pragma unroll
      for (int e = 0; e < 3; ++e) ub[i][e] = __ldcs(ubuf + ((size_t)i * 3 + e) * N + j);
    const double v0 = __ldcs(d.vj + j), v1 = __ldcs(d.vj + N + j), v2 = __ldcs(d.vj + 2 * N + j);
    const double p0 = p[0], p1 = p[1], p2 = p[2];
    if (d.sc->flags & FLAG_PEER_TIMEOUT) return;    // incomplete cross-rank sums: the step is not applied
#pragma unroll
    for (int i = 0; i < NIT; ++i) {
      const double alpha = d.sc->alpha_hist[i], beta = d.sc->beta_hist[i];
#pragma unroll
      for (int e = 0; e < 3; ++e) {
        P[e] = (double)ub[i][e] + beta * P[e];
        u[e] += alpha * P[e];
      }
    }
    p[0] = p0 + (v0 - u[0]);
    p[1] = p1 + (v1 - u[1]);
    p[2] = p2 + (v2 - u[2]);
    return;
  }
  if (d.sc->flags & FLAG_PEER_TIMEOUT) return;      // incomplete cross-rank sums: the step is not applied
  const int iters = d.sc->pcg_iters;
  for (int i = 0; i < iters && i < PCG_UBUF; ++i) {
    const double alpha = d.sc->alpha_hist[i], beta = d.sc->beta_hist[i];
    const UT* ub = ubuf + (size_t)i * 3 * N + j;
#pragma unroll
    for (int e = 0; e < 3; ++e) {
      P[e] = (double)ub[e * N] + beta * P[e];
      u[e] += alpha * P[e];
    }
  }
#pragma unroll
  for (int e = 0; e < 3; ++e) p[e] = p[e] + (d.vj[e * N + j] - u[e]);
}

// cams += delta_p; q /= |q|                      (ba_processor.py:383-392); and the camera-table row of the result
__global__ void k_update_cams(BaDev d) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d.V) return;
  if (d.sc->flags & FLAG_PEER_TIMEOUT) return;      // incomplete cross-rank sums: the step is not applied
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
  cam_table_row(d, c, v);                           // the table row of the new parameters (no k_cam_prepare launch)
}

// ---- adaptive LM (extension; the reference has fixed lambda and no accept/reject) -----------------
__global__ void k_lm_backup(BaDev d) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (size_t)7 * d.V) d.cams_backup[i] = d.cams[i];
  if (i < (size_t)3 * d.N) d.pts_backup[i] = d.pts[i];
}

// Predicted decrease of the damped Gauss-Newton model for the step (delta_p, delta_x) that solves
// (J^T J + lambda I) delta = J^T r:   |r|^2 - |r - J delta|^2 = delta . J^T r + lambda |delta|^2
//   = delta_p . (g + lambda delta_p)  +  sum_j e_x,j^T D_j^-1 e_x,j  +  lambda |delta_x|^2
// with g = e_p - B D^-1 e_x the reduced right-hand side (delta_x = D^-1 (e_x - B^T delta_p)).  Runs after the update:
// delta_x = pts - pts_backup; the e_x^T D^-1 e_x sums come from the linearisation (grp_exv).  The camera part goes to
// sc->model_decrease, this rank's point part to sc->md_pts (summed over ranks with cost_new).
__global__ void __launch_bounds__(256) k_lm_model_decrease(BaDev d) {
  __shared__ double red[32];
  Scalars* sc = d.sc;
  const double lam = sc->lambda;
  const size_t stride = (size_t)gridDim.x * blockDim.x, t0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  double a = 0.0;
  for (size_t i = t0; i < (size_t)3 * d.N; i += stride) {
    const double dx = d.pts[i] - d.pts_backup[i];
    a += lam * dx * dx;
  }
  for (size_t g = t0; g < (size_t)d.G; g += stride) a += d.grp_exv[g];
  a = block_sum(a, red);
  if (threadIdx.x == 0) d.red_a[blockIdx.x] = a;
  if (!last_cta_arrives(&sc->ticket_a)) return;
  double b = 0.0;
  for (int k = threadIdx.x; k < (int)gridDim.x; k += blockDim.x) b += __ldcg(d.red_a + k);
  b = block_sum(b, red);
  double c = 0.0;
  for (int k = threadIdx.x; k < 7 * d.V; k += blockDim.x) c += d.x[k] * (lam * d.x[k] + d.gvec[k]);
  c = block_sum(c, red);
  if (threadIdx.x == 0) {
    sc->md_pts = b;
    sc->model_decrease = c;
  }
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
      const double md = sc->model_decrease + sc->md_pts;
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


static MvShape matvec_shape(const TileShape& ts, bool backsub, bool robust) {
  MvShape sh;
  sh.slot_cap = ((ts.max_chunk_len + 15) & ~15) + 1;
  sh.pt_cap = (ts.max_group_pts + 1) & ~1;
  sh.seg_cap = ts.max_nseg;
  sh.k_cap = ts.max_k;
  const int a = 3 * sh.slot_cap, bb = RUN_LANES * MV_STRIDE;
  sh.un_doubles = ((a > bb ? a : bb) + 1) & ~1;
  sh.vn_doubles = backsub ? 3 * sh.pt_cap : 0;
  sh.ws_doubles = robust ? sh.k_cap * RUN_LANES : 0;
  return sh;
}

// grid of the vector kernels: one warp per camera, never more CTAs than SMs (k_pcg_step holds a grid-wide barrier,
// so all of its CTAs must be resident at once; one 512-thread CTA per SM always is)
static int vec_grid(const BaDev& d) {
  static int sms[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  int& n = sms[dev & 63];
  if (n == 0) {
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 1;
  }
  int g = (d.V + VEC_WARPS - 1) / VEC_WARPS;
  if (g < 1) g = 1;
  return g < n ? g : n;
}

void launch_pcg_init(const BaDev& d, unsigned epoch, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc) {
  launch_pdl(k_pcg_init, dim3(vec_grid(d)), dim3(VEC_THREADS), 0, st, d, epoch, dvdc_quirk ? 1 : 0);
  lc.launches += 1;
}

void launch_matvec(const BaDev& d, const TileShape& ts, bool backsub, bool dvdc_quirk, cudaStream_t st,
                   LaunchCounter& lc) {
  if (d.G <= 0) return;
  const bool robust = d.huber_delta > 0.0;
  const MvShape sh = matvec_shape(ts, backsub, robust);
  const size_t smem = mv_smem_bytes(sh);
  // the opt-in shared memory size is sticky per function and device: raise it only when needed
  static std::mutex granted_mu;                       // contexts of different host threads share this table
  static size_t granted[4][64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  size_t& g = granted[(backsub ? 1 : 0) + (robust ? 2 : 0)][dev & 63];
  auto go = [&](auto kernel) {
    {
      std::lock_guard<std::mutex> lock(granted_mu);
      if (smem > g) {
        cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        g = smem;
      }
    }
    launch_pdl(kernel, dim3(d.G), dim3(RUN_LANES), smem, st, d, sh, dvdc_quirk ? 1 : 0);
  };
  if (backsub) {
    if (robust) go(k_matvec<true, true>); else go(k_matvec<true, false>);
  } else {
    if (robust) go(k_matvec<false, true>); else go(k_matvec<false, false>);
  }
  lc.launches++;
}

void launch_camsum12(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  k_camsum12<<<vec_grid(d), VEC_THREADS, 0, st>>>(d);
  lc.launches++;
}

// the vector part of a PCG iteration (one launch); mode 1: y from this GPU's rows, 2: from svec, 3: peer windows
void launch_pcg_tail(const BaDev& d, int mode, double tol, int max_iter, unsigned epoch, bool dvdc_quirk, bool rows_f32,
                     bool abort_slow, cudaStream_t st, LaunchCounter& lc) {
  const int rf = rows_f32 ? 1 : 0;
  const int ab = abort_slow ? 1 : 0;
  const int q = dvdc_quirk ? 1 : 0;
  const int g1 = vec_grid(d);                          // <= SM count
  static const bool step1_enabled = !(getenv("SFMBA_STEP1") && atoi(getenv("SFMBA_STEP1")) == 0);   // (A/B switch)
  const bool wide = (d.V + VEC_WARPS - 1) / VEC_WARPS > g1;      // more cameras than that grid has warps
  static const bool step1_peer = getenv("SFMBA_STEP1_PEER") && atoi(getenv("SFMBA_STEP1_PEER")) != 0;
  if (!wide && step1_enabled && (mode == 1 || (mode == 3 && step1_peer))) {
    // one camera per warp on one GPU: the latency-trimmed kernel.  With the peer exchange inside it measured SLOWER than
    // k_pcg_step<3> (4 GPUs, same box: 33 against 26 us between the matvecs, profiles/r2_ab32_*), so several GPUs keep
    // the older kernel unless SFMBA_STEP1_PEER=1
    if (mode == 1) launch_pdl(k_pcg_step1<1>, dim3(g1), dim3(VEC_THREADS), 0, st, d, tol, max_iter, epoch, q, rf, ab);
    else launch_pdl(k_pcg_step1<3>, dim3(g1), dim3(VEC_THREADS), 0, st, d, tol, max_iter, epoch, q, rf, ab);
  } else if (!wide) {
    if (mode == 1) launch_pdl(k_pcg_step<1, false>, dim3(g1), dim3(VEC_THREADS), 0, st, d, tol, max_iter, epoch, q, rf, ab);
    else if (mode == 2) launch_pdl(k_pcg_step<2, false>, dim3(g1), dim3(VEC_THREADS), 0, st, d, tol, max_iter, epoch, q, rf, ab);
    else launch_pdl(k_pcg_step<3, false>, dim3(g1), dim3(VEC_THREADS), 0, st, d, tol, max_iter, epoch, q, rf, ab);
  } else {
    // two CTAs per SM are resident at 64 registers: twice the warps, still every CTA of the grid on the chip (the kernel
    // holds a grid-wide barrier)
    int g2 = (d.V + VEC_WARPS - 1) / VEC_WARPS;
    if (g2 > 2 * g1) g2 = 2 * g1;
    if (mode == 1) launch_pdl(k_pcg_step<1, true>, dim3(g2), dim3(VEC_THREADS), 0, st, d, tol, max_iter, epoch, q, rf, ab);
    else if (mode == 2) launch_pdl(k_pcg_step<2, true>, dim3(g2), dim3(VEC_THREADS), 0, st, d, tol, max_iter, epoch, q, rf, ab);
    else launch_pdl(k_pcg_step<3, true>, dim3(g2), dim3(VEC_THREADS), 0, st, d, tol, max_iter, epoch, q, rf, ab);
  }
  lc.launches += 1;
}

// x keeps its value, the residual is recomputed from it: camit rows of x, matvec on x (the caller's, FP64), refresh
void launch_pcg_reopen(const BaDev& d, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc) {
  k_cam_rows<<<(d.V + 127) / 128, 128, 0, st>>>(d, d.x, dvdc_quirk ? 1 : 0);
  k_pcg_reopen<<<1, 1, 0, st>>>(d);
  lc.launches += 2;
}

void launch_pcg_refresh(const BaDev& d, int mode, double tol, unsigned epoch, bool dvdc_quirk, cudaStream_t st,
                        LaunchCounter& lc) {
  const int g = vec_grid(d);
  const int q = dvdc_quirk ? 1 : 0;
  if (mode == 1) {
    launch_pdl(k_pcg_refresh<1>, dim3(g), dim3(VEC_THREADS), 0, st, d, tol, epoch, q);
  } else if (mode == 2) {
    launch_pdl(k_pcg_refresh<2>, dim3(g), dim3(VEC_THREADS), 0, st, d, tol, epoch, q);
  } else {
    launch_pdl(k_pcg_refresh<3>, dim3(g), dim3(VEC_THREADS), 0, st, d, tol, epoch, q);
  }
  lc.launches += 1;
}

// Sum of n <= 32 device scalars over the ranks through the peer windows (flag-in-data words in the PCG area): what the
// cost of a candidate step and the final cost need.  One warp; every rank ends with the same bits.
__global__ void k_peer_sum_scalars(BaDev d, double* vals, int n) {
  const PeerWindow& xw = d.xw;
  const unsigned long long seq64 = *xw.counter() + 1ull;
  const unsigned seq = (unsigned)seq64;
  const int par = (int)(seq64 & 1ull), i = threadIdx.x;
  if (i < n) {
    const double v = vals[i];
    for (int w = 0; w < xw.world; ++w) ll_store(ll_slot(xw, w, par, xw.rank, xw.ll_y) + 2 * i, v, seq);
    vals[i] = ll_sum_ranks(ll_slot(xw, xw.rank, par, 0, xw.ll_y) + 2 * i, xw.cap, xw.world, seq, xw.timeout_ns, d.sc);
  }
  __syncwarp();
  if (i == 0) *xw.counter() = seq64;
}

void launch_peer_sum_scalars(const BaDev& d, double* vals, int n, cudaStream_t st, LaunchCounter& lc) {
  k_peer_sum_scalars<<<1, 32, 0, st>>>(d, vals, n);
  lc.launches++;
}

void launch_cam_rows(const BaDev& d, const double* vec, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc) {
  k_cam_rows<<<(d.V + 127) / 128, 128, 0, st>>>(d, vec, dvdc_quirk ? 1 : 0);
  lc.launches++;
}

template <typename UT>
static void backsub_dispatch(const BaDev& d, const UT* ubuf, int iters, cudaStream_t st) {
  const dim3 g((unsigned)((d.N + 255) / 256)), b(256);
  switch (iters) {
    case 1: k_backsub_points<UT, 1><<<g, b, 0, st>>>(d, ubuf); break;
    case 2: k_backsub_points<UT, 2><<<g, b, 0, st>>>(d, ubuf); break;
    case 3: k_backsub_points<UT, 3><<<g, b, 0, st>>>(d, ubuf); break;
    case 4: k_backsub_points<UT, 4><<<g, b, 0, st>>>(d, ubuf); break;
    case 5: k_backsub_points<UT, 5><<<g, b, 0, st>>>(d, ubuf); break;
    case 6: k_backsub_points<UT, 6><<<g, b, 0, st>>>(d, ubuf); break;
    case 7: k_backsub_points<UT, 7><<<g, b, 0, st>>>(d, ubuf); break;
    case 8: k_backsub_points<UT, 8><<<g, b, 0, st>>>(d, ubuf); break;
    default: k_backsub_points<UT, 0><<<g, b, 0, st>>>(d, ubuf); break;
  }
}

// `iters`: PCG iterations of the solve as the host read them from its last signal (1..PCG_UBUF), or 0 = ask the device
void launch_backsub_points(const BaDev& d, bool from_f32, int iters, cudaStream_t st, LaunchCounter& lc) {
  if (d.N <= 0) return;
  static_assert(PCG_UBUF == 8, "backsub_dispatch lists the iteration counts 1..PCG_UBUF");
  if (from_f32) backsub_dispatch<float>(d, d.ubuf32, iters, st);
  else backsub_dispatch<double>(d, d.ubuf, iters, st);
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
  size_t g = ((size_t)3 * d.N + 4 * 256 - 1) / (4 * 256);
  if (g < 1) g = 1;
  if (g > 592) g = 592;                    // red_a holds one partial per CTA
  k_lm_model_decrease<<<(unsigned)g, 256, 0, st>>>(d);
  lc.launches++;
}

void launch_lm_decide(const BaDev& d, double lam_min, double lam_max, cudaStream_t st, LaunchCounter& lc) {
  size_t n = (size_t)3 * d.N;
  if ((size_t)7 * d.V > n) n = (size_t)7 * d.V;
  k_lm_decide<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d, lam_min, lam_max);
  lc.launches++;
}

}  // namespace sfmba

#ifdef SFMBA_STEP_TRACE
extern "C" __attribute__((visibility("default"))) int sfmba_debug_step_trace(unsigned long long* out_host) {
  return (int)cudaMemcpyFromSymbol(out_host, sfmba::g_step_trace, sizeof(sfmba::g_step_trace));
}
#endif
