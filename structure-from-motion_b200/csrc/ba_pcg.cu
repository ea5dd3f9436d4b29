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
#include <cooperative_groups.h>

#include <stdio.h>
#include <stdlib.h>

#include "tile.cuh"

namespace cg = cooperative_groups;

namespace sfmba {

constexpr int MV_STRIDE = 13;   // doubles per run row in shared memory (odd: conflict-free flushes)

// Q[j][3 c + k] = d R[k][c] / d q_j for q2 = 2 (w,x,y,z)        (campose_processor.py:636-702)
__device__ __forceinline__ void quat_dRt(const double q2[4], double Q[4][9]) {
  const double w2 = q2[0], x2 = q2[1], y2 = q2[2], z2 = q2[3];
  Q[0][0] = 0.0;       Q[0][1] = z2;        Q[0][2] = -y2;
  Q[0][3] = -z2;       Q[0][4] = 0.0;       Q[0][5] = x2;
  Q[0][6] = y2;        Q[0][7] = -x2;       Q[0][8] = 0.0;
  Q[1][0] = 0.0;       Q[1][1] = y2;        Q[1][2] = z2;
  Q[1][3] = y2;        Q[1][4] = -2.0 * x2; Q[1][5] = w2;
  Q[1][6] = z2;        Q[1][7] = -w2;       Q[1][8] = -2.0 * x2;
  Q[2][0] = -2.0 * y2; Q[2][1] = x2;        Q[2][2] = -w2;
  Q[2][3] = x2;        Q[2][4] = 0.0;       Q[2][5] = z2;
  Q[2][6] = w2;        Q[2][7] = z2;        Q[2][8] = -2.0 * y2;
  Q[3][0] = -2.0 * z2; Q[3][1] = w2;        Q[3][2] = x2;
  Q[3][3] = -w2;       Q[3][4] = -2.0 * z2; Q[3][5] = y2;
  Q[3][6] = x2;        Q[3][7] = y2;        Q[3][8] = 0.0;
}

// per camera and per matvec: (PCG) p = z + beta p, then E = sum_j p_q[j] Q_j, w = R^T p_C
__global__ void __launch_bounds__(128) k_cam_iter(BaDev d, const double* __restrict__ vec, int pcg_direction,
                                                  int dvdc_quirk) {
  const int cam = blockIdx.x * blockDim.x + threadIdx.x;
  if (cam >= d.V) return;
  if (pcg_direction && d.sc->pcg_done) return;
  double dl[7];
  if (pcg_direction) {
    const double beta = d.sc->beta;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      const double v = d.z[7 * (size_t)cam + k] + beta * d.p[7 * (size_t)cam + k];
      d.p[7 * (size_t)cam + k] = v;
      dl[k] = v;
    }
  } else {
#pragma unroll
    for (int k = 0; k < 7; ++k) dl[k] = vec[7 * (size_t)cam + k];
  }
  const double* t = d.camtab + (size_t)CAMTAB_STRIDE * cam;
  double q2[4] = {2.0 * t[12], 2.0 * t[13], 2.0 * t[14], 2.0 * t[15]};
  double Q[4][9];
  quat_dRt(q2, Q);
  double* o = d.camit + (size_t)CAMIT_STRIDE * cam;
#pragma unroll
  for (int e = 0; e < 9; ++e) o[e] = dl[3] * Q[0][e] + dl[4] * Q[1][e] + dl[5] * Q[2][e] + dl[6] * Q[3][e];
  double w[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) w[c] = t[c] * dl[0] + t[3 + c] * dl[1] + t[6 + c] * dl[2];
  o[9] = w[0];
  o[10] = w[1];
  o[11] = w[2];
  o[12] = dvdc_quirk ? 2.0 * w[2] : 0.0;
  o[13] = dvdc_quirk ? 1.0 : -1.0;
  o[14] = 0.0;
  o[15] = 0.0;
}

// ---- K3: matrix-free implicit-Schur matvec / back-substitution -----------------------------------
// Everything a group needs is staged into shared memory up front with cp.async (points, D^-1, row
// pointers, record words, segment cameras, run tables), then the camera constants of the chunk's
// segments are gathered; the two passes touch global memory only to write their results.
// BACKSUB: pass 1 with delta_p, then pts_j += v_j - D_j^-1 t_j        (ba_processor.py:405-406)
struct MvShape {
  int slot_cap, pt_cap, seg_cap, k_cap, un_doubles, vn_doubles;
};
constexpr int CAMC = 26;   // doubles per staged camera row: R[9] C[3] E[9] w[3] wq pad

__host__ __device__ inline size_t mv_smem_bytes(const MvShape& s) {
  size_t dbl = (size_t)3 * s.pt_cap + 3 * s.pt_cap + 6 * s.pt_cap + s.vn_doubles + s.un_doubles + (size_t)CAMC * s.seg_cap;
  size_t b = dbl * 8;
  b += (size_t)4 * (s.pt_cap + 4);                 // rowp
  b += (size_t)4 * s.k_cap * RUN_LANES;            // ixs
  b += (size_t)4 * ((s.seg_cap + 3) / 4 * 4);      // segc
  b += (size_t)2 * (SEG_RUN_STRIDE + 6);           // srun (padded to 16 B)
  return b;
}

// One group (blockIdx-independent, so a persistent kernel can walk over groups).  Ends with a
// block-wide barrier unless BACKSUB, so the shared memory can be restaged right away.
template <bool BACKSUB>
__device__ __forceinline__ void matvec_group(const BaDev& d, const MvShape& sh, int dvdc_quirk, int grp, double* smem) {
  // The shape words and the lane id pass through an empty asm: when this body sits inside a persistent
  // kernel's group loop, the compiler would otherwise hoist every address derived from them out of the
  // loop and keep them live across it, which costs the inner loops their registers.
  int pt_cap = sh.pt_cap, slot_cap = sh.slot_cap, seg_cap = sh.seg_cap, k_cap = sh.k_cap;
  int un_doubles = sh.un_doubles, vn_doubles = sh.vn_doubles;
  int tid = threadIdx.x;
  asm volatile("" : "+r"(pt_cap), "+r"(slot_cap), "+r"(seg_cap), "+r"(k_cap), "+r"(un_doubles), "+r"(vn_doubles),
               "+r"(tid));
  double* Xa = smem;                               // AoS [np][3]
  double* tacc = Xa + 3 * pt_cap;                  // SoA [3][pt_cap]
  double* Dn = tacc + 3 * pt_cap;                  // SoA [6][pt_cap]
  double* Vn = Dn + 6 * pt_cap;                    // SoA [3][pt_cap] v_j (back-substitution only)
  double* un = Vn + vn_doubles;
  double* camc = un + un_doubles;
  ChunkStage cs;
  cs.ixs = reinterpret_cast<uint32_t*>(camc + (size_t)CAMC * seg_cap);
  cs.segc = reinterpret_cast<int32_t*>(cs.ixs + (size_t)k_cap * RUN_LANES);
  cs.srun = reinterpret_cast<uint16_t*>(cs.segc + (seg_cap + 3) / 4 * 4);
  cs.bs = nullptr;
  int32_t* rowp = reinterpret_cast<int32_t*>(cs.srun + SEG_RUN_STRIDE + 6);
  const GroupMeta gm = d.grp_meta[grp];
  const int np = gm.np, pt0 = gm.pt0, c0 = gm.c0, c1 = gm.c0 + gm.nc;
  {
    const size_t N = (size_t)d.N;
    const double* src = d.pts + 3 * (size_t)pt0;
    for (int i = tid; i < 3 * np; i += RUN_LANES) cp_async8(Xa + i, src + i);
    for (int i = tid; i <= np; i += RUN_LANES) cp_async4(rowp + i, d.rowptr + pt0 + i);
#pragma unroll
    for (int e = 0; e < 6; ++e) {
      for (int j = tid; j < np; j += RUN_LANES) cp_async8(Dn + e * pt_cap + j, d.dinv + e * N + pt0 + j);
    }
    if (BACKSUB) {
#pragma unroll
      for (int e = 0; e < 3; ++e) {
        for (int j = tid; j < np; j += RUN_LANES) cp_async8(Vn + e * pt_cap + j, d.vj + e * N + pt0 + j);
      }
    }
    for (int i = tid; i < 3 * pt_cap; i += RUN_LANES) tacc[i] = 0.0;
  }
  ChunkInfo ci;
  ci.obs0 = gm.obs0;
  ci.len = gm.len0;
  ci.nseg = gm.nseg0;
  ci.K = gm.K0;
  // stage one chunk: async tables, then gather the camera constants of its segments
  auto stage = [&](int c, bool with_iter) {
    stage_chunk_async(d, c, ci, cs);
    cp_async_wait_all();
    __syncthreads();
    const int parts = with_iter ? 13 : 6;
    for (int i = tid; i < ci.nseg * parts; i += RUN_LANES) {
      const int s = i / parts, part = i - s * parts;
      const int cam = cs.segc[s];
      const double2 v = part < 6 ? reinterpret_cast<const double2*>(d.camtab + (size_t)CAMTAB_STRIDE * cam)[part]
                                 : reinterpret_cast<const double2*>(d.camit + (size_t)CAMIT_STRIDE * cam)[part - 6];
      reinterpret_cast<double2*>(camc + (size_t)CAMC * s)[part] = v;
    }
    __syncthreads();
  };
  // ---- pass 1: t_j = sum_o J_x^T (J_p p_cam)
  for (int c = c0; c < c1; ++c) {
    if (c > c0) ci = d.chunk_info[c];
    stage(c, true);
    const int K = ci.K;
    double R[9], C[3], E[9], w[3], wq = 0.0;
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
    auto pass1 = [&](uint32_t ix, double t[3]) {
      const int ptl = rec_ptl(ix);
      const Proj pr = project(R, C, Xa[3 * ptl], Xa[3 * ptl + 1], Xa[3 * ptl + 2]);
      const double g0 = E[0] * pr.d0 + E[1] * pr.d1 + E[2] * pr.d2 - w[0];
      const double g1 = E[3] * pr.d0 + E[4] * pr.d1 + E[5] * pr.d2 - w[1];
      const double g2 = E[6] * pr.d0 + E[7] * pr.d1 + E[8] * pr.d2 - w[2];
      const double a0 = pr.iw * (g0 - pr.u * g2);
      const double a1 = pr.iw * (g1 - pr.v * (g2 + wq));
      const double q0 = pr.iw * a0, q1 = pr.iw * a1;
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
      pass1(ixa, ta);
      pass1(ixb, tb);
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
    for (int idx = tid; idx < 3 * np; idx += RUN_LANES) {
      const int j = idx / 3, e = idx - 3 * j;
      const int lo = max(rowp[j], ci.obs0) - ci.obs0;
      const int hi = min(rowp[j + 1], ci.obs0 + ci.len) - ci.obs0;
      if (lo < hi) {
        const double* sl = un + e * slot_cap;
        double a = tacc[e * pt_cap + j];
        for (int s = lo; s < hi; ++s) a += sl[s];
        tacc[e * pt_cap + j] = a;
      }
    }
    __syncthreads();
  }
  // ---- u_j = D_j^-1 t_j
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
    }
  }
  if (BACKSUB) return;
  __syncthreads();
  // ---- pass 2: per camera sums of c~ d^T and c'
  const double sgn = dvdc_quirk ? 1.0 : -1.0;
  for (int c = c0; c < c1; ++c) {
    if (gm.nc > 1) {                       // single-chunk groups still hold their chunk from pass 1
      ci = d.chunk_info[c];
      stage(c, false);
    }
    const int K = ci.K;
    double R[9], C[3];
    double acc[MV_ACC];
#pragma unroll
    for (int e = 0; e < MV_ACC; ++e) acc[e] = 0.0;
    const uint32_t ix_first = cs.ixs[tid];
    const bool any = ix_first != REC_INVALID;
    if (any) {
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
      pass2(ixa, 1.0);
      pass2(ixb, vb ? 1.0 : 0.0);
    }
    if (any) {
      double* dst = un + tid * MV_STRIDE;
#pragma unroll
      for (int e = 0; e < MV_ACC; ++e) dst[e] = acc[e];
    }
    __syncthreads();
    for (int idx = tid; idx < ci.nseg * MV_ACC; idx += RUN_LANES) {
      const int s = idx / MV_ACC, e = idx - s * MV_ACC;
      const int ra = cs.srun[s], rb = cs.srun[s + 1];
      double a = 0.0;
      for (int rr = ra; rr < rb; ++rr) a += un[rr * MV_STRIDE + e];
      d.part12[((size_t)c * SEG_CAP + s) * MV_ACC + e] = a;
    }
    __syncthreads();
  }
}

template <bool BACKSUB>
__global__ void __launch_bounds__(RUN_LANES, 4) k_matvec(BaDev d, MvShape sh, int dvdc_quirk) {
  extern __shared__ __align__(16) double smem[];
  if (!BACKSUB && d.sc->pcg_done) return;
  matvec_group<BACKSUB>(d, sh, dvdc_quirk, blockIdx.x, smem);
}

// ---- "last block" finish of a grid-wide sum: per-block partials are summed in block order ----------
// Returns true in every thread of the block that finished last; `total` is then valid in all of them.
template <int NV>
__device__ __forceinline__ bool grid_sum_finish(double (&v)[NV], double* __restrict__ partials, uint32_t* ticket,
                                                double* red, double (&total)[NV]) {
  __shared__ bool s_last;
  __threadfence();   // this block's earlier global writes (q) must be visible to the block that finishes last
#pragma unroll
  for (int i = 0; i < NV; ++i) v[i] = block_sum(v[i], red);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) partials[(size_t)blockIdx.x * NV + i] = v[i];
    __threadfence();
    const uint32_t t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return false;
  __threadfence();
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    double a = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) a += ((volatile double*)partials)[(size_t)b * NV + i];
    total[i] = block_sum(a, red);
  }
  if (threadIdx.x == 0) *ticket = 0u;
  return true;
}

// camit row of one camera for direction dl = (dC, dq): E = sum_j dq[j] Q_j, w = R^T dC, quirk term
__device__ __forceinline__ void write_camit(const BaDev& d, int cam, const double dl[7], int dvdc_quirk) {
  const double* t = d.camtab + (size_t)CAMTAB_STRIDE * cam;
  const double q2[4] = {2.0 * t[12], 2.0 * t[13], 2.0 * t[14], 2.0 * t[15]};
  double Q[4][9];
  quat_dRt(q2, Q);
  double* o = d.camit + (size_t)CAMIT_STRIDE * cam;
#pragma unroll
  for (int e = 0; e < 9; ++e) o[e] = dl[3] * Q[0][e] + dl[4] * Q[1][e] + dl[5] * Q[2][e] + dl[6] * Q[3][e];
  double w[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) w[c] = t[c] * dl[0] + t[3 + c] * dl[1] + t[6 + c] * dl[2];
  o[9] = w[0];
  o[10] = w[1];
  o[11] = w[2];
  o[12] = dvdc_quirk ? 2.0 * w[2] : 0.0;
  o[13] = 0.0;
  o[14] = 0.0;
  o[15] = 0.0;
}

// y_i from the camera's segment rows, one warp per camera: lane = (row slot, component), two row slots
// of 16 lanes each take alternate rows with eight loads in flight, slot sums are combined in fixed
// order.  y_C = -R s, y_q[j] = <Q_j, T>.  Returns y in lanes 0..6 (0 elsewhere).
__device__ __forceinline__ double cam_y_from_rows(const BaDev& d, int cam, int lane, bool live) {
  const int sub = lane & 15, slot = lane >> 4;
  double a = 0.0;
  if (live && sub < MV_ACC) {
    const int r0 = d.camrow_ptr[cam], r1 = d.camrow_ptr[cam + 1];
    int k = r0 + slot;
    for (; k + 14 < r1; k += 16) {                     // rows k, k+2, ..., k+14 of this slot
      int idx[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) idx[u] = d.camrow_idx[k + 2 * u];
      double v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = d.part12[(size_t)idx[u] * MV_ACC + sub];
#pragma unroll
      for (int u = 0; u < 8; ++u) a += v[u];
    }
    for (; k < r1; k += 2) a += d.part12[(size_t)d.camrow_idx[k] * MV_ACC + sub];
  }
  a += __shfl_down_sync(0xffffffffu, a, 16);           // slot 0 + slot 1
  double T[MV_ACC];
#pragma unroll
  for (int e = 0; e < MV_ACC; ++e) T[e] = __shfl_sync(0xffffffffu, a, e);
  double y = 0.0;
  if (live && lane < 7) {
    const double* t = d.camtab + (size_t)CAMTAB_STRIDE * cam;
    if (lane < 3) {
      y = -(t[3 * lane] * T[9] + t[3 * lane + 1] * T[10] + t[3 * lane + 2] * T[11]);
    } else {
      const double q2[4] = {2.0 * t[12], 2.0 * t[13], 2.0 * t[14], 2.0 * t[15]};
      double Q[4][9];
      quat_dRt(q2, Q);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (j == lane - 3) {
#pragma unroll
          for (int e = 0; e < 9; ++e) y += Q[j][e] * T[e];
        }
      }
    }
  }
  return y;
}

// svec = y only (multi-GPU: all-reduced before k_pcg_tail<2>)
__global__ void __launch_bounds__(256) k_camsum12(BaDev d) {
  if (d.sc->pcg_done) return;
  const int sub = threadIdx.x & 31;
  const int cam = blockIdx.x * 8 + (threadIdx.x >> 5);
  const bool live = cam < d.V;
  const double y = cam_y_from_rows(d, cam, sub, live);
  if (live && sub < 7) d.svec[7 * (size_t)cam + sub] = y;
}

// The vector part of one PCG iteration in ONE cooperative launch (grid-wide barriers between phases).
//   MODE 0  init: x = 0, r = g, z = Minv g, rz, gg, p = z, camit(p)
//   MODE 1  phase 1: y from the segment rows, q = A p - y, p.q            (single GPU)
//   MODE 2  phase 1: y = svec (already all-reduced), q = A p - y, p.q     (several GPUs, NCCL)
//   MODE 3  phase 0: this rank's partial y from its rows, stored into every rank's peer window
//           over NVLink; flags; phase 1: y = sum over ranks in rank order (several GPUs, fused)
//   phase 2: alpha = rz / p.q;  x += alpha p;  r' = r - alpha q;  z = Minv r';  r'.z, r'.r
//   phase 3: beta = rz' / rz;  convergence test and host signal;  r = r';  p = z + beta p;  camit(p)
// Every block sums the per-block partials itself, in block order, so all blocks hold identical scalars.
__device__ __forceinline__ double all_blocks_sum(const double* __restrict__ partials, int stride, int which,
                                                 double* red) {
  double a = 0.0;
  for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) a += __ldcg(partials + (size_t)b * stride + which);
  return block_sum(a, red);
}

// ---- system-scope flag accesses for the peer exchange (flags live in peer HBM, reached over NVLink)
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
constexpr unsigned long long PEER_TIMEOUT_NS = 20ull * 1000ull * 1000ull * 1000ull;

// Signal "this rank's slot of exchange `seq` is complete" to every rank, then wait for all ranks.
// Call after a grid-wide barrier that follows the (system-fenced) slot stores; every block waits
// for itself.  A peer that never arrives (crashed rank) raises FLAG_PEER_TIMEOUT instead of hanging.
__device__ __forceinline__ void peer_signal_and_wait(const PeerWindow& xw, unsigned long long seq, Scalars* sc) {
  const int par = (int)(seq & 1ull);
  if (blockIdx.x == 0 && (int)threadIdx.x < xw.world)
    st_release_sys(xw.flags(threadIdx.x) + par * MAX_PEERS + xw.rank, seq);
  if ((int)threadIdx.x < xw.world) {
    const unsigned long long* f = xw.flags(xw.rank) + par * MAX_PEERS + threadIdx.x;
    const unsigned long long t0 = global_timer_ns();
    unsigned spins = 0;
    while (ld_acquire_sys(f) < seq) {
      if ((++spins & 0x3ffu) == 0 && global_timer_ns() - t0 > PEER_TIMEOUT_NS) {
        atomicOr(&sc->flags, FLAG_PEER_TIMEOUT);
        break;
      }
    }
  }
  __syncthreads();
}

template <int MODE>
__global__ void __launch_bounds__(256) k_pcg_tail(BaDev d, double tol, int max_iter, int dvdc_quirk) {
  __shared__ double red[32];
  cg::grid_group grid = cg::this_grid();
  Scalars* sc = d.sc;
  if (MODE != 0 && sc->pcg_done) return;            // uniform over the grid: nobody reaches a barrier
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int gsize = gridDim.x * blockDim.x;
  const int n7 = 7 * d.V;
  double alpha = 0.0;
  const double rz_old = sc->rz;
  if (MODE != 0) {
    double pq = 0.0;
    const int sub = threadIdx.x & 31;
    const int ngroups = gsize >> 5;
    for (int c0 = 0; c0 < d.V; c0 += ngroups) {     // uniform trip count (shuffles inside)
      const int cam = c0 + (gtid >> 5);
      const bool live = cam < d.V;
      double y = 0.0;
      if (live && sub < 7) {
        const size_t k = 7 * (size_t)cam + sub;
        if (MODE == 2) y = d.svec[k];
        const double* row = d.ablk + (size_t)cam * 49 + sub * 7;
        const double* pv = d.p + 7 * (size_t)cam;
        double ap = 0.0;
#pragma unroll
        for (int b = 0; b < 7; ++b) ap += row[b] * pv[b];
        const double qv = ap - y;
        d.q[k] = qv;
        pq += pv[sub] * qv;
      }
    }
    pq = block_sum(pq, red);
    if (threadIdx.x == 0) d.red_a[blockIdx.x] = pq;
    grid.sync();
    const double pq_all = all_blocks_sum(d.red_a, 1, 0, red);
    alpha = rz_old / pq_all;
    if (gtid == 0) {
      sc->pq = pq_all;
      sc->alpha = alpha;
    }
  }
  // ---- phase 2: one thread per (camera, row); r' goes to svec (r is still being read by the row's peers)
  double v0 = 0.0, v1 = 0.0;
  for (int k = gtid; k < n7; k += gsize) {
    const int cam = k / 7, a = k - 7 * cam;
    const size_t base = 7 * (size_t)cam;
    double rn[7];
    if (MODE == 0) {
#pragma unroll
      for (int b = 0; b < 7; ++b) rn[b] = d.gvec[base + b];
      d.x[k] = 0.0;
    } else {
#pragma unroll
      for (int b = 0; b < 7; ++b) rn[b] = d.r[base + b] - alpha * __ldcg(d.q + base + b);
      d.x[k] += alpha * d.p[k];
    }
    const double* row = d.minv + (size_t)cam * 49 + a * 7;
    double z = 0.0;
#pragma unroll
    for (int b = 0; b < 7; ++b) z += row[b] * rn[b];
    d.z[k] = z;
    d.svec[k] = rn[a];
    v0 += rn[a] * z;
    v1 += rn[a] * rn[a];
  }
  v0 = block_sum(v0, red);
  v1 = block_sum(v1, red);
  if (threadIdx.x == 0) {
    d.red_b[2 * blockIdx.x] = v0;
    d.red_b[2 * blockIdx.x + 1] = v1;
  }
  grid.sync();
  const double rz_new = all_blocks_sum(d.red_b, 2, 0, red);
  const double rr = all_blocks_sum(d.red_b, 2, 1, red);
  const double beta = MODE == 0 ? 0.0 : rz_new / rz_old;
  // convergence: every thread evaluates the same scalars; thread 0 of the grid publishes them
  int done = 0, it = 0;
  const double gg = MODE == 0 ? rr : sc->gg;
  uint32_t newflags = 0;
  if (MODE == 0) {
    done = (rr == 0.0 || !(rr == rr)) ? 1 : 0;
    if (!(rr == rr)) newflags |= FLAG_NONFINITE;
  } else {
    it = sc->pcg_iters + 1;
    if (!(rz_old / alpha > 0.0) || !(rr == rr)) {    // p.q <= 0 or NaN
      newflags |= FLAG_NONFINITE;
      done = 1;
    }
    if (rr <= tol * tol * gg) done = 1;
    if (!done && it >= max_iter) {
      newflags |= FLAG_PCG_MAXITER;
      done = 1;
    }
  }
  // ---- phase 3
  for (int k = gtid; k < n7; k += gsize) {
    const double pn = __ldcg(d.z + k) + beta * (MODE == 0 ? 0.0 : d.p[k]);
    d.r[k] = __ldcg(d.svec + k);
    d.p[k] = pn;
  }
  grid.sync();   // p complete (camit needs the 7 entries of a camera), pcg_iters read by everyone
  if (!done) {
    for (int cam = gtid; cam < d.V; cam += gsize) {
      double dl[7];
#pragma unroll
      for (int k = 0; k < 7; ++k) dl[k] = __ldcg(d.p + 7 * (size_t)cam + k);
      write_camit(d, cam, dl, dvdc_quirk);
    }
  }
  if (gtid == 0) {
    if (MODE == 0) sc->gg = rr;
    if (MODE != 0) sc->pcg_iters_total += 1;
    sc->pcg_iters = it;
    if (newflags) atomicOr(&sc->flags, newflags);
    sc->beta = beta;
    sc->rz = rz_new;
    sc->rr = rr;
    sc->pcg_done = done;
  }
}

// ---- the whole PCG solve of one LM iteration in ONE persistent cooperative launch -------------------
// Grid = as many CTAs as are co-resident (4 per SM); a CTA walks over groups g = blockIdx, + gridDim, ...
// for the matvec, then the grid meets at a barrier and runs the vector part with one warp per camera
// (lane a < 7 owns entry a of the camera's 7-vectors; x, r, z, p, q are touched by their owner only,
// so only camit, the segment rows and the dot-product partials cross a grid barrier).
//   init      x = 0, r = g, z = Minv g, p = z, camit(p); rz, gg
//   A         matvec over the groups -> segment rows (part12)
//   [PEER]    partial y of this rank -> every rank's window over NVLink; flags
//   B         y per camera, q = A p - y, p.q
//   C         alpha; x += alpha p; r -= alpha q; z = Minv r; r.z, r.r
//   D         beta; convergence; p = z + beta p; camit(p)
// At exit camit holds the rows of x (the back-substitution direction), so k_matvec<true> follows directly.
__device__ __forceinline__ double row7_dot(const double* __restrict__ M49, int lane, double v_lane) {
  double a = 0.0;
#pragma unroll
  for (int b = 0; b < 7; ++b) {
    const double vb = __shfl_sync(0xffffffffu, v_lane, b);
    if (lane < 7) a += M49[lane * 7 + b] * vb;
  }
  return a;
}

// lanes 0..6 hold a direction (dC, dq) of camera `cam`: lane 0 writes its camit row
__device__ __forceinline__ void camit_from_lanes(const BaDev& d, int cam, int lane, double v_lane, int dvdc_quirk) {
  double dl[7];
#pragma unroll
  for (int b = 0; b < 7; ++b) dl[b] = __shfl_sync(0xffffffffu, v_lane, b);
  if (lane == 0) write_camit(d, cam, dl, dvdc_quirk);
}

// Loop-carried state of the solve lives in shared memory (every CTA holds the same values): nothing
// but addresses stays live in registers across the matvec, whose inner loops need all 128 of them.
struct SolveCarry {
  double rz, gg, rr, alpha, beta;
  unsigned long long xseq, stamp, ns[5];
  int it, done, grp;
  uint32_t newflags;
};

#ifndef SOLVE_CTAS
#define SOLVE_CTAS 3
#endif
template <bool PEER>
__global__ void __launch_bounds__(RUN_LANES, SOLVE_CTAS)
k_pcg_solve(BaDev d, MvShape sh, double tol, int max_iter, int dvdc_quirk) {
  extern __shared__ __align__(16) double smem[];
  __shared__ double red[32];
  __shared__ SolveCarry cy;
  cg::grid_group grid = cg::this_grid();
  const int tid = threadIdx.x;
  const bool clock_thread = blockIdx.x == 0 && tid == 0;
  auto lap = [&](int which) {
    if (clock_thread) {
      const unsigned long long now = global_timer_ns();
      cy.ns[which] += now - cy.stamp;
      cy.stamp = now;
    }
  };
  // ---- init
  {
    const int lane = tid & 31;
    const int gwarp = (blockIdx.x * RUN_LANES + tid) >> 5, nwarps = gridDim.x * (RUN_LANES / 32);
    const bool l7 = lane < 7;
    double v0 = 0.0, v1 = 0.0;
    for (int cam = gwarp; cam < d.V; cam += nwarps) {
      const size_t k = 7 * (size_t)cam + lane;
      const double g = l7 ? d.gvec[k] : 0.0;
      const double z = row7_dot(d.minv + 49 * (size_t)cam, lane, g);
      if (l7) {
        d.x[k] = 0.0;
        d.r[k] = g;
        d.p[k] = z;
        v0 += g * z;
        v1 += g * g;
      }
      camit_from_lanes(d, cam, lane, z, dvdc_quirk);
    }
    v0 = block_sum(v0, red);
    v1 = block_sum(v1, red);
    if (tid == 0) {
      d.red_b[2 * blockIdx.x] = v0;
      d.red_b[2 * blockIdx.x + 1] = v1;
    }
    grid.sync();
    const double rz = all_blocks_sum(d.red_b, 2, 0, red);
    const double gg = all_blocks_sum(d.red_b, 2, 1, red);
    if (tid == 0) {
      cy.rz = rz;
      cy.gg = gg;
      cy.rr = gg;
      cy.alpha = 0.0;
      cy.beta = 0.0;
      cy.it = 0;
      cy.newflags = !(gg == gg) ? FLAG_NONFINITE : 0u;
      cy.done = (gg == 0.0 || !(gg == gg)) ? 1 : 0;
      cy.xseq = PEER ? *d.xw.counter() : 0ull;       // rewritten by one thread at the very end
#pragma unroll
      for (int k = 0; k < 5; ++k) cy.ns[k] = 0ull;
      cy.stamp = global_timer_ns();
    }
    __syncthreads();
  }
  while (!cy.done) {
    // ---- A: matvec.  Groups are handed out by a ticket counter (one per iteration parity; the idle
    // one is re-armed in phase C), fetched one group ahead so the atomic's latency hides behind work.
    {
      uint32_t* ctr = (cy.it & 1) ? &d.sc->ticket_b : &d.sc->ticket_a;
      if (tid == 0) cy.grp = (int)atomicAdd(ctr, 1u);
      __syncthreads();
      int g = cy.grp;
      while (g < d.G) {
        unsigned nxt = 0;
        if (tid == 0) nxt = atomicAdd(ctr, 1u);
        matvec_group<false>(d, sh, dvdc_quirk, g, smem);
        if (tid == 0) cy.grp = (int)nxt;
        __syncthreads();
        g = cy.grp;
      }
    }
    grid.sync();
    lap(0);
    // ---- vector part: one warp per camera
    const int lane = tid & 31;
    const int gwarp = (blockIdx.x * RUN_LANES + tid) >> 5, nwarps = gridDim.x * (RUN_LANES / 32);
    const bool l7 = lane < 7;
    const int V = d.V;
    Scalars* sc = d.sc;
    unsigned long long xseq = cy.xseq;
    if (PEER) {
      const PeerWindow& xw = d.xw;
      ++xseq;
      const int par = (int)(xseq & 1ull);
      for (int cam = gwarp; cam < V; cam += nwarps) {
        const double y = cam_y_from_rows(d, cam, lane, true);
        if (l7) {
          const size_t k = 7 * (size_t)cam + lane;
          for (int w = 0; w < xw.world; ++w) xw.slot(w, par, xw.rank)[k] = y;
        }
      }
      __threadfence_system();
      grid.sync();
      peer_signal_and_wait(xw, xseq, sc);
      lap(1);
    }
    // ---- B: q = A p - y, p.q
    double pq = 0.0;
    for (int cam = gwarp; cam < V; cam += nwarps) {
      const size_t k = 7 * (size_t)cam + lane;
      double y = 0.0;
      if (PEER) {
        if (l7) {
          const int par = (int)(xseq & 1ull);
          for (int w = 0; w < d.xw.world; ++w) y += __ldcg(d.xw.slot(d.xw.rank, par, w) + k);
        }
      } else {
        y = cam_y_from_rows(d, cam, lane, true);
      }
      const double pv = l7 ? d.p[k] : 0.0;
      const double ap = row7_dot(d.ablk + 49 * (size_t)cam, lane, pv);
      if (l7) {
        const double qv = ap - y;
        d.q[k] = qv;
        pq += pv * qv;
      }
    }
    pq = block_sum(pq, red);
    if (tid == 0) d.red_a[blockIdx.x] = pq;
    grid.sync();
    lap(2);
    int it = cy.it;
    if (clock_thread) *((it & 1) ? &sc->ticket_a : &sc->ticket_b) = 0u;   // the next iteration's counter
    const double rz = cy.rz, gg = cy.gg;
    const double pq_all = all_blocks_sum(d.red_a, 1, 0, red);
    const double alpha = rz / pq_all;
    // ---- C: x += alpha p, r -= alpha q, z = Minv r
    double v0 = 0.0, v1 = 0.0;
    for (int cam = gwarp; cam < V; cam += nwarps) {
      const size_t k = 7 * (size_t)cam + lane;
      double rn = 0.0;
      if (l7) {
        d.x[k] += alpha * d.p[k];
        rn = d.r[k] - alpha * d.q[k];
      }
      const double z = row7_dot(d.minv + 49 * (size_t)cam, lane, rn);
      if (l7) {
        d.r[k] = rn;
        d.z[k] = z;
        v0 += rn * z;
        v1 += rn * rn;
      }
    }
    v0 = block_sum(v0, red);
    v1 = block_sum(v1, red);
    if (tid == 0) {
      d.red_b[2 * blockIdx.x] = v0;
      d.red_b[2 * blockIdx.x + 1] = v1;
    }
    grid.sync();
    lap(3);
    const double rz_new = all_blocks_sum(d.red_b, 2, 0, red);
    const double rr = all_blocks_sum(d.red_b, 2, 1, red);
    const double beta = rz_new / rz;
    ++it;
    uint32_t newflags = cy.newflags;
    int done = 0;
    if (!(pq_all > 0.0) || !(rr == rr)) {            // p.q <= 0 or NaN
      newflags |= FLAG_NONFINITE;
      done = 1;
    }
    if (rr <= tol * tol * gg) done = 1;
    if (PEER && (*(volatile uint32_t*)&sc->flags & FLAG_PEER_TIMEOUT)) done = 1;
    if (!done && it >= max_iter) {
      newflags |= FLAG_PCG_MAXITER;
      done = 1;
    }
    // ---- D: next direction (or, when done, the camit rows of x for the back-substitution)
    for (int cam = gwarp; cam < V; cam += nwarps) {
      const size_t k = 7 * (size_t)cam + lane;
      double pn = 0.0;
      if (l7) {
        pn = d.z[k] + beta * d.p[k];
        d.p[k] = pn;
      }
      const double dir = done ? (l7 ? d.x[k] : 0.0) : pn;
      camit_from_lanes(d, cam, lane, dir, dvdc_quirk);
    }
    __syncthreads();                                  // every thread has read the carry of this iteration
    if (tid == 0) {
      cy.rz = rz_new;
      cy.rr = rr;
      cy.alpha = alpha;
      cy.beta = beta;
      cy.it = it;
      cy.done = done;
      cy.newflags = newflags;
      cy.xseq = xseq;
    }
    if (!done) grid.sync();
    else __syncthreads();
    lap(4);
  }
  if (cy.it == 0) {                                   // zero right-hand side: x = 0
    const int lane = tid & 31;
    const int gwarp = (blockIdx.x * RUN_LANES + tid) >> 5, nwarps = gridDim.x * (RUN_LANES / 32);
    for (int cam = gwarp; cam < d.V; cam += nwarps) camit_from_lanes(d, cam, lane, 0.0, dvdc_quirk);
  }
  if (clock_thread) {
    Scalars* sc = d.sc;
    if (PEER) *d.xw.counter() = cy.xseq;
    sc->ticket_a = 0u;
    sc->ticket_b = 0u;
#pragma unroll
    for (int k = 0; k < 5; ++k) sc->pcg_ns[k] += cy.ns[k];
    sc->gg = cy.gg;
    sc->pcg_iters = cy.it;
    sc->pcg_iters_total += cy.it;
    if (cy.newflags) atomicOr(&sc->flags, cy.newflags);
    sc->alpha = cy.alpha;
    sc->beta = cy.beta;
    sc->rz = cy.rz;
    sc->rr = cy.rr;
    sc->pcg_done = 1;
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


static MvShape matvec_shape(const TileShape& ts, bool backsub) {
  MvShape sh;
  sh.slot_cap = ((ts.max_chunk_len + 15) & ~15) + 1;
  sh.pt_cap = (ts.max_group_pts + 1) & ~1;
  sh.seg_cap = ts.max_nseg;
  sh.k_cap = ts.max_k;
  const int a = 3 * sh.slot_cap, bb = RUN_LANES * MV_STRIDE;
  sh.un_doubles = ((a > bb ? a : bb) + 1) & ~1;
  sh.vn_doubles = backsub ? 3 * sh.pt_cap : 0;
  return sh;
}

static int tail_grid(const BaDev& d) {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 1;
  }
  int g = (d.V + 7) / 8;                  // phase 1 wants one warp per camera
  if (g < 1) g = 1;
  return g < sms ? g : sms;               // cooperative launch: every block must be resident
}

template <int MODE>
static void launch_tail(const BaDev& d, double tol, int max_iter, bool dvdc_quirk, cudaStream_t st) {
  BaDev dd = d;
  int quirk = dvdc_quirk ? 1 : 0;
  void* args[] = {&dd, &tol, &max_iter, &quirk};
  cudaLaunchCooperativeKernel((const void*)k_pcg_tail<MODE>, dim3(tail_grid(d)), dim3(256), args, 0, st);
}

void launch_pcg_init(const BaDev& d, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc) {
  launch_tail<0>(d, 0.0, 0, dvdc_quirk, st);
  lc.launches++;
}

void launch_cam_iter(const BaDev& d, const double* vec, bool pcg_direction, bool dvdc_quirk, cudaStream_t st,
                     LaunchCounter& lc) {
  k_cam_iter<<<(d.V + 127) / 128, 128, 0, st>>>(d, vec, pcg_direction ? 1 : 0, dvdc_quirk ? 1 : 0);
  lc.launches++;
}

void launch_matvec(const BaDev& d, const TileShape& ts, bool backsub, bool dvdc_quirk, cudaStream_t st,
                   LaunchCounter& lc) {
  if (d.G <= 0) return;
  const MvShape sh = matvec_shape(ts, backsub);
  const size_t smem = mv_smem_bytes(sh);
  if (backsub) {
    cudaFuncSetAttribute(k_matvec<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_matvec<true><<<d.G, RUN_LANES, smem, st>>>(d, sh, dvdc_quirk ? 1 : 0);
  } else {
    cudaFuncSetAttribute(k_matvec<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_matvec<false><<<d.G, RUN_LANES, smem, st>>>(d, sh, dvdc_quirk ? 1 : 0);
  }
  lc.launches++;
}

// persistent PCG solve: grid = co-resident CTAs (cooperative launch)
template <bool PEER>
static int launch_solve_t(const BaDev& d, const MvShape& sh, size_t smem, double tol, int max_iter, int quirk,
                          cudaStream_t st) {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  cudaFuncSetAttribute(k_pcg_solve<PEER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int per_sm = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg_solve<PEER>, RUN_LANES, smem);
  if (per_sm < 1) return 0;
  int grid = per_sm * sms;
  if (grid > d.G) grid = d.G;
  if (grid > 1024) grid = 1024;                     // red_a / red_b hold one partial per CTA
  static bool told = false;
  if (!told && getenv("SFMBA_DEBUG")) {
    told = true;
    fprintf(stderr, "[sfmba] k_pcg_solve: %d CTAs/SM x %d SMs, grid %d for %d groups, %zu B dynamic smem "
            "(slot_cap %d pt_cap %d seg_cap %d k_cap %d)\n", per_sm, sms, grid, d.G, smem, sh.slot_cap, sh.pt_cap,
            sh.seg_cap, sh.k_cap);
  }
  BaDev dd = d;
  MvShape s2 = sh;
  void* args[] = {&dd, &s2, &tol, &max_iter, &quirk};
  cudaLaunchCooperativeKernel((const void*)k_pcg_solve<PEER>, dim3(grid), dim3(RUN_LANES), args, smem, st);
  return grid;
}

bool launch_pcg_solve(const BaDev& d, const TileShape& ts, bool peer, double tol, int max_iter, bool dvdc_quirk,
                      cudaStream_t st, LaunchCounter& lc) {
  if (d.G <= 0) return false;
  const MvShape sh = matvec_shape(ts, false);
  const size_t smem = mv_smem_bytes(sh);
  const int grid = peer ? launch_solve_t<true>(d, sh, smem, tol, max_iter, dvdc_quirk ? 1 : 0, st)
                        : launch_solve_t<false>(d, sh, smem, tol, max_iter, dvdc_quirk ? 1 : 0, st);
  if (grid <= 0) return false;
  lc.launches++;
  return true;
}

void launch_camsum12(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  k_camsum12<<<(d.V + 7) / 8, 256, 0, st>>>(d);
  lc.launches++;
}

// the vector part of a PCG iteration; from_rows: y from the segment rows (one GPU), else from svec
void launch_pcg_tail(const BaDev& d, double tol, int max_iter, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc) {
  launch_tail<2>(d, tol, max_iter, dvdc_quirk, st);
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
