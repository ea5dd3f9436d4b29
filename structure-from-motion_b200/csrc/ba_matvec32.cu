// K3 in single precision: y = W D^-1 W^T z with FP32 per-observation arithmetic (BASELINE.json north_star: "FP32 with
// FP64 accumulation of the normal equations").  Same tile structure and the same index as k_matvec (ba_pcg.cu): one CTA
// per group, one lane per run of K camera-sorted observations, pass 1 -> per-point solve -> pass 2 -> segment rows.
//
// What makes FP32 safe here: every quantity a lane touches is small and local.  The linearisation (k_linearize, FP64)
// leaves, per LM iteration, the points relative to their group's origin O_g (xrel = X - O_g, subtracted in FP64, stored
// FP32), the camera centres relative to the same origin per (chunk, segment) (segst: R | C - O_g), and D^-1 in FP32; so
// d = (X - O_g) - (C - O_g) carries no large-coordinate cancellation.  The normal equations themselves (A, D, g, the
// preconditioner, every PCG vector and dot product) stay FP64; the segment rows this kernel writes are FP64 sums of FP32
// run sums.  The operator the PCG sees is therefore S (1 + O(1e-7)) against an exact right-hand side.
//
// Staging: the contiguous arrays of a group (xrel, the record words, the segment rows) arrive by bulk copies
// (cp.async.bulk -> UBLKCP, completion on an mbarrier) issued by one thread; the per-iteration camera row (E, w of z)
// is read by each lane straight from global memory.
#include "ba_internal.h"
#include "sfmba_device.cuh"

#include <mutex>

#include "launch.cuh"
#include "tile.cuh"

namespace sfmba {

// ---- mbarrier + bulk copy (shared::cta addresses are valid shared::cluster addresses of the own CTA) --------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// bytes: a multiple of 16; both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, void* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

struct Mv32Shape {
  int slot_cap, pt_cap, seg_cap, k_cap;
};
constexpr int SEGST = 12;          // floats per static segment row: R[9] (row-major, camera-to-world), C - O_g [3]
constexpr int SEGIT = 16;          // floats per per-iteration camera row: E[9], w[3], wq, pad[3]
#ifndef MV32_MINB
#define MV32_MINB 8    // resident CTAs per SM the register budget is compiled for
#endif
#ifndef MV32_PAIR
#define MV32_PAIR 0    // 1 = two observations of a lane per trip
#endif
constexpr int MV32_STRIDE = 13;    // floats per lane in the run-sum flush (odd: conflict-free)

__host__ __device__ inline size_t mv32_smem_bytes(const Mv32Shape& s) {
  size_t b = 16;                                                   // mbarrier
  b += (size_t)16 * s.pt_cap;                                      // xs   float4[pt_cap]
  b += (size_t)16 * s.pt_cap;                                      // us   float4[pt_cap]  (t_j, then u_j)
  const size_t slots = (size_t)16 * s.slot_cap, flush = (size_t)4 * RUN_LANES * MV32_STRIDE;
  b += ((slots > flush ? slots : flush) + 15) & ~(size_t)15;       // un   float4 slots | run-sum flush
  b += (size_t)4 * SEGST * s.seg_cap;                              // segst
  b += (size_t)4 * s.k_cap * RUN_LANES;                            // ixs
  b += (size_t)8 * ((s.seg_cap + 3) / 4 * 4);                      // segc, rpos
  b += (size_t)2 * (SEG_RUN_STRIDE + 6);                           // srun (padded to 16 B)
  b += (size_t)4 * (s.pt_cap + 4);                                 // rowp
  return b;
}

__global__ void __launch_bounds__(RUN_LANES, MV32_MINB) k_matvec32(BaDev d, Mv32Shape sh, int dvdc_quirk) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  pdl_trigger();                                   // the next kernel may queue up behind this one's last wave
  const int tid = threadIdx.x;
  const int pt_cap = sh.pt_cap, seg_cap = sh.seg_cap, k_cap = sh.k_cap;
  unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem_raw);
  float4* xs = reinterpret_cast<float4*>(smem_raw + 16);
  float4* us = xs + pt_cap;
  float4* un = us + pt_cap;
  const size_t slots_b = (size_t)16 * sh.slot_cap, flush_b = (size_t)4 * RUN_LANES * MV32_STRIDE;
  float* segst = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(un) +
                                          (((slots_b > flush_b ? slots_b : flush_b) + 15) & ~(size_t)15));
  uint32_t* ixs = reinterpret_cast<uint32_t*>(segst + (size_t)SEGST * seg_cap);
  int32_t* segc = reinterpret_cast<int32_t*>(ixs + (size_t)k_cap * RUN_LANES);
  int32_t* rpos = segc + (seg_cap + 3) / 4 * 4;
  uint16_t* srun = reinterpret_cast<uint16_t*>(rpos + (seg_cap + 3) / 4 * 4);
  int32_t* rowp = reinterpret_cast<int32_t*>(srun + SEG_RUN_STRIDE + 6);
  float* flush = reinterpret_cast<float*>(un);

  const int grp = blockIdx.x;
  const GroupMeta gm = d.grp_meta[grp];
  const int np = gm.np, pt0 = gm.pt0, c0 = gm.c0, c1 = gm.c0 + gm.nc;
  ChunkInfo ci;
  ci.obs0 = gm.obs0;
  ci.len = gm.len0;
  ci.nseg = gm.nseg0;
  ci.K = gm.K0;
  unsigned phase = 0;
  if (tid == 0) mbar_init(bar, 1);
  __syncthreads();
  // bulk copies of one chunk's contiguous tables (and, with the first chunk, of the group's points)
  auto issue_chunk = [&](int c, const ChunkInfo& cc, bool with_points) {
    if (tid == 0) {
      // the tables were last touched through the generic proxy (previous chunk); order that before the bulk writes
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      const unsigned b_ix = (unsigned)cc.K * RUN_LANES * 4u;
      const unsigned b_st = (unsigned)cc.nseg * SEGST * 4u;
      const unsigned b_x = with_points ? (unsigned)np * 16u : 0u;
      mbar_expect_tx(bar, b_ix + b_st + b_x);
      if (b_x) bulk_g2s(xs, d.xrel + pt0, b_x, bar);
      if (b_ix) bulk_g2s(ixs, d.rec_ix + (size_t)c * CHUNK_CAP, b_ix, bar);
      if (b_st) bulk_g2s(segst, d.segst + (size_t)c * SEG_CAP * SEGST, b_st, bar);
    }
    // small tables: per-thread asynchronous copies
    const int32_t* sc = d.seg_cam + (size_t)c * SEG_CAP;
    for (int i = tid; i < (cc.nseg + 3) / 4; i += RUN_LANES) cp_async16(segc + 4 * i, sc + 4 * i);
    const int32_t* rp = d.rowpos + (size_t)c * SEG_CAP;
    for (int i = tid; i < (cc.nseg + 3) / 4; i += RUN_LANES) cp_async16(rpos + 4 * i, rp + 4 * i);
    const uint16_t* sr = d.seg_lane0 + (size_t)c * SEG_RUN_STRIDE;
    for (int i = tid; i < (cc.nseg + 2) / 2; i += RUN_LANES) cp_async4(srun + 2 * i, sr + 2 * i);
  };
  if (gm.nc > 0) issue_chunk(c0, ci, true);
  for (int i = tid; i <= np; i += RUN_LANES) cp_async4(rowp + i, d.rowptr + pt0 + i);
  for (int i = tid; i < pt_cap; i += RUN_LANES) us[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  // Everything staged so far is constant during the PCG solve, so under a programmatic dependent launch it overlaps
  // the previous kernel's tail; camit32 and the done flag are that kernel's results.
  pdl_wait();
  if (d.sc->pcg_done) {
    cp_async_wait_all();
    if (gm.nc > 0) mbar_wait(bar, 0);
    return;
  }
  float* ukeep = nullptr;                            // u_j of this matvec, kept for the back-substitution (FP32 planes)
  {
    const int pcg_it = d.sc->pcg_iters;
    if (pcg_it < PCG_UBUF) ukeep = d.ubuf32 + (size_t)pcg_it * 3 * (size_t)d.N + pt0;
  }
  // wait for a chunk's tables
  auto finish_chunk = [&]() {
    cp_async_wait_all();
    mbar_wait(bar, phase);
    phase ^= 1u;
    __syncthreads();
  };
  const bool single = gm.nc == 1;
  float R[9], C[3];
#pragma unroll
  for (int e = 0; e < 9; ++e) R[e] = 0.f;
  C[0] = C[1] = C[2] = 0.f;
  auto load_rc = [&](int seg) {
    const float4* row = reinterpret_cast<const float4*>(segst + (size_t)SEGST * seg);
    const float4 a = row[0], b = row[1], c = row[2];
    R[0] = a.x; R[1] = a.y; R[2] = a.z; R[3] = a.w; R[4] = b.x; R[5] = b.y; R[6] = b.z; R[7] = b.w; R[8] = c.x;
    C[0] = c.y; C[1] = c.z; C[2] = c.w;
  };
  struct P32 {
    float d0, d1, d2, iw, u, v;
  };
  auto project32 = [&](const float4 x) {
    P32 p;
    p.d0 = x.x - C[0];
    p.d1 = x.y - C[1];
    p.d2 = x.z - C[2];
    const float pu = R[0] * p.d0 + R[3] * p.d1 + R[6] * p.d2;
    const float pv = R[1] * p.d0 + R[4] * p.d1 + R[7] * p.d2;
    const float pw = R[2] * p.d0 + R[5] * p.d1 + R[8] * p.d2;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(p.iw) : "f"(pw));   // MUFU.RCP: 1 ulp, one instruction
    p.u = pu * p.iw;
    p.v = pv * p.iw;
    return p;
  };
  // ---- pass 1: t_j = sum_o J_x^T (J_p z_cam)
  for (int c = c0; c < c1; ++c) {
    if (c > c0) {
      ci = d.chunk_info[c];
      issue_chunk(c, ci, false);
    }
    finish_chunk();
    const int K = ci.K;
    float E[9], w[3], wq = 0.f;
    const uint32_t ix_first = ixs[tid];
    if (ix_first != REC_INVALID) {                         // one camera per lane
      const int seg = rec_seg(ix_first);
      load_rc(seg);
      // the per-iteration row of the lane's camera comes straight from global memory (L2): the lanes of a segment
      // read the same 64 bytes
      const float4* row = reinterpret_cast<const float4*>(d.camit32) + (size_t)segc[seg] * (SEGIT / 4);
      const float4 a = __ldcg(row), b = __ldcg(row + 1), cc = __ldcg(row + 2), e = __ldcg(row + 3);
      E[0] = a.x; E[1] = a.y; E[2] = a.z; E[3] = a.w; E[4] = b.x; E[5] = b.y; E[6] = b.z; E[7] = b.w; E[8] = cc.x;
      w[0] = cc.y; w[1] = cc.z; w[2] = cc.w;
      wq = e.x;
    }
    auto pass1 = [&](uint32_t ix) {
      const P32 pr = project32(xs[rec_ptl(ix)]);
      const float g0 = E[0] * pr.d0 + E[1] * pr.d1 + E[2] * pr.d2 - w[0];
      const float g1 = E[3] * pr.d0 + E[4] * pr.d1 + E[5] * pr.d2 - w[1];
      const float g2 = E[6] * pr.d0 + E[7] * pr.d1 + E[8] * pr.d2 - w[2];
      const float a0 = pr.iw * (g0 - pr.u * g2);
      const float a1 = pr.iw * (g1 - pr.v * (g2 + wq));
      const float q0 = pr.iw * a0, q1 = pr.iw * a1;
      const float q2 = -(pr.u * q0 + pr.v * q1);
      float4 t;
      t.x = R[0] * q0 + R[1] * q1 + R[2] * q2;
      t.y = R[3] * q0 + R[4] * q1 + R[5] * q2;
      t.z = R[6] * q0 + R[7] * q1 + R[8] * q2;
      t.w = 0.f;
      return t;
    };
#if MV32_PAIR
    // two observations of the lane per trip: their dependent chains (gather, projection, reciprocal) interleave
    for (int k = 0; k < K; k += 2) {
      const uint32_t ixa = ixs[k * RUN_LANES + tid];
      if (ixa == REC_INVALID) break;
      const uint32_t ixb = (k + 1 < K) ? ixs[(k + 1) * RUN_LANES + tid] : REC_INVALID;
      const bool vb = ixb != REC_INVALID;
      const float4 ta = pass1(ixa);
      const float4 tb = pass1(vb ? ixb : ixa);
      un[rec_pos(ixa)] = ta;
      if (vb) un[rec_pos(ixb)] = tb;
    }
#else
    for (int k = 0; k < K; ++k) {
      const uint32_t ix = ixs[k * RUN_LANES + tid];
      if (ix == REC_INVALID) break;
      un[rec_pos(ix)] = pass1(ix);
    }
#endif
    __syncthreads();
    // one thread per point: its slots are summed in observation order; a single-chunk group has the whole t_j here,
    // so the same thread goes on to u_j = D_j^-1 t_j
    for (int j = tid; j < np; j += RUN_LANES) {
      const int lo = max(rowp[j], ci.obs0) - ci.obs0;
      const int hi = min(rowp[j + 1], ci.obs0 + ci.len) - ci.obs0;
      // D^-1 of the point: asked for before the slot sums, used after them
      const float4* dg = reinterpret_cast<const float4*>(d.dinv32 + 8 * ((size_t)pt0 + j));
      float4 d0 = make_float4(0.f, 0.f, 0.f, 0.f), d1 = d0;  // 00 01 02 11 | 12 22 - -
      if (single) {
        d0 = dg[0];
        d1 = dg[1];
      }
      float4 a = us[j];
      for (int s = lo; s < hi; s += 4) {                     // up to four loads in flight, sums in observation order
        const int n = hi - s;
        const float4 v0 = un[s];
        const float4 v1 = n > 1 ? un[s + 1] : make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 v2 = n > 2 ? un[s + 2] : make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 v3 = n > 3 ? un[s + 3] : make_float4(0.f, 0.f, 0.f, 0.f);
        a.x += v0.x; a.y += v0.y; a.z += v0.z;
        a.x += v1.x; a.y += v1.y; a.z += v1.z;
        a.x += v2.x; a.y += v2.y; a.z += v2.z;
        a.x += v3.x; a.y += v3.y; a.z += v3.z;
      }
      if (single) {
        float4 u;
        u.x = d0.x * a.x + d0.y * a.y + d0.z * a.z;
        u.y = d0.y * a.x + d0.w * a.y + d1.x * a.z;
        u.z = d0.z * a.x + d1.x * a.y + d1.y * a.z;
        u.w = 0.f;
        a = u;
        if (ukeep) {
          ukeep[j] = u.x;
          ukeep[(size_t)d.N + j] = u.y;
          ukeep[2 * (size_t)d.N + j] = u.z;
        }
      }
      us[j] = a;
    }
    __syncthreads();
  }
  if (!single) {
    cp_async_wait_all();
    __syncthreads();
    for (int j = tid; j < np; j += RUN_LANES) {
      const float4 a = us[j];
      const float4* dg = reinterpret_cast<const float4*>(d.dinv32 + 8 * ((size_t)pt0 + j));
      const float4 d0 = dg[0], d1 = dg[1];
      float4 u;
      u.x = d0.x * a.x + d0.y * a.y + d0.z * a.z;
      u.y = d0.y * a.x + d0.w * a.y + d1.x * a.z;
      u.z = d0.z * a.x + d1.x * a.y + d1.y * a.z;
      u.w = 0.f;
      us[j] = u;
      if (ukeep) {
        ukeep[j] = u.x;
        ukeep[(size_t)d.N + j] = u.y;
        ukeep[2 * (size_t)d.N + j] = u.z;
      }
    }
    __syncthreads();
  }
  // ---- pass 2: per camera sums of c~ d^T and c'
  for (int c = c0; c < c1; ++c) {
    if (gm.nc > 1) {                       // single-chunk groups still hold their chunk from pass 1
      ci = d.chunk_info[c];
      issue_chunk(c, ci, false);
      finish_chunk();
    }
    const int K = ci.K;
    float acc[MV_ACC];
#pragma unroll
    for (int e = 0; e < MV_ACC; ++e) acc[e] = 0.f;
    const uint32_t ix_first = ixs[tid];
    const bool any = ix_first != REC_INVALID;
    const float sgn = dvdc_quirk ? 1.f : -1.f;             // the reference's dv/dC sign quirk (campose_processor.py:802-804)
    if (any && gm.nc > 1) load_rc(rec_seg(ix_first));
    auto pass2 = [&](uint32_t ix, float scale) {
      const int ptl = rec_ptl(ix);
      const P32 pr = project32(xs[ptl]);
      const float4 uu = us[ptl];
      const float e0 = R[0] * uu.x + R[3] * uu.y + R[6] * uu.z;
      const float e1 = R[1] * uu.x + R[4] * uu.y + R[7] * uu.z;
      const float e2 = R[2] * uu.x + R[5] * uu.y + R[8] * uu.z;
      const float iws = pr.iw * scale;
      const float b0 = pr.iw * (e0 - pr.u * e2);
      const float b1 = pr.iw * (e1 - pr.v * e2);
      const float ct0 = iws * b0, ct1 = iws * b1;
      const float ct2 = -(pr.u * ct0 + pr.v * ct1);
      acc[0] += ct0 * pr.d0; acc[1] += ct0 * pr.d1; acc[2] += ct0 * pr.d2;
      acc[3] += ct1 * pr.d0; acc[4] += ct1 * pr.d1; acc[5] += ct1 * pr.d2;
      acc[6] += ct2 * pr.d0; acc[7] += ct2 * pr.d1; acc[8] += ct2 * pr.d2;
      acc[9] += ct0;
      acc[10] += ct1;
      acc[11] += sgn * pr.v * ct1 - pr.u * ct0;
    };
#if MV32_PAIR
    for (int k = 0; k < K; k += 2) {
      const uint32_t ixa = ixs[k * RUN_LANES + tid];
      if (ixa == REC_INVALID) break;
      const uint32_t ixb = (k + 1 < K) ? ixs[(k + 1) * RUN_LANES + tid] : REC_INVALID;
      const bool vb = ixb != REC_INVALID;
      pass2(ixa, 1.f);
      pass2(vb ? ixb : ixa, vb ? 1.f : 0.f);
    }
#else
    for (int k = 0; k < K; ++k) {
      const uint32_t ix = ixs[k * RUN_LANES + tid];
      if (ix == REC_INVALID) break;
      pass2(ix, 1.f);
    }
#endif
    if (any) {                                             // (the slots of pass 1 were consumed before the last barrier)
      float* dst = flush + tid * MV32_STRIDE;
#pragma unroll
      for (int e = 0; e < MV_ACC; ++e) dst[e] = acc[e];
    }
    __syncthreads();
    // one thread per (segment, component): its runs are added in lane order; consecutive threads write consecutive
    // floats of a row (the per-camera sums over the rows are FP64, cam_y_gather32)
    // -- the padding floats 12..15 are written too (zeros, by the threads of components 8..11): a row is two whole
    // 32-byte sectors, so L2 never holds a partially written sector that the reader's load would first have to complete
    // from DRAM
    for (int i = tid; i < ci.nseg * MV_ACC; i += RUN_LANES) {
      const int s = i / MV_ACC, e = i - s * MV_ACC;
      const int ra = srun[s], rb = srun[s + 1];
      float a = 0.f;
      for (int rr = ra; rr < rb; rr += 4) {                  // up to four loads in flight, sums in lane order
        const int n = rb - rr;
        const float* q = flush + rr * MV32_STRIDE + e;
        const float v0 = q[0], v1 = n > 1 ? q[MV32_STRIDE] : 0.f, v2 = n > 2 ? q[2 * MV32_STRIDE] : 0.f;
        const float v3 = n > 3 ? q[3 * MV32_STRIDE] : 0.f;
        a += v0;
        a += v1;
        a += v2;
        a += v3;
      }
      float* row = d.part12f + (size_t)rpos[s] * MV_ROW32;
      row[e] = a;
      if (e >= 8) row[e + 4] = 0.f;
    }
    __syncthreads();
  }
}

void launch_matvec32(const BaDev& d, const TileShape& ts, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc) {
  if (d.G <= 0) return;
  Mv32Shape sh;
  sh.slot_cap = (ts.max_chunk_len + 3) & ~3;
  sh.pt_cap = (ts.max_group_pts + 3) & ~3;
  sh.seg_cap = (ts.max_nseg + 3) & ~3;
  sh.k_cap = ts.max_k;
  const size_t smem = mv32_smem_bytes(sh);
  static std::mutex granted_mu;
  static size_t granted[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  {
    std::lock_guard<std::mutex> lock(granted_mu);
    size_t& g = granted[dev & 63];
    if (smem > g) {
      cudaFuncSetAttribute(k_matvec32, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      g = smem;
    }
  }
  launch_pdl(k_matvec32, dim3(d.G), dim3(RUN_LANES), smem, st, d, sh, dvdc_quirk ? 1 : 0);
  lc.launches++;
}

}  // namespace sfmba
