// Device side of the NVLink peer exchange (PeerWindow in ba_internal.h, host side in peer_window.cu).
// An exchange with sequence number `seq` (the window's counter + 1) goes
//   1. every CTA stores this rank's partial values into slot (seq & 1, rank) of EVERY rank's window,
//   2. peer_publish: the CTA that finishes last raises this rank's flag in every window,
//   3. (next kernel) peer_wait: every CTA waits for all ranks' flags, then sums the slots of its own
//      window in rank order, so all ranks obtain bit-identical sums,
//   4. one thread advances the counter once every CTA of step 3 has read it.
// The only waiting is on flags written by OTHER GPUs; it is bounded by PEER_TIMEOUT_NS.
#pragma once
#include "ba_internal.h"
#include "sfmba_device.cuh"

namespace sfmba {

// system-scope flag accesses (flags live in peer HBM, reached over NVLink)
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

// true in every thread of the CTA that arrives last (its view of the other CTAs' writes is complete)
__device__ __forceinline__ bool last_cta_arrives(uint32_t* ticket) {
  __shared__ bool s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
    if (s_last) *ticket = 0u;
  }
  __syncthreads();
  if (s_last) __threadfence();
  return s_last;
}

// call after the slot stores of this CTA: the last CTA to arrive raises the flags
__device__ __forceinline__ void peer_publish(const PeerWindow& xw, unsigned long long seq, uint32_t* ticket) {
  __threadfence_system();
  if (last_cta_arrives(ticket)) {
    __threadfence_system();
    const int par = (int)(seq & 1ull);
    if ((int)threadIdx.x < xw.world) st_release_sys(xw.flags(threadIdx.x) + par * MAX_PEERS + xw.rank, seq);
  }
}

// every CTA waits until all ranks have raised the flag of exchange `seq` (peers run on other GPUs;
// a peer that never arrives raises FLAG_PEER_TIMEOUT instead of hanging this GPU)
__device__ __forceinline__ void peer_wait(const PeerWindow& xw, unsigned long long seq, Scalars* sc) {
  const int par = (int)(seq & 1ull);
  if ((int)threadIdx.x < xw.world) {
    const unsigned long long* f = xw.flags(xw.rank) + par * MAX_PEERS + threadIdx.x;
    const unsigned long long t0 = global_timer_ns();
    unsigned spins = 0;
    while (ld_acquire_sys(f) < seq) {
      if ((++spins & 0x3ffu) == 0 && global_timer_ns() - t0 > xw.timeout_ns) {
        atomicOr(&sc->flags, FLAG_PEER_TIMEOUT);
        break;
      }
    }
  }
  __syncthreads();
}

// After a warp's words have been stored: push them out (SFMBA_LL_FENCE 1 = membar.sys by the storing lanes).
#ifndef SFMBA_LL_FENCE
#define SFMBA_LL_FENCE 0
#endif
__device__ __forceinline__ void ll_flush() {
#if SFMBA_LL_FENCE == 1
  __threadfence_system();
#elif SFMBA_LL_FENCE == 2
  __threadfence();
#endif
}
#ifndef SFMBA_LL_SLEEP
#define SFMBA_LL_SLEEP 0
#endif
// ---- flag-in-data ("LL") exchange: every double travels as one 16-byte word {lo, seq, hi, seq}; the reader spins until
// both sequence tags match.  No fence, no ticket, no separate flag: a reader waits only for the numbers it needs.
__device__ __forceinline__ void ll_store(double* slot16, double v, unsigned seq) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(slot16), "r"((unsigned)b), "r"(seq),
               "r"((unsigned)(b >> 32)), "r"(seq)
               : "memory");
}
__device__ __forceinline__ double ll_load(const double* slot16, unsigned seq, unsigned long long timeout_ns, Scalars* sc) {
  unsigned lo, f0, hi, f1;
  unsigned spins = 0;
  unsigned long long t0 = 0;
  for (;;) {
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(lo), "=r"(f0), "=r"(hi), "=r"(f1) : "l"(slot16) : "memory");
    if (f0 == seq && f1 == seq) break;
#if SFMBA_LL_SLEEP > 0
    __nanosleep(SFMBA_LL_SLEEP);     // thousands of lanes poll: back off, or the polling loads delay the very stores they wait for
#endif
    if ((++spins & 0xffu) == 0) {
      const unsigned long long now = global_timer_ns();
      if (t0 == 0) t0 = now;
      if (now - t0 > timeout_ns) {
        atomicOr(&sc->flags, FLAG_PEER_TIMEOUT);
        break;
      }
    }
  }
  return __longlong_as_double((long long)(((unsigned long long)hi << 32) | lo));
}
// Sum over the ranks' words of one number, in rank order (all ranks get the same bits).  The loads of all ranks' words go
// out together and only the words whose tags do not match yet are waited for: a loop of ll_load pays one L2 round trip per
// rank even when every word has long arrived.  `first` = the number's word in writer 0's slot of this rank's window,
// `writer_stride` = doubles between the slots of consecutive writers (PeerWindow::cap).
__device__ __forceinline__ double ll_sum_ranks(const double* first, unsigned long long writer_stride, int world, unsigned seq,
                                               unsigned long long timeout_ns, Scalars* sc) {
  unsigned lo[MAX_PEERS], hi[MAX_PEERS];
  unsigned missing = 0;
#pragma unroll
  for (int w = 0; w < MAX_PEERS; ++w) {
    lo[w] = hi[w] = 0u;
    if (w < world) {
      unsigned f0, f1;
      asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                   : "=r"(lo[w]), "=r"(f0), "=r"(hi[w]), "=r"(f1)
                   : "l"(first + (size_t)w * writer_stride)
                   : "memory");
      if (f0 != seq || f1 != seq) missing |= 1u << w;
    }
  }
  double sum = 0.0;
#pragma unroll
  for (int w = 0; w < MAX_PEERS; ++w) {
    if (w < world) {
      double v = __longlong_as_double((long long)(((unsigned long long)hi[w] << 32) | lo[w]));
      if (missing & (1u << w)) v = ll_load(first + (size_t)w * writer_stride, seq, timeout_ns, sc);
      sum += v;
    }
  }
  return sum;
}
// a flag-in-data area of a window slot (offset in doubles from the slot's start); element k at 16 k bytes
__device__ __forceinline__ double* ll_slot(const PeerWindow& xw, int window, int parity, int writer, unsigned long long off) {
  return xw.slot(window, parity, writer) + off;
}

}  // namespace sfmba
