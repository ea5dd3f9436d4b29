// Index construction for one BA problem (runs once per solve, on the device).
//
// Input: the point-major CSR (obs_pt / obs_cam / pt_rowptr) -- the bit-exact visibility index of
// ba_processor.py:304-310 + key_tracker.py:198-204.  Output: groups, chunks, the camera-sorted and
// lane-transposed observation records, run / segment tables and the camera -> segment-row CSR
// (layout described in ba_internal.h).  Observations are normalised here once
// (b = K^-1 [u,v,1]^T / z, ba_processor.py:339-342) and stored as f64 in the records.
#include <cub/cub.cuh>

#include "ba_internal.h"
#include "sfmba_device.cuh"

namespace sfmba {

// pt_rowptr must be a CSR over [0, m] and obs_cam must lie in [0, V): everything below indexes the observation
// arrays and the camera tables through them.  The host reads the flag after the counting phase.
__global__ void k_validate_index(BaDev d) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  bool bad = false;
  if (j <= d.N) {
    const int a = d.rowptr[j];
    bad = a < 0 || a > d.m;
    if (j == 0 && a != 0) bad = true;
    if (j == d.N && a != d.m) bad = true;
    if (j < d.N && a > d.rowptr[j + 1]) bad = true;
  }
  const int stride = gridDim.x * blockDim.x;
  for (int o = j; o < d.m; o += stride)
    if ((uint32_t)d.obs_cam[o] >= (uint32_t)d.V) bad = true;
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(&d.sc->flags, FLAG_BAD_INDEX);
}

// group g owns the points j with rowptr[j] + POINT_WEIGHT * j in [g W, (g+1) W).  One warp per group: 32 probes per round
// cut the range 33-fold (four dependent rounds for a million points instead of twenty bisection steps)
__global__ void __launch_bounds__(256) k_group_bounds(BaDev d) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (g > d.G) return;
  if (g == d.G) {
    if (lane == 0) d.grp_pt[g] = d.N;
    return;
  }
  const long long target = (long long)g * GROUP_WEIGHT;
  int lo = 0, hi = d.N;   // first j in [0,N] with weight prefix >= target
  while (lo < hi) {
    const long long span = hi - lo;
    const int p = lo + (int)(((long long)(lane + 1) * span) / 33);           // lo <= p < hi, non-decreasing in the lane
    const bool below = (long long)d.rowptr[p] + (long long)POINT_WEIGHT * p < target;
    const unsigned m = __ballot_sync(0xffffffffu, below);
    const int c = __popc(m);                                                 // the predicate is monotone: lanes 0..c-1
    const int p_last_below = __shfl_sync(0xffffffffu, p, c > 0 ? c - 1 : 0);
    const int p_first_not = __shfl_sync(0xffffffffu, p, c < 32 ? c : 31);
    if (c > 0) lo = p_last_below + 1;
    if (c < 32) hi = p_first_not;
  }
  if (lane == 0) d.grp_pt[g] = lo;
}

// Cameras of one group are counted in a small open-addressing table: only groups with <= SEG_CAP distinct
// cameras become a single chunk, so the insertion stops as soon as more have been seen (every thread can
// add at most one more camera before it notices, which keeps the table below half full).
constexpr int HASH_SLOTS = 512;
static_assert(HASH_SLOTS >= 2 * (SEG_CAP + RUN_LANES), "table must hold SEG_CAP cameras plus one per thread");

// lanes a chunk needs when every camera's observations are cut into runs of K
__device__ __forceinline__ int lanes_needed(const int* cnt, int K, int* s_sum) {
  int mine = 0;
  for (int i = threadIdx.x; i < HASH_SLOTS; i += blockDim.x) mine += (cnt[i] + K - 1) / K;
  __syncthreads();
  if (threadIdx.x == 0) *s_sum = 0;
  __syncthreads();
  if (mine) atomicAdd(s_sum, mine);
  __syncthreads();
  return *s_sum;
}

// chunks per group: one chunk when some K <= CHUNK_K_MAX fits the group into RUN_LANES lanes and
// SEG_CAP cameras, otherwise SEG_CAP-observation chunks with one observation per lane
__global__ void __launch_bounds__(RUN_LANES) k_group_count(BaDev d, int fine_chunks) {
  __shared__ int s_key[HASH_SLOTS];
  __shared__ int s_cnt[HASH_SLOTS];
  __shared__ int s_distinct, s_sum;
  const int g = blockIdx.x;
  if (d.sc->flags & FLAG_BAD_INDEX) {               // k_validate_index: the host rejects the problem after this phase
    if (threadIdx.x == 0) {
      d.grp_nchunk[g] = 0;
      d.grp_K[g] = 1;
    }
    return;
  }
  const int pt0 = d.grp_pt[g], pt1 = d.grp_pt[g + 1];
  const int o0 = d.rowptr[pt0], o1 = d.rowptr[pt1];
  const int L = o1 - o0;
  int n, K = 1;
  if (L == 0) {
    n = 0;
  } else if (fine_chunks || L > CHUNK_CAP) {
    n = (L + SEG_CAP - 1) / SEG_CAP;
  } else {
    for (int i = threadIdx.x; i < HASH_SLOTS; i += blockDim.x) {
      s_key[i] = -1;
      s_cnt[i] = 0;
    }
    if (threadIdx.x == 0) s_distinct = 0;
    __syncthreads();
    // the loads of a thread's cameras are issued together, the table work follows
    for (int i0 = threadIdx.x; i0 < L; i0 += 4 * RUN_LANES) {
      int cams[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * RUN_LANES;
        cams[u] = i < L ? d.obs_cam[o0 + i] : -1;
      }
      if (*(volatile int*)&s_distinct > SEG_CAP) break;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int cam = cams[u];
        if (cam < 0) continue;
        unsigned h = ((unsigned)cam * 2654435761u) >> 20;
        for (;;) {
          const int slot = h & (HASH_SLOTS - 1);
          const int prev = atomicCAS(&s_key[slot], -1, cam);
          if (prev == -1 || prev == cam) {
            if (prev == -1) atomicAdd(&s_distinct, 1);
            atomicAdd(&s_cnt[slot], 1);
            break;
          }
          ++h;
        }
        if (*(volatile int*)&s_distinct > SEG_CAP) break;
      }
    }
    __syncthreads();
    n = (L + SEG_CAP - 1) / SEG_CAP;
    if (s_distinct <= SEG_CAP) {
      for (K = (L + RUN_LANES - 1) / RUN_LANES; K <= CHUNK_K_MAX; ++K) {
        if (lanes_needed(s_cnt, K, &s_sum) <= RUN_LANES) {
          n = 1;
          break;
        }
      }
    }
    if (n != 1) K = 1;
    if (threadIdx.x == 0 && n == 1) {
      atomicMax(&d.sc->max_chunk_len, L);
      atomicMax(&d.sc->max_nseg, s_distinct);
      atomicMax(&d.sc->max_k, K);
    }
  }
  if (threadIdx.x == 0) {
    d.grp_nchunk[g] = n;
    d.grp_K[g] = K;
    atomicMax(&d.sc->max_group_pts, pt1 - pt0);
    if (n > 1 || (n == 1 && K == 1 && L <= SEG_CAP)) {
      atomicMax(&d.sc->max_chunk_len, L < SEG_CAP ? L : SEG_CAP);
      atomicMax(&d.sc->max_nseg, L < SEG_CAP ? L : SEG_CAP);
      atomicMax(&d.sc->max_k, 1);
    }
  }
}

__device__ __forceinline__ void finish_scan(const BaDev& d, int total, IndexStats* host_stats, unsigned epoch) {
  d.grp_chunk[d.G] = total;
  d.sc->num_chunks = total;
  if (host_stats) {
    // what the host needs to size the chunk arrays and the tiles' shared memory, straight into pinned host memory: the
    // host spins on the epoch word instead of synchronising the stream
    host_stats->num_chunks = total;
    host_stats->max_chunk_len = d.sc->max_chunk_len;
    host_stats->max_group_pts = d.sc->max_group_pts;
    host_stats->max_nseg = d.sc->max_nseg;
    host_stats->max_k = d.sc->max_k;
    host_stats->flags = d.sc->flags;
    __threadfence_system();
    *(volatile unsigned*)&host_stats->epoch = epoch;
  }
}

// (after cub's device scan) grp_chunk holds the exclusive scan of grp_nchunk over [0, G); close it
__global__ void k_finish_scan(BaDev d, IndexStats* host_stats, unsigned epoch) {
  const int total = d.G > 0 ? d.grp_chunk[d.G - 1] + d.grp_nchunk[d.G - 1] : 0;
  finish_scan(d, total, host_stats, epoch);
}

// the same in one launch for the usual group counts: one CTA scans grp_nchunk tile by tile, then closes the scan
constexpr int SCAN1_THREADS = 1024;
constexpr int SCAN1_ITEMS = 8;                       // consecutive groups per thread: one tile covers 8192 groups
constexpr int SCAN1_MAX_GROUPS = 64 * SCAN1_THREADS * SCAN1_ITEMS;
__global__ void __launch_bounds__(SCAN1_THREADS) k_scan_chunks(BaDev d, IndexStats* host_stats, unsigned epoch) {
  using Scan = cub::BlockScan<int, SCAN1_THREADS>;
  __shared__ typename Scan::TempStorage tmp;
  int carry = 0;
  for (int base = 0; base < d.G; base += SCAN1_THREADS * SCAN1_ITEMS) {
    const int i0 = base + threadIdx.x * SCAN1_ITEMS;
    int v[SCAN1_ITEMS], mine = 0;
#pragma unroll
    for (int u = 0; u < SCAN1_ITEMS; ++u) {
      v[u] = i0 + u < d.G ? d.grp_nchunk[i0 + u] : 0;
      mine += v[u];
    }
    int ex, tot;
    Scan(tmp).ExclusiveSum(mine, ex, tot);
    int run = carry + ex;
#pragma unroll
    for (int u = 0; u < SCAN1_ITEMS; ++u) {
      if (i0 + u < d.G) d.grp_chunk[i0 + u] = run;
      run += v[u];
    }
    carry += tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) finish_scan(d, carry, host_stats, epoch);
}

struct BuildSmem {
  uint32_t key[CHUNK_CAP];
  uint16_t seg[CHUNK_CAP];          // segment of each sorted position
  int seg_start[SEG_CAP + 1];       // first sorted position of each segment (+ end)
  int seg_lane0[SEG_CAP + 1];       // first lane of each segment (+ end)
  uint16_t lane_seg[RUN_LANES];
  int rowp[GROUP_PT_CAP + 1];       // pt_rowptr of the group's points (observation -> point by bisection)
};

// one CTA per group: sort each chunk by camera, cut segments into runs of K, emit records and tables
__global__ void __launch_bounds__(RUN_LANES) k_chunk_build(BaDev d, int cam_bits, int g0) {
  using Sort = cub::BlockRadixSort<uint32_t, RUN_LANES, CHUNK_K_MAX>;
  using Scan = cub::BlockScan<int, RUN_LANES>;
  __shared__ union {
    typename Sort::TempStorage sort;
    typename Scan::TempStorage scan;
  } tmp;
  __shared__ BuildSmem s;
  const int g = g0 + blockIdx.x;
  const int pt0 = d.grp_pt[g], pt1 = d.grp_pt[g + 1];
  const int c0 = d.grp_chunk[g], c1 = d.grp_chunk[g + 1];
  const int o_begin = d.rowptr[pt0], o_end = d.rowptr[pt1];
  const int nc = c1 - c0;
  const int K = nc == 1 ? d.grp_K[g] : 1;
  const int tid = threadIdx.x;
  for (int i = tid; i <= pt1 - pt0; i += RUN_LANES) s.rowp[i] = d.rowptr[pt0 + i];
  if (tid == 0) {
    GroupMeta gm;
    gm.pt0 = pt0;
    gm.np = pt1 - pt0;
    gm.c0 = c0;
    gm.nc = nc;
    gm.len0 = gm.nseg0 = gm.obs0 = 0;
    gm.K0 = 1;
    d.grp_meta[g] = gm;
  }
  for (int c = c0; c < c1; ++c) {
    const int obs0 = nc == 1 ? o_begin : o_begin + (c - c0) * SEG_CAP;
    const int len = nc == 1 ? (o_end - o_begin) : min(SEG_CAP, o_end - obs0);
    uint32_t keys[CHUNK_K_MAX];
    bool bad_cam = false;
#pragma unroll
    for (int i = 0; i < CHUNK_K_MAX; ++i) {
      const int idx = tid * CHUNK_K_MAX + i;
      uint32_t cam = idx < len ? (uint32_t)d.obs_cam[obs0 + idx] : 0u;
      if (cam >= (uint32_t)d.V) {                    // (k_validate_index has flagged it; stay inside the tables)
        bad_cam = true;
        cam = 0u;
      }
      keys[i] = idx < len ? ((cam << 11) | (uint32_t)idx) : 0xffffffffu;
    }
    if (bad_cam) atomicOr(&d.sc->flags, FLAG_BAD_INDEX);
    // the keys arrive in position order (blocked arrangement) and the radix sort is stable, so sorting on
    // the camera bits alone leaves equal cameras in position order; padding keys (all ones) stay last
    Sort(tmp.sort).Sort(keys, 11, 11 + cam_bits);
#pragma unroll
    for (int i = 0; i < CHUNK_K_MAX; ++i) s.key[tid * CHUNK_K_MAX + i] = keys[i];
    __syncthreads();
    // segment heads among my consecutive sorted items
    uint32_t prev_cam = tid == 0 ? 0xffffffffu : (s.key[tid * CHUNK_K_MAX - 1] >> 11);
    int heads = 0;
    uint32_t flags = 0;
#pragma unroll
    for (int i = 0; i < CHUNK_K_MAX; ++i) {
      const int idx = tid * CHUNK_K_MAX + i;
      if (idx < len) {
        const uint32_t cam = keys[i] >> 11;
        if (idx == 0 || cam != prev_cam) {
          ++heads;
          flags |= 1u << i;
        }
        prev_cam = cam;
      }
    }
    int excl, nseg;
    Scan(tmp.scan).ExclusiveSum(heads, excl, nseg);
    int seg = excl - 1;
#pragma unroll
    for (int i = 0; i < CHUNK_K_MAX; ++i) {
      const int idx = tid * CHUNK_K_MAX + i;
      if (idx < len) {
        if (flags & (1u << i)) {
          ++seg;
          s.seg_start[seg] = idx;
          d.seg_cam[(size_t)c * SEG_CAP + seg] = (int32_t)(keys[i] >> 11);
        }
        s.seg[idx] = (uint16_t)seg;
      }
    }
    if (tid == 0) s.seg_start[nseg] = len;
    for (int sidx = nseg + tid; sidx < SEG_CAP; sidx += RUN_LANES) d.seg_cam[(size_t)c * SEG_CAP + sidx] = d.V;
    // the identity permutation the camera sort of the rows starts from (rowpos is that buffer until the sort is done)
    for (int sidx = tid; sidx < SEG_CAP; sidx += RUN_LANES) d.rowpos[(size_t)c * SEG_CAP + sidx] = c * SEG_CAP + sidx;
    __syncthreads();
    // lanes per segment -> first lane of each segment
    const int nl = tid < nseg ? (s.seg_start[tid + 1] - s.seg_start[tid] + K - 1) / K : 0;
    int lane0, nlanes;
    __syncthreads();
    Scan(tmp.scan).ExclusiveSum(nl, lane0, nlanes);
    if (tid < nseg) {
      s.seg_lane0[tid] = lane0;
      d.seg_lane0[(size_t)c * SEG_RUN_STRIDE + tid] = (uint16_t)lane0;
      for (int j = 0; j < nl; ++j) s.lane_seg[lane0 + j] = (uint16_t)tid;
    }
    if (tid == 0) {
      s.seg_lane0[nseg] = nlanes;
      d.seg_lane0[(size_t)c * SEG_RUN_STRIDE + nseg] = (uint16_t)nlanes;
      ChunkInfo ci;
      ci.obs0 = obs0;
      ci.len = len;
      ci.nseg = nseg;
      ci.K = K;
      d.chunk_info[c] = ci;
      if (c == c0) {
        GroupMeta* gm = d.grp_meta + g;
        gm->len0 = len;
        gm->nseg0 = nseg;
        gm->obs0 = obs0;
        gm->K0 = K;
      }
    }
    __syncthreads();
    // records, lane-transposed: slot = k * RUN_LANES + lane; lane = run (K sorted observations of ONE camera)
    for (int slot = tid; slot < K * RUN_LANES; slot += RUN_LANES) {
      const int k = slot / RUN_LANES, lane = slot % RUN_LANES;
      uint32_t ix = REC_INVALID;
      double2 b = make_double2(0.0, 0.0);
      if (lane < nlanes) {
        const int sg = s.lane_seg[lane];
        // the lanes of a segment take its sorted observations round-robin: at trip k they read neighbouring
        // observations (neighbouring points: fewer shared-memory bank conflicts) and their lengths differ by <= 1
        const int nl = s.seg_lane0[sg + 1] - s.seg_lane0[sg];
        const int idx = s.seg_start[sg] + k * nl + (lane - s.seg_lane0[sg]);
        if (idx < s.seg_start[sg + 1]) {
          const uint32_t key = s.key[idx];
          const int pos = (int)(key & 2047u);
          const int cam = (int)(key >> 11);
          const int o = obs0 + pos;
          int lo = 0, hi = pt1 - pt0;                       // last point j with rowp[j] <= o
          while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (s.rowp[mid] <= o) lo = mid; else hi = mid;
          }
          const int ptl = lo;
          ix = (uint32_t)ptl | ((uint32_t)pos << 10) | ((uint32_t)sg << 21);
          double bb[2];
          normalise_obs(d.kinv ? d.kinv + 9 * (size_t)cam : nullptr, d.obs_uv[o], d.obs_uv[d.m + o], bb);
          b = make_double2(bb[0], bb[1]);
        }
      }
      d.rec_ix[(size_t)c * CHUNK_CAP + slot] = ix;
      d.rec_b[(size_t)c * CHUNK_CAP + slot] = b;
    }
    __syncthreads();
  }
}

// rowpos[slot] = position of (chunk, segment) slot in the camera-sorted order; camrow_ptr[c] = first sorted row whose
// camera is >= c, written where the sorted camera changes (cameras without rows included)
__global__ void k_rowpos_camptr(const int32_t* __restrict__ sorted_slot, const int32_t* __restrict__ sorted_cam, int n, int V,
                                int32_t* __restrict__ rowpos, int32_t* __restrict__ camrow_ptr) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  rowpos[sorted_slot[k]] = k;
  const int cam = sorted_cam[k], prev = k > 0 ? sorted_cam[k - 1] : -1;      // ascending, <= V (V = unused slot)
  for (int c = prev + 1; c <= cam; ++c) camrow_ptr[c] = k;
  if (k == n - 1)
    for (int c = cam + 1; c <= V; ++c) camrow_ptr[c] = n;
}

// camrow_ptr[c] = first sorted row whose camera is >= c
__global__ void k_cam_rowptr(const int32_t* __restrict__ sorted_cam, int n, int V, int32_t* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c > V) return;
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (sorted_cam[mid] < c) lo = mid + 1; else hi = mid;
  }
  out[c] = lo;
}

__global__ void k_widen_u16(const uint16_t* __restrict__ in, int32_t* __restrict__ out, size_t n) {
  // eight indices per thread: one 16-byte load, two 16-byte stores
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (i + 8 <= n) {
    const uint4 v = *reinterpret_cast<const uint4*>(in + i);
    int4 a, b;
    a.x = v.x & 0xffff; a.y = v.x >> 16; a.z = v.y & 0xffff; a.w = v.y >> 16;
    b.x = v.z & 0xffff; b.y = v.z >> 16; b.z = v.w & 0xffff; b.w = v.w >> 16;
    *reinterpret_cast<int4*>(out + i) = a;
    *reinterpret_cast<int4*>(out + i + 4) = b;
  } else {
    for (size_t k = i; k < n; ++k) out[k] = in[k];
  }
}

void launch_widen_u16(const uint16_t* in, int32_t* out, size_t n, cudaStream_t st, LaunchCounter& lc) {
  if (n == 0) return;
  const size_t threads = (n + 7) / 8;
  k_widen_u16<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(in, out, n);
  lc.launches++;
}

static int bits_for(int n) {   // bits needed to represent values in [0, n]
  int b = 1;
  while ((1ll << b) <= (long long)n) ++b;
  return b;
}

size_t index_temp_bytes(int G, int chunk_bound, int V) {
  size_t a = 0, b = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, a, (const int32_t*)nullptr, (int32_t*)nullptr, G > 0 ? G : 1);
  const int n = chunk_bound * SEG_CAP;
  cub::DeviceRadixSort::SortPairs(nullptr, b, (const int32_t*)nullptr, (int32_t*)nullptr, (const int32_t*)nullptr,
                                  (int32_t*)nullptr, n > 0 ? n : 1, 0, bits_for(V));
  return a > b ? a : b;
}

// phase 1: groups and chunk counts (the host then reads sc->num_chunks and the shape statistics)
void launch_index_count(const BaDev& d, bool fine_chunks, void* temp, size_t temp_bytes, IndexStats* host_stats,
                        unsigned epoch, cudaStream_t st, LaunchCounter& lc) {
  k_validate_index<<<(d.N + 1 + 255) / 256, 256, 0, st>>>(d);
  k_group_bounds<<<(unsigned)(((size_t)(d.G + 1) * 32 + 255) / 256), 256, 0, st>>>(d);
  lc.launches += 2;
  if (d.G > 0) {
    k_group_count<<<d.G, RUN_LANES, 0, st>>>(d, fine_chunks ? 1 : 0);
    lc.launches++;
  }
  if (d.G <= SCAN1_MAX_GROUPS) {
    k_scan_chunks<<<1, SCAN1_THREADS, 0, st>>>(d, host_stats, epoch);
    lc.launches++;
  } else {
    cub::DeviceScan::ExclusiveSum(temp, temp_bytes, d.grp_nchunk, d.grp_chunk, d.G, st);
    k_finish_scan<<<1, 1, 0, st>>>(d, host_stats, epoch);
    lc.launches += 3;
  }
}

// phase 2: records, tables and the camera CSR (chunk arrays allocated for `nchunks`)
void launch_chunk_build(const BaDev& d, int g0, int g1, cudaStream_t st, LaunchCounter& lc) {
  if (g1 > g0) {
    k_chunk_build<<<g1 - g0, RUN_LANES, 0, st>>>(d, bits_for(d.V > 0 ? d.V - 1 : 0), g0);
    lc.launches++;
  }
}

void launch_index_build(const BaDev& d, int nchunks, void* temp, size_t temp_bytes, int32_t* sort_keys_out,
                        int32_t* iota, cudaStream_t st, LaunchCounter& lc) {
  const int n = nchunks * SEG_CAP;
  if (n > 0) {
    // (iota = d.rowpos: filled by k_chunk_build)
    cub::DeviceRadixSort::SortPairs(temp, temp_bytes, (const int32_t*)d.seg_cam, sort_keys_out, (const int32_t*)iota,
                                    d.camrow_idx, n, 0, bits_for(d.V), st);
    // (rowpos is iota's buffer: free after the sort)
    k_rowpos_camptr<<<(n + 255) / 256, 256, 0, st>>>(d.camrow_idx, sort_keys_out, n, d.V, d.rowpos, d.camrow_ptr);
    lc.launches += 4;
  } else {
    k_cam_rowptr<<<(d.V + 1 + 255) / 256, 256, 0, st>>>(sort_keys_out, n, d.V, d.camrow_ptr);   // no rows: all zeros
    lc.launches++;
  }
}

}  // namespace sfmba
