// Index construction for one BA problem: point-major tiles, the camera-major permutation and the
// one-camera-per-chunk table.  Runs once per solve; everything stays on the device (no host sync).
//
// The point-major CSR itself (obs_pt / obs_cam / pt_rowptr) is the bit-exact visibility index of
// ba_processor.py:304-310 + key_tracker.py:198-204 and is an INPUT here.
#include <cub/cub.cuh>

#include "ba_internal.h"

namespace sfmba {

// tile t owns the points whose first observation lies in [t*TILE_OBS, (t+1)*TILE_OBS)
__global__ void k_tile_bounds(BaDev d) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t > d.ntiles) return;
  if (t == d.ntiles) {
    d.tile_pt[t] = d.N;
    return;
  }
  const int target = t * TILE_OBS;
  int lo = 0, hi = d.N;   // first j in [0,N] with rowptr[j] >= target
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (d.rowptr[mid] < target) lo = mid + 1; else hi = mid;
  }
  d.tile_pt[t] = lo;
}

__global__ void k_iota(int32_t* v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = i;
}

__global__ void k_invert_perm(const int32_t* __restrict__ perm, int32_t* __restrict__ inv, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) inv[perm[i]] = i;
}

// cam_rowptr[c] = first camera-major slot whose camera is >= c (sorted keys)
__global__ void k_cam_rowptr(const int32_t* __restrict__ sorted_cam, int m, int V, int32_t* __restrict__ cam_rowptr) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c > V) return;
  int lo = 0, hi = m;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (sorted_cam[mid] < c) lo = mid + 1; else hi = mid;
  }
  cam_rowptr[c] = lo;
}

// Single block: exclusive scan of ceil(count/CHUNK_OBS) over cameras, then every camera writes its
// chunk rows.  V is small (<= ~1e5), a sequential-by-block scan is plenty.
__global__ void k_chunk_table(BaDev d) {
  __shared__ int s_carry;
  __shared__ int s_scan[1024];
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < d.V; base += blockDim.x) {
    const int c = base + threadIdx.x;
    int n = 0;
    if (c < d.V) {
      const int cnt = d.cam_rowptr[c + 1] - d.cam_rowptr[c];
      n = (cnt + CHUNK_OBS - 1) / CHUNK_OBS;
    }
    s_scan[threadIdx.x] = n;
    __syncthreads();
    for (int off = 1; off < blockDim.x; off <<= 1) {   // Hillis-Steele inclusive scan
      int v = 0;
      if ((int)threadIdx.x >= off) v = s_scan[threadIdx.x - off];
      __syncthreads();
      s_scan[threadIdx.x] += v;
      __syncthreads();
    }
    const int excl = s_carry + s_scan[threadIdx.x] - n;
    if (c < d.V) {
      d.cam_chunk_ptr[c] = excl;
      const int beg = d.cam_rowptr[c], end = d.cam_rowptr[c + 1];
      for (int k = 0; k < n; ++k) {
        d.chunk_cam[excl + k] = c;
        d.chunk_beg[excl + k] = beg + k * CHUNK_OBS;
      }
      (void)end;
    }
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) s_carry += s_scan[threadIdx.x];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    d.cam_chunk_ptr[d.V] = s_carry;
    d.chunk_beg[s_carry] = d.m;
    d.sc->num_chunks = s_carry;
  }
}

static int key_bits(int V) {
  int b = 1;
  while ((1ll << b) < (long long)V) ++b;
  return b;
}

size_t index_sort_temp_bytes(int m, int V) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const int32_t*)nullptr, (int32_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, m > 0 ? m : 1, 0, key_bits(V));
  return bytes;
}

void launch_build_index(const BaDev& d, void* sort_temp, size_t sort_temp_bytes, int32_t* sort_keys_out,
                        int32_t* iota, cudaStream_t st, LaunchCounter& lc) {
  k_tile_bounds<<<(d.ntiles + 1 + 255) / 256, 256, 0, st>>>(d);
  lc.launches++;
  if (d.m > 0) {
    k_iota<<<(d.m + 255) / 256, 256, 0, st>>>(iota, d.m);
    lc.launches++;
    // stable LSD radix sort by camera: ties keep point-major order, so the camera-major order (and
    // with it every floating-point summation order downstream) is deterministic
    cub::DeviceRadixSort::SortPairs(sort_temp, sort_temp_bytes, d.obs_cam, sort_keys_out, (const int32_t*)iota,
                                    d.cm_perm, d.m, 0, key_bits(d.V), st);
    lc.launches += 4;
    k_invert_perm<<<(d.m + 255) / 256, 256, 0, st>>>(d.cm_perm, d.pm2cm, d.m);
    lc.launches++;
  }
  k_cam_rowptr<<<(d.V + 1 + 255) / 256, 256, 0, st>>>(sort_keys_out, d.m, d.V, d.cam_rowptr);
  lc.launches++;
  k_chunk_table<<<1, 1024, 0, st>>>(d);
  lc.launches++;
}

}  // namespace sfmba
