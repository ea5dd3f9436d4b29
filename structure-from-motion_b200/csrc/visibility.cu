// Visibility / track index on the device, bit-exact with the reference's read-out.
//
// Reference: BA asks KeyTracker.is_visible(view, tri_idx) for every (point, view) pair
// (ba_processor.py:304-310).  is_visible (key_tracker.py:198-204) scans the diagonal row
// table[view, :] for tri_idx and applies np.any() to the *indices* found, so
//   - no match                          -> invisible
//   - the only match is key index 0     -> invisible  (index-0 quirk)
//   - otherwise                         -> the smallest matching key index
// Here every used key emits (tri_idx * V + view, key_idx); one stable radix sort groups them in the
// reference's loop order (point-major, view-minor, smallest key first), a flag pass applies the
// quirk, a scan compacts, and a binary search per point yields the CSR row pointer.
#include <cub/cub.cuh>

#include "ba_internal.h"

namespace sfmba {

__global__ void k_vis_keys(int V, int N, int64_t total, const int64_t* __restrict__ diag,
                           const int64_t* __restrict__ key_off, uint64_t* __restrict__ keys,
                           int32_t* __restrict__ vals) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= total) return;
  int lo = 0, hi = V;           // last v with key_off[v] <= g
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (key_off[mid] <= g) lo = mid; else hi = mid;
  }
  const int64_t t = diag[g];
  const bool valid = t >= 0 && t < N;
  keys[g] = valid ? (uint64_t)t * (uint64_t)V + (uint64_t)lo : ~0ull;
  vals[g] = (int32_t)(g - key_off[lo]);
}

__global__ void k_vis_flags(int64_t total, const uint64_t* __restrict__ keys, const int32_t* __restrict__ vals,
                            int32_t* __restrict__ flag) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const uint64_t k = keys[i];
  bool vis = false;
  if (k != ~0ull && (i == 0 || keys[i - 1] != k)) {
    const bool several = (i + 1 < total) && keys[i + 1] == k;
    vis = vals[i] > 0 || several;
  }
  flag[i] = vis ? 1 : 0;
}

__global__ void k_vis_scatter(int V, int64_t total, const uint64_t* __restrict__ keys,
                              const int32_t* __restrict__ vals, const int32_t* __restrict__ flag,
                              const int32_t* __restrict__ pos, int32_t* __restrict__ obs_pt,
                              int32_t* __restrict__ obs_cam, int32_t* __restrict__ obs_key,
                              int32_t* __restrict__ count) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  if (flag[i]) {
    const int p = pos[i];
    const uint64_t k = keys[i];
    obs_pt[p] = (int32_t)(k / (uint64_t)V);
    obs_cam[p] = (int32_t)(k % (uint64_t)V);
    obs_key[p] = vals[i];
  }
  if (i == total - 1) *count = pos[i] + flag[i];
}

__global__ void k_vis_rowptr(int N, const int32_t* __restrict__ obs_pt, const int32_t* __restrict__ count,
                             int32_t* __restrict__ rowptr) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j > N) return;
  const int m = *count;
  int lo = 0, hi = m;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (obs_pt[mid] < j) lo = mid + 1; else hi = mid;
  }
  rowptr[j] = lo;
}

int build_visibility_device(int V, int N, const int64_t* diag, const int64_t* key_off_host, int64_t total,
                            int32_t* obs_pt, int32_t* obs_cam, int32_t* obs_key, int32_t* pt_rowptr,
                            int32_t* host_num_obs, cudaStream_t st, LaunchCounter& lc, std::string& err) {
  if (V <= 0 || N < 0 || total < 0) {
    err = "bad sizes";
    return SFMBA_ERR_INVALID_ARG;
  }
  if (total >= (1ll << 31)) {
    err = "more than 2^31 keys";
    return SFMBA_ERR_UNSUPPORTED;
  }
  int32_t* d_count = nullptr;
  cudaError_t e = cudaMalloc(&d_count, sizeof(int32_t));
  if (e != cudaSuccess) {
    err = cudaGetErrorString(e);
    return SFMBA_ERR_CUDA;
  }
  cudaMemsetAsync(d_count, 0, sizeof(int32_t), st);
  void* scratch = nullptr;
  if (total > 0) {
    int bits = 1;
    while (bits < 64 && (1ull << bits) <= (uint64_t)N * (uint64_t)V) ++bits;
    bits = bits + 1 > 64 ? 64 : bits + 1;   // the all-ones sentinel must sort last
    bits = 64;                                // sentinel uses the top bits: sort on all of them
    size_t sort_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (uint64_t*)nullptr, (uint64_t*)nullptr, (int32_t*)nullptr,
                                    (int32_t*)nullptr, (int)total, 0, bits);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (int32_t*)nullptr, (int32_t*)nullptr, (int)total);
    const size_t tmp_bytes = (sort_bytes > scan_bytes ? sort_bytes : scan_bytes) + 256;
    const size_t n = (size_t)total;
    const size_t bytes = 2 * n * sizeof(uint64_t) + 4 * n * sizeof(int32_t) + (size_t)(V + 1) * sizeof(int64_t) +
                         tmp_bytes + 1024;
    e = cudaMalloc(&scratch, bytes);
    if (e != cudaSuccess) {
      cudaFree(d_count);
      err = cudaGetErrorString(e);
      return SFMBA_ERR_CUDA;
    }
    char* p = (char*)scratch;
    uint64_t* keys_a = (uint64_t*)p; p += n * sizeof(uint64_t);
    uint64_t* keys_b = (uint64_t*)p; p += n * sizeof(uint64_t);
    int64_t* d_off = (int64_t*)p; p += (size_t)(V + 1) * sizeof(int64_t);
    int32_t* vals_a = (int32_t*)p; p += n * sizeof(int32_t);
    int32_t* vals_b = (int32_t*)p; p += n * sizeof(int32_t);
    int32_t* flag = (int32_t*)p; p += n * sizeof(int32_t);
    int32_t* pos = (int32_t*)p; p += n * sizeof(int32_t);
    p = (char*)(((uintptr_t)p + 255) & ~(uintptr_t)255);
    void* tmp = p;
    cudaMemcpyAsync(d_off, key_off_host, (size_t)(V + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st);
    const unsigned g = (unsigned)((total + 255) / 256);
    k_vis_keys<<<g, 256, 0, st>>>(V, N, total, diag, d_off, keys_a, vals_a);
    size_t sb = tmp_bytes;
    cub::DeviceRadixSort::SortPairs(tmp, sb, keys_a, keys_b, vals_a, vals_b, (int)total, 0, bits, st);
    k_vis_flags<<<g, 256, 0, st>>>(total, keys_b, vals_b, flag);
    sb = tmp_bytes;
    cub::DeviceScan::ExclusiveSum(tmp, sb, flag, pos, (int)total, st);
    k_vis_scatter<<<g, 256, 0, st>>>(V, total, keys_b, vals_b, flag, pos, obs_pt, obs_cam, obs_key, d_count);
    lc.launches += 12;
  }
  k_vis_rowptr<<<(N + 1 + 255) / 256, 256, 0, st>>>(N, obs_pt, d_count, pt_rowptr);
  lc.launches++;
  int32_t m = 0;
  cudaMemcpyAsync(&m, d_count, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
  e = cudaStreamSynchronize(st);
  if (scratch) cudaFree(scratch);
  cudaFree(d_count);
  if (e != cudaSuccess) {
    err = cudaGetErrorString(e);
    return SFMBA_ERR_CUDA;
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    err = cudaGetErrorString(e);
    return SFMBA_ERR_CUDA;
  }
  if (host_num_obs) *host_num_obs = m;
  return SFMBA_OK;
}

}  // namespace sfmba
