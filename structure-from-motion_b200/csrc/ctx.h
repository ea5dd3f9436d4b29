// Host-side context of libsfmba: device workspace arena, NCCL communicator (dlopen'ed), profiling.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "ba_internal.h"

struct ProfileEntry {
  double ms = 0.0;
  int64_t launches = 0;
};

// Grow-only device buffer.
struct DevBuf {
  void* ptr = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes);
  void release();
  template <typename T>
  T* as() const { return reinterpret_cast<T*>(ptr); }
};

// Host-buffer solves (sfmba_ba_solve_host): the uploads run on a copy stream in the order the solve
// needs them and the solve waits for each group at its first use, so the index build overlaps the
// uploads; the downloads start as soon as the parameters are final, beside the final cost pass.
struct SolveIo {
  cudaStream_t copy = nullptr;
  cudaEvent_t ev_index = nullptr;    // pt_rowptr, obs_cam resident
  cudaEvent_t ev_obs = nullptr;      // obs_uv, cam_kinv resident
  cudaEvent_t ev_params = nullptr;   // cams, pts resident
  cudaEvent_t ev_final = nullptr;    // recorded by the solve when cams / pts are final
  double* host_cams = nullptr;
  double* host_pts = nullptr;
  // upload ranges of the host-buffer solve: range r = groups [range_grp[r], range_grp[r+1]); ev_range_obs[r] /
  // ev_range_pts[r] fire when the observations / the points of the range are resident
  static constexpr int MAX_RANGES = 8;
  int num_ranges = 0;
  int range_grp[MAX_RANGES + 1] = {};
  cudaEvent_t ev_range_obs[MAX_RANGES] = {};
  cudaEvent_t ev_range_pts[MAX_RANGES] = {};
};

// what the workspace's index was built from (sfmba_options.reuse_index)
struct IndexCache {
  bool valid = false;
  int V = 0, N = 0, m = 0, fine = 0, nchunks = 0;
  const void *obs_uv = nullptr, *obs_cam = nullptr, *rowptr = nullptr, *kinv = nullptr;
  sfmba::TileShape ts;
};

struct sfmba_ctx {
  int device = 0;
  int world = 1;
  int rank = 0;
  std::string err;
  sfmba::LaunchCounter lc;
  // NCCL (only when world > 1)
  void* nccl_lib = nullptr;
  void* nccl_comm = nullptr;
  // NVLink peer exchange windows (peer_window.cu); peer.world == 0 until mapped
  void* peer_local = nullptr;
  void* peer_mapped[sfmba::MAX_PEERS] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  sfmba::PeerWindow peer = {};
  bool peer_disabled = false;
  DevBuf misc;
  cudaError_t reserve_misc(size_t bytes, void** out) {
    const cudaError_t e = misc.reserve(bytes);
    *out = misc.ptr;
    return e;
  }
  // profiling
  bool profile = false;          // per-kernel spans (event records between the launches)
  bool phase_timing = false;     // per-phase spans of sfmba_result.ms_*
  std::map<std::string, ProfileEntry> prof;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaEvent_t phase_ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  // workspace
  std::vector<DevBuf> bufs;      // indexed by slot
  DevBuf host_stage[12];         // device copies of host inputs for *_host entry points
  void* pinned = nullptr;        // pinned staging for results/scalars
  size_t pinned_cap = 0;
  cudaStream_t own_stream = nullptr;
  SolveIo io;                    // copy stream and events of the host-buffer entry point
  sfmba::DenseStreams dense_streams;   // dense-Cholesky look-ahead (dense.cu)
  IndexCache index_cache;
  unsigned pcg_epoch = 0;        // id of the current PCG solve (see host_sig)
  int pcg_hint = 0;              // iterations the previous PCG solve needed
  std::vector<cudaEvent_t> timer_events;   // phase-timer events, created once and reused by every solve
  size_t timer_next = 0;

  cudaError_t reserve(int slot, size_t bytes, void** out);
};

extern std::string g_last_error;

int sfm_set_error(sfmba_ctx* ctx, int code, const std::string& msg);
int sfm_check_cuda(sfmba_ctx* ctx, cudaError_t e, const char* what);

// NCCL helpers (nccl_dl.cu)
int sfm_nccl_unique_id(void* out128, std::string& err);
int sfm_nccl_init(sfmba_ctx* ctx, const void* unique_id);
int sfm_nccl_allreduce_f64(sfmba_ctx* ctx, double* buf, size_t count, cudaStream_t st);
int sfm_nccl_allgather_bytes(sfmba_ctx* ctx, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st);
int sfm_nccl_barrier(sfmba_ctx* ctx, cudaStream_t st);
void sfm_nccl_destroy(sfmba_ctx* ctx);

// peer windows (peer_window.cu)
int sfm_peer_setup(sfmba_ctx* ctx, size_t doubles_cap, cudaStream_t st);
void sfm_peer_destroy(sfmba_ctx* ctx);
