// Peer exchange windows: every rank allocates one window, exports it with CUDA IPC and maps the
// windows of all other ranks, so kernels can store partial sums straight into peer HBM over
// NVLink / NVSwitch (see PeerWindow in ba_internal.h and k_pcg_tail<3> in ba_pcg.cu).
// The handles travel through the NCCL communicator the context already owns (one all-gather).
// All ranks call sfm_peer_setup with the same capacity at the same point of the solve, so every
// collective below is matched.  If any rank cannot map a peer, all ranks fall back to NCCL.
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "ctx.h"

using sfmba::MAX_PEERS;
using sfmba::PEER_HEADER_BYTES;

static void close_peers(sfmba_ctx* ctx) {
  for (int r = 0; r < MAX_PEERS; ++r) {
    if (ctx->peer_mapped[r]) cudaIpcCloseMemHandle(ctx->peer_mapped[r]);
    ctx->peer_mapped[r] = nullptr;
  }
}

void sfm_peer_destroy(sfmba_ctx* ctx) {
  close_peers(ctx);
  if (ctx->peer_local) cudaFree(ctx->peer_local);
  ctx->peer_local = nullptr;
  ctx->peer.world = 0;
  ctx->peer.cap = 0;
}

// Returns SFMBA_OK with ctx->peer.world == world when the windows are usable, SFMBA_OK with
// ctx->peer.world == 0 when the ranks agreed to use NCCL instead, or an error.
int sfm_peer_setup(sfmba_ctx* ctx, size_t doubles_cap, cudaStream_t st) {
  if (ctx->world <= 1 || ctx->peer_disabled) return SFMBA_OK;
  if (ctx->world > MAX_PEERS) {
    ctx->peer_disabled = true;
    return SFMBA_OK;
  }
  if (ctx->peer.world == ctx->world && ctx->peer.cap >= doubles_cap) return SFMBA_OK;
  const char* env = getenv("SFMBA_PEER");
  if (env && env[0] == '0') {      // same environment on every rank (torchrun)
    ctx->peer_disabled = true;
    return SFMBA_OK;
  }
  int rc;
  const int W = ctx->world;
  // ---- drop the old windows: everyone unmaps, then (barrier) everyone frees
  if (ctx->peer_local) {
    if ((rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "peer regrow sync")) != 0) return rc;
    close_peers(ctx);
    if ((rc = sfm_nccl_barrier(ctx, st)) != 0) return rc;
    cudaFree(ctx->peer_local);
    ctx->peer_local = nullptr;
    ctx->peer.world = 0;
  }
  const size_t cap = (doubles_cap + 255) & ~(size_t)255;
  const size_t bytes = PEER_HEADER_BYTES + 2 * (size_t)W * cap * sizeof(double);
  if ((rc = sfm_check_cuda(ctx, cudaMalloc(&ctx->peer_local, bytes), "cudaMalloc(peer window)")) != 0) return rc;
  if ((rc = sfm_check_cuda(ctx, cudaMemsetAsync(ctx->peer_local, 0, bytes, st), "memset(peer window)")) != 0) return rc;
  // ---- exchange the IPC handles
  cudaIpcMemHandle_t mine;
  int ok = cudaIpcGetMemHandle(&mine, ctx->peer_local) == cudaSuccess ? 1 : 0;
  if (!ok) {
    cudaGetLastError();
    memset(&mine, 0, sizeof(mine));
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  const size_t rec = 64 + 8;                      // handle + ok word
  void* dev = nullptr;
  if ((rc = sfm_check_cuda(ctx, ctx->reserve_misc(4096, &dev), "cudaMalloc(peer handles)")) != 0) return rc;
  std::vector<char> host(rec * (W + 1), 0);
  memcpy(host.data(), &mine, 64);
  memcpy(host.data() + 64, &ok, sizeof(int));
  char* dsend = reinterpret_cast<char*>(dev);
  char* drecv = dsend + rec;
  cudaMemcpyAsync(dsend, host.data(), rec, cudaMemcpyHostToDevice, st);
  if ((rc = sfm_nccl_allgather_bytes(ctx, dsend, drecv, rec, st)) != 0) return rc;
  cudaMemcpyAsync(host.data() + rec, drecv, rec * W, cudaMemcpyDeviceToHost, st);
  if ((rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "peer handle exchange")) != 0) return rc;
  // ---- map the peers
  int all_ok = 1;
  for (int r = 0; r < W; ++r) {
    int okr;
    memcpy(&okr, host.data() + rec * (r + 1) + 64, sizeof(int));
    all_ok &= okr;
  }
  if (all_ok) {
    for (int r = 0; r < W && all_ok; ++r) {
      if (r == ctx->rank) continue;
      cudaIpcMemHandle_t h;
      memcpy(&h, host.data() + rec * (r + 1), 64);
      void* p = nullptr;
      if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        all_ok = 0;
      } else {
        ctx->peer_mapped[r] = p;
      }
    }
  }
  // ---- agree: the sum of the per-rank verdicts must be W
  double verdict = all_ok ? 1.0 : 0.0;
  double* dverdict = reinterpret_cast<double*>(dsend);
  cudaMemcpyAsync(dverdict, &verdict, sizeof(double), cudaMemcpyHostToDevice, st);
  if ((rc = sfm_nccl_allreduce_f64(ctx, dverdict, 1, st)) != 0) return rc;
  cudaMemcpyAsync(&verdict, dverdict, sizeof(double), cudaMemcpyDeviceToHost, st);
  if ((rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "peer verdict")) != 0) return rc;
  if ((int)(verdict + 0.5) != W) {
    close_peers(ctx);
    if ((rc = sfm_nccl_barrier(ctx, st)) != 0) return rc;
    cudaFree(ctx->peer_local);
    ctx->peer_local = nullptr;
    ctx->peer_disabled = true;
    ctx->peer.world = 0;
    return SFMBA_OK;
  }
  ctx->peer.world = W;
  ctx->peer.rank = ctx->rank;
  ctx->peer.cap = cap;
  {
    // 20 s by default; SFMBA_PEER_TIMEOUT_MS shortens it (tests of the failure path)
    const char* ms = getenv("SFMBA_PEER_TIMEOUT_MS");
    const long long v = ms ? atoll(ms) : 0;
    ctx->peer.timeout_ns = (v > 0 ? (unsigned long long)v : 20000ull) * 1000000ull;
  }
  for (int r = 0; r < MAX_PEERS; ++r)
    ctx->peer.base[r] = r == ctx->rank ? reinterpret_cast<char*>(ctx->peer_local)
                                       : reinterpret_cast<char*>(r < W ? ctx->peer_mapped[r] : nullptr);
  return SFMBA_OK;
}
