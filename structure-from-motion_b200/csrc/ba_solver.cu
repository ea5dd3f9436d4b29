// Host driver of the BA solve: workspace carving, the LM / PCG launch sequence, collectives.
//
// Reference loop: ba_processor.py:297-406 (fixed lambda, `iteration` always-accepted steps).
// SFMBA_MODE_ADAPTIVE_LM adds cost evaluation, gain ratio and accept/reject, all decided on the
// device (the reference has none of these).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "ctx.h"

using namespace sfmba;

std::string g_last_error;

int sfm_set_error(sfmba_ctx* ctx, int code, const std::string& msg) {
  g_last_error = msg;
  if (ctx) ctx->err = msg;
  return code;
}

int sfm_check_cuda(sfmba_ctx* ctx, cudaError_t e, const char* what) {
  if (e == cudaSuccess) return SFMBA_OK;
  return sfm_set_error(ctx, SFMBA_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

cudaError_t DevBuf::reserve(size_t bytes) {
  if (bytes <= cap) return cudaSuccess;
  if (ptr) cudaFree(ptr);
  ptr = nullptr;
  cap = 0;
  size_t want = bytes + bytes / 8 + 256;
  cudaError_t e = cudaMalloc(&ptr, want);
  if (e != cudaSuccess) {
    want = bytes;
    e = cudaMalloc(&ptr, want);
  }
  if (e == cudaSuccess) cap = want;
  return e;
}

void DevBuf::release() {
  if (ptr) cudaFree(ptr);
  ptr = nullptr;
  cap = 0;
}

cudaError_t sfmba_ctx::reserve(int slot, size_t bytes, void** out) {
  if ((int)bufs.size() <= slot) bufs.resize(slot + 1);
  cudaError_t e = bufs[slot].reserve(bytes ? bytes : 16);
  *out = bufs[slot].ptr;
  return e;
}

namespace {

enum Slot {
  S_SCALARS = 0, S_CAMTAB, S_GRP_COST, S_GRP_EXV, S_GRP_PT, S_GRP_NCHUNK, S_GRP_K, S_GRP_CHUNK, S_GRP_META, S_DINV, S_VJ, S_UBUF, S_CAM_ACC, S_ABLK, S_MINV,
  S_CAMIT, S_CAMIT32, S_UBUF32, S_PART12F, S_PART64F, S_XREL, S_DINV32, S_SEGST, S_SVEC, S_X, S_R, S_Z, S_P, S_Q, S_S, S_GVEC, S_RED_A, S_RED_B, S_CAMS_BAK, S_PTS_BAK,
  S_CHUNK_INFO, S_REC_B, S_REC_IX, S_REC_W, S_SEG_CAM, S_SEG_LANE0, S_CAMROW_PTR, S_CAMROW_IDX, S_PART64, S_PART12,
  S_TEMP, S_SORT_KEYS, S_IOTA, S_DENSE, S_DENSE_LINV, S_DENSE_VEC, S_COUNT
};

template <typename T>
int carve(sfmba_ctx* ctx, int slot, size_t count, T** out) {
  void* p = nullptr;
  cudaError_t e = ctx->reserve(slot, count * sizeof(T), &p);
  *out = reinterpret_cast<T*>(p);
  return sfm_check_cuda(ctx, e, "cudaMalloc(workspace)");
}

}  // namespace

#define CARVE(slot, count, field) \
  if ((rc = carve(ctx, slot, (size_t)(count), &d.field)) != SFMBA_OK) return rc;

// Fills the BaDev view.  `full` = false carves only what K1 / the index-free cost need.
int sfm_setup_problem(sfmba_ctx* ctx, const sfmba_problem* pr, bool full, BaDev* out) {
  if (!pr || pr->num_cams < 0 || pr->num_pts < 0 || pr->num_obs < 0)
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "bad problem sizes");
  if (pr->num_cams == 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "num_cams == 0");
  if (!pr->cams || (pr->num_pts > 0 && !pr->pts) || !pr->pt_rowptr ||
      (pr->num_obs > 0 && (!pr->obs_uv || !pr->obs_cam || (!full && !pr->obs_pt))))
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "null pointer in problem");
  BaDev d;
  memset(&d, 0, sizeof(d));
  d.V = pr->num_cams;
  d.N = pr->num_pts;
  d.m = pr->num_obs;
  d.G = (int)(((long long)d.m + (long long)POINT_WEIGHT * d.N + GROUP_WEIGHT - 1) / GROUP_WEIGHT);
  d.obs_uv = pr->obs_uv;
  d.obs_cam = pr->obs_cam;
  d.obs_pt = pr->obs_pt;
  d.rowptr = pr->pt_rowptr;
  d.kinv = pr->cam_kinv;
  d.cams = pr->cams;
  d.pts = pr->pts;
  int rc;
  const size_t V = (size_t)d.V, N = (size_t)std::max(d.N, 1);
  CARVE(S_SCALARS, 1, sc);
  CARVE(S_CAMTAB, V * CAMTAB_STRIDE, camtab);
  CARVE(S_GRP_COST, std::max(d.G, 148 * 8) + 1, grp_cost);
  if (full) {
    CARVE(S_GRP_PT, d.G + 1, grp_pt);
    CARVE(S_GRP_NCHUNK, d.G + 1, grp_nchunk);
    CARVE(S_GRP_K, d.G + 1, grp_K);
    CARVE(S_GRP_CHUNK, d.G + 1, grp_chunk);
    CARVE(S_GRP_META, d.G + 1, grp_meta);
    CARVE(S_DINV, 6 * N, dinv);
    CARVE(S_VJ, 3 * N, vj);
    CARVE(S_UBUF, (size_t)PCG_UBUF * 3 * N, ubuf);
    CARVE(S_CAM_ACC, V * CAM_ACC + 8, cam_acc);
    CARVE(S_ABLK, V * 49, ablk);
    CARVE(S_MINV, V * 49, minv);
    CARVE(S_CAMIT, V * CAMIT_STRIDE, camit);
    CARVE(S_SVEC, 7 * V, svec);
    CARVE(S_X, 7 * V, x);
    CARVE(S_R, 7 * V, r);
    CARVE(S_Z, 7 * V, z);
    CARVE(S_P, 7 * V, p);
    CARVE(S_Q, 7 * V, q);
    CARVE(S_S, 7 * V, s);
    CARVE(S_GVEC, 7 * V, gvec);
    CARVE(S_RED_A, 2048, red_a);
    CARVE(S_RED_B, 2048, red_b);
    CARVE(S_CAMS_BAK, 7 * V, cams_backup);
    CARVE(S_PTS_BAK, 3 * N, pts_backup);
    CARVE(S_CAMROW_PTR, V + 1, camrow_ptr);
  }
  *out = d;
  return SFMBA_OK;
}

// chunk-sized arrays, once the number of chunks is known
static int sfm_setup_chunks(sfmba_ctx* ctx, int nchunks, BaDev* io) {
  BaDev& d = *io;
  int rc;
  const size_t nc = (size_t)std::max(nchunks, 1);
  d.chunk_bound = (int)nc;
  CARVE(S_CHUNK_INFO, nc, chunk_info);
  CARVE(S_REC_B, nc * CHUNK_CAP, rec_b);
  CARVE(S_REC_IX, nc * CHUNK_CAP, rec_ix);
  if (d.huber_delta > 0.0) CARVE(S_REC_W, nc * CHUNK_CAP, rec_w);
  CARVE(S_SEG_CAM, nc * SEG_CAP, seg_cam);
  CARVE(S_SEG_LANE0, nc * SEG_RUN_STRIDE, seg_lane0);
  CARVE(S_CAMROW_IDX, nc * SEG_CAP, camrow_idx);
  CARVE(S_PART64, nc * SEG_CAP * CAM_ACC, part64);
  CARVE(S_PART12, nc * SEG_CAP * MV_ACC, part12);
  return SFMBA_OK;
}
#undef CARVE

namespace {

struct PhaseTimer {
  sfmba_ctx* ctx;
  cudaStream_t st;
  float acc[6] = {0, 0, 0, 0, 0, 0};
  // phase timing needs event pairs; to stay asynchronous we record (begin,end) pairs lazily and
  // resolve them after the final synchronisation
  std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> spans;
  // events come from a pool the context keeps: creating and destroying ~40 of them per solve was host time spent
  // between two solves, with the GPU idle
  cudaEvent_t get() {
    if (ctx->timer_next == ctx->timer_events.size()) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      ctx->timer_events.push_back(e);
    }
    return ctx->timer_events[ctx->timer_next++];
  }
  // phase spans are recorded only on request (sfmba_profile_enable): ~40 event records and as many elapsed-time queries
  // per solve are host time a tight loop of solves would pay for numbers nobody reads.  The total is always timed.
  cudaEvent_t begin(bool always = false) {
    if (!always && !ctx->phase_timing && !ctx->profile) return nullptr;
    cudaEvent_t e = get();
    cudaEventRecord(e, st);
    return e;
  }
  void end(int phase, cudaEvent_t b) {
    if (!b) return;
    cudaEvent_t e = get();
    cudaEventRecord(e, st);
    spans.push_back({phase, {b, e}});
  }
  // named per-kernel spans, recorded only when ctx->profile is on (sfmba_profile_enable)
  struct KSpan { const char* name; cudaEvent_t b, e; };
  std::vector<KSpan> kspans;
  cudaEvent_t kbegin() { return ctx->profile ? begin(true) : nullptr; }
  void kend(const char* name, cudaEvent_t b) {
    if (!ctx->profile) return;
    cudaEvent_t e = get();
    cudaEventRecord(e, st);
    kspans.push_back({name, b, e});
  }
  void resolve() {
    for (auto& s : spans) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, s.second.first, s.second.second) == cudaSuccess) acc[s.first] += ms;
    }
    for (auto& k : kspans) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, k.b, k.e) == cudaSuccess) {
        ProfileEntry& pe = ctx->prof[k.name];
        pe.ms += ms;
        pe.launches += 1;
      }
    }
    kspans.clear();
    spans.clear();
    ctx->timer_next = 0;
  }
};

enum Phase { PH_INDEX = 0, PH_LINEARIZE, PH_CAMREDUCE, PH_PCG, PH_UPDATE, PH_TOTAL };

}  // namespace


int sfm_ba_solve_device(sfmba_ctx* ctx, const sfmba_problem* pr, const sfmba_options* opt_in,
                        sfmba_result* res, cudaStream_t st, const SolveIo* io) {
  sfmba_options opt;
  if (opt_in) opt = *opt_in; else sfmba_default_options(&opt);
  if (opt.iterations < 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "iterations < 0");
  if (opt.solver != SFMBA_SOLVER_PCG && opt.solver != SFMBA_SOLVER_DENSE_CHOLESKY)
    return sfm_set_error(ctx, SFMBA_ERR_UNSUPPORTED, "unknown solver");
  const bool dense = opt.solver == SFMBA_SOLVER_DENSE_CHOLESKY;
  if (!(opt.huber_delta >= 0.0)) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "huber_delta < 0");
  if (dense && opt.huber_delta > 0.0)
    return sfm_set_error(ctx, SFMBA_ERR_UNSUPPORTED, "the Huber loss is implemented for the PCG solver only");
  BaDev d;
  int rc = sfm_setup_problem(ctx, pr, true, &d);
  if (rc != SFMBA_OK) return rc;
  d.huber_delta = opt.huber_delta;
  if (opt.mode == SFMBA_MODE_ADAPTIVE_LM && (rc = carve(ctx, S_GRP_EXV, (size_t)d.G + 1, &d.grp_exv)) != 0) return rc;
  if (opt.precond != SFMBA_PRECOND_JACOBI && opt.precond != SFMBA_PRECOND_SCHUR_JACOBI)
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "precond is not an sfmba_precond");
  d.schur_diag = (dense || opt.precond == SFMBA_PRECOND_SCHUR_JACOBI) ? 1 : 0;
  const int n7 = 7 * d.V;
  DenseWork dw;
  memset(&dw, 0, sizeof(dw));
  if (dense) {
    dw.n = n7;
    dw.npad = (int)dense_padded_dim(d.V);
    dw.ld = (size_t)dw.npad;
    const size_t bytes = dw.ld * (size_t)dw.npad * sizeof(double);
    if (bytes > ((size_t)96 << 30)) {
      char buf[128];
      snprintf(buf, sizeof(buf), "dense solver: the %d x %d reduced matrix needs %.1f GB", n7, n7, bytes / 1e9);
      return sfm_set_error(ctx, SFMBA_ERR_UNSUPPORTED, buf);
    }
    if ((rc = carve(ctx, S_DENSE, dw.ld * (size_t)dw.npad, &dw.S)) != 0) return rc;
    if ((rc = carve(ctx, S_DENSE_LINV, (size_t)dw.npad * 128, &dw.linv)) != 0) return rc;
    if ((rc = carve(ctx, S_DENSE_VEC, (size_t)3 * dw.npad, &dw.vec)) != 0) return rc;
  }
  const double tol = opt.pcg_tol > 0 ? opt.pcg_tol : 1e-8;
  const int max_iter = opt.pcg_max_iter > 0 ? opt.pcg_max_iter : std::max(200, std::min(4 * n7, 5000));
  const bool quirk = opt.dvdc_sign_quirk != 0;
  const bool fine = opt.force_slow_path != 0;
  if (opt.matvec_precision < SFMBA_MATVEC_AUTO || opt.matvec_precision > SFMBA_MATVEC_F32)
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "matvec_precision is not an sfmba_matvec_precision");
  // FP32 matvec: PCG solver, plain least squares, peers or one GPU.  AUTO takes it when the requested tolerance is
  // within its reach (the operator itself is only good to ~1e-7) and falls back to FP64 iterations for slow solves.
  const bool f32_ok = !dense && opt.huber_delta == 0.0 && !(ctx->world > 1 && ctx->peer_disabled);
  const bool auto_prec = opt.matvec_precision == SFMBA_MATVEC_AUTO;
  const bool use_f32 = f32_ok && (opt.matvec_precision == SFMBA_MATVEC_F32 || (auto_prec && tol >= 1e-8));
  constexpr int F32_SAFE_ITERS = 8;
  // FP32 Jacobian arithmetic in the linearisation (ba_linearize32.cu): where the FP32 matvec runs, no Schur-diagonal sums
  static const int lin32_env = getenv("SFMBA_LIN32") ? atoi(getenv("SFMBA_LIN32")) : -1;
  int relinearized = 0;
  int f32_cond = 0;     // what the previous PCG solve OF THIS CALL said about the reduced system: 1 = converged fast, -1 = slow,
                        // 0 = none yet (nothing is carried over from earlier calls: the result never depends on them)
  int refine_iters = 0;
  const bool adaptive = opt.mode == SFMBA_MODE_ADAPTIVE_LM;
  const bool multi = ctx->world > 1;
  const int64_t launches0 = ctx->lc.launches;
  LaunchCounter& lc = ctx->lc;

  // several GPUs: map the peer windows (first solve, or more cameras than before); NCCL otherwise
  memset(&d.xw, 0, sizeof(d.xw));
  if (multi) {
    // a slot: plain doubles of the camera-sum exchange of the unfused path (+ the cost), then the flag-in-data areas
    // (16 bytes per entry) of the camera sums (+ the cost) and of the PCG matvec result
    const size_t ll_cam = ((size_t)d.V * CAM_ACC + 64 + 1) & ~(size_t)1;
    const size_t ll_y = ll_cam + 2 * ((size_t)d.V * CAM_ACC + 2);
    if ((rc = sfm_peer_setup(ctx, ll_y + 2 * (size_t)n7 + 16, st)) != 0) return rc;
    d.xw = ctx->peer;
    d.xw.ll_cam = ll_cam;
    d.xw.ll_y = ll_y;
  }
  const bool peer = multi && d.xw.world == ctx->world;
  // PCG progress word in pinned memory: polled by the host when the ranks need no matched collectives
  d.host_sig = (!multi || peer) ? reinterpret_cast<unsigned long long*>(ctx->pinned) : nullptr;

  PhaseTimer pt{ctx, st};
  ctx->timer_next = 0;                     // (a solve that failed half-way left its events checked out)
  cudaEvent_t t_total = pt.begin(true);

  // scalars (pageable source: the copy is staged by the driver before the call returns)
  Scalars h0;
  memset(&h0, 0, sizeof(h0));
  h0.lambda = opt.lambda;
  h0.nu = 2.0;
  if ((rc = sfm_check_cuda(ctx, cudaMemcpyAsync(d.sc, &h0, sizeof(h0), cudaMemcpyHostToDevice, st), "memcpy(scalars)")) != 0)
    return rc;

  // ---- index: count chunks (one host round trip), then build
  char* pin = reinterpret_cast<char*>(ctx->pinned) + 64;
  Scalars* hs = reinterpret_cast<Scalars*>(pin);
  double* hfinal = reinterpret_cast<double*>(pin + sizeof(Scalars));
  cudaEvent_t t0 = pt.begin();
  const int chunk_upper = d.G + d.m / SEG_CAP + 1;
  void* temp = nullptr;
  size_t temp_bytes = index_temp_bytes(d.G, 1, d.V);
  if ((rc = sfm_check_cuda(ctx, ctx->reserve(S_TEMP, temp_bytes, &temp), "cudaMalloc(temp)")) != 0) return rc;
  if (io) cudaStreamWaitEvent(st, io->ev_index, 0);
  volatile IndexStats* ist = reinterpret_cast<volatile IndexStats*>(reinterpret_cast<char*>(ctx->pinned) + 32);
  // sfmba_options.reuse_index: the caller vouches that the structure and the observations are those of the previous
  // device-buffer solve on this context; the index (groups, chunks, records, camera rows) is then still in the workspace
  IndexCache& ic = ctx->index_cache;
  const bool reuse = opt.reuse_index != 0 && !io && ic.valid && ic.V == d.V && ic.N == d.N && ic.m == d.m &&
                     ic.obs_uv == d.obs_uv && ic.obs_cam == d.obs_cam && ic.rowptr == d.rowptr && ic.kinv == d.kinv &&
                     ic.fine == (fine ? 1 : 0);
  ic.valid = false;
  const unsigned idx_epoch = ++ctx->pcg_epoch;
  if (!reuse) launch_index_count(d, fine, temp, temp_bytes, const_cast<IndexStats*>(ist), idx_epoch, st, lc);
  if (!reuse) {
    // the counting kernels' last act is to write the statistics into pinned memory: spin on their epoch word (a stream
    // synchronisation costs several times the latency); the stream is only queried to notice a failed launch
    long spins = 0;
    while (ist->epoch != idx_epoch) {
      if ((++spins & 0x3fff) == 0 && cudaStreamQuery(st) != cudaErrorNotReady) {
        if (ist->epoch == idx_epoch) break;
        rc = sfm_check_cuda(ctx, cudaGetLastError(), "index count");
        return rc ? rc : sfm_set_error(ctx, SFMBA_ERR_CUDA, "index count made no progress");
      }
    }
  }
  const int nchunks = reuse ? ic.nchunks : ist->num_chunks;
  if (!reuse && ((ist->flags & SFMBA_FLAG_BAD_INDEX) || nchunks < 0 || nchunks > chunk_upper))
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG,
                         "pt_rowptr is not a CSR over [0, num_obs] or obs_cam has an index outside [0, num_cams)");
  TileShape ts;
  if (reuse) {
    ts = ic.ts;
  } else {
    ts.max_chunk_len = std::max((int)ist->max_chunk_len, 1);
    ts.max_group_pts = std::max((int)ist->max_group_pts, 1);
    ts.max_nseg = std::max((int)ist->max_nseg, 1);
    ts.max_k = std::max((int)ist->max_k, 1);
  }
  if ((rc = sfm_setup_chunks(ctx, nchunks, &d)) != 0) return rc;
  if (use_f32 && !(multi && !peer)) {
    if ((rc = carve(ctx, S_XREL, (size_t)std::max(d.N, 1), &d.xrel)) != 0) return rc;
    if ((rc = carve(ctx, S_DINV32, (size_t)8 * std::max(d.N, 1), &d.dinv32)) != 0) return rc;
    if ((rc = carve(ctx, S_SEGST, (size_t)d.chunk_bound * SEG_CAP * 12, &d.segst)) != 0) return rc;
    if ((rc = carve(ctx, S_CAMIT32, (size_t)d.V * CAMIT_STRIDE, &d.camit32)) != 0) return rc;
    if ((rc = carve(ctx, S_PART64F, (size_t)d.chunk_bound * SEG_CAP * ROW_F32, &d.part64f)) != 0) return rc;
    if ((rc = carve(ctx, S_PART12F, (size_t)d.chunk_bound * SEG_CAP * MV_ROW32, &d.part12f)) != 0) return rc;
    if ((rc = carve(ctx, S_UBUF32, (size_t)PCG_UBUF * 3 * std::max(d.N, 1), &d.ubuf32)) != 0) return rc;
  }
  int32_t *sort_keys = nullptr, *iota = nullptr;
  temp_bytes = index_temp_bytes(d.G, d.chunk_bound, d.V);
  if ((rc = sfm_check_cuda(ctx, ctx->reserve(S_TEMP, temp_bytes, &temp), "cudaMalloc(temp)")) != 0) return rc;
  if ((rc = carve(ctx, S_SORT_KEYS, (size_t)d.chunk_bound * SEG_CAP, &sort_keys)) != 0) return rc;
  if ((rc = carve(ctx, S_IOTA, (size_t)d.chunk_bound * SEG_CAP, &iota)) != 0) return rc;
  d.rowpos = iota;   // the identity permutation fed to the sort; overwritten with the inverse permutation after it
  const bool ranged = io && io->num_ranges > 1 && io->range_grp[io->num_ranges] == d.G;
  if (reuse) {
    // nothing to build
  } else if (ranged) {
    // records of a range are built as soon as its observations have arrived
    for (int r = 0; r < io->num_ranges; ++r) {
      cudaStreamWaitEvent(st, io->ev_range_obs[r], 0);
      launch_chunk_build(d, io->range_grp[r], io->range_grp[r + 1], st, lc);
    }
  } else {
    if (io) cudaStreamWaitEvent(st, io->ev_obs, 0);
    launch_chunk_build(d, 0, d.G, st, lc);
  }
  if (!reuse) launch_index_build(d, nchunks, temp, temp_bytes, sort_keys, iota, st, lc);
  if (!io) {                               // (host-buffer solves index their own staging copies)
    ic.V = d.V; ic.N = d.N; ic.m = d.m;
    ic.obs_uv = d.obs_uv; ic.obs_cam = d.obs_cam; ic.rowptr = d.rowptr; ic.kinv = d.kinv;
    ic.fine = fine ? 1 : 0;
    ic.nchunks = nchunks;
    ic.ts = ts;
    ic.valid = true;
  }
  pt.end(PH_INDEX, t0);
  if (io) cudaStreamWaitEvent(st, io->ev_params, 0);      // ranged: the cameras; otherwise cameras and points
  if (ranged && adaptive) {
    for (int r = 0; r < io->num_ranges; ++r) cudaStreamWaitEvent(st, io->ev_range_pts[r], 0);   // (the backup reads every point)
  }

  static const bool fused_finish = !(getenv("SFMBA_FUSED_FINISH") && atoi(getenv("SFMBA_FUSED_FINISH")) == 0);
  // the FP32 linearisation writes FP32 segment rows, which only the fused camera finish reads
  const bool lin32_ok = use_f32 && d.xrel != nullptr && opt.precond == SFMBA_PRECOND_JACOBI && fused_finish && (!multi || peer);
  int cost_slot = 0;
  bool camtab_fresh = false;              // camtab holds the rows of the current cameras (k_update_cams leaves them so)
  for (int it = 0; it < opt.iterations; ++it) {
    t0 = pt.begin();
    if (adaptive) launch_lm_backup(d, st, lc);
    // FP32 Jacobian arithmetic in this linearisation?  F32: always.  AUTO: unless a PCG solve of this call has shown an
    // ill-conditioned reduced system (more than F32_SAFE_ITERS iterations).  The solve on an FP32 linearisation is watched:
    // the device gives it up as soon as its convergence rate predicts more than F32_SAFE_ITERS iterations (pcg_slow,
    // ba_pcg.cu) or, failing that, the host sees it take that many -- the LM iteration is then redone from an FP64
    // linearisation (the parameters are still untouched) and the rest of the call stays FP64.
    bool lin32 = lin32_ok && (lin32_env >= 0 ? lin32_env != 0
                                             : (opt.matvec_precision == SFMBA_MATVEC_F32 || (auto_prec && f32_cond >= 0)));
    int pcg_finished = -1;                 // iterations of this PCG solve when the host knows the number
    bool mv32 = false;                     // arithmetic of the matvecs being launched
   for (int attempt = 0;; ++attempt) {
    if (attempt > 0) t0 = pt.begin();
    if (!camtab_fresh) launch_cam_prepare(d, st, lc);
    camtab_fresh = true;
    // (the fused camera finish also sums the cost partials of the linearisation: no k_sum_cost launch)
    const bool finish_sums_cost = fused_finish && !dense && (!multi || peer);
    if (it == 0 && ranged) {
      // first linearisation of a host-buffer solve: range by range, behind the upload of the points
      for (int r = 0; r < io->num_ranges; ++r) {
        cudaStreamWaitEvent(st, io->ev_range_pts[r], 0);
        const bool close = r + 1 == io->num_ranges && !finish_sums_cost;
        if (lin32) launch_linearize32(d, ts, quirk, cost_slot, st, lc, io->range_grp[r], io->range_grp[r + 1], close);
        else launch_linearize(d, ts, quirk, cost_slot, st, lc, io->range_grp[r], io->range_grp[r + 1], close);
      }
    } else {
      cudaEvent_t kb = pt.kbegin();
      if (lin32) launch_linearize32(d, ts, quirk, cost_slot, st, lc, 0, -1, !finish_sums_cost);
      else launch_linearize(d, ts, quirk, cost_slot, st, lc, 0, -1, !finish_sums_cost);
      pt.kend(lin32 ? "linearize32" : "linearize", kb);
    }
    pt.end(PH_LINEARIZE, t0);

    t0 = pt.begin();
    const unsigned pcg_epoch = ++ctx->pcg_epoch;
    if (fused_finish && !dense && (!multi || peer)) {
      // PCG on one GPU or over peer windows: camera sums, blocks and the start of the PCG solve in one kernel per camera
      launch_cam_finish(d, peer, lin32, true, cost_slot, pcg_epoch, quirk, st, lc);
    } else {
      if (peer) {
        launch_camsum64_peer(d, cost_slot, st, lc);
      } else {
        launch_camsum64(d, st, lc);
      }
      if (multi && !peer) {
        // one message: the camera accumulators plus this rank's cost in the trailing slot
        cudaMemcpyAsync(d.cam_acc + (size_t)d.V * CAM_ACC, &d.sc->cost_cur, sizeof(double),
                        cudaMemcpyDeviceToDevice, st);
        if ((rc = sfm_nccl_allreduce_f64(ctx, d.cam_acc, (size_t)d.V * CAM_ACC + 1, st)) != 0) return rc;
        cudaMemcpyAsync(&d.sc->cost_cur, d.cam_acc + (size_t)d.V * CAM_ACC, sizeof(double),
                        cudaMemcpyDeviceToDevice, st);
        if (cost_slot < SFMBA_MAX_COST_HISTORY)
          cudaMemcpyAsync(&d.sc->costs[cost_slot], d.cam_acc + (size_t)d.V * CAM_ACC, sizeof(double),
                          cudaMemcpyDeviceToDevice, st);
      }
      launch_cam_blocks(d, st, lc);
      if (!dense) launch_pcg_init(d, pcg_epoch, quirk, st, lc);
    }
    pt.end(PH_CAMREDUCE, t0);

    // ---- PCG.  Kernels of an already-converged solve exit at their first instruction, so launching
    // past convergence is harmless but not free.  One GPU, or several with peer windows (the exchange
    // happens inside k_pcg_tail, and launches past convergence exit before it, so the ranks need not
    // agree on their number): k_pcg_tail publishes (epoch, iteration, done) in one 8-byte store to
    // pinned host memory; the host keeps LOOKAHEAD iterations queued and reads that word without any
    // synchronisation.  NCCL fallback: every rank must issue the same number of collectives, so
    // iterations go out in fixed batches and the (rank-identical) done flag of batch b is inspected
    // after batch b+1 has been queued.
    t0 = pt.begin();
    if (dense) {
      // direct solve: explicit S (this rank's off-diagonal blocks, summed over ranks), Cholesky, substitution
      cudaEvent_t kb = pt.kbegin();
      launch_dense_assemble(d, dw, quirk, st, lc);
      if (multi && (rc = sfm_nccl_allreduce_f64(ctx, dw.S, dw.ld * (size_t)dw.npad, st)) != 0) return rc;
      pt.kend("dense_assemble", kb);
      kb = pt.kbegin();
      launch_dense_factor_solve(d, dw, ctx->dense_streams, st, lc);
      pt.kend("dense_factor_solve", kb);
      launch_cam_rows(d, d.x, quirk, st, lc);
    }
    const int tail_mode = !multi ? 1 : (peer ? 3 : 2);
    // a solve on an FP32 linearisation that AUTO chose is on probation (see above)
    const bool watch_solve = lin32 && auto_prec && lin32_env < 0 && attempt == 0;
    int pcg_count = -1;                    // iterations of the solve's first run (before any FP64 continuation)
    pcg_finished = -1;
    mv32 = use_f32 && tail_mode != 2;
    unsigned cur_epoch = pcg_epoch;
    auto launch_iteration = [&]() -> int {
      cudaEvent_t kb = pt.kbegin();
      if (mv32) launch_matvec32(d, ts, quirk, st, lc); else launch_matvec(d, ts, false, quirk, st, lc);
      pt.kend("matvec", kb);
      if (tail_mode == 2) {
        launch_camsum12(d, st, lc);
        const int r2 = sfm_nccl_allreduce_f64(ctx, d.svec, (size_t)n7, st);
        if (r2 != 0) return r2;
      }
      kb = pt.kbegin();
      launch_pcg_tail(d, tail_mode, tol, max_iter, cur_epoch, quirk, mv32, watch_solve, st, lc);
      pt.kend("pcg_tail", kb);
      return 0;
    };
    if (dense) {
    } else if (tail_mode != 2) {
      // The iteration count of the previous solve is the hint: that many iterations are queued without
      // looking at the flag (consecutive LM iterations need the same count almost always), then the host
      // waits for the last of them; only if the solve is not done does it fall back to keeping LOOKAHEAD
      // iterations in flight.  A right hint costs no no-op launches and no GPU idle time.
      static const int LOOKAHEAD = getenv("SFMBA_LOOKAHEAD") ? std::max(1, atoi(getenv("SFMBA_LOOKAHEAD"))) : 2;
      volatile unsigned long long* sig = reinterpret_cast<volatile unsigned long long*>(ctx->pinned);
      int launched = 0, finished = 0;
      bool done = false, gave_up = false;    // gave_up: the device stopped a solve it predicts to be slow (abort_slow)
      // wait until iteration `want` (1-based) has finished or the solve is done; updates finished/done
      auto wait_for = [&](int want) -> int {
        long spins = 0;
        for (;;) {
          const unsigned long long w = *sig;
          if ((unsigned)(w >> 32) == cur_epoch) {
            finished = (int)((w & 0x7fffffffull) >> 1);
            if (w & 1ull) { done = true; gave_up = (w & 0x80000000ull) != 0; return 0; }
            if (finished >= want) return 0;
          }
          if ((++spins & 0x3fff) == 0 && cudaStreamQuery(st) != cudaErrorNotReady) {
            const unsigned long long w2 = *sig;       // stream drained: final state or an error
            if ((unsigned)(w2 >> 32) == cur_epoch) {
              finished = (int)((w2 & 0x7fffffffull) >> 1);
              if (w2 & 1ull) { done = true; gave_up = (w2 & 0x80000000ull) != 0; return 0; }
              if (finished >= want) return 0;
            }
            int r2 = sfm_check_cuda(ctx, cudaGetLastError(), "pcg kernels");
            return r2 ? r2 : sfm_set_error(ctx, SFMBA_ERR_CUDA, "PCG made no progress");
          }
        }
      };
      auto run = [&](int hint) -> int {
        for (; launched < hint; ++launched)
          if ((rc = launch_iteration()) != 0) return rc;
        if (hint > 0 && (rc = wait_for(hint)) != 0) return rc;
        while (launched < max_iter && !done) {
          if (launched >= LOOKAHEAD + finished) {
            if ((rc = wait_for(launched - LOOKAHEAD + 1)) != 0) return rc;
            if (done) break;
          }
          if ((rc = launch_iteration()) != 0) return rc;
          ++launched;
        }
        if (!done && (rc = wait_for(launched)) != 0) return rc;   // max_iter reached: the kernel sets done itself
        return 0;
      };
      if ((rc = run(std::min(ctx->pcg_hint, max_iter))) != 0) return rc;
      if (gave_up) {
        // (the hint for the next solve stays what it was)
        pcg_count = F32_SAFE_ITERS + 1;                      // counts as a slow solve below
        pcg_finished = -1;
      } else {
      ctx->pcg_hint = finished;
      pcg_finished = finished;
      pcg_count = finished;
      }
      if (!gave_up && mv32 && auto_prec && finished > F32_SAFE_ITERS && finished < max_iter && !watch_solve) {
        // AUTO: a solve this slow amplifies the 1e-7 operator error of the FP32 matvec too much to accept unseen.  Its
        // x is the starting point of an FP64 continuation: true residual from one FP64 matvec on x, then FP64 iterations
        // until that residual meets the tolerance (usually very few)
        mv32 = false;
        cur_epoch = ++ctx->pcg_epoch;
        launch_pcg_reopen(d, quirk, st, lc);
        cudaEvent_t kb = pt.kbegin();
        launch_matvec(d, ts, false, quirk, st, lc);
        pt.kend("matvec64", kb);
        launch_pcg_refresh(d, tail_mode, tol, cur_epoch, quirk, st, lc);
        launched = 0;
        finished = 0;
        done = false;
        if ((rc = wait_for(0)) != 0) return rc;              // the refresh itself may find the solve converged
        if (!done && (rc = run(0)) != 0) return rc;
        refine_iters += finished;
        pcg_finished = -1;                                   // x is no longer sum alpha_i p_i of one run: tile back-substitution
      }
    } else {
      const int batch = 4;
      int32_t* done_host = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(ctx->pinned) + 16);
      cudaEvent_t ev_prev = nullptr;
      bool have_prev = false;
      int launched = 0, which = 0;
      done_host[0] = done_host[1] = 0;
      while (launched < max_iter) {
        const int nb = std::min(batch, max_iter - launched);
        for (int k = 0; k < nb; ++k)
          if ((rc = launch_iteration()) != 0) return rc;
        launched += nb;
        cudaMemcpyAsync(&done_host[which], &d.sc->pcg_done, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
        cudaEvent_t ev = pt.get();
        cudaEventRecord(ev, st);
        if (have_prev) {
          cudaEventSynchronize(ev_prev);
          if (done_host[which ^ 1]) break;
        }
        ev_prev = ev;
        have_prev = true;
        which ^= 1;
      }
    }
    pt.end(PH_PCG, t0);
    if (pcg_count >= 0) f32_cond = pcg_count <= F32_SAFE_ITERS ? 1 : -1;
    if (lin32 && auto_prec && lin32_env < 0 && pcg_count > F32_SAFE_ITERS && attempt == 0) {
      lin32 = false;                       // mispredicted: this reduced system is not one to perturb by 1e-7
      ++relinearized;
      continue;
    }
    break;
   }

    t0 = pt.begin();
    {
      cudaEvent_t kb = pt.kbegin();
      if (pcg_finished >= 0 && pcg_finished <= PCG_UBUF) {
        launch_backsub_points(d, mv32, pcg_finished, st, lc);             // the PCG matvecs have left u_j = Dinv B^T p_i per point
      } else {
        // dense solve, NCCL fallback or a long PCG solve: one tile pass forms pts += v - Dinv B^T dp from the camit
        // rows of x
        if (!dense) launch_cam_rows(d, d.x, quirk, st, lc);
        launch_matvec(d, ts, true, quirk, st, lc);
      }
      pt.kend("backsub", kb);
    }
    launch_update_cams(d, st, lc);                      // (and their camtab rows)
    if (adaptive) {
      launch_cost_tiles(d, ts, nullptr, st, lc);       // -> sc->cost_new
      launch_lm_model_decrease(d, st, lc);             // -> sc->model_decrease, sc->md_pts (the next word)
      if (multi) {
        if (peer) launch_peer_sum_scalars(d, &d.sc->cost_new, 2, st, lc);
        else if ((rc = sfm_nccl_allreduce_f64(ctx, &d.sc->cost_new, 2, st)) != 0) return rc;
      }
      launch_lm_decide(d, opt.lm_lambda_min > 0 ? opt.lm_lambda_min : 1e-12,
                       opt.lm_lambda_max > 0 ? opt.lm_lambda_max : 1e12, st, lc);
      camtab_fresh = false;                            // a rejected step puts the old cameras back
    }
    pt.end(PH_UPDATE, t0);
    ++cost_slot;
  }

  if (ranged) {
    // (with zero iterations nothing above has waited for the points of the later upload ranges)
    for (int r = 0; r < io->num_ranges; ++r) cudaStreamWaitEvent(st, io->ev_range_pts[r], 0);
  }
  if (io) {
    // parameters are final: their download runs on the copy stream beside the final cost pass
    cudaEventRecord(io->ev_final, st);
    cudaStreamWaitEvent(io->copy, io->ev_final, 0);
    cudaMemcpyAsync(io->host_cams, d.cams, (size_t)7 * d.V * sizeof(double), cudaMemcpyDeviceToHost, io->copy);
    if (d.N) cudaMemcpyAsync(io->host_pts, d.pts, (size_t)3 * d.N * sizeof(double), cudaMemcpyDeviceToHost, io->copy);
  }
  // final cost
  double* final_cost_dev = d.cam_acc + (size_t)d.V * CAM_ACC + 1;
  if (opt.compute_costs || adaptive) {
    if (!camtab_fresh) launch_cam_prepare(d, st, lc);
    cudaEvent_t kb = pt.kbegin();
    launch_cost_tiles(d, ts, final_cost_dev, st, lc);
    pt.kend("cost", kb);
    if (peer) launch_peer_sum_scalars(d, final_cost_dev, 1, st, lc);
    else if (multi && (rc = sfm_nccl_allreduce_f64(ctx, final_cost_dev, 1, st)) != 0) return rc;
  }
  pt.end(PH_TOTAL, t_total);

  // results
  *hfinal = 0.0;
  cudaMemcpyAsync(hs, d.sc, sizeof(Scalars), cudaMemcpyDeviceToHost, st);
  if (opt.compute_costs || adaptive)
    cudaMemcpyAsync(hfinal, final_cost_dev, sizeof(double), cudaMemcpyDeviceToHost, st);
  if ((rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "solve sync")) != 0) return rc;
  if (io && (rc = sfm_check_cuda(ctx, cudaStreamSynchronize(io->copy), "D2H sync")) != 0) return rc;
  if ((rc = sfm_check_cuda(ctx, cudaGetLastError(), "solve kernels")) != 0) return rc;
  pt.resolve();
  if (res) {
    memset(res, 0, sizeof(*res));
    res->iterations_done = opt.iterations;
    res->iterations_accepted = adaptive ? hs->lm_accepted : opt.iterations;
    res->pcg_iterations = hs->pcg_iters_total;
    res->flags = hs->flags;
    res->final_lambda = hs->lambda;
    const int nc = std::min(opt.iterations, SFMBA_MAX_COST_HISTORY - 1);
    for (int k = 0; k < nc; ++k) res->costs[k] = hs->costs[k];
    res->num_costs = nc;
    res->initial_cost = nc > 0 ? hs->costs[0] : *hfinal;
    res->final_cost = *hfinal;
    if (opt.compute_costs || adaptive) res->costs[res->num_costs++] = *hfinal;
    res->ms_index = pt.acc[PH_INDEX];
    res->ms_linearize = pt.acc[PH_LINEARIZE];
    res->ms_camreduce = pt.acc[PH_CAMREDUCE];
    res->ms_pcg = pt.acc[PH_PCG];
    res->ms_update = pt.acc[PH_UPDATE];
    res->ms_total = pt.acc[PH_TOTAL];
    res->kernel_launches = (int32_t)(ctx->lc.launches - launches0);
    res->pcg_refine_iterations = refine_iters;
    res->relinearized = relinearized;
    res->matvec_precision = (use_f32 && d.xrel) ? SFMBA_MATVEC_F32 : SFMBA_MATVEC_F64;
  }
  if (hs->flags & SFMBA_FLAG_PEER_TIMEOUT)
    return sfm_set_error(ctx, SFMBA_ERR_COMM,
                         "a peer rank never delivered its partial sums (20 s): the cross-rank sums of that iteration are "
                         "incomplete and its update was not applied");
  if (hs->flags & (SFMBA_FLAG_QW_NEAR_ZERO | SFMBA_FLAG_NONFINITE | SFMBA_FLAG_DEPTH_ZERO | SFMBA_FLAG_SINGULAR)) {
    char buf[96];
    snprintf(buf, sizeof(buf), "device numeric flags 0x%x", hs->flags);
    return sfm_set_error(ctx, SFMBA_ERR_NUMERIC, buf);
  }
  return SFMBA_OK;
}
