// extern "C" surface of libsfmba (include/sfmba.h).  No exceptions cross this file.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "ctx.h"

using namespace sfmba;

int sfm_setup_problem(sfmba_ctx* ctx, const sfmba_problem* pr, bool full, BaDev* out);
int sfm_ba_solve_device(sfmba_ctx* ctx, const sfmba_problem* pr, const sfmba_options* opt, sfmba_result* res,
                        cudaStream_t st, const SolveIo* io);

namespace {

constexpr size_t PINNED_BYTES = 1 << 16;

// device copy of a host array in ctx->host_stage[slot]
template <typename T>
int stage_in(sfmba_ctx* ctx, int slot, const T* host, size_t count, T** dev, cudaStream_t st) {
  *dev = nullptr;
  if (!host || count == 0) {
    // keep a valid non-null pointer for zero-sized arrays
    cudaError_t e = ctx->host_stage[slot].reserve(16);
    if (e != cudaSuccess) return sfm_check_cuda(ctx, e, "cudaMalloc(stage)");
    *dev = ctx->host_stage[slot].as<T>();
    return SFMBA_OK;
  }
  cudaError_t e = ctx->host_stage[slot].reserve(count * sizeof(T));
  if (e != cudaSuccess) return sfm_check_cuda(ctx, e, "cudaMalloc(stage)");
  *dev = ctx->host_stage[slot].as<T>();
  e = cudaMemcpyAsync(*dev, host, count * sizeof(T), cudaMemcpyHostToDevice, st);
  return sfm_check_cuda(ctx, e, "cudaMemcpyAsync(H2D)");
}

int read_flags(sfmba_ctx* ctx, uint32_t* dflags, uint32_t* host_flags, cudaStream_t st) {
  uint32_t* pin = reinterpret_cast<uint32_t*>(ctx->pinned);
  cudaMemcpyAsync(pin, dflags, sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
  int rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "refine sync");
  if (rc) return rc;
  rc = sfm_check_cuda(ctx, cudaGetLastError(), "refine kernels");
  if (rc) return rc;
  if (host_flags) *host_flags = *pin;
  if (*pin & SFMBA_FLAG_BAD_INDEX)
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "rowptr is not a CSR over [0, num_obs] or an index is out of range");
  if (*pin & (SFMBA_FLAG_QW_NEAR_ZERO | SFMBA_FLAG_NONFINITE)) {
    char buf[64];
    snprintf(buf, sizeof(buf), "device numeric flags 0x%x", *pin);
    return sfm_set_error(ctx, SFMBA_ERR_NUMERIC, buf);
  }
  return SFMBA_OK;
}

}  // namespace

extern "C" {

int sfmba_version(void) { return SFMBA_VERSION; }

void sfmba_default_options(sfmba_options* opt) {
  if (!opt) return;
  memset(opt, 0, sizeof(*opt));
  opt->lambda = 5.0;            // ba_processor.py:24
  opt->iterations = 3;          // ba_processor.py:24
  opt->mode = SFMBA_MODE_REFERENCE_FIXED;
  opt->solver = SFMBA_SOLVER_PCG;
  opt->pcg_max_iter = 0;
  opt->pcg_tol = 1e-8;
  opt->dvdc_sign_quirk = 1;
  opt->force_slow_path = 0;
  opt->compute_costs = 1;
  opt->precond = SFMBA_PRECOND_JACOBI;
  opt->lm_lambda_min = 1e-12;
  opt->lm_lambda_max = 1e12;
  opt->matvec_precision = SFMBA_MATVEC_AUTO;
}

const char* sfmba_last_error(const sfmba_ctx* ctx) { return ctx ? ctx->err.c_str() : g_last_error.c_str(); }

int sfmba_nccl_unique_id(void* out128) {
  if (!out128) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null id buffer");
  std::string err;
  const int rc = sfm_nccl_unique_id(out128, err);
  if (rc) return sfm_set_error(nullptr, rc, err);
  return SFMBA_OK;
}

int sfmba_create(sfmba_ctx** out, int device, int world_size, int rank, const void* nccl_unique_id) {
  if (!out) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null out");
  *out = nullptr;
  if (world_size < 1 || rank < 0 || rank >= world_size)
    return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "bad world_size / rank");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return sfm_set_error(nullptr, SFMBA_ERR_NO_DEVICE,
                         "no CUDA device: libsfmba has no CPU fallback (sm_100a kernels only)");
  }
  if (device < 0 || device >= ndev) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "bad device index");
  e = cudaSetDevice(device);
  if (e != cudaSuccess) return sfm_check_cuda(nullptr, e, "cudaSetDevice");
  sfmba_ctx* ctx = new (std::nothrow) sfmba_ctx();
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_CUDA, "out of host memory");
  ctx->device = device;
  ctx->world = world_size;
  ctx->rank = rank;
  e = cudaMallocHost(&ctx->pinned, PINNED_BYTES);
  if (e != cudaSuccess) {
    delete ctx;
    return sfm_check_cuda(nullptr, e, "cudaMallocHost");
  }
  ctx->pinned_cap = PINNED_BYTES;
  memset(ctx->pinned, 0, PINNED_BYTES);
  cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking);
  if (world_size > 1) {
    if (!nccl_unique_id) {
      sfmba_destroy(ctx);
      return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "world_size > 1 needs an ncclUniqueId");
    }
    const int rc = sfm_nccl_init(ctx, nccl_unique_id);
    if (rc) {
      g_last_error = ctx->err;
      sfmba_destroy(ctx);
      return rc;
    }
  }
  *out = ctx;
  return SFMBA_OK;
}

int sfmba_destroy(sfmba_ctx* ctx) {
  if (!ctx) return SFMBA_OK;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  sfm_peer_destroy(ctx);
  sfm_nccl_destroy(ctx);
  ctx->misc.release();
  for (auto& b : ctx->bufs) b.release();
  for (auto& b : ctx->host_stage) b.release();
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  for (auto e : ctx->timer_events) cudaEventDestroy(e);
  ctx->timer_events.clear();
  ctx->dense_streams.release();
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  if (ctx->io.copy) {
    cudaStreamDestroy(ctx->io.copy);
    cudaEventDestroy(ctx->io.ev_index);
    cudaEventDestroy(ctx->io.ev_obs);
    cudaEventDestroy(ctx->io.ev_params);
    cudaEventDestroy(ctx->io.ev_final);
    for (int r = 0; r < SolveIo::MAX_RANGES; ++r) {
      if (ctx->io.ev_range_obs[r]) cudaEventDestroy(ctx->io.ev_range_obs[r]);
      if (ctx->io.ev_range_pts[r]) cudaEventDestroy(ctx->io.ev_range_pts[r]);
    }
  }
  delete ctx;
  return SFMBA_OK;
}

int sfmba_ba_solve(sfmba_ctx* ctx, const sfmba_problem* problem, const sfmba_options* opt, sfmba_result* result,
                   void* stream) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (problem && problem->flags != 0)
    return sfm_set_error(ctx, SFMBA_ERR_UNSUPPORTED, "sfmba_problem.flags: the compact layouts are for sfmba_ba_solve_host");
  cudaSetDevice(ctx->device);
  return sfm_ba_solve_device(ctx, problem, opt, result, (cudaStream_t)stream, nullptr);
}

int sfmba_ba_solve_host(sfmba_ctx* ctx, const sfmba_problem* pr, const sfmba_options* opt, sfmba_result* result) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (!pr) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "null problem");
  cudaSetDevice(ctx->device);
  cudaStream_t st = ctx->own_stream;
  SolveIo& io = ctx->io;
  if (!io.copy) {
    cudaStreamCreateWithFlags(&io.copy, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&io.ev_index, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&io.ev_obs, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&io.ev_params, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&io.ev_final, cudaEventDisableTiming);
  }
  sfmba_problem dp = *pr;
  const size_t m = (size_t)std::max(pr->num_obs, 0), N = (size_t)std::max(pr->num_pts, 0),
               V = (size_t)std::max(pr->num_cams, 0);
  int rc;
  float* d_uv;
  int32_t *d_cam, *d_row;
  double *d_kinv = nullptr, *d_cams, *d_pts;
  // uploads in the order of first use: index inputs, cameras, observations, points.  Large problems are cut into
  // ranges of whole groups (the solver's own partition: group g starts at the first point j with
  // pt_rowptr[j] + 2 j >= 1024 g, ba_index.cu) so that the record build and the first linearisation of a range run
  // while the later ranges are still on the bus.
  if ((rc = stage_in(ctx, 3, pr->pt_rowptr, N + 1, &d_row, io.copy))) return rc;
  if (pr->flags & SFMBA_PROBLEM_CAM_U16) {
    // half-width camera indices on the bus, widened on the device behind the copy
    if (pr->num_cams > 65536) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "SFMBA_PROBLEM_CAM_U16 needs num_cams <= 65536");
    uint16_t* d_cam16;
    if ((rc = stage_in(ctx, 8, reinterpret_cast<const uint16_t*>(pr->obs_cam), m, &d_cam16, io.copy))) return rc;
    cudaError_t e = ctx->host_stage[1].reserve(std::max<size_t>(16, m * sizeof(int32_t)));
    if (e != cudaSuccess) return sfm_check_cuda(ctx, e, "cudaMalloc(stage)");
    d_cam = ctx->host_stage[1].as<int32_t>();
    launch_widen_u16(d_cam16, d_cam, m, io.copy, ctx->lc);
  } else {
    if ((rc = stage_in(ctx, 1, pr->obs_cam, m, &d_cam, io.copy))) return rc;
  }
  cudaEventRecord(io.ev_index, io.copy);
  if (pr->cam_kinv && (rc = stage_in(ctx, 4, pr->cam_kinv, 9 * V, &d_kinv, io.copy))) return rc;
  if ((rc = stage_in(ctx, 5, pr->cams, 7 * V, &d_cams, io.copy))) return rc;
  io.num_ranges = 0;
  const long long G = ((long long)m + 2ll * (long long)N + 1023) / 1024;
  // (measured at the Venice shape: 5.87 ms per call in one piece, 5.64 ms in two ranges, 5.67 in four, 5.78 in eight:
  // profiles/prof_e2e.py)
  int R = (m >= ((size_t)1 << 20) && pr->pt_rowptr && pr->obs_uv && pr->pts) ? 2 : 1;
  if (G < 64) R = 1;
  if (const char* ev = getenv("SFMBA_UPLOAD_RANGES")) R = std::max(1, std::min(atoi(ev), (int)SolveIo::MAX_RANGES));
  if (R > 1) {
    // range boundaries in groups, points and observations; clamped and monotone whatever pt_rowptr holds
    size_t rp[SolveIo::MAX_RANGES + 1], ro[SolveIo::MAX_RANGES + 1];
    for (int r = 0; r <= R; ++r) {
      const long long g = r == R ? G : G * r / R;
      io.range_grp[r] = (int)g;
      size_t lo = 0, hi = N;
      if (r == R) {
        lo = N;
      } else {
        const long long target = g * 1024;
        while (lo < hi) {
          const size_t mid = (lo + hi) >> 1;
          if ((long long)pr->pt_rowptr[mid] + 2ll * (long long)mid < target) lo = mid + 1; else hi = mid;
        }
      }
      rp[r] = lo;
      long long o = r == R ? (long long)m : (long long)pr->pt_rowptr[lo];
      if (o < 0) o = 0;
      if (o > (long long)m) o = (long long)m;
      ro[r] = (size_t)o;
      if (r > 0 && ro[r] < ro[r - 1]) ro[r] = ro[r - 1];
      if (r > 0 && rp[r] < rp[r - 1]) rp[r] = rp[r - 1];
    }
    ro[0] = 0;
    rp[0] = 0;
    for (int r = 0; r < R; ++r) {
      if (!io.ev_range_obs[r]) {
        cudaEventCreateWithFlags(&io.ev_range_obs[r], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&io.ev_range_pts[r], cudaEventDisableTiming);
      }
    }
    cudaError_t e = ctx->host_stage[0].reserve(std::max<size_t>(16, 2 * m * sizeof(float)));
    if (e == cudaSuccess) e = ctx->host_stage[6].reserve(std::max<size_t>(16, 3 * N * sizeof(double)));
    if (e != cudaSuccess) return sfm_check_cuda(ctx, e, "cudaMalloc(stage)");
    d_uv = ctx->host_stage[0].as<float>();
    d_pts = ctx->host_stage[6].as<double>();
    cudaEventRecord(io.ev_params, io.copy);                   // ranged: "cameras resident"
    for (int r = 0; r < R; ++r) {
      const size_t o0 = ro[r], cnt = ro[r + 1] - ro[r];
      if (cnt) {
        cudaMemcpyAsync(d_uv + o0, pr->obs_uv + o0, cnt * sizeof(float), cudaMemcpyHostToDevice, io.copy);
        cudaMemcpyAsync(d_uv + m + o0, pr->obs_uv + m + o0, cnt * sizeof(float), cudaMemcpyHostToDevice, io.copy);
      }
      cudaEventRecord(io.ev_range_obs[r], io.copy);
    }
    cudaEventRecord(io.ev_obs, io.copy);
    for (int r = 0; r < R; ++r) {
      const size_t p0 = rp[r], cnt = rp[r + 1] - rp[r];
      if (cnt)
        cudaMemcpyAsync(d_pts + 3 * p0, pr->pts + 3 * p0, 3 * cnt * sizeof(double), cudaMemcpyHostToDevice, io.copy);
      cudaEventRecord(io.ev_range_pts[r], io.copy);
    }
    if ((rc = sfm_check_cuda(ctx, cudaGetLastError(), "ranged upload")) != 0) return rc;
    io.num_ranges = R;
  } else {
    if ((rc = stage_in(ctx, 0, pr->obs_uv, 2 * m, &d_uv, io.copy))) return rc;
    cudaEventRecord(io.ev_obs, io.copy);
    if ((rc = stage_in(ctx, 6, pr->pts, 3 * N, &d_pts, io.copy))) return rc;
    cudaEventRecord(io.ev_params, io.copy);
  }
  dp.obs_uv = d_uv;
  dp.obs_cam = d_cam;
  dp.obs_pt = nullptr;   // the solve derives the point of an observation from pt_rowptr
  dp.pt_rowptr = d_row;
  dp.cam_kinv = pr->cam_kinv ? d_kinv : nullptr;
  dp.cams = d_cams;
  dp.pts = d_pts;
  io.host_cams = pr->cams;
  io.host_pts = pr->pts;
  rc = sfm_ba_solve_device(ctx, &dp, opt, result, st, &io);   // downloads cams / pts and waits for both streams
  if (rc != SFMBA_OK && rc != SFMBA_ERR_NUMERIC) {
    cudaStreamSynchronize(io.copy);                           // nothing of this call stays in flight
    cudaStreamSynchronize(st);
  }
  return rc;
}

namespace {
// K1 alone; reads the device flag word back (one synchronisation) so that bad indices and zero depths are reported
int residual_jacobian_impl(sfmba_ctx* ctx, const sfmba_problem* problem, int dvdc_sign_quirk, float* out_r, float* out_jp,
                           float* out_jx, double* out_r64, double* out_jp64, double* out_jx64, double* out_cost,
                           void* stream) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  cudaSetDevice(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  BaDev d;
  int rc = sfm_setup_problem(ctx, problem, false, &d);
  if (rc) return rc;
  cudaMemsetAsync(d.sc, 0, sizeof(Scalars), st);
  launch_cam_prepare(d, st, ctx->lc);
  launch_k1(d, dvdc_sign_quirk != 0, out_r, out_jp, out_jx, out_r64, out_jp64, out_jx64, out_cost, st, ctx->lc);
  uint32_t* pin_flags = reinterpret_cast<uint32_t*>(ctx->pinned);
  cudaMemcpyAsync(pin_flags, &d.sc->flags, sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
  if ((rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "k1 sync")) != 0) return rc;
  if ((rc = sfm_check_cuda(ctx, cudaGetLastError(), "k1 launch")) != 0) return rc;
  if (*pin_flags & SFMBA_FLAG_BAD_INDEX)
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG,
                         "obs_cam / obs_pt has an index outside its array (the outputs of those observations were not written)");
  if (*pin_flags & (SFMBA_FLAG_DEPTH_ZERO | SFMBA_FLAG_QW_NEAR_ZERO | SFMBA_FLAG_NONFINITE)) {
    char buf[64];
    snprintf(buf, sizeof(buf), "device numeric flags 0x%x", *pin_flags);
    return sfm_set_error(ctx, SFMBA_ERR_NUMERIC, buf);
  }
  return SFMBA_OK;
}
}  // namespace

int sfmba_residual_jacobian(sfmba_ctx* ctx, const sfmba_problem* problem, int dvdc_sign_quirk, float* out_r,
                            float* out_jp, float* out_jx, double* out_r64, double* out_cost, void* stream) {
  return residual_jacobian_impl(ctx, problem, dvdc_sign_quirk, out_r, out_jp, out_jx, out_r64, nullptr, nullptr, out_cost,
                                stream);
}

int sfmba_residual_jacobian_f64(sfmba_ctx* ctx, const sfmba_problem* problem, int dvdc_sign_quirk, double* out_r64,
                                double* out_jp64, double* out_jx64, double* out_cost, void* stream) {
  return residual_jacobian_impl(ctx, problem, dvdc_sign_quirk, nullptr, nullptr, nullptr, out_r64, out_jp64, out_jx64,
                                out_cost, stream);
}

int sfmba_ba_cost(sfmba_ctx* ctx, const sfmba_problem* problem, double* host_cost, void* stream) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  cudaSetDevice(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  BaDev d;
  int rc = sfm_setup_problem(ctx, problem, false, &d);
  if (rc) return rc;
  cudaMemsetAsync(d.sc, 0, sizeof(Scalars), st);
  launch_cam_prepare(d, st, ctx->lc);
  double* dev_cost = d.grp_cost;    // k_sum_cost reads the partials before thread 0 writes slot 0
  launch_cost(d, dev_cost, st, ctx->lc);
  double* pin = reinterpret_cast<double*>(ctx->pinned);
  uint32_t* pin_flags = reinterpret_cast<uint32_t*>(pin + 1);
  cudaMemcpyAsync(pin, dev_cost, sizeof(double), cudaMemcpyDeviceToHost, st);
  cudaMemcpyAsync(pin_flags, &d.sc->flags, sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
  rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "cost sync");
  if (rc) return rc;
  if (*pin_flags & SFMBA_FLAG_BAD_INDEX)
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "obs_cam / obs_pt has an index outside its array");
  if (host_cost) *host_cost = *pin;
  return sfm_check_cuda(ctx, cudaGetLastError(), "cost kernels");
}

int sfmba_triangulate_refine_batched(sfmba_ctx* ctx, int32_t num_pts, int32_t num_obs, int32_t num_projs,
                                     const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                                     const double* obs_uv, double* pts, double lambda, int32_t iterations,
                                     uint32_t* host_flags, void* stream) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (num_pts < 0 || num_obs < 0 || num_projs < 0 || iterations < 0)
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "negative size");
  cudaSetDevice(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  void* fl = nullptr;
  int rc = sfm_check_cuda(ctx, ctx->reserve(0, sizeof(Scalars), &fl), "cudaMalloc(flags)");
  if (rc) return rc;
  uint32_t* dflags = reinterpret_cast<uint32_t*>(fl);
  cudaMemsetAsync(dflags, 0, sizeof(uint32_t), st);
  launch_validate_csr(num_pts, num_obs, rowptr, obs_view, num_projs, dflags, st, ctx->lc);
  launch_triangulate(num_pts, projs, rowptr, obs_view, obs_uv, pts, lambda, iterations, dflags, st, ctx->lc);
  return read_flags(ctx, dflags, host_flags, st);
}

int sfmba_triangulate_refine_batched_host(sfmba_ctx* ctx, int32_t num_pts, int32_t num_obs, int32_t num_projs,
                                          const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                                          const double* obs_uv, double* pts, double lambda, int32_t iterations,
                                          uint32_t* host_flags) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (num_pts < 0 || num_obs < 0 || num_projs < 0)
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "negative size");
  cudaSetDevice(ctx->device);
  cudaStream_t st = ctx->own_stream;
  double *d_projs, *d_uv, *d_pts;
  int32_t *d_row, *d_view;
  int rc;
  if ((rc = stage_in(ctx, 0, projs, (size_t)12 * num_projs, &d_projs, st))) return rc;
  if ((rc = stage_in(ctx, 1, rowptr, (size_t)num_pts + 1, &d_row, st))) return rc;
  if ((rc = stage_in(ctx, 2, obs_view, (size_t)num_obs, &d_view, st))) return rc;
  if ((rc = stage_in(ctx, 3, obs_uv, (size_t)2 * num_obs, &d_uv, st))) return rc;
  if ((rc = stage_in(ctx, 4, pts, (size_t)3 * num_pts, &d_pts, st))) return rc;
  rc = sfmba_triangulate_refine_batched(ctx, num_pts, num_obs, num_projs, d_projs, d_row, d_view, d_uv, d_pts,
                                        lambda, iterations, host_flags, st);
  if (rc != SFMBA_OK && rc != SFMBA_ERR_NUMERIC) return rc;
  if (num_pts) cudaMemcpyAsync(pts, d_pts, (size_t)3 * num_pts * sizeof(double), cudaMemcpyDeviceToHost, st);
  const int rc2 = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "D2H sync");
  return rc2 ? rc2 : rc;
}

int sfmba_triangulate_linear_batched(sfmba_ctx* ctx, int32_t num_pts, int32_t num_obs, int32_t num_projs,
                                     const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                                     const double* obs_uv, double* pts, int32_t* front, uint32_t* host_flags,
                                     void* stream) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (num_pts < 0 || num_obs < 0 || num_projs < 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "negative size");
  cudaSetDevice(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  void* fl = nullptr;
  int rc = sfm_check_cuda(ctx, ctx->reserve(0, sizeof(Scalars), &fl), "cudaMalloc(flags)");
  if (rc) return rc;
  uint32_t* dflags = reinterpret_cast<uint32_t*>(fl);
  cudaMemsetAsync(dflags, 0, sizeof(uint32_t), st);
  launch_validate_csr(num_pts, num_obs, rowptr, obs_view, num_projs, dflags, st, ctx->lc);
  launch_triangulate_linear(num_pts, projs, rowptr, obs_view, obs_uv, pts, front, dflags, st, ctx->lc);
  uint32_t* pin = reinterpret_cast<uint32_t*>(ctx->pinned);
  cudaMemcpyAsync(pin, dflags, sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
  if ((rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "triangulate sync")) != 0) return rc;
  if ((rc = sfm_check_cuda(ctx, cudaGetLastError(), "triangulate kernel")) != 0) return rc;
  if (host_flags) *host_flags = *pin;
  if (*pin & SFMBA_FLAG_BAD_INDEX)
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "rowptr is not a CSR over [0, num_obs] or an index is out of range");
  if (*pin & SFMBA_FLAG_NONFINITE) return sfm_set_error(ctx, SFMBA_ERR_NUMERIC, "linear triangulation: point at infinity");
  return SFMBA_OK;
}

int sfmba_triangulate_linear_batched_host(sfmba_ctx* ctx, int32_t num_pts, int32_t num_obs, int32_t num_projs,
                                          const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                                          const double* obs_uv, double* pts, int32_t* front, uint32_t* host_flags) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (num_pts < 0 || num_obs < 0 || num_projs < 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "negative size");
  cudaSetDevice(ctx->device);
  cudaStream_t st = ctx->own_stream;
  double *d_projs, *d_uv;
  int32_t *d_row, *d_view;
  int rc;
  if ((rc = stage_in(ctx, 0, projs, (size_t)12 * num_projs, &d_projs, st))) return rc;
  if ((rc = stage_in(ctx, 1, rowptr, (size_t)num_pts + 1, &d_row, st))) return rc;
  if ((rc = stage_in(ctx, 2, obs_view, (size_t)num_obs, &d_view, st))) return rc;
  if ((rc = stage_in(ctx, 3, obs_uv, (size_t)2 * num_obs, &d_uv, st))) return rc;
  cudaError_t e = ctx->host_stage[4].reserve(std::max<size_t>(16, (size_t)3 * num_pts * sizeof(double)));
  if (e != cudaSuccess) return sfm_check_cuda(ctx, e, "cudaMalloc(stage)");
  e = ctx->host_stage[7].reserve(std::max<size_t>(16, (size_t)num_pts * sizeof(int32_t)));
  if (e != cudaSuccess) return sfm_check_cuda(ctx, e, "cudaMalloc(stage)");
  double* d_pts = ctx->host_stage[4].as<double>();
  int32_t* d_front = front ? ctx->host_stage[7].as<int32_t>() : nullptr;
  rc = sfmba_triangulate_linear_batched(ctx, num_pts, num_obs, num_projs, d_projs, d_row, d_view, d_uv, d_pts, d_front,
                                        host_flags, st);
  if (rc != SFMBA_OK && rc != SFMBA_ERR_NUMERIC) return rc;
  if (num_pts) {
    cudaMemcpyAsync(pts, d_pts, (size_t)3 * num_pts * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (front) cudaMemcpyAsync(front, d_front, (size_t)num_pts * sizeof(int32_t), cudaMemcpyDeviceToHost, st);
  }
  const int rc2 = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "D2H sync");
  return rc2 ? rc2 : rc;
}

int sfmba_pnp_refine_batched(sfmba_ctx* ctx, int32_t num_views, int32_t num_obs, const int32_t* rowptr,
                             const double* obs_xy, const double* obs_X, double* rot, double* loc, double lambda,
                             int32_t iterations, int32_t row_packing_quirk, int32_t dvdc_sign_quirk,
                             uint32_t* host_flags, void* stream) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (num_views < 0 || num_obs < 0 || iterations < 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "negative size");
  cudaSetDevice(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  void* fl = nullptr;
  int rc = sfm_check_cuda(ctx, ctx->reserve(0, sizeof(Scalars), &fl), "cudaMalloc(flags)");
  if (rc) return rc;
  uint32_t* dflags = reinterpret_cast<uint32_t*>(fl);
  cudaMemsetAsync(dflags, 0, sizeof(uint32_t), st);
  launch_validate_csr(num_views, num_obs, rowptr, nullptr, 0, dflags, st, ctx->lc);
  launch_pnp(num_views, rowptr, obs_xy, obs_X, rot, loc, lambda, iterations, row_packing_quirk != 0,
             dvdc_sign_quirk != 0, dflags, st, ctx->lc);
  return read_flags(ctx, dflags, host_flags, st);
}

int sfmba_pnp_refine_batched_host(sfmba_ctx* ctx, int32_t num_views, int32_t num_obs, const int32_t* rowptr,
                                  const double* obs_xy, const double* obs_X, double* rot, double* loc,
                                  double lambda, int32_t iterations, int32_t row_packing_quirk,
                                  int32_t dvdc_sign_quirk, uint32_t* host_flags) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (num_views < 0 || num_obs < 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "negative size");
  cudaSetDevice(ctx->device);
  cudaStream_t st = ctx->own_stream;
  int32_t* d_row;
  double *d_xy, *d_X, *d_rot, *d_loc;
  int rc;
  if ((rc = stage_in(ctx, 0, rowptr, (size_t)num_views + 1, &d_row, st))) return rc;
  if ((rc = stage_in(ctx, 1, obs_xy, (size_t)2 * num_obs, &d_xy, st))) return rc;
  if ((rc = stage_in(ctx, 2, obs_X, (size_t)3 * num_obs, &d_X, st))) return rc;
  if ((rc = stage_in(ctx, 3, rot, (size_t)9 * num_views, &d_rot, st))) return rc;
  if ((rc = stage_in(ctx, 4, loc, (size_t)3 * num_views, &d_loc, st))) return rc;
  rc = sfmba_pnp_refine_batched(ctx, num_views, num_obs, d_row, d_xy, d_X, d_rot, d_loc, lambda, iterations,
                                row_packing_quirk, dvdc_sign_quirk, host_flags, st);
  if (rc != SFMBA_OK && rc != SFMBA_ERR_NUMERIC) return rc;
  if (num_views) {
    cudaMemcpyAsync(rot, d_rot, (size_t)9 * num_views * sizeof(double), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(loc, d_loc, (size_t)3 * num_views * sizeof(double), cudaMemcpyDeviceToHost, st);
  }
  const int rc2 = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "D2H sync");
  return rc2 ? rc2 : rc;
}

// ---- f3: RANSAC hypotheses ---------------------------------------------------------------------------
namespace {
bool inv3_host(const double* a, double* o) {
  const double c00 = a[4] * a[8] - a[5] * a[7], c01 = a[5] * a[6] - a[3] * a[8], c02 = a[3] * a[7] - a[4] * a[6];
  const double det = a[0] * c00 + a[1] * c01 + a[2] * c02;
  if (det == 0.0 || det != det) return false;
  o[0] = c00 / det;
  o[1] = (a[2] * a[7] - a[1] * a[8]) / det;
  o[2] = (a[1] * a[5] - a[2] * a[4]) / det;
  o[3] = c01 / det;
  o[4] = (a[0] * a[8] - a[2] * a[6]) / det;
  o[5] = (a[2] * a[3] - a[0] * a[5]) / det;
  o[6] = c02 / det;
  o[7] = (a[1] * a[6] - a[0] * a[7]) / det;
  o[8] = (a[0] * a[4] - a[1] * a[3]) / det;
  return true;
}

// scratch layout inside ctx->host_stage[11]: flags(16 B) | best i32[2] (16 B) | K, Kinv (18 doubles) | counts[H] |
// models[H][stride] | projs[H][12]
struct RansacScratch {
  uint32_t* flags;
  int32_t* best;
  double* kk;
  int32_t* counts;
  double* models;
  double* projs;
};
int ransac_scratch(sfmba_ctx* ctx, int H, int stride, bool want_projs, RansacScratch* r) {
  const size_t hc = ((size_t)H * sizeof(int32_t) + 15) & ~(size_t)15;
  const size_t bytes = 32 + 18 * sizeof(double) + hc + (size_t)H * stride * sizeof(double) +
                       (want_projs ? (size_t)H * 12 * sizeof(double) : 0) + 64;
  cudaError_t e = ctx->host_stage[11].reserve(bytes);
  if (e != cudaSuccess) return sfm_check_cuda(ctx, e, "cudaMalloc(ransac scratch)");
  char* b = ctx->host_stage[11].as<char>();
  r->flags = reinterpret_cast<uint32_t*>(b);
  r->best = reinterpret_cast<int32_t*>(b + 16);
  r->kk = reinterpret_cast<double*>(b + 32);
  r->counts = reinterpret_cast<int32_t*>(b + 32 + 18 * sizeof(double));
  r->models = reinterpret_cast<double*>(b + 32 + 18 * sizeof(double) + hc);
  r->projs = r->models + (size_t)H * stride;
  return SFMBA_OK;
}
}  // namespace

int sfmba_pnp_ransac(sfmba_ctx* ctx, int32_t num_pts, int32_t num_hyp, const double* key, const double* X,
                     const double* K, const int32_t* samples, const int32_t* loc_sign, double inlier_threshold,
                     double* out_rot, double* out_loc, int32_t* out_best, uint8_t* out_mask, int32_t* out_counts,
                     double* out_models, uint32_t* host_flags, void* stream) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (num_pts < 0 || num_hyp < 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "negative size");
  if (!K || !out_best || (num_pts > 0 && (!key || !X || !out_mask)) || (num_hyp > 0 && !samples))
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "null pointer in pnp_ransac");
  if (num_hyp > 0 && num_pts < 6) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "pnp_ransac needs >= 6 points");
  cudaSetDevice(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  double kk[18];
  for (int i = 0; i < 9; ++i) kk[i] = K[i];
  if (!inv3_host(K, kk + 9)) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "singular intrinsic matrix");
  RansacScratch rs;
  int rc = ransac_scratch(ctx, std::max(num_hyp, 1), 12, true, &rs);
  if (rc) return rc;
  cudaMemsetAsync(rs.flags, 0, 32, st);
  cudaMemcpyAsync(rs.kk, kk, sizeof(kk), cudaMemcpyHostToDevice, st);
  launch_pnp_ransac(num_pts, num_hyp, key, X, rs.kk, rs.kk + 9, samples, loc_sign, inlier_threshold, rs.models,
                    rs.projs, rs.counts, rs.best, out_mask, rs.flags, st, ctx->lc);
  // pinned read-back: flags | best[2] | winner's model
  char* pin = reinterpret_cast<char*>(ctx->pinned);
  cudaMemcpyAsync(pin, rs.flags, 32, cudaMemcpyDeviceToHost, st);
  if ((rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "pnp_ransac sync")) != 0) return rc;
  if ((rc = sfm_check_cuda(ctx, cudaGetLastError(), "pnp_ransac kernels")) != 0) return rc;
  const uint32_t fl = *reinterpret_cast<uint32_t*>(pin);
  if (fl & SFMBA_FLAG_BAD_INDEX) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "sample index out of range");
  int32_t best[2] = {-1, 0};
  if (num_hyp > 0) {
    best[0] = reinterpret_cast<int32_t*>(pin + 16)[0];
    best[1] = reinterpret_cast<int32_t*>(pin + 16)[1];
  }
  out_best[0] = best[0];
  out_best[1] = best[1];
  if (host_flags) *host_flags = fl;
  double model[12] = {1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0};       // campose_processor.py:526-527
  if (best[0] >= 0) {
    cudaMemcpyAsync(pin + 64, rs.models + 12 * (size_t)best[0], sizeof(model), cudaMemcpyDeviceToHost, st);
    if ((rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "pnp_ransac model")) != 0) return rc;
    memcpy(model, pin + 64, sizeof(model));
  } else if (num_hyp <= 0 && num_pts > 0) {
    cudaMemsetAsync(out_mask, 0, (size_t)num_pts, st);
  }
  if (out_rot) memcpy(out_rot, model, 9 * sizeof(double));
  if (out_loc) memcpy(out_loc, model + 9, 3 * sizeof(double));
  if (num_hyp > 0 && out_counts)
    cudaMemcpyAsync(out_counts, rs.counts, (size_t)num_hyp * sizeof(int32_t), cudaMemcpyDeviceToDevice, st);
  if (num_hyp > 0 && out_models)
    cudaMemcpyAsync(out_models, rs.models, (size_t)num_hyp * 12 * sizeof(double), cudaMemcpyDeviceToDevice, st);
  return sfm_check_cuda(ctx, cudaStreamSynchronize(st), "pnp_ransac outputs");
}

int sfmba_pnp_ransac_host(sfmba_ctx* ctx, int32_t num_pts, int32_t num_hyp, const double* key, const double* X,
                          const double* K, const int32_t* samples, const int32_t* loc_sign, double inlier_threshold,
                          double* out_rot, double* out_loc, int32_t* out_best, uint8_t* out_mask,
                          int32_t* out_counts, double* out_models, uint32_t* host_flags) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (num_pts < 0 || num_hyp < 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "negative size");
  cudaSetDevice(ctx->device);
  cudaStream_t st = ctx->own_stream;
  double *d_key, *d_X;
  int32_t *d_s, *d_sign = nullptr;
  int rc;
  if ((rc = stage_in(ctx, 0, key, (size_t)3 * num_pts, &d_key, st))) return rc;
  if ((rc = stage_in(ctx, 1, X, (size_t)4 * num_pts, &d_X, st))) return rc;
  if ((rc = stage_in(ctx, 2, samples, (size_t)6 * num_hyp, &d_s, st))) return rc;
  if (loc_sign && (rc = stage_in(ctx, 3, loc_sign, (size_t)num_hyp, &d_sign, st))) return rc;
  cudaError_t e = ctx->host_stage[4].reserve(std::max<size_t>(16, (size_t)num_pts));
  if (e == cudaSuccess) e = ctx->host_stage[5].reserve(std::max<size_t>(16, (size_t)num_hyp * sizeof(int32_t)));
  if (e == cudaSuccess) e = ctx->host_stage[6].reserve(std::max<size_t>(16, (size_t)num_hyp * 12 * sizeof(double)));
  if (e != cudaSuccess) return sfm_check_cuda(ctx, e, "cudaMalloc(stage)");
  uint8_t* d_mask = ctx->host_stage[4].as<uint8_t>();
  int32_t* d_counts = out_counts ? ctx->host_stage[5].as<int32_t>() : nullptr;
  double* d_models = out_models ? ctx->host_stage[6].as<double>() : nullptr;
  rc = sfmba_pnp_ransac(ctx, num_pts, num_hyp, d_key, d_X, K, d_s, loc_sign ? d_sign : nullptr, inlier_threshold,
                        out_rot, out_loc, out_best, d_mask, d_counts, d_models, host_flags, st);
  if (rc != SFMBA_OK) return rc;
  if (num_pts && out_mask) cudaMemcpyAsync(out_mask, d_mask, (size_t)num_pts, cudaMemcpyDeviceToHost, st);
  if (num_hyp && out_counts)
    cudaMemcpyAsync(out_counts, d_counts, (size_t)num_hyp * sizeof(int32_t), cudaMemcpyDeviceToHost, st);
  if (num_hyp && out_models)
    cudaMemcpyAsync(out_models, d_models, (size_t)num_hyp * 12 * sizeof(double), cudaMemcpyDeviceToHost, st);
  return sfm_check_cuda(ctx, cudaStreamSynchronize(st), "D2H sync");
}

int sfmba_fund_ransac(sfmba_ctx* ctx, int32_t num_pairs, int32_t num_hyp, const double* pairs,
                      const int32_t* samples, double inlier_threshold, double* out_F, int32_t* out_best,
                      uint8_t* out_mask, int32_t* out_counts, double* out_models, uint32_t* host_flags,
                      void* stream) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (num_pairs < 0 || num_hyp < 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "negative size");
  if (!out_best || (num_pairs > 0 && (!pairs || !out_mask)) || (num_hyp > 0 && !samples))
    return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "null pointer in fund_ransac");
  if (num_hyp > 0 && num_pairs < 8) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "fund_ransac needs >= 8 pairs");
  cudaSetDevice(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  RansacScratch rs;
  int rc = ransac_scratch(ctx, std::max(num_hyp, 1), 9, false, &rs);
  if (rc) return rc;
  cudaMemsetAsync(rs.flags, 0, 32, st);
  launch_fund_ransac(num_pairs, num_hyp, pairs, samples, inlier_threshold, rs.models, rs.counts, rs.best, out_mask,
                     rs.flags, st, ctx->lc);
  char* pin = reinterpret_cast<char*>(ctx->pinned);
  cudaMemcpyAsync(pin, rs.flags, 32, cudaMemcpyDeviceToHost, st);
  if ((rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "fund_ransac sync")) != 0) return rc;
  if ((rc = sfm_check_cuda(ctx, cudaGetLastError(), "fund_ransac kernels")) != 0) return rc;
  const uint32_t fl = *reinterpret_cast<uint32_t*>(pin);
  if (fl & SFMBA_FLAG_BAD_INDEX) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "sample index out of range");
  int32_t best[2] = {-1, 0};
  if (num_hyp > 0) {
    best[0] = reinterpret_cast<int32_t*>(pin + 16)[0];
    best[1] = reinterpret_cast<int32_t*>(pin + 16)[1];
  }
  out_best[0] = best[0];
  out_best[1] = best[1];
  if (host_flags) *host_flags = fl;
  double model[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};                  // epipolar_processor.py:213
  if (best[0] >= 0) {
    cudaMemcpyAsync(pin + 64, rs.models + 9 * (size_t)best[0], sizeof(model), cudaMemcpyDeviceToHost, st);
    if ((rc = sfm_check_cuda(ctx, cudaStreamSynchronize(st), "fund_ransac model")) != 0) return rc;
    memcpy(model, pin + 64, sizeof(model));
  } else if (num_hyp <= 0 && num_pairs > 0) {
    cudaMemsetAsync(out_mask, 0, (size_t)num_pairs, st);
  }
  if (out_F) memcpy(out_F, model, sizeof(model));
  if (num_hyp > 0 && out_counts)
    cudaMemcpyAsync(out_counts, rs.counts, (size_t)num_hyp * sizeof(int32_t), cudaMemcpyDeviceToDevice, st);
  if (num_hyp > 0 && out_models)
    cudaMemcpyAsync(out_models, rs.models, (size_t)num_hyp * 9 * sizeof(double), cudaMemcpyDeviceToDevice, st);
  return sfm_check_cuda(ctx, cudaStreamSynchronize(st), "fund_ransac outputs");
}

int sfmba_fund_ransac_host(sfmba_ctx* ctx, int32_t num_pairs, int32_t num_hyp, const double* pairs,
                           const int32_t* samples, double inlier_threshold, double* out_F, int32_t* out_best,
                           uint8_t* out_mask, int32_t* out_counts, double* out_models, uint32_t* host_flags) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (num_pairs < 0 || num_hyp < 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "negative size");
  cudaSetDevice(ctx->device);
  cudaStream_t st = ctx->own_stream;
  double* d_pairs;
  int32_t* d_s;
  int rc;
  if ((rc = stage_in(ctx, 0, pairs, (size_t)4 * num_pairs, &d_pairs, st))) return rc;
  if ((rc = stage_in(ctx, 2, samples, (size_t)8 * num_hyp, &d_s, st))) return rc;
  cudaError_t e = ctx->host_stage[4].reserve(std::max<size_t>(16, (size_t)num_pairs));
  if (e == cudaSuccess) e = ctx->host_stage[5].reserve(std::max<size_t>(16, (size_t)num_hyp * sizeof(int32_t)));
  if (e == cudaSuccess) e = ctx->host_stage[6].reserve(std::max<size_t>(16, (size_t)num_hyp * 9 * sizeof(double)));
  if (e != cudaSuccess) return sfm_check_cuda(ctx, e, "cudaMalloc(stage)");
  uint8_t* d_mask = ctx->host_stage[4].as<uint8_t>();
  int32_t* d_counts = out_counts ? ctx->host_stage[5].as<int32_t>() : nullptr;
  double* d_models = out_models ? ctx->host_stage[6].as<double>() : nullptr;
  rc = sfmba_fund_ransac(ctx, num_pairs, num_hyp, d_pairs, d_s, inlier_threshold, out_F, out_best, d_mask, d_counts,
                         d_models, host_flags, st);
  if (rc != SFMBA_OK) return rc;
  if (num_pairs && out_mask) cudaMemcpyAsync(out_mask, d_mask, (size_t)num_pairs, cudaMemcpyDeviceToHost, st);
  if (num_hyp && out_counts)
    cudaMemcpyAsync(out_counts, d_counts, (size_t)num_hyp * sizeof(int32_t), cudaMemcpyDeviceToHost, st);
  if (num_hyp && out_models)
    cudaMemcpyAsync(out_models, d_models, (size_t)num_hyp * 9 * sizeof(double), cudaMemcpyDeviceToHost, st);
  return sfm_check_cuda(ctx, cudaStreamSynchronize(st), "D2H sync");
}

int sfmba_build_visibility(sfmba_ctx* ctx, int32_t num_views, int32_t num_pts, const int64_t* diag,
                           const int64_t* key_off, int32_t* obs_pt, int32_t* obs_cam, int32_t* obs_key,
                           int32_t* pt_rowptr, int32_t* host_num_obs, void* stream) {
  if (!ctx) return sfm_set_error(nullptr, SFMBA_ERR_INVALID_ARG, "null ctx");
  if (!key_off || num_views <= 0) return sfm_set_error(ctx, SFMBA_ERR_INVALID_ARG, "bad key_off / num_views");
  cudaSetDevice(ctx->device);
  std::string err;
  const int rc = build_visibility_device(num_views, num_pts, diag, key_off, key_off[num_views] - key_off[0], obs_pt,
                                         obs_cam, obs_key, pt_rowptr, host_num_obs, (cudaStream_t)stream, ctx->lc,
                                         err);
  if (rc) return sfm_set_error(ctx, rc, err);
  return SFMBA_OK;
}

int sfmba_profile_reset(sfmba_ctx* ctx) {
  if (!ctx) return SFMBA_ERR_INVALID_ARG;
  ctx->prof.clear();
  return SFMBA_OK;
}

int sfmba_profile_enable(sfmba_ctx* ctx, int enable) {
  if (!ctx) return SFMBA_ERR_INVALID_ARG;
  ctx->profile = (enable & 1) != 0;
  ctx->phase_timing = (enable & 2) != 0;
  return SFMBA_OK;
}

int sfmba_profile_get(sfmba_ctx* ctx, const char* name, double* ms_total, int64_t* launches) {
  if (!ctx || !name) return SFMBA_ERR_INVALID_ARG;
  auto it = ctx->prof.find(name);
  if (it == ctx->prof.end()) {
    if (ms_total) *ms_total = 0.0;
    if (launches) *launches = 0;
    return SFMBA_OK;
  }
  if (ms_total) *ms_total = it->second.ms;
  if (launches) *launches = it->second.launches;
  return SFMBA_OK;
}

int64_t sfmba_total_launches(const sfmba_ctx* ctx) { return ctx ? ctx->lc.launches : 0; }

}  // extern "C"
