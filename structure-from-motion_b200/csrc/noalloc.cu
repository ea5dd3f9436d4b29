// sfmba_ba_cost_noalloc: the index-free reprojection cost of sfmba_ba_cost for callers that cannot allocate
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has no counterpart in
// the reference.
//
// The two scratch areas sfmba_ba_cost carves out of its context -- the camera table (R, C, q per camera) and the
// per-CTA partial sums with the flag word -- live here in ONE file-scope device array of fixed size.  Every GPU that
// loads the library holds it until the process exits.  All calls on a device share it and it keeps the last call's
// contents: the flag word is cleared by every call (the table and the partials are fully rewritten before they are
// read); two calls that overlap on the same GPU (different streams) race on it.
//
// The array's name is a host-side stand-in in host code, so its device address is looked up once per device with
// cudaGetSymbolAddress (with that device current) and handed to the kernels through the BaDev view.  Nothing on this
// path calls cudaMalloc / cudaFree or any allocating library routine.
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>

#include "ctx.h"

using namespace sfmba;

namespace {

constexpr int NOALLOC_MAX_CAMS = SFMBA_NOALLOC_MAX_CAMS;
constexpr int NOALLOC_PARTIALS = 148 * 8 + 8;           // cost grid of launch_cost (<= 1184 CTAs) + the result slot
constexpr size_t NOALLOC_CAMTAB_DOUBLES = (size_t)NOALLOC_MAX_CAMS * CAMTAB_STRIDE;
constexpr size_t NOALLOC_SCALARS_DOUBLES = (sizeof(Scalars) + 7) / 8;
constexpr size_t NOALLOC_DOUBLES = NOALLOC_CAMTAB_DOUBLES + NOALLOC_PARTIALS + NOALLOC_SCALARS_DOUBLES;

// camtab[MAX_CAMS][16] | partials[1192] | Scalars (only .flags is used)                      (2.1 MB)
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_DOUBLES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
double* g_scratch_dev[MAX_DEVICES] = {};                 // device address of the array, per device

int fail(int code, const std::string& msg) { return sfm_set_error(nullptr, code, msg); }

// device address of this device's copy of the array (looked up once, with the device current)
int scratch_address(int device, double** out) {
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_dev[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(SFMBA_ERR_CUDA, std::string("cudaGetSymbolAddress(noalloc scratch): ") + cudaGetErrorString(e));
    }
    g_scratch_dev[device] = reinterpret_cast<double*>(p);
  }
  *out = g_scratch_dev[device];
  return SFMBA_OK;
}

}  // namespace

extern "C" int sfmba_ba_cost_noalloc(int device, const sfmba_problem* pr, double* out_cost, uint32_t* out_flags,
                                     void* stream) {
  if (!pr || !out_cost) return fail(SFMBA_ERR_INVALID_ARG, "sfmba_ba_cost_noalloc: null problem / out_cost");
  if (device < 0 || device >= MAX_DEVICES) return fail(SFMBA_ERR_INVALID_ARG, "sfmba_ba_cost_noalloc: bad device index");
  if (pr->num_cams <= 0 || pr->num_pts < 0 || pr->num_obs < 0)
    return fail(SFMBA_ERR_INVALID_ARG, "sfmba_ba_cost_noalloc: bad problem sizes");
  if (pr->num_cams > NOALLOC_MAX_CAMS) {
    char buf[160];
    snprintf(buf, sizeof(buf),
             "sfmba_ba_cost_noalloc: %d cameras exceed the fixed scratch of SFMBA_NOALLOC_MAX_CAMS = %d; use sfmba_ba_cost",
             pr->num_cams, NOALLOC_MAX_CAMS);
    return fail(SFMBA_ERR_UNSUPPORTED, buf);
  }
  if (!pr->cams || (pr->num_pts > 0 && !pr->pts) ||
      (pr->num_obs > 0 && (!pr->obs_uv || !pr->obs_cam || !pr->obs_pt)))
    return fail(SFMBA_ERR_INVALID_ARG, "sfmba_ba_cost_noalloc: null pointer in problem (obs_pt is required)");
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return sfm_check_cuda(nullptr, e, "cudaSetDevice");
  double* scratch = nullptr;
  int rc = scratch_address(device, &scratch);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  BaDev d;
  memset(&d, 0, sizeof(d));
  d.V = pr->num_cams;
  d.N = pr->num_pts;
  d.m = pr->num_obs;
  d.obs_uv = pr->obs_uv;
  d.obs_cam = pr->obs_cam;
  d.obs_pt = pr->obs_pt;
  d.rowptr = pr->pt_rowptr;
  d.kinv = pr->cam_kinv;
  d.cams = pr->cams;
  d.pts = pr->pts;
  d.camtab = scratch;
  d.grp_cost = scratch + NOALLOC_CAMTAB_DOUBLES;
  d.sc = reinterpret_cast<Scalars*>(scratch + NOALLOC_CAMTAB_DOUBLES + NOALLOC_PARTIALS);
  // the flag word must start at zero (the kernels only OR into it); everything else is written before it is read
  e = cudaMemsetAsync(d.sc, 0, sizeof(Scalars), st);
  if (e != cudaSuccess) return sfm_check_cuda(nullptr, e, "cudaMemsetAsync(noalloc flags)");
  LaunchCounter lc;
  launch_cam_prepare(d, st, lc);                          // k_cam_prepare: camtab rows of the scratch
  launch_cost(d, out_cost, st, lc);                       // k_cost -> partials of the scratch; k_sum_cost -> *out_cost
  if (out_flags) {
    e = cudaMemcpyAsync(out_flags, &d.sc->flags, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return sfm_check_cuda(nullptr, e, "cudaMemcpyAsync(noalloc flags)");
  }
  return sfm_check_cuda(nullptr, cudaGetLastError(), "sfmba_ba_cost_noalloc launch");
}
