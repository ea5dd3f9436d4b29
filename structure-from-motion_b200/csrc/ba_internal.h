// Internal (non-ABI) declarations shared by the translation units of libsfmba.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/sfmba.h"

namespace sfmba {

constexpr int CAMTAB_STRIDE = 16;  // doubles per camera-table row: R[9] row-major, C[3], qj[4]
constexpr int TILE_OBS = 256;      // observations per point-major tile (by first observation)
constexpr int TILE_CAP = 512;      // shared-memory slots of the fast path (tile + straddling tail)
constexpr int TILE_THREADS = 256;
constexpr int CHUNK_OBS = 256;     // observations per camera-major chunk (one camera per chunk)
constexpr int CAM_ACC = 64;        // doubles per camera accumulator row: A(28) S(28) g(7) pad(1)

// Device-resident scalars of one solve (one allocation, zeroed per call).
struct Scalars {
  double rz, pq, rr, gg, alpha, beta;
  double lambda;         // current damping (device copy; adaptive mode changes it on device)
  double cost_cur;       // cost at the current linearisation point
  double cost_new;       // cost at the candidate point (adaptive mode)
  double nu;             // LM rejection growth factor
  double model_decrease; // predicted decrease of the last step (adaptive mode)
  int32_t pcg_done;
  int32_t pcg_iters;     // iterations of the current PCG solve
  int32_t pcg_iters_total;
  int32_t num_chunks;
  int32_t lm_accepted;   // number of accepted steps
  int32_t lm_last_accepted;
  uint32_t flags;
  int32_t pad;
  double costs[SFMBA_MAX_COST_HISTORY];
};

// Device view of one BA problem plus its workspace.  Plain pointers so kernels take it by value.
struct BaDev {
  int V, N, m, mp;                    // mp: SoA row stride (m rounded up to 32)
  int ntiles, nchunk_bound;
  // inputs
  const float* obs_uv;
  const int32_t* obs_cam;
  const int32_t* obs_pt;
  const int32_t* rowptr;
  const double* kinv;
  double* cams;
  double* pts;
  // index
  int32_t* tile_pt;                   // [ntiles+1]
  int32_t* cm_perm;                   // [m] camera-major slot -> point-major observation
  int32_t* pm2cm;                     // [m] inverse
  int32_t* cam_rowptr;                // [V+1] offsets into camera-major order
  int32_t* cam_chunk_ptr;             // [V+1] offsets into the chunk table
  int32_t* chunk_cam;                 // [nchunk_bound]
  int32_t* chunk_beg;                 // [nchunk_bound+1] camera-major slot range of each chunk
  // linearisation
  double* camtab;                     // [V][16]
  float* jp_pm;                       // [14][mp] point-major SoA
  float* jx_pm;                       // [6][mp]
  float4* rec_cm;                     // [m][4]: Jp(14) rho(2), camera-major AoS (64 B records)
  float4* y_cm;                       // [m]: (yuu, yuv, yvv, 0), Y = Jx D^-1 Jx^T
  float* dinv;                        // [6][N] symmetric packing 00 01 02 11 12 22
  float* vj;                          // [3][N] D^-1 e_x
  double* tile_cost;                  // [ntiles]
  // camera side
  double* chunk_acc;                  // [nchunk_bound][CAM_ACC]
  double* cam_acc;                    // [V][CAM_ACC] (+ 8 trailing doubles: local cost at [V*CAM_ACC])
  double* ablk;                       // [V][49] A_i + lambda I
  double* minv;                       // [V][49] block-Jacobi inverse of S_ii
  // PCG
  double2* b_cm;                      // [m]
  double* chunk_y;                    // [nchunk_bound][8]
  double* svec;                       // [7V] sum_j W D^-1 W^T p (local; allreduced when world > 1)
  double* x;                          // [7V] solution (delta p)
  double* r;
  double* z;
  double* p;
  double* gvec;                       // [7V] right-hand side
  double* cams_backup;                // [V][7]  adaptive mode
  double* pts_backup;                 // [N][3]
  Scalars* sc;
  unsigned long long* host_sig;       // pinned, device-visible: (epoch << 32) | (pcg_iters << 1) | done
};

struct LaunchCounter {
  int64_t launches = 0;
};

// ---- index (ba_index.cu) ------------------------------------------------------------------------
size_t index_sort_temp_bytes(int m, int V);
void launch_build_index(const BaDev& d, void* sort_temp, size_t sort_temp_bytes, int32_t* sort_keys_out,
                        int32_t* iota, cudaStream_t st, LaunchCounter& lc);

// ---- linearisation (ba_linearize.cu) ------------------------------------------------------------
void launch_cam_prepare(const BaDev& d, cudaStream_t st, LaunchCounter& lc);
void launch_linearize(const BaDev& d, bool dvdc_quirk, bool force_slow, int cost_slot, cudaStream_t st,
                      LaunchCounter& lc);
void launch_cost(const BaDev& d, double* out_cost_dev, cudaStream_t st, LaunchCounter& lc);
void launch_k1(const BaDev& d, bool dvdc_quirk, float* out_r, float* out_jp, float* out_jx, double* out_r64,
               double* out_cost, cudaStream_t st, LaunchCounter& lc);

// ---- camera-side reduction (ba_schur.cu) --------------------------------------------------------
void launch_camreduce(const BaDev& d, cudaStream_t st, LaunchCounter& lc);
void launch_cam_finalize(const BaDev& d, cudaStream_t st, LaunchCounter& lc);   // chunk_acc -> cam_acc
void launch_cam_blocks(const BaDev& d, cudaStream_t st, LaunchCounter& lc);     // ablk, minv, gvec

// ---- PCG (ba_pcg.cu) -----------------------------------------------------------------------------
void launch_pcg_init(const BaDev& d, unsigned epoch, cudaStream_t st, LaunchCounter& lc);
void launch_matvec_a(const BaDev& d, const double* xvec, bool backsub, bool force_slow, cudaStream_t st,
                     LaunchCounter& lc);
void launch_matvec_b(const BaDev& d, cudaStream_t st, LaunchCounter& lc);
void launch_matvec_finalize(const BaDev& d, cudaStream_t st, LaunchCounter& lc);
void launch_pcg_step(const BaDev& d, double tol, int max_iter, unsigned epoch, cudaStream_t st, LaunchCounter& lc);
void launch_update_cams(const BaDev& d, cudaStream_t st, LaunchCounter& lc);
void launch_lm_decide(const BaDev& d, double lam_min, double lam_max, cudaStream_t st, LaunchCounter& lc);
void launch_lm_backup(const BaDev& d, cudaStream_t st, LaunchCounter& lc);
void launch_lm_model_decrease(const BaDev& d, cudaStream_t st, LaunchCounter& lc);

// ---- refiners (refine.cu) ------------------------------------------------------------------------
void launch_triangulate(int num_pts, const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                        const double* obs_uv, double* pts, double lambda, int iterations, uint32_t* flags,
                        cudaStream_t st, LaunchCounter& lc);
void launch_pnp(int num_views, const int32_t* rowptr, const double* obs_xy, const double* obs_X, double* rot,
                double* loc, double lambda, int iterations, bool row_packing_quirk, bool dvdc_quirk,
                uint32_t* flags, cudaStream_t st, LaunchCounter& lc);

// ---- visibility (visibility.cu) ------------------------------------------------------------------
int build_visibility_device(int num_views, int num_pts, const int64_t* diag, const int64_t* key_off_host,
                            int64_t total_keys, int32_t* obs_pt, int32_t* obs_cam, int32_t* obs_key,
                            int32_t* pt_rowptr, int32_t* host_num_obs, cudaStream_t st, LaunchCounter& lc,
                            std::string& err);

}  // namespace sfmba
