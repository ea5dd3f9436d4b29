/*
 * sfmba.h -- C ABI of libsfmba.so, the B200 (sm_100a) bundle-adjustment / nonlinear-refinement engine.
 *
 * The reference (willSapgreen/structure-from-motion) is pure Python and has NO FFI for this path; its
 * boundary is the Python class surface.  Each entry point below names the
 * reference interface it replaces.  The reference-side binding a maintainer would add is the ctypes
 * stub shown in INTEGRATION.md (and implemented in structure-from-motion_b200/_capi.py).
 *
 * Rules of the ABI
 *   - plain C: pointers, sizes, POD structs; no C++ exceptions cross it; no torch types.
 *   - every function returns 0 (SFMBA_OK) or a negative sfmba_status; sfmba_last_error() gives text.
 *   - "device" pointers are CUDA device pointers on the ctx's device (e.g. torch.Tensor.data_ptr());
 *     "host" pointers are ordinary host memory.  `stream` is a cudaStream_t passed as void*
 *     (torch.cuda.current_stream().cuda_stream); NULL = the default stream.
 *   - a ctx owns only its workspace; it is bound to one device and is not thread-safe.
 *   - *_host entry points take HOST buffers, stage them through pinned memory, run the same device
 *     path and copy the results back before returning (the end-to-end path bench.py times as `e2e`).
 *
 * Conventions (reference: view_processor.py:57, utils.py:58, ba_processor.py:285-294)
 *   world->camera  p = R^T (X - C);  quaternion order (w,x,y,z);  pose = [Cx,Cy,Cz,qw,qx,qy,qz];
 *   observations sorted point-major / view-minor (ba_processor.py:304-306).
 */
#ifndef SFMBA_H_
#define SFMBA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SFMBA_API __attribute__((visibility("default")))
#else
#define SFMBA_API
#endif

#define SFMBA_VERSION 100
#define SFMBA_MAX_COST_HISTORY 64

typedef struct sfmba_ctx sfmba_ctx;

typedef enum {
  SFMBA_OK = 0,
  SFMBA_ERR_INVALID_ARG = -1,
  SFMBA_ERR_CUDA = -2,
  SFMBA_ERR_NCCL = -3,
  SFMBA_ERR_NUMERIC = -4,   /* device raised a numeric flag, see sfmba_result.flags              */
  SFMBA_ERR_UNSUPPORTED = -5,
  SFMBA_ERR_NO_DEVICE = -6,
  SFMBA_ERR_COMM = -7       /* a peer rank never arrived at an exchange (SFMBA_FLAG_PEER_TIMEOUT): the    */
                            /* parameters of the failed iteration were not updated                        */
} sfmba_status;

/* bits of sfmba_result.flags / refine `flags` outputs (device-side error word)                     */
#define SFMBA_FLAG_QW_NEAR_ZERO   1u  /* |qw| < 1e-6 when q is re-derived from R (utils.py:49-51)  */
#define SFMBA_FLAG_DEPTH_ZERO     2u  /* p_w == 0 for some observation                             */
#define SFMBA_FLAG_NONFINITE      4u  /* NaN/Inf in cost or update                                 */
#define SFMBA_FLAG_PCG_MAXITER    8u  /* PCG stopped on pcg_max_iter, not on pcg_tol               */
#define SFMBA_FLAG_SINGULAR      16u  /* a 3x3 / 7x7 block was not positive definite               */
#define SFMBA_FLAG_PEER_TIMEOUT  32u  /* several GPUs: a rank waited 20 s for a peer's partial sums (SFMBA_ERR_COMM) */
#define SFMBA_FLAG_BAD_INDEX     64u  /* a RANSAC sample index outside [0, n): the call returns INVALID_ARG */

typedef enum { SFMBA_MODE_REFERENCE_FIXED = 0, SFMBA_MODE_ADAPTIVE_LM = 1 } sfmba_mode;
/* PCG: matrix-free block-Jacobi preconditioned conjugate gradients on the implicit Schur complement.
 * DENSE_CHOLESKY: the reduced camera matrix (7V x 7V, f64) is assembled and factored on the FP64 tensor
 * cores -- the direct solve of ba_processor.py:382; needs 8 (7V)^2 bytes, replicated on every rank.     */
typedef enum { SFMBA_SOLVER_PCG = 0, SFMBA_SOLVER_DENSE_CHOLESKY = 1 } sfmba_solver;
/* Block-Jacobi preconditioner of the PCG solve.  JACOBI: the 7x7 blocks of A + lambda I (the sums the matvec needs
 * anyway).  SCHUR_JACOBI: the 7x7 diagonal blocks of the Schur complement S itself; costs 28 more running sums and the
 * 2x2 point term Y = J_x D^-1 J_x^T per observation in the linearisation.  On the BAL-shaped problems both need the
 * same number of iterations at every damping tried, so JACOBI is the default.  The solution is
 * the same to the PCG tolerance either way.                                                                        */
typedef enum { SFMBA_PRECOND_JACOBI = 0, SFMBA_PRECOND_SCHUR_JACOBI = 1 } sfmba_precond;
/* Arithmetic of the per-observation work (the name is historical: it started with the PCG matvec K3 and now also governs
 * the Jacobian arithmetic of the linearisation K1 + K2).  The normal equations (A, D, the right-hand side, the
 * preconditioner), every PCG vector and dot product, the cost and the LM bookkeeping are FP64 in all modes.
 *   F64:  everything FP64.
 *   F32:  the matvec runs in FP32 on group-local coordinates (FP32 run sums, FP64 above them); the linearisation keeps
 *         the projection, the residual and the cost in FP64 and builds the Jacobian blocks and their products in FP32
 *         (BASELINE.json north_star: "FP32 with FP64 accumulation").  The reduced system is perturbed by ~1e-7 relative:
 *         the step moves by ~1e-7 cond(Minv S).  Plain least squares with the A + lambda I preconditioner only.
 *   AUTO (default): FP32 matvec when pcg_tol >= 1e-8, a solve that needs more than 8 iterations is finished in FP64 from
 *         a true FP64 residual; FP32 Jacobians as long as no PCG solve of the call has been that slow.  A solve on an
 *         FP32 linearisation is watched: the device gives it up when its convergence rate predicts more than 8
 *         iterations (or the host sees it take them), the LM iteration is redone from an FP64 linearisation
 *         (sfmba_result.relinearized counts these) and the rest of the call stays FP64.  Depends on nothing but the
 *         call itself (bit-reproducible).  DESIGN.md section 4 has the measured distances.                            */
typedef enum { SFMBA_MATVEC_AUTO = 0, SFMBA_MATVEC_F64 = 1, SFMBA_MATVEC_F32 = 2 } sfmba_matvec_precision;

/*
 * One BA problem (this rank's shard when world_size > 1: cameras replicated, points and their
 * observations partitioned by point).  Replaces the packing at ba_processor.py:285-294 and the
 * per-observation gathers at ba_processor.py:309-342.
 */
/* sfmba_problem.flags.  CAM_U16 (sfmba_ba_solve_host only): obs_cam points to uint16_t[m] instead of int32_t[m]
 * (num_cams <= 65536) -- half the bytes of the largest index array on the bus; widened on the device, bit-exact.   */
#define SFMBA_PROBLEM_CAM_U16 1

typedef struct {
  int32_t num_cams;          /* V                                                                   */
  int32_t num_pts;           /* N (local)                                                           */
  int32_t num_obs;           /* m (local)                                                           */
  int32_t flags;             /* SFMBA_PROBLEM_* bits, 0 = the plain layout below                    */
  const float*   obs_uv;     /* f32 [2][m] SoA: all u then all v.  Pixel coordinates when cam_kinv  */
                             /* is given (cv2.KeyPoint.pt is float32), else already normalised      */
  const int32_t* obs_cam;    /* i32 [m] camera index of each observation (u16 [m] with SFMBA_PROBLEM_CAM_U16) */
  const int32_t* obs_pt;     /* i32 [m] point index, non-decreasing (point-major order).  Read only */
                             /* by sfmba_residual_jacobian / sfmba_ba_cost; the solve derives it   */
                             /* from pt_rowptr and accepts NULL                                    */
  const int32_t* pt_rowptr;  /* i32 [N+1] CSR offsets into the observation arrays                   */
  const double*  cam_kinv;   /* f64 [V][9] row-major K^-1 per camera, or NULL = identity            */
                             /* (ba_processor.py:339-342: b = K^-1 [u,v,1]^T / z)                   */
  double*        cams;       /* f64 [V][7] in/out: [C(3), q(w,x,y,z)]                               */
  double*        pts;        /* f64 [N][3] in/out                                                   */
} sfmba_problem;

typedef struct {
  double  lambda;            /* damping factor (ba_processor.py:24 default 5)                       */
  int32_t iterations;        /* number of (reference mode) / max (adaptive mode) LM iterations      */
  int32_t mode;              /* sfmba_mode                                                          */
  int32_t solver;            /* sfmba_solver                                                        */
  int32_t pcg_max_iter;      /* <= 0: default                                                       */
  double  pcg_tol;           /* relative residual ||r|| <= tol ||g||;  <= 0: default 1e-8 (at Venice the */
                             /* final cost then lies 2e-9 and the poses 9e-9 from a 1e-13 solve, three   */
                             /* orders inside the parity bar: profiles/r1_sweep_pcg_tol.jsonl)           */
  int32_t dvdc_sign_quirk;   /* 1 = reproduce campose_processor.py:802-804 (reference), 0 = exact   */
  int32_t force_slow_path;   /* testing: route every tile through the long-track code path          */
  int32_t compute_costs;     /* 1 = fill sfmba_result.costs (one extra cost pass at the end)        */
  int32_t precond;           /* sfmba_precond (PCG solver)                                          */
  double  lm_lambda_min;     /* adaptive mode clamps                                                */
  double  lm_lambda_max;
  double  huber_delta;       /* > 0: Huber loss on every observation's residual norm, in the units of the */
                             /* residual (normalised camera coordinates): rho(s) = s for s = |r|^2 <=      */
                             /* delta^2, 2 delta sqrt(s) - delta^2 beyond.  Each linearisation re-weights   */
                             /* the observation by rho'(s) (IRLS); costs reported are sum rho.  0 = plain   */
                             /* least squares (the reference has no robust loss). */
                             /* PCG solver only.                                                          */
  int32_t matvec_precision;  /* sfmba_matvec_precision (PCG solver)                                          */
  int32_t reuse_index;       /* sfmba_ba_solve only.  1 = the caller vouches that pt_rowptr, obs_cam, obs_uv and    */
                             /* cam_kinv (same device pointers, same contents) are those of the previous           */
                             /* sfmba_ba_solve on this context: the observation index is not rebuilt (0.25 ms at   */
                             /* Venice).  Ignored when anything the library can check differs.                     */
} sfmba_options;

typedef struct {
  double   initial_cost;     /* sum ||b - f||^2 at entry (normalised camera coordinates)            */
  double   final_cost;       /* at exit                                                             */
  double   final_lambda;
  int32_t  iterations_done;
  int32_t  iterations_accepted;
  int32_t  pcg_iterations;   /* total over all LM iterations                                        */
  uint32_t flags;            /* SFMBA_FLAG_*                                                        */
  int32_t  num_costs;
  int32_t  pcg_refine_iterations;  /* FP64 iterations that finished solves begun with the FP32 matvec (AUTO)   */
  double   costs[SFMBA_MAX_COST_HISTORY];  /* cost at the start of iteration k, then the final cost */
  /* device time of the call (ms_total, always) and of its phases (after sfmba_profile_enable(ctx, 2), else 0),
   * milliseconds, CUDA events on `stream`                                                           */
  float    ms_total, ms_index, ms_linearize, ms_camreduce, ms_pcg, ms_update;
  int32_t  kernel_launches;  /* kernels launched by the call                                        */
  int32_t  matvec_precision; /* SFMBA_MATVEC_F64 or SFMBA_MATVEC_F32: what the PCG matvecs of the call ran     */
  int32_t  relinearized;     /* AUTO: LM iterations redone from an FP64 linearisation (slow PCG after an FP32 one) */
  int32_t  reserved1;
} sfmba_result;

/* ---- lifetime -------------------------------------------------------------------------------- */
/* nccl_unique_id: NULL when world_size == 1, else the 128-byte ncclUniqueId made by
 * sfmba_nccl_unique_id() on rank 0 and distributed by the caller (torch.distributed broadcast).   */
SFMBA_API int sfmba_create(sfmba_ctx** out, int device, int world_size, int rank, const void* nccl_unique_id);
SFMBA_API int sfmba_destroy(sfmba_ctx* ctx);
SFMBA_API const char* sfmba_last_error(const sfmba_ctx* ctx);      /* ctx may be NULL: last global error      */
SFMBA_API int sfmba_version(void);
SFMBA_API int sfmba_nccl_unique_id(void* out128);
SFMBA_API void sfmba_default_options(sfmba_options* opt);

/* ---- a1..a17: BaProcessor.__execute_bundle_adjustment  (ba_processor.py:274-440) -------------- */
SFMBA_API int sfmba_ba_solve(sfmba_ctx* ctx, const sfmba_problem* problem, const sfmba_options* opt,
                   sfmba_result* result, void* stream);
/* same with HOST buffers in `problem` (all pointers host); cams/pts updated in place on the host  */
SFMBA_API int sfmba_ba_solve_host(sfmba_ctx* ctx, const sfmba_problem* problem, const sfmba_options* opt,
                        sfmba_result* result);

/* ---- a5,a7..a12 alone: residual + Jacobian blocks per observation (K1, for parity tests and for
 * CamposeProcessor.construct_jacobian_matrix, campose_processor.py:462-482 /
 * TriangulationProcessor.construct_jacobian_matrix, triangulation_processor.py:237-271) ---------
 * out_r  f32 [2][m], out_jp f32 [14][m] (row-major 2x7 block, entry e = 7*row+col), out_jx f32
 * [6][m] (entry e = 3*row+col), all SoA; any of them may be NULL.  out_cost: device f64 [1] or NULL.
 * out_r64: optional device f64 [2][m] full-precision residuals.                                    */
SFMBA_API int sfmba_residual_jacobian(sfmba_ctx* ctx, const sfmba_problem* problem, int dvdc_sign_quirk,
                            float* out_r, float* out_jp, float* out_jx, double* out_r64,
                            double* out_cost, void* stream);
/* the same blocks in full precision: out_r64 f64 [2][m], out_jp64 f64 [14][m], out_jx64 f64 [6][m] (any may be NULL) --
 * what CamposeProcessor.construct_jacobian_matrix hands to callers that build their own normal equations.
 * Both variants synchronise and report bad indices (SFMBA_ERR_INVALID_ARG) and zero depths (SFMBA_ERR_NUMERIC).   */
SFMBA_API int sfmba_residual_jacobian_f64(sfmba_ctx* ctx, const sfmba_problem* problem, int dvdc_sign_quirk,
                                double* out_r64, double* out_jp64, double* out_jx64, double* out_cost, void* stream);
SFMBA_API int sfmba_ba_cost(sfmba_ctx* ctx, const sfmba_problem* problem, double* host_cost, void* stream);

/* ---- sfmba_ba_cost without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid it).  Same kernels and same bits as sfmba_ba_cost, but the scratch they need (camera
 * table, per-CTA partial sums, flag word) is a fixed file-scope device array (2.1 MB, held by every GPU that
 * loads the library until the process exits) instead of context workspace, so no ctx is needed and nothing on
 * the path calls cudaMalloc / cudaFree.  Limit: num_cams <= SFMBA_NOALLOC_MAX_CAMS (any num_pts / num_obs);
 * more cameras return SFMBA_ERR_UNSUPPORTED with a message in sfmba_last_error(NULL).
 *   device     CUDA device index (made current); every pointer is a device pointer on it; obs_pt is required
 *   out_cost   device f64 [1]: sum ||b - f||^2
 *   out_flags  device u32 [1] or NULL: SFMBA_FLAG_* raised by the kernels (SFMBA_FLAG_BAD_INDEX: an index outside
 *              its array; that observation is skipped)
 * Asynchronous on `stream`; the caller synchronises before reading the outputs.  All calls on one device share
 * the scratch, which keeps the last call's contents (each call clears what must start at zero): calls that
 * overlap on the GPU -- two streams, or two threads -- race on it and must be serialised by the caller.        */
#define SFMBA_NOALLOC_MAX_CAMS 16384
SFMBA_API int sfmba_ba_cost_noalloc(int device, const sfmba_problem* problem, double* out_cost, uint32_t* out_flags,
                          void* stream);

/* ---- a18: TriangulationProcessor.nonlinear_triangulate (triangulation_processor.py:160-234),
 * batched over points with a CSR of (view, observation) per point.
 *   projs   f64 [P][12] row-major 3x4 projection matrices (K included by the caller)
 *   rowptr  i32 [M+1], obs_view i32 [m] index into projs, obs_uv f64 [m][2]
 *   pts     f64 [M][3] in/out;  X -= inv(J^T J + lambda I3) J^T (proj - obs), `iterations` times    */
SFMBA_API int sfmba_triangulate_refine_batched(sfmba_ctx* ctx, int32_t num_pts, int32_t num_obs, int32_t num_projs,
                                     const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                                     const double* obs_uv, double* pts, double lambda, int32_t iterations,
                                     uint32_t* host_flags, void* stream);
SFMBA_API int sfmba_triangulate_refine_batched_host(sfmba_ctx* ctx, int32_t num_pts, int32_t num_obs, int32_t num_projs,
                                          const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                                          const double* obs_uv, double* pts, double lambda, int32_t iterations,
                                          uint32_t* host_flags);

/* ---- f2: TriangulationProcessor.linear_triangulate (triangulation_processor.py:91-157)
 * and the depth test of CamposeProcessor.evalulate_cam_pose_cheirality (campose_processor.py:139-189),
 * batched over points with the same CSR as above (the reference takes exactly two views per point; any
 * number >= 2 is accepted here).
 *   pts    f64 [M][3] out: right singular vector of the smallest singular value of the 2n x 4 DLT system,
 *          divided by its last entry
 *   front  i32 [M] out or NULL: 1 when the point has positive depth (third projected coordinate) in every one
 *          of its views, else 0                                                                          */
SFMBA_API int sfmba_triangulate_linear_batched(sfmba_ctx* ctx, int32_t num_pts, int32_t num_obs, int32_t num_projs,
                                     const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                                     const double* obs_uv, double* pts, int32_t* front, uint32_t* host_flags,
                                     void* stream);
SFMBA_API int sfmba_triangulate_linear_batched_host(sfmba_ctx* ctx, int32_t num_pts, int32_t num_obs, int32_t num_projs,
                                          const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                                          const double* obs_uv, double* pts, int32_t* front, uint32_t* host_flags);

/* ---- a19: CamposeProcessor.nonlinear_estimate_cam_pose_pnp (campose_processor.py:308-459),
 * batched over views with a CSR of observations per view.
 *   rowptr   i32 [P+1];  obs_xy f64 [m][2] normalised observations (K^-1 x / z);  obs_X f64 [m][3]
 *   world points (homogeneous coordinate 1);  rot f64 [P][9] in/out (row-major, camera-to-world);  loc f64 [P][3]
 *   in/out.  row_packing_quirk = 1 reproduces campose_processor.py:404-405.                        */
SFMBA_API int sfmba_pnp_refine_batched(sfmba_ctx* ctx, int32_t num_views, int32_t num_obs, const int32_t* rowptr,
                             const double* obs_xy, const double* obs_X, double* rot, double* loc,
                             double lambda, int32_t iterations, int32_t row_packing_quirk,
                             int32_t dvdc_sign_quirk, uint32_t* host_flags, void* stream);
SFMBA_API int sfmba_pnp_refine_batched_host(sfmba_ctx* ctx, int32_t num_views, int32_t num_obs, const int32_t* rowptr,
                                  const double* obs_xy, const double* obs_X, double* rot, double* loc,
                                  double lambda, int32_t iterations, int32_t row_packing_quirk,
                                  int32_t dvdc_sign_quirk, uint32_t* host_flags);

/* ---- f3: RANSAC hypotheses evaluated in parallel.  The sample indices are drawn
 * by the caller (the Python layer uses random.sample exactly as the reference does), the library evaluates
 * all H minimal-sample models, counts inliers over all n points per hypothesis and returns the first
 * hypothesis with the strictly largest count (the reference's update rule) together with its inlier mask.
 *
 * sfmba_pnp_ransac: CamposeProcessor.__linear_determine_cam_pos / __estimate_six_pts
 * (campose_processor.py:485-633).
 *   key      f64 [n][3] homogeneous image points (rows of key_2d_pts^T)
 *   X        f64 [n][4] homogeneous world points (rows of tri_3d_pts^T)
 *   K        f64 [9] intrinsic matrix, row-major (host pointer in both variants)
 *   samples  i32 [H][6] point indices of every hypothesis
 *   loc_sign i32 [H] or NULL: +1 / -1 per hypothesis.  campose_processor.py:620-631 returns (R, C) when LAPACK
 *            hands back the last right singular vector with det(rot_t_noise) > 0 and (R, -C) otherwise; that
 *            sign is an implementation detail of the host's LAPACK, so the caller passes it when it wants
 *            the reference's outcome bit for bit.  NULL = +1 everywhere (the geometrically consistent pose).
 *   out_rot f64 [9] (row-major, camera-to-world), out_loc f64 [3]: winner's pose (host pointers; identity / 0
 *            when no hypothesis has an inlier, like the reference's initial best_rot / best_loc)
 *   out_best i32 [2] host: {winning hypothesis or -1, its inlier count}
 *   out_mask u8 [n]: 1 for the winner's inliers;  out_counts i32 [H] or NULL;  out_models f64 [H][12] or NULL
 *            (rot | loc of every hypothesis)
 * sfmba_fund_ransac: EpipolarProcessor.__estimate_ransac / __estimate_eight_pts (epipolar_processor.py:140-247)
 *   pairs    f64 [n][4] normalised matched pairs [xl, yl, xr, yr];  samples i32 [H][8]
 *   out_F    f64 [9] host: winner's fundamental matrix (f__ / f__[2][2]); zeros when best = -1
 *   out_models f64 [H][9] or NULL.  host_flags carries SFMBA_FLAG_SINGULAR when some hypothesis's f__ does not
 *   have rank 2 (the reference raises ValueError there, epipolar_processor.py:179-183).
 * The plain variants take device pointers for key / X / pairs / samples / loc_sign / out_mask / out_counts /
 * out_models; the _host variants take host pointers for everything.  Both synchronise before returning.   */
SFMBA_API int sfmba_pnp_ransac(sfmba_ctx* ctx, int32_t num_pts, int32_t num_hyp, const double* key, const double* X,
                     const double* K, const int32_t* samples, const int32_t* loc_sign, double inlier_threshold,
                     double* out_rot, double* out_loc, int32_t* out_best, uint8_t* out_mask, int32_t* out_counts,
                     double* out_models, uint32_t* host_flags, void* stream);
SFMBA_API int sfmba_pnp_ransac_host(sfmba_ctx* ctx, int32_t num_pts, int32_t num_hyp, const double* key, const double* X,
                          const double* K, const int32_t* samples, const int32_t* loc_sign, double inlier_threshold,
                          double* out_rot, double* out_loc, int32_t* out_best, uint8_t* out_mask,
                          int32_t* out_counts, double* out_models, uint32_t* host_flags);
SFMBA_API int sfmba_fund_ransac(sfmba_ctx* ctx, int32_t num_pairs, int32_t num_hyp, const double* pairs,
                      const int32_t* samples, double inlier_threshold, double* out_F, int32_t* out_best,
                      uint8_t* out_mask, int32_t* out_counts, double* out_models, uint32_t* host_flags, void* stream);
SFMBA_API int sfmba_fund_ransac_host(sfmba_ctx* ctx, int32_t num_pairs, int32_t num_hyp, const double* pairs,
                           const int32_t* samples, double inlier_threshold, double* out_F, int32_t* out_best,
                           uint8_t* out_mask, int32_t* out_counts, double* out_models, uint32_t* host_flags);

/* ---- a3/a4: visibility index from the diagonal rows of KeyTrack.table (key_tracker.py:29-59,
 * 198-204), bit-exact including the index-0 quirk.  diag i64 [total_keys] is the concatenation of
 * table[v, :] for v = 0..V-1, key_off i64 [V+1] the offsets.  Outputs (device i32): obs_pt, obs_cam,
 * obs_key sized >= min(total_keys, V*N); pt_rowptr [N+1]; *host_num_obs receives m.                */
SFMBA_API int sfmba_build_visibility(sfmba_ctx* ctx, int32_t num_views, int32_t num_pts, const int64_t* diag,
                           const int64_t* key_off, int32_t* obs_pt, int32_t* obs_cam, int32_t* obs_key,
                           int32_t* pt_rowptr, int32_t* host_num_obs, void* stream);

/* ---- introspection used by bench.py ---------------------------------------------------------- */
/* device time (ms) and launch count of named kernels accumulated since the last reset; names:
 * "linearize", "linearize32", "matvec", "pcg_tail", "backsub", "cost"                                            */
SFMBA_API int sfmba_profile_reset(sfmba_ctx* ctx);
/* enable: bit 0 = per-kernel event timing (records between the launches: the kernels no longer overlap), bit 1 = the
 * per-phase times of sfmba_result.ms_* (ms_total is always measured).  Both off by default.                          */
SFMBA_API int sfmba_profile_enable(sfmba_ctx* ctx, int enable);
SFMBA_API int sfmba_profile_get(sfmba_ctx* ctx, const char* name, double* ms_total, int64_t* launches);
SFMBA_API int64_t sfmba_total_launches(const sfmba_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* SFMBA_H_ */
