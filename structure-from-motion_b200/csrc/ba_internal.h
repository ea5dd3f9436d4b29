// Internal (non-ABI) declarations shared by the translation units of libsfmba.
//
// Engine layout (v1, matrix-free):
//   group  = consecutive points whose weight prefix (observations + 2 per point) falls into one
//            GROUP_WEIGHT bucket; one CTA owns one group in every tile kernel.
//   chunk  = <= CHUNK_CAP consecutive (point-major) observations of a group.  A group is one chunk
//            unless it overflows (a very long track) or touches more than SEG_CAP cameras, in which
//            case it is cut into SEG_CAP-observation chunks.
//   Inside a chunk the observations are sorted by camera ("segments").  Every segment is cut into
//   runs of K consecutive sorted observations (the last run of a segment may be shorter) and every
//   run belongs to exactly one lane of the CTA; K is the smallest value for which the chunk needs at
//   most RUN_LANES lanes.  A lane therefore sees ONE camera: its constants are loaded once into
//   registers, the camera-side sums of the run accumulate in registers, and the per-observation work
//   is pure FP64 arithmetic without divergence.  Records are stored lane-transposed ([k][lane]) so
//   every load is coalesced.
//   Camera-side sums: lane -> (shared memory) -> segment row in HBM -> per-camera sum, all in a
//   fixed order (bitwise reproducible, no atomics on data).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/sfmba.h"

namespace sfmba {

constexpr int CAMTAB_STRIDE = 16;   // doubles per camera-table row: R[9] row-major, C[3], qj[4]
constexpr int CAMIT_STRIDE = 16;    // doubles per PCG-iteration camera row: E[9], w[3], wq, pad[3]
constexpr int RUN_LANES = 128;      // threads per tile CTA = lanes per chunk
constexpr int CHUNK_K_MAX = 12;     // sorted observations per lane
constexpr int CHUNK_CAP = RUN_LANES * CHUNK_K_MAX;   // 1536 record slots per chunk
constexpr int GROUP_WEIGHT = 1024;  // bucket width of (rowptr[j] + 2 j)
constexpr int POINT_WEIGHT = 2;
constexpr int GROUP_PT_CAP = GROUP_WEIGHT / POINT_WEIGHT;   // <= 512 points per group
constexpr int SEG_CAP = 128;        // distinct cameras per chunk
constexpr int SEG_RUN_STRIDE = SEG_CAP + 2;             // u16 entries per chunk in seg_lane0 (4-byte aligned rows)
constexpr int CAM_ACC = 64;         // doubles per camera accumulator row: A(28) S(28) g(7) pad(1)
constexpr int ROW_COMPACT = 36;     // doubles per segment row of the Gram sums without the S entries: A(28) g(7) pad(1)
constexpr int MV_ROW32 = 16;        // floats per segment row of the FP32 matvec: 12 sums + pad = 64 bytes (two whole sectors)
constexpr int ROW_F32 = 40;         // floats per segment row of the FP32 linearisation: A(28) g(7) pad(5) = 160 bytes, whole
                                    // 32-byte sectors (144-byte rows made neighbouring CTAs share sectors: slower to sum)
constexpr int MV_ACC = 12;          // doubles per matvec row: T[3][3] (c~ d^T), s[3]

// record index word: ptl[0:10) pos[10:21) seg[21:28)
constexpr uint32_t REC_INVALID = 0xffffffffu;
__host__ __device__ __forceinline__ uint32_t rec_ptl(uint32_t ix) { return ix & 1023u; }
__host__ __device__ __forceinline__ uint32_t rec_pos(uint32_t ix) { return (ix >> 10) & 2047u; }
__host__ __device__ __forceinline__ uint32_t rec_seg(uint32_t ix) { return (ix >> 21) & 127u; }

// Device-resident scalars of one solve (one allocation, zeroed per call).
struct Scalars {
  double rz, pq, rr, gg, alpha, beta;
  double lambda;         // current damping (device copy; adaptive mode changes it on device)
  double cost_cur;       // cost at the current linearisation point
  double cost_new;       // cost at the candidate point (adaptive mode)
  double md_pts;         // point part of the predicted decrease, this rank's points (adjacent to cost_new: one all-reduce)
  double nu;             // LM rejection growth factor
  double model_decrease; // camera part of the predicted decrease of the last step (adaptive mode)
  int32_t pcg_done;
  int32_t pcg_iters;     // iterations of the current PCG solve
  int32_t pcg_iters_total;
  int32_t num_chunks;
  int32_t lm_accepted;   // number of accepted steps
  int32_t lm_last_accepted;
  uint32_t flags;
  int32_t max_chunk_len; // index statistics (read back once per solve to size shared memory)
  int32_t max_group_pts;
  int32_t max_nseg;
  int32_t max_k;
  uint32_t ticket_a;     // "last block done" tickets of the multi-CTA PCG kernels
  uint32_t ticket_b;
  double costs[SFMBA_MAX_COST_HISTORY];
  double alpha_hist[8];  // step lengths of the first PCG_UBUF iterations of the current solve
  double beta_hist[8];   // and the direction coefficients (p_i = z_i + beta_i p_(i-1))
  uint32_t bar_a;        // grid-wide barrier of k_pcg_step (arrival counter, zeroed by the CTA that finishes last)
  uint32_t pad0;
};
constexpr int PCG_UBUF = 8;            // PCG iterations whose per-point u_j are kept for the back-substitution

// One 32-byte load tells a tile CTA everything it needs to start staging its group.
struct GroupMeta {
  int32_t pt0, np;       // points [pt0, pt0 + np)
  int32_t c0, nc;        // chunks [c0, c0 + nc)
  int32_t len0, nseg0;   // first chunk: observations, distinct cameras
  int32_t obs0, K0;      // first chunk: first observation, observations per lane
};

struct ChunkInfo {
  int32_t obs0;          // first (point-major) observation
  int32_t len;           // observations
  int32_t nseg;          // distinct cameras
  int32_t K;             // observations per lane
};

// Device view of one BA problem plus its workspace.  Plain pointers so kernels take it by value.
// Peer exchange window, one per rank, every rank maps every other rank's window (CUDA IPC over
// NVLink / NVSwitch).  Layout of a window: 256-byte header = flags[2][MAX_PEERS] (u64, written by the
// peers) + this rank's exchange counter; then data[2][world][cap] doubles: slot (parity, r) is
// written by rank r only.  An exchange with counter value s uses parity s & 1; a rank can enter
// exchange s + 1 only after every peer has signalled s, and a peer signals s only after it has left
// exchange s - 1 (stream order), so two parities suffice.
constexpr int MAX_PEERS = 8;
constexpr int PEER_HEADER_BYTES = 256;
struct PeerWindow {
  int world, rank;
  unsigned long long cap;             // doubles per (parity, rank) slot
  unsigned long long ll_cam;          // doubles from the start of a slot to the flag-in-data area of the camera sums
  unsigned long long ll_y;            // ... and of the PCG matvec result
  unsigned long long timeout_ns;      // how long a rank waits for a peer's flag before it gives up (SFMBA_ERR_COMM)
  char* base[MAX_PEERS];              // window of every rank in this rank's address space
  __host__ __device__ unsigned long long* flags(int r) const { return reinterpret_cast<unsigned long long*>(base[r]); }
  __host__ __device__ unsigned long long* counter() const { return flags(rank) + 2 * MAX_PEERS; }
  __host__ __device__ double* slot(int window, int parity, int writer) const {
    return reinterpret_cast<double*>(base[window] + PEER_HEADER_BYTES) + ((size_t)parity * world + writer) * cap;
  }
};

struct BaDev {
  int V, N, m;
  int G;                              // groups
  int chunk_bound;                    // allocated chunks
  // inputs
  const float* obs_uv;
  const int32_t* obs_cam;
  const int32_t* obs_pt;
  const int32_t* rowptr;
  const double* kinv;
  double* cams;
  double* pts;
  // index
  int32_t* grp_pt;                    // [G+1]
  int32_t* grp_nchunk;                // [G]   chunks per group (scan input)
  int32_t* grp_K;                     // [G]   observations per lane of a single-chunk group
  int32_t* grp_chunk;                 // [G+1] first chunk of each group
  GroupMeta* grp_meta;                // [G]
  ChunkInfo* chunk_info;              // [chunk_bound]
  double2* rec_b;                     // [chunk_bound][CHUNK_CAP] normalised observation, lane-transposed
  uint32_t* rec_ix;                   // [chunk_bound][CHUNK_CAP]
  double* rec_w;                      // [chunk_bound][CHUNK_CAP] Huber weight of each record at the linearisation
                                      // point (robust solves only: written by k_linearize, read by the matvec)
  double huber_delta;                 // 0 = plain least squares
  int schur_diag;                     // 1 = the linearisation also sums the diagonal blocks of B D^-1 B^T (Schur-Jacobi
                                      // preconditioner, dense solver); 0 = those 28 entries of every row are zero
  int32_t* seg_cam;                   // [chunk_bound][SEG_CAP]   camera of each segment (V = unused)
  uint16_t* seg_lane0;                // [chunk_bound][SEG_RUN_STRIDE] first lane of each segment (+ end)
  int32_t* camrow_ptr;                // [V+1]  CSR: camera -> segment rows
  int32_t* camrow_idx;                // [chunk_bound*SEG_CAP] (chunk, segment) slots sorted by camera (index build only)
  int32_t* rowpos;                    // [chunk_bound*SEG_CAP] inverse: row of slot (chunk, segment) in part12 / part64,
                                      // so that the rows of one camera are contiguous (camrow_ptr[cam] .. camrow_ptr[cam+1])
  // linearisation
  double* camtab;                     // [V][16]
  double* dinv;                       // [6][N] symmetric packing 00 01 02 11 12 22
  double* vj;                         // [3][N] D^-1 e_x
  double* ubuf;                       // [PCG_UBUF][3][N] u_j = D_j^-1 W_j^T p_i of PCG iteration i < PCG_UBUF: with the step
                                      // lengths alpha_i (Scalars::alpha_hist) their sum is D^-1 B^T delta_p, so the
                                      // back-substitution of a short solve needs no pass over the observations
  double* grp_cost;                   // [max(G, cost grid)]
  double* grp_exv;                    // [G] sum_j e_x^T D^-1 e_x of the group (adaptive LM only, else null)
  double* part64;                     // [chunk_bound*SEG_CAP][CAM_ACC or ROW_COMPACT] segment rows of the Gram sums
  double* cam_acc;                    // [V][CAM_ACC] (+ 8 trailing doubles: local cost at [V*CAM_ACC])
  double* ablk;                       // [V][49] A_i + lambda I
  double* minv;                       // [V][49] block-Jacobi inverse of S_ii
  // PCG
  double* camit;                      // [V][16]
  // single-precision matvec (ba_matvec32.cu); null when the solve runs the FP64 matvec only
  float4* xrel;                       // [N] point - origin of its group (first point of the group), FP64 difference stored FP32
  float* dinv32;                      // [N][8] D^-1 (00 01 02 11 12 22 - -)
  float* segst;                       // [chunk_bound*SEG_CAP][12] per (chunk, segment): R (9), C - group origin (3)
  float* camit32;                     // [V][16] FP32 copy of the camit rows
  float* part64f;                     // [chunk_bound*SEG_CAP][ROW_F32] segment rows of the FP32 linearisation
  float* part12f;                     // [chunk_bound*SEG_CAP][MV_ROW32] segment rows of the FP32 matvec
  float* ubuf32;                      // [PCG_UBUF][3][N] the u_j the FP32 matvecs keep (as ubuf)
  double* part12;                     // [chunk_bound*SEG_CAP][MV_ACC]
  double* svec;                       // [7V] sum_j W D^-1 W^T p (local; allreduced when world > 1)
  double* x;                          // [7V] solution (delta p)
  double* r;
  double* z;
  double* p;
  double* q;                          // [7V] w = S z of the current iteration
  double* s;                          // [7V] S p (by recurrence)
  double* gvec;                       // [7V] right-hand side
  double* red_a;                      // [<= 2048] per-block partials of the PCG dot products (eight per CTA in k_pcg_step1)
  double* red_b;                      // [<= 2048]
  double* cams_backup;                // [V][7]  adaptive mode
  double* pts_backup;                 // [N][3]
  Scalars* sc;
  unsigned long long* host_sig;       // pinned, device-visible: (epoch << 32) | (pcg_iters << 1) | done; or null
  PeerWindow xw;                      // NVLink peer exchange window (world > 1 and peers mapped), else world = 0
};

// index statistics of one solve as the device hands them to the host (pinned memory)
struct IndexStats {
  unsigned epoch;                     // written last: the record belongs to this solve
  int32_t num_chunks, max_chunk_len, max_group_pts, max_nseg, max_k;
  uint32_t flags;
};

struct LaunchCounter {
  int64_t launches = 0;
};

struct TileShape {                    // from the index statistics; sizes dynamic shared memory
  int max_chunk_len = CHUNK_CAP;
  int max_group_pts = GROUP_PT_CAP;
  int max_nseg = SEG_CAP;
  int max_k = CHUNK_K_MAX;
};

// ---- index (ba_index.cu) ------------------------------------------------------------------------
size_t index_temp_bytes(int G, int chunk_bound, int V);
// phase 1: groups + chunk counts + shape statistics in sc (host reads them back once)
void launch_index_count(const BaDev& d, bool fine_chunks, void* temp, size_t temp_bytes, IndexStats* host_stats,
                        unsigned epoch, cudaStream_t st, LaunchCounter& lc);
// phase 2a: records and run / segment tables of the groups [g0, g1) (needs obs_uv of their observations only)
void launch_chunk_build(const BaDev& d, int g0, int g1, cudaStream_t st, LaunchCounter& lc);
// phase 2b: camera -> segment-row CSR over all chunks
void launch_index_build(const BaDev& d, int nchunks, void* temp, size_t temp_bytes, int32_t* sort_keys_out,
                        int32_t* iota, cudaStream_t st, LaunchCounter& lc);

void launch_widen_u16(const uint16_t* in, int32_t* out, size_t n, cudaStream_t st, LaunchCounter& lc);

// ---- linearisation (ba_linearize.cu) ------------------------------------------------------------
void launch_cam_prepare(const BaDev& d, cudaStream_t st, LaunchCounter& lc);
// per-point Schur elimination: D_j^-1, v_j, cost -> sc->costs[cost_slot] / cost_cur
// cost only at the current parameters -> sc->cost_new and (optionally) *out_cost_dev
void launch_cost_tiles(const BaDev& d, const TileShape& ts, double* out_cost_dev, cudaStream_t st, LaunchCounter& lc);
void launch_cost(const BaDev& d, double* out_cost_dev, cudaStream_t st, LaunchCounter& lc);   // index-free
void launch_k1(const BaDev& d, bool dvdc_quirk, float* out_r, float* out_jp, float* out_jx, double* out_r64,
               double* out_jp64, double* out_jx64, double* out_cost, cudaStream_t st, LaunchCounter& lc);

// ---- camera-side reduction (ba_schur.cu) --------------------------------------------------------
// fused linearisation: point blocks (D^-1, v, cost -> sc->costs[cost_slot]) + camera-side Gram rows
void launch_linearize(const BaDev& d, const TileShape& ts, bool dvdc_quirk, int cost_slot, cudaStream_t st,
                      LaunchCounter& lc, int g0 = 0, int g1 = -1, bool sum_cost = true);
// the same with FP32 Jacobian arithmetic (ba_linearize32.cu); plain least squares without the Schur-diagonal sums only
void launch_linearize32(const BaDev& d, const TileShape& ts, bool dvdc_quirk, int cost_slot, cudaStream_t st,
                        LaunchCounter& lc, int g0 = 0, int g1 = -1, bool sum_cost = true);
void launch_sum_cost(const BaDev& d, int cost_slot, cudaStream_t st, LaunchCounter& lc);
void launch_camsum64(const BaDev& d, cudaStream_t st, LaunchCounter& lc);
// several GPUs with peer windows: segment rows -> every rank's window -> cam_acc and cost summed over ranks
void launch_camsum64_peer(const BaDev& d, int cost_slot, cudaStream_t st, LaunchCounter& lc);       // part64 -> cam_acc
void launch_cam_blocks(const BaDev& d, cudaStream_t st, LaunchCounter& lc);     // ablk, minv, gvec
// PCG solver: camera sums (this GPU's rows, or all ranks' through the peer windows), blocks and the PCG start in one
// kernel per camera
void launch_cam_finish(const BaDev& d, bool peer, bool rows_f32, bool sum_cost, int cost_slot, unsigned epoch,
                       bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc);

// ---- PCG (ba_pcg.cu) -----------------------------------------------------------------------------
void launch_pcg_init(const BaDev& d, unsigned epoch, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc);
void launch_matvec(const BaDev& d, const TileShape& ts, bool backsub, bool dvdc_quirk, cudaStream_t st,
                   LaunchCounter& lc);
void launch_matvec32(const BaDev& d, const TileShape& ts, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc);   // FP32 arithmetic
void launch_camsum12(const BaDev& d, cudaStream_t st, LaunchCounter& lc);   // part12 -> svec (multi-GPU)
// vector part of one PCG iteration.  mode 1: y from this GPU's rows; 2: y = svec (NCCL all-reduced);
// 3: partial y exchanged through the peer windows over NVLink inside the kernel
void launch_pcg_tail(const BaDev& d, int mode, double tol, int max_iter, unsigned epoch, bool dvdc_quirk, bool rows_f32,
                     bool abort_slow, cudaStream_t st, LaunchCounter& lc);
void launch_pcg_reopen(const BaDev& d, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc);
void launch_pcg_refresh(const BaDev& d, int mode, double tol, unsigned epoch, bool dvdc_quirk, cudaStream_t st,
                        LaunchCounter& lc);
void launch_peer_sum_scalars(const BaDev& d, double* vals, int n, cudaStream_t st, LaunchCounter& lc);
void launch_update_cams(const BaDev& d, cudaStream_t st, LaunchCounter& lc);
void launch_backsub_points(const BaDev& d, bool from_f32, int iters, cudaStream_t st, LaunchCounter& lc);
void launch_lm_decide(const BaDev& d, double lam_min, double lam_max, cudaStream_t st, LaunchCounter& lc);
void launch_lm_backup(const BaDev& d, cudaStream_t st, LaunchCounter& lc);
void launch_lm_model_decrease(const BaDev& d, cudaStream_t st, LaunchCounter& lc);

// ---- dense direct solve of the reduced system (dense.cu) -------------------------------------------
struct DenseWork {
  double* S;        // [npad][ld] column-major, lower triangle used
  size_t ld;        // = npad
  int n, npad;      // 7 V and its padding to a multiple of 128
  double* linv;     // [npad / 128][128][128] inverses of the diagonal factor tiles
  double* vec;      // [3][npad] rhs | y | x
};
// the look-ahead of the factorisation: a high-priority stream for the critical path and the events that order it
// against the caller's stream.  Owned by the context (created on first use, destroyed with it).
struct DenseStreams {
  cudaStream_t crit = nullptr;
  std::vector<cudaEvent_t> ev_panel, ev_upd;
  void release() {
    for (cudaEvent_t e : ev_panel) cudaEventDestroy(e);
    for (cudaEvent_t e : ev_upd) cudaEventDestroy(e);
    ev_panel.clear();
    ev_upd.clear();
    if (crit) cudaStreamDestroy(crit);
    crit = nullptr;
  }
};
size_t dense_padded_dim(int V);
// S = -(off-diagonal blocks) of this rank's points (zeroed first); all-reduced by the caller when world > 1
void launch_dense_assemble(const BaDev& d, const DenseWork& w, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc);
// adds the diagonal blocks (cam_acc, lambda), factors, solves S x = gvec -> d.x
void launch_dense_factor_solve(const BaDev& d, const DenseWork& w, DenseStreams& ds, cudaStream_t st, LaunchCounter& lc);
// camit rows of an arbitrary 7V-vector (the back-substitution direction)
void launch_cam_rows(const BaDev& d, const double* vec, bool dvdc_quirk, cudaStream_t st, LaunchCounter& lc);

// ---- refiners (refine.cu) ------------------------------------------------------------------------
void launch_validate_csr(int n_items, int num_obs, const int32_t* rowptr, const int32_t* idx, int bound,
                         uint32_t* flags, cudaStream_t st, LaunchCounter& lc);
void launch_triangulate(int num_pts, const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                        const double* obs_uv, double* pts, double lambda, int iterations, uint32_t* flags,
                        cudaStream_t st, LaunchCounter& lc);
void launch_triangulate_linear(int num_pts, const double* projs, const int32_t* rowptr, const int32_t* obs_view,
                               const double* obs_uv, double* pts, int32_t* front, uint32_t* flags, cudaStream_t st,
                               LaunchCounter& lc);
void launch_pnp(int num_views, const int32_t* rowptr, const double* obs_xy, const double* obs_X, double* rot,
                double* loc, double lambda, int iterations, bool row_packing_quirk, bool dvdc_quirk,
                uint32_t* flags, cudaStream_t st, LaunchCounter& lc);

// ---- RANSAC hypotheses (ransac.cu) -----------------------------------------------------------------
void launch_pnp_ransac(int n, int H, const double* key, const double* X, const double* Kmat, const double* Kinv,
                       const int32_t* samples, const int32_t* loc_sign, double thr, double* models, double* projs,
                       int32_t* counts, int32_t* best, uint8_t* mask, uint32_t* flags, cudaStream_t st,
                       LaunchCounter& lc);
void launch_fund_ransac(int n, int H, const double* pairs, const int32_t* samples, double thr, double* models,
                        int32_t* counts, int32_t* best, uint8_t* mask, uint32_t* flags, cudaStream_t st,
                        LaunchCounter& lc);

// ---- visibility (visibility.cu) ------------------------------------------------------------------
int build_visibility_device(int num_views, int num_pts, const int64_t* diag, const int64_t* key_off_host,
                            int64_t total_keys, int32_t* obs_pt, int32_t* obs_cam, int32_t* obs_key,
                            int32_t* pt_rowptr, int32_t* host_num_obs, cudaStream_t st, LaunchCounter& lc,
                            std::string& err);

}  // namespace sfmba
