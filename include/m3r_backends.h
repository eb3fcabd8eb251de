/*
 * m3r_backends.h -- C ABI of the B200-native geometric backend (libm3r_b200.so).
 *
 * This is the drop-in boundary for the reference's one native extension,
 * `mast3r_slam_backends` (pybind11 module, /root/reference/mast3r_slam/backend/src/gn.cpp:116-123,
 * declarations in backend/include/gn.h).  Each entry point below is what a binding of the
 * corresponding reference export would call; the reference interface it replaces is cited
 * per function.  Plain pointers, sizes and scalars only: no torch types.
 *
 * Conventions
 *  - every tensor argument is a DEVICE pointer to a dense row-major (contiguous) buffer on
 *    the current CUDA device, laid out exactly as the reference's tensors are
 *    (the reference checks contiguity only: gn.h:5, gn.cpp:14-21);
 *  - inputs are borrowed for the duration of the call; outputs are caller-allocated;
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  All work is
 *    enqueued on it; no entry point synchronises the device or the stream;
 *  - return value: M3R_OK (0) or an M3R_ERR_* code; m3r_last_error() gives the message of
 *    the last failure on the calling thread.  Numerical failure of the Gauss-Newton linear
 *    solve is NOT an error: the step is zero, as in the reference (gn_kernels.cu:147-150).
 */
#ifndef M3R_BACKENDS_H
#define M3R_BACKENDS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define M3R_ABI_VERSION 1

#define M3R_OK 0
#define M3R_ERR_INVALID_ARGUMENT 1
#define M3R_ERR_CUDA 2
#define M3R_ERR_UNSUPPORTED 3
#define M3R_ERR_NO_MEMORY 4

/* dtype codes for refine_matches descriptors (AT_DISPATCH_FLOATING_TYPES_AND_HALF,
 * matching_kernels.cu:103) */
#define M3R_DTYPE_F16 0
#define M3R_DTYPE_F32 1
#define M3R_DTYPE_F64 2

/* Gauss-Newton residual kinds */
#define M3R_GN_RAYS 0   /* ray_align_kernel,  gn_kernels.cu:813-1138  */
#define M3R_GN_CALIB 1  /* calib_proj_kernel, gn_kernels.cu:1231-1543 */
#define M3R_GN_POINTS 2 /* point_align_kernel, gn_kernels.cu:455-723  */

int m3r_abi_version(void);
const char *m3r_last_error(void);

/* Number of kernels this library has launched in the calling process since load (all
 * streams); bench.py reads it around the timed region for its `gpu_launches` claim. */
uint64_t m3r_launch_count(void);

/* ---------------------------------------------------------------------------------------
 * iter_proj -- replaces `iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter,
 * lambda_init, cost_thresh) -> [p_new, converged]`  (gn.h:89-95, gn.cpp:84-99,
 * matching_kernels.cu:119-316).
 *   rays_img_with_grad [b,h,w,9] f32   unit ray (3) + d/du (3) + d/dv (3) per pixel
 *   pts_3d_norm        [b,n,3]   f32   unit rays to match
 *   p_init             [b,n,2]   f32   initial (u,v)
 *   p_new              [b,n,2]   f32   out
 *   converged          [b,n]     u8    out (torch.bool storage)
 * `variant`: 0 = library default; other values select alternative kernels with identical
 * results (used by the benchmarks).
 * ------------------------------------------------------------------------------------- */
int m3r_iter_proj(const float *rays_img_with_grad, const float *pts_3d_norm,
                  const float *p_init, float *p_new, uint8_t *converged, int b, int h, int w,
                  int n, int max_iter, float lambda_init, float cost_thresh, void *stream);
int m3r_iter_proj_v(const float *rays_img_with_grad, const float *pts_3d_norm,
                    const float *p_init, float *p_new, uint8_t *converged, int b, int h, int w,
                    int n, int max_iter, float lambda_init, float cost_thresh, int variant,
                    void *stream);

/* ---------------------------------------------------------------------------------------
 * refine_matches -- replaces `refine_matches(D11, D21, p1, radius, dilation_max) ->
 * [p1_new]`  (gn.h:112-117, gn.cpp:101-114, matching_kernels.cu:25-116).
 *   D11    [b,h,w,fdim]  dtype     descriptors of view 1
 *   D21    [b,n,fdim]    dtype     descriptors of the query points
 *   p1     [b,n,2]       i64       (u,v) from iter_proj
 *   p1_new [b,n,2]       i64       out
 * Scores are accumulated in `dtype` exactly as the reference does (for f16: every multiply
 * and every add rounds to binary16), so the result is bit-exact with the reference.
 * ------------------------------------------------------------------------------------- */
int m3r_refine_matches(const void *D11, const void *D21, const int64_t *p1, int64_t *p1_new,
                       int b, int h, int w, int n, int fdim, int dtype, int radius,
                       int dilation_max, void *stream);
int m3r_refine_matches_v(const void *D11, const void *D21, const int64_t *p1, int64_t *p1_new,
                         int b, int h, int w, int n, int fdim, int dtype, int radius,
                         int dilation_max, int variant, void *stream);

/* Allocation-free variant of m3r_refine_matches.  It has NO counterpart in the reference: it was
 * added for callers that cannot allocate (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same arguments, same
 * results (bit-exact), but the tile plan the kernel needs as scratch lives in one file-scope
 * device array of fixed size inside the library, so the call performs no cudaMalloc /
 * cudaMallocAsync / cudaFree and no other allocation.
 *   limits: f16 descriptors with fdim 24, radius 3, one query per pixel (n = h*w), and at most
 *           m3r_refine_matches_static_max_tiles() = 3200 query tiles of 16x16 per call
 *           (b <= 4 frames at 512x384); anything else returns M3R_ERR_INVALID_ARGUMENT with a
 *           message that names the limit;
 *   cost:   every GPU that loads the library holds the 55 296-byte array until process exit;
 *   sharing: all calls on a device share the array and it keeps the last call's contents
 *           (nothing in it needs to start at zero); calls that overlap on the GPU -- different
 *           streams, parallel graph branches -- race on it and must be serialised by the caller. */
int m3r_refine_matches_static(const void *D11, const void *D21, const int64_t *p1, int64_t *p1_new,
                              int b, int h, int w, int n, int fdim, int dtype, int radius,
                              int dilation_max, void *stream);
int64_t m3r_refine_matches_static_max_tiles(void);

/* ---------------------------------------------------------------------------------------
 * Matching front end -- the torch code either side of the two kernels above, fused
 * (SURVEY.md 8f rows N1 and N3).
 *
 * m3r_prep_rays replaces the ray-image part of `prep_for_iter_proj(X11, X21,
 * idx_1_to_2_init)` (mast3r_slam/matching.py:25-36) and `img_gradient(img)`
 * (mast3r_slam/image.py:5-38): F.normalize over xyz (eps 1e-12), the two depthwise 3x3
 * filters on the reflect-padded ray image, and the concat/permute into iter_proj's layout.
 *   X11                [b,h,w,3] f32   canonical pointmap of view 1
 *   rays_img_with_grad [b,h,w,9] f32   out
 *
 * m3r_match replaces `match(X11, X21, D11, D21, idx_1_to_2_init) -> (idx_1_to_2,
 * valid_match2)` (matching.py:8-10, 52-88): prep, iter_proj, `.long()`, the 3D-distance
 * occlusion test, refine_matches (when radius > 0) and pixel_to_lin, in three launches.
 *   X21        [b,h,w,3]    f32    view-2 points in view 1's frame (n = h*w queries)
 *   D11, D21   [b,h,w,fdim] dtype  descriptors, ALREADY in the dtype the reference would
 *                                  refine in (it casts to half: matching.py:79-80); may be
 *                                  NULL when radius == 0
 *   idx_init   [b,n] i64 or NULL   previous idx_1_to_2 (NULL = identity, matching.py:41-44)
 *   idx_1_to_2 [b,n] i64           out: u + w*v of the refined pixel
 *   valid      [b,n] u8            out: converged && |X11[p1] - X21| < dist_thresh
 *   rays_ws    b*h*w*12 floats     caller-owned scratch, 16-byte aligned (the ray image in an
 *                                  internal 48-byte-texel layout read with 128-bit loads)
 *   p1_ws      2*b*n*2 int64       caller-owned scratch, 16-byte aligned
 * ------------------------------------------------------------------------------------- */
int m3r_prep_rays(const float *X11, float *rays_img_with_grad, int b, int h, int w, void *stream);
int m3r_match(const float *X11, const float *X21, const void *D11, const void *D21, int fdim,
              int dtype, const int64_t *idx_init, int64_t *idx_1_to_2, uint8_t *valid,
              float *rays_ws, int64_t *p1_ws, int b, int h, int w, int max_iter,
              float lambda_init, float cost_thresh, float dist_thresh, int radius,
              int dilation_max, void *stream);

/* ---------------------------------------------------------------------------------------
 * Match confidences and fractions -- replaces the torch gathers and
 * reductions that follow every match: `Qk = sqrt(Qff[idx_f2k] * Qkf)`, `valid_opt`,
 * `valid_kf`, `match_frac`, `match_frac_k`, `unique_frac_f` (mast3r_slam/tracker.py:41,
 * 59-67, 104-110) and `Qj = sqrt(Qii[batch_inds, idx_i2j] * Qji)`, `valid_j`,
 * `match_frac_j` (mast3r_slam/global_opt.py:55-68), one pass over idx.
 *   idx         [b,n]     i64   match index into the source view (n_src points per batch row)
 *   Q_src       [b,n_src] f32   gathered through idx;  Q_dst [b,n] f32
 *   valid_match [b,n]     u8
 *   C_src       [b,n_src] f32 or NULL, C_dst [b,n] f32 or NULL   averaged confidences
 *   Q_out       [b,n]     f32   sqrt(Q_src[idx] * Q_dst), bit-identical to torch
 *   valid_q     [b,n]     u8    valid_match && Q > Q_conf
 *   valid_opt   [b,n]     u8 or NULL (with C): valid_q && C_src[idx] > C_conf && C_dst > C_conf
 *   counts      [b,3]     i32   sum(valid_q), sum(valid_opt), #distinct idx among valid_match
 *   scratch     m3r_match_confidence_scratch_bytes(b, n_src) bytes, or NULL to skip the
 *               distinct count (counts[:,2] = 0)
 * ------------------------------------------------------------------------------------- */
int64_t m3r_match_confidence_scratch_bytes(int b, int n_src);
int m3r_match_confidence(const int64_t *idx, const float *Q_src, const float *Q_dst,
                         const uint8_t *valid_match, const float *C_src, const float *C_dst,
                         float Q_conf, float C_conf, float *Q_out, uint8_t *valid_q,
                         uint8_t *valid_opt, int32_t *counts, void *scratch, int b, int n, int n_src,
                         void *stream);

/* ---------------------------------------------------------------------------------------
 * Tracking pose solve -- replaces the bodies of
 * `FrameTracker.opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid)` and
 * `FrameTracker.opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K,
 * img_size)` (mast3r_slam/tracker.py:173-215, 217-266) including `solve` (:157-171), the
 * geometry helpers they call (geometry.py:12-33, 45-53, 64-107), `check_convergence` and
 * `huber` (nonlinear_optimizer.py:5-35) and the lietorch Sim3 algebra around them: one
 * cooperative kernel runs all iterations on the device, no host synchronisation.
 *   Xf          [n_f,3] f32   frame pointmap.  With idx_f2k == NULL it is the already gathered
 *                             `Xf[idx_f2k]` of tracker.py:155 (n_f == n); otherwise the
 *                             canonical pointmap, gathered inside the kernel
 *   idx_f2k     [n] i64|NULL  for keyframe point k, its match in the frame
 *   Xk          [n,3]  f32    keyframe pointmap (rays only; may be NULL for calib)
 *   Qk          [n]    f32    match confidence sqrt(Qff[idx] * Qkf) (tracker.py:41)
 *   valid       [n]    u8     valid_opt (tracker.py:62-66)
 *   meas_k      [n,3]  f32    calib only: (u, v, log z) of the keyframe points (:145-150)
 *   valid_meas_k[n]    u8     calib only
 *   K           [3,3]  f32    calib only: intrinsics, row-major, on the device
 *   T_WCf,T_WCk [8]    f32    lietorch.Sim3.data rows (t xyz, q xyzw, s), 16-byte aligned
 *   result      device struct, see below; status CHOLESKY_FAILED is where the reference's
 *               torch.linalg.cholesky raises (tracker.py:168, caught at :90-92)
 *   scratch     m3r_track_scratch_doubles() doubles of caller-owned device memory, ZEROED once
 *               when allocated and then passed to every call (it carries the grid-barrier
 *               counter from launch to launch); one scratch per stream
 * ------------------------------------------------------------------------------------- */
#define M3R_TRACK_OK 0
#define M3R_TRACK_CHOLESKY_FAILED 1

typedef struct m3r_track_params {
  int kind;                /* M3R_GN_RAYS (ray + distance) or M3R_GN_CALIB (pixel + log depth) */
  int max_iters;           /* config tracking.max_iters */
  float rel_error;         /* tracking.rel_error  */
  float delta_norm;        /* tracking.delta_norm */
  float huber_k;           /* tracking.huber      */
  float sigma_a, sigma_b;  /* rays: sigma_ray, sigma_dist; calib: sigma_pixel, sigma_depth */
  int height, width;       /* calib: img_size */
  float pixel_border;      /* calib: tracking.pixel_border */
  float z_eps;             /* calib: tracking.depth_eps */
} m3r_track_params;

typedef struct m3r_track_result {
  float T_WCf[8];      /* T_WCk * T_CkCf (tracker.py:213) */
  float T_CkCf[8];     /* the optimised relative pose */
  int32_t status;      /* M3R_TRACK_* */
  int32_t iterations;  /* executed iterations */
  double cost;         /* 0.5 * b^T b of the last linearisation (tracker.py:166) */
} m3r_track_result;

int64_t m3r_track_scratch_doubles(void);
int m3r_track_pose(const m3r_track_params *params, const float *Xf, int64_t n_f,
                   const int64_t *idx_f2k, const float *Xk, const float *Qk, const uint8_t *valid,
                   const float *meas_k, const uint8_t *valid_meas_k, const float *K, int64_t n,
                   const float *T_WCf, const float *T_WCk, m3r_track_result *result, double *scratch,
                   void *stream);

/* ---------------------------------------------------------------------------------------
 * Pointmap producers.
 *
 * m3r_update_pointmap replaces the in-place fusion branches of `Frame.update_pointmap(X, C)`
 * (mast3r_slam/frame.py:41-102; the `first` / `recent` / `best_score` branches are plain
 * copies decided on the host and stay there), optionally preceded by the Sim(3) transport
 * `Xkk = T_CkCf.act(Xkf)` of mast3r_slam/tracker.py:97-98 when T != NULL -- T may point
 * straight at `m3r_track_result.T_CkCf`, so no host round trip sits between the pose solve
 * and the keyframe update.  weighted_pointmap and indep_conf are bit-identical to torch.
 *   X_canon [n,3] f32 in/out, C_canon [n] f32 in/out (the running SUM of confidences, frame.py:76)
 *   X [n,3] f32, C [n] f32 the new observation;  T [8] f32 or NULL, 16-byte aligned
 *
 * m3r_constrain_points_to_ray replaces `constrain_points_to_ray(img_size, Xs, K)`
 * (mast3r_slam/geometry.py:37-42 with backproject :107-115 and get_pixel_coords :118-123):
 * X <- z * ((u - cx)/fx, (v - cy)/fy, 1).  Xs, out [b,h,w,3] f32 (out may alias Xs); K [3,3]
 * f32 on the device.  Bit-identical to torch.
 * ------------------------------------------------------------------------------------- */
#define M3R_FILTER_WEIGHTED_POINTMAP 0  /* frame.py:74-77 */
#define M3R_FILTER_WEIGHTED_SPHERICAL 1 /* frame.py:78-101 */
#define M3R_FILTER_INDEP_CONF 2         /* frame.py:69-73 */
int m3r_update_pointmap(int mode, float *X_canon, float *C_canon, const float *X, const float *C,
                        const float *T, int64_t n, void *stream);
int m3r_constrain_points_to_ray(const float *Xs, float *out, const float *K, int b, int h, int w,
                                void *stream);

/* ---------------------------------------------------------------------------------------
 * Gauss-Newton on the Sim(3) keyframe graph.
 *
 * A workspace caches device scratch (per-edge blocks, the assembled system, index maps)
 * across calls; create one per device and per host thread that calls the solver.
 * ------------------------------------------------------------------------------------- */
typedef struct m3r_gn_workspace m3r_gn_workspace;
int m3r_gn_workspace_create(m3r_gn_workspace **out);
void m3r_gn_workspace_destroy(m3r_gn_workspace *ws);

/* Common tensor arguments (reference layouts, gn.h:22-87):
 *   Twc         [N,8]   f32  in/out  (t xyz, q xyzw, s); rows in sorted-unique keyframe order
 *   Xs          [N,n,3] f32          canonical pointmaps
 *   Cs          [N,n,1] f32          confidences
 *   ii, jj      [E]     i64          keyframe ids of each directed edge i->j
 *   idx_ii2jj   [E,n]   i64          for point k of j, its match in i
 *   valid_match [E,n,1] u8 (bool)
 *   Q           [E,n,1] f32          match confidence
 *   dx          [N-1,7] f32  out     step of the last executed iteration (tau, omega, sigma)
 * Pose 0 is held fixed (num_fix = 1, gn_kernels.cu:741,1157,1566).  The loop runs up to
 * max_iter iterations and stops after the first one whose ||dx||_F < delta_thresh
 * (gn_kernels.cu:1219-1222); the test happens on the device, the call never blocks.
 */

/* replaces gauss_newton_rays (gn.h:45-55, gn.cpp:28-52, gn_kernels.cu:1140-1228) */
int m3r_gauss_newton_rays(m3r_gn_workspace *ws, float *Twc, const float *Xs, const float *Cs,
                          const int64_t *ii, const int64_t *jj, const int64_t *idx_ii2jj,
                          const uint8_t *valid_match, const float *Q, float *dx, int num_poses,
                          int num_points, int num_edges, float sigma_ray, float sigma_dist,
                          float C_thresh, float Q_thresh, int max_iter, float delta_thresh,
                          void *stream);

/* replaces gauss_newton_calib (gn.h:74-87, gn.cpp:54-82, gn_kernels.cu:1546-1638);
 * K is a DEVICE pointer to the [3,3] f32 intrinsics (fx=K[0], fy=K[4], cx=K[2], cy=K[5]) */
int m3r_gauss_newton_calib(m3r_gn_workspace *ws, float *Twc, const float *Xs, const float *Cs,
                           const float *K, const int64_t *ii, const int64_t *jj,
                           const int64_t *idx_ii2jj, const uint8_t *valid_match, const float *Q,
                           float *dx, int num_poses, int num_points, int num_edges, int height,
                           int width, int pixel_border, float z_eps, float sigma_pixel,
                           float sigma_depth, float C_thresh, float Q_thresh, int max_iter,
                           float delta_thresh, void *stream);

/* replaces gauss_newton_points (gn.h:22-31, gn.cpp:3-26, gn_kernels.cu:725-811) */
int m3r_gauss_newton_points(m3r_gn_workspace *ws, float *Twc, const float *Xs, const float *Cs,
                            const int64_t *ii, const int64_t *jj, const int64_t *idx_ii2jj,
                            const uint8_t *valid_match, const float *Q, float *dx,
                            int num_poses, int num_points, int num_edges, float sigma_point,
                            float C_thresh, float Q_thresh, int max_iter, float delta_thresh,
                            void *stream);

/* ---------------------------------------------------------------------------------------
 * Staged Gauss-Newton (what the three calls above are made of).  Used by the multi-GPU
 * path, where each rank linearises its own shard of the directed edges and the assembled
 * systems are summed across ranks before every rank solves and retracts its replica of
 * Twc, and by the parity tests, which check each stage against the oracle.
 * ------------------------------------------------------------------------------------- */
typedef struct {
  int kind;                 /* M3R_GN_RAYS | M3R_GN_CALIB | M3R_GN_POINTS */
  int num_poses;            /* N */
  int num_points;           /* n */
  int num_edges;            /* E: directed edges of the WHOLE graph (length of ii, jj) */
  int edge_begin, edge_end; /* this call's shard [begin,end) of the directed edges */
  float *Twc;               /* [N,8] in/out (replicated) */
  const float *Xs;          /* [N,n,3] (replicated) */
  const float *Cs;          /* [N,n]   (replicated) */
  const int64_t *ii;        /* [E] whole graph */
  const int64_t *jj;        /* [E] whole graph */
  const int64_t *idx_ii2jj; /* [edge_end-edge_begin, n] rows of this shard only */
  const uint8_t *valid_match; /* [edge_end-edge_begin, n] */
  const float *Q;             /* [edge_end-edge_begin, n] */
  /* residual parameters (unused ones ignored) */
  float sigma_ray, sigma_dist;           /* rays   */
  const float *K;                        /* calib: device [3,3] */
  int height, width, pixel_border;       /* calib  */
  float z_eps, sigma_pixel, sigma_depth; /* calib  */
  float sigma_point;                     /* points */
  float C_thresh, Q_thresh;
} m3r_gn_problem;

/* Size in doubles of the assembled system buffer: D*D (dense, row-major, lower triangle
 * meaningful) followed by D right-hand-side entries, D = 7*(N-1). */
int64_t m3r_gn_system_size(int num_poses);

/* Resets the device-side stop flag and remaps keyframe ids to packed pose indices
 * (replaces get_unique_kf_idx/create_inds, gn_kernels.cu:161-170, without the host sync). */
int m3r_gn_begin(m3r_gn_workspace *ws, const m3r_gn_problem *prob, void *stream);

/* One accumulation pass over the shard's edges and deterministic assembly into `system`
 * (caller-allocated device buffer of m3r_gn_system_size doubles; fully overwritten). */
int m3r_gn_linearize(m3r_gn_workspace *ws, const m3r_gn_problem *prob, double *system,
                     void *stream);

/* Cholesky-solve `system` (destroyed), dx = -H^{-1} b (zeros if the factorisation fails),
 * left-retract poses 1..N-1 in place, set the stop flag if ||dx||_F < delta_thresh.
 * After the stop flag is set, m3r_gn_linearize / m3r_gn_solve_retract become no-ops on the
 * device, which is how the max_iter loop ends early without a host round trip. */
int m3r_gn_solve_retract(m3r_gn_workspace *ws, const m3r_gn_problem *prob, double *system,
                         float *dx, float delta_thresh, void *stream);

/* Per-edge world-frame blocks of the last m3r_gn_linearize, for the parity tests:
 * Hjj [E_shard,7,7] and gj [E_shard,7] in double, with H_ii = H_jj, H_ij = H_ji = -H_jj,
 * g_i = -g_j (the identities the reference's Ji = -Jj implies, gn_kernels.cu:1000).
 * Device pointers owned by the workspace, valid until the next call on it. */
int m3r_gn_edge_blocks(m3r_gn_workspace *ws, const double **Hjj, const double **gj);

/* Number of iterations the last m3r_gauss_newton_* / staged loop executed.  This call
 * synchronises `stream` (it reads a device counter); it is for tests and benchmarks. */
int m3r_gn_iterations_done(m3r_gn_workspace *ws, void *stream, int *iters);

/* ---------------------------------------------------------------------------------------
 * Multi-GPU exchange of the assembled systems over NVLink peer memory (new; the reference
 * is single-GPU).  One process per GPU.  Every rank owns an exchange area allocated by this
 * library (two system buffers, used alternately, plus one arrival flag per peer) and maps
 * the areas of all peers through CUDA IPC.  Per Gauss-Newton iteration a rank
 *   1. linearises its edge shard straight into its own exchange buffer
 *      (m3r_gn_linearize with system = m3r_gn_comm_system()),
 *   2. m3r_gn_comm_signal: publishes "my buffer of this round is complete" into every
 *      peer's flag array (system-scope release store after a system fence),
 *   3. m3r_gn_comm_reduce: waits for all peers' flags of this round, then sums the W buffers
 *      in rank order (bitwise identical result on every rank, so the replicated solves stay
 *      in lockstep with no broadcast) into `out`, reading the peers' buffers directly over
 *      NVLink,
 *   4. m3r_gn_solve_retract on `out`.
 * Two buffers suffice: a rank reaches round t+2 only after every peer has signalled round
 * t+1, i.e. after it finished reading round t.
 * Both kernels are no-ops once the workspace's stop flag is set (all ranks agree on it).
 * ------------------------------------------------------------------------------------- */
typedef struct m3r_gn_comm m3r_gn_comm;
#define M3R_IPC_HANDLE_BYTES 64

/* Allocates the exchange area for systems of up to `max_system_doubles` doubles. */
int m3r_gn_comm_create(m3r_gn_comm **out, int rank, int world_size, int64_t max_system_doubles);
void m3r_gn_comm_destroy(m3r_gn_comm *comm);
/* IPC handle of this rank's exchange area (M3R_IPC_HANDLE_BYTES bytes), to be all-gathered. */
int m3r_gn_comm_handle(m3r_gn_comm *comm, void *handle_out);
/* `handles`: world_size consecutive handles in rank order (own entry ignored). */
int m3r_gn_comm_connect(m3r_gn_comm *comm, const void *handles);
/* Single-process variant for tests: connect communicators that live in the same process
 * (one per device, peer access enabled by the call). */
int m3r_gn_comm_connect_local(m3r_gn_comm **comms, int world_size);
/* Starts the next exchange round and returns this rank's buffer for it. */
int m3r_gn_comm_next_round(m3r_gn_comm *comm, double **system_out);
int m3r_gn_comm_signal(m3r_gn_comm *comm, m3r_gn_workspace *ws, void *stream);
/* out (m3r_gn_system_size(num_poses) doubles): block-lower triangle and right-hand side of
 * the summed system, i.e. everything m3r_gn_solve_retract reads. */
int m3r_gn_comm_reduce(m3r_gn_comm *comm, m3r_gn_workspace *ws, double *out, int num_poses,
                       void *stream);
/* m3r_gn_comm_reduce in two halves, for a host thread that drives SEVERAL ranks itself (ranks that
 * share a process: m3r_gn_multi_*, the loop-back tests).  Such a host must not rely on the reduce
 * kernels' spin-wait: a kernel that waits for a flag is only safe when the kernel that sets it is
 * guaranteed to run beside it, which CUDA does not promise for two streams of one device, nor when
 * the setter has not been enqueued yet.  Instead it enqueues, for every rank, signal -> [event],
 * then phase 1 behind the events of all ranks -> [event], then phase 2 behind those, so that every
 * wait finds its flags set.  phase 1: reduce-scatter + second signal (nothing in the one-shot form);
 * phase 2: all-gather (the whole reduce in the one-shot form). */
int m3r_gn_comm_reduce_phase(m3r_gn_comm *comm, m3r_gn_workspace *ws, double *out, int num_poses,
                             int phase, void *stream);
/* Exchange algorithm of m3r_gn_comm_reduce: 0 = chosen by size (one-shot below 8 MB or fewer
 * than 4 ranks, else reduce-scatter + all-gather), 1 = always one-shot, 2 = always two-phase.
 * Process-wide; for tests and benchmarks. */
void m3r_debug_set_exchange_mode(int mode);

/* ---------------------------------------------------------------------------------------
 * Single-process multi-GPU Gauss-Newton (new).  The reference's caller
 * (global_opt.py:142-143, one backend process, main.py:268) hands over tensors on ONE device;
 * this handle lets that unchanged call use several devices of the box.
 *
 *   devices[0] is the device the caller's tensors live on.  The other devices read those
 *   tensors in place over NVLink peer access during their per-call preparation (packing Xs / Cs
 *   and compacting their shard's rows of idx_ii2jj / valid_match / Q into local memory: that
 *   pass is the scatter, nothing is staged), then run every iteration from local memory:
 *   linearise the shard -> m3r_gn_comm exchange (rank order, bitwise identical sums) ->
 *   replicated solve and retraction.  Rank 0 updates the caller's Twc in place.
 *   A device may be listed more than once (loop-back ranks on one GPU, each with its own
 *   workspace, exchange area and stream): the 1-GPU tests run the protocol that way.
 *
 * m3r_gn_multi_solve: `prob` as for m3r_gn_begin with every pointer on devices[0] and the WHOLE
 * edge arrays (edge_begin / edge_end are ignored); `stream` is a stream of devices[0].  The other
 * devices' streams are ordered after the work already enqueued on `stream`, and `stream` is
 * ordered after their completion, so the caller keeps ordinary stream semantics.  Never blocks
 * the host (except when the graph outgrows the exchange areas, which are then re-allocated).
 * ------------------------------------------------------------------------------------- */
typedef struct m3r_gn_multi m3r_gn_multi;
int m3r_gn_multi_create(m3r_gn_multi **out, const int *devices, int num_devices);
void m3r_gn_multi_destroy(m3r_gn_multi *multi);
int m3r_gn_multi_solve(m3r_gn_multi *multi, const m3r_gn_problem *prob, float *dx, int max_iter,
                       float delta_thresh, void *stream);
/* After a solve (synchronises every rank; tests and benchmarks): iterations executed, the time
 * each rank's preparation took in ms (begin_ms[num_devices]; for ranks > 0 this is the scatter:
 * remote reads + packing + compaction), and whether every rank's pose replica equals `Twc`
 * (device pointer on devices[0]) bit for bit.  Any of the outputs may be null. */
int m3r_gn_multi_report(m3r_gn_multi *multi, const float *Twc, int num_poses, float *begin_ms,
                        int *iters, int *replicas_equal);

#ifdef __cplusplus
}
#endif
#endif /* M3R_BACKENDS_H */
