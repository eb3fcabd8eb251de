/*
 * m3r_oracle.h -- CPU restatement of the reference `mast3r_slam_backends` hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (lightweight-mast3r-slam_b200/)
 * may include, link or call this.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker / the timed CPU arm.
 *
 * Every function cites the reference lines it restates (paths relative to
 * /root/reference/mast3r_slam/backend/src/).  The reference has no CPU implementation
 * (setup.py:47-48) and no tests or golden vectors for this path, so
 * the oracle is pinned against outputs of the reference's own CUDA kernels compiled for
 * sm_100 (oracle/build_ref.py -> oracle/_ref/) and run on the B200 box; the vectors are
 * committed under tests/golden/ with the script that made them (tests/golden/make_golden.py).
 *
 * Arithmetic notes kept from the reference:
 *  - double-promoted sub-expressions (the `1.0`, `0.1`, `10.0`, `1.345` literals) are
 *    restated literally, so C's usual arithmetic conversions reproduce them;
 *  - refine_matches accumulates in the tensor dtype; for half every `*` and `+=` rounds to
 *    binary16 through float exactly as c10::Half does (torch/headeronly/util/Half.h);
 *  - compiled with -ffp-contract=off: each operation rounds once (nvcc contracts some
 *    mul+add pairs into FMA in the reference build; that is a <=1 ulp per-op difference
 *    and is covered by the stated tolerances, never by refine_matches which has none).
 */
#ifndef M3R_ORACLE_H
#define M3R_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* number of OpenMP threads the oracle will use (1 when built without OpenMP) */
int m3r_oracle_num_threads(void);
void m3r_oracle_set_num_threads(int n);

/* matching_kernels.cu:119-275 (iter_proj_kernel) + :279-316 (launcher, zero-init outputs) */
void m3r_oracle_iter_proj(const float *rays_img_with_grad, /* [b,h,w,9] */
                          const float *pts_3d_norm,        /* [b,n,3]   */
                          const float *p_init,             /* [b,n,2]   */
                          float *p_new,                    /* [b,n,2]   */
                          uint8_t *converged,              /* [b,n]     */
                          int b, int h, int w, int n, int max_iter, float lambda_init,
                          float cost_thresh);

/* matching_kernels.cu:25-81 (refine_matches_kernel<scalar_t>), dtype: 0=f16 1=f32 2=f64 */
void m3r_oracle_refine_matches(const void *D11, /* [b,h,w,fdim] */
                               const void *D21, /* [b,n,fdim]   */
                               const int64_t *p1, /* [b,n,2] */
                               int64_t *p1_new,   /* [b,n,2] */
                               int b, int h, int w, int n, int fdim, int dtype, int radius,
                               int dilation_max);

/* binary16 helpers (c10::Half semantics: round-to-nearest-even through float) */
uint16_t m3r_oracle_f2h(float f);
float m3r_oracle_h2f(uint16_t h);

/* gn_kernels.cu:161-170 (get_unique_kf_idx + create_inds): sorted unique of cat(ii,jj),
 * searchsorted -> packed index in [0,N).  Returns the number of unique ids. */
int m3r_oracle_edge_indices(const int64_t *ii, const int64_t *jj, int num_edges,
                            int32_t *ii_edge, int32_t *jj_edge);

/* kind: 0 = rays (ray_align_kernel :813-1138), 1 = calib (calib_proj_kernel :1231-1543),
 *       2 = points (point_align_kernel :455-723). */
typedef struct {
  int kind;
  /* rays */
  float sigma_ray, sigma_dist;
  /* calib */
  float fx, fy, cx, cy;
  int height, width, pixel_border;
  float z_eps, sigma_pixel, sigma_depth;
  /* points */
  float sigma_point;
  /* common */
  float C_thresh, Q_thresh;
} m3r_oracle_gn_params;

/* One launch of the accumulation kernel: per directed edge the reference's block of 256
 * threads (thread t visits points t, t+256, ...; gn_kernels.cu:31-32) followed by the
 * shared-memory tree of blockReduce/warpReduce (:36-55), restated in that exact summation
 * order.  Hs [4,E,7,7] (ii,ij,ji,jj) and gs [2,E,7] as the kernel stores them (:1097-1137).
 * ii_edge/jj_edge are packed pose indices. */
void m3r_oracle_gn_accumulate(const m3r_oracle_gn_params *p, const float *Twc, /* [N,8] */
                              const float *Xs,                                 /* [N,n,3] */
                              const float *Cs,                                 /* [N,n]   */
                              const int32_t *ii_edge, const int32_t *jj_edge,
                              const int64_t *idx_ii2jj,   /* [E,n] */
                              const uint8_t *valid_match, /* [E,n] */
                              const float *Q,             /* [E,n] */
                              int num_poses, int num_points, int num_edges, float *Hs,
                              float *gs);

/* SparseBlock::update_lhs/update_rhs/solve (gn_kernels.cu:57-159) with num_fix = 1:
 * assemble the (N-1)*7 system in double from the float blocks, Cholesky, x.  A dense LLT
 * replaces Eigen's SimplicialLLT (same factorisation up to the fill-reducing permutation;
 * both in double).  Returns 0 on success and fills x [(N-1)*7] doubles; returns 1 on
 * factorisation failure (the reference then uses dx = 0, :147-150). */
int m3r_oracle_gn_solve(const float *Hs, const float *gs, const int32_t *ii_edge,
                        const int32_t *jj_edge, int num_poses, int num_edges, double *x);

/* pose_retr_kernel (:415-453) with retrSim3/expSim3/expSO3 (:299-413): in-place left
 * retraction of poses k >= num_fix by dx[k-num_fix]. */
void m3r_oracle_pose_retr(float *Twc, const float *dx, int num_poses, int num_fix);

/* The driver loops gauss_newton_{rays,calib,points}_cuda (:1140-1228, :1546-1638,
 * :725-811).  ii/jj are the caller's global keyframe ids.  Twc is updated in place; dx
 * [(N-1),7] receives the last executed iteration's step.  Returns the number of
 * iterations executed. */
int m3r_oracle_gauss_newton(const m3r_oracle_gn_params *p, float *Twc, const float *Xs,
                            const float *Cs, const int64_t *ii, const int64_t *jj,
                            const int64_t *idx_ii2jj, const uint8_t *valid_match,
                            const float *Q, int num_poses, int num_points, int num_edges,
                            int max_iter, float delta_thresh, float *dx);

#ifdef __cplusplus
}
#endif
#endif
