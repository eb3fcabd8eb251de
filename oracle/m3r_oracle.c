/*
 * m3r_oracle.c -- CPU restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY
 * (see m3r_oracle.h).  Build: `make -C oracle` -> oracle/libm3r_oracle.so
 * (gcc -O2 -ffp-contract=off -fopenmp).
 *
 * Reference files restated (relative to /root/reference/mast3r_slam/backend/src/):
 *   matching_kernels.cu  iter_proj_kernel :119-275, refine_matches_kernel :25-81
 *   gn_kernels.cu        Sim(3) helpers :172-413, pose_retr_kernel :415-453,
 *                        point_align_kernel :455-723, ray_align_kernel :813-1138,
 *                        calib_proj_kernel :1231-1543, SparseBlock :57-159,
 *                        get_unique_kf_idx/create_inds :161-170, driver loops :725-811,
 *                        :1140-1228, :1546-1638
 */
#include "m3r_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#if defined(__x86_64__)
#include <immintrin.h>
#endif

/* ------------------------------------------------------------------------------------ */
/* threads                                                                              */
/* ------------------------------------------------------------------------------------ */
int m3r_oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void m3r_oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* ------------------------------------------------------------------------------------ */
/* binary16 <-> binary32, round-to-nearest-even (what c10::Half / __float2half do)        */
/* ------------------------------------------------------------------------------------ */
static inline float h2f_sw(uint16_t h) {
  uint32_t sgn = (uint32_t)(h & 0x8000u) << 16;
  uint32_t em = h & 0x7fffu;
  uint32_t x;
  float f;
  if (em >= 0x7c00u) {
    x = sgn | 0x7f800000u | ((em & 0x3ffu) << 13);
  } else if (em >= 0x0400u) {
    x = sgn | ((em << 13) + 0x38000000u);
  } else {
    float v = (float)em * 5.9604644775390625e-8f; /* em * 2^-24, exact */
    memcpy(&x, &v, 4);
    x |= sgn;
  }
  memcpy(&f, &x, 4);
  return f;
}

static inline uint16_t f2h_sw(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  uint32_t sgn = (x >> 16) & 0x8000u;
  uint32_t a = x & 0x7fffffffu;
  if (a > 0x7f800000u) return (uint16_t)(sgn | 0x7e00u); /* NaN */
  if (a >= 0x47800000u) return (uint16_t)(sgn | 0x7c00u); /* >= 65536 or inf */
  if (a < 0x38800000u) {
    /* |f| < 2^-14: result is m * 2^-24; adding 0.5f rounds (RNE) to a multiple of 2^-24 */
    float af, r;
    uint32_t rb;
    memcpy(&af, &a, 4);
    r = af + 0.5f;
    memcpy(&rb, &r, 4);
    return (uint16_t)(sgn | (rb - 0x3f000000u));
  }
  {
    uint32_t odd = (a >> 13) & 1u;
    a += 0xC8000000u + 0xfffu + odd; /* rebias exponent by -112, round half to even */
    return (uint16_t)(sgn | (a >> 13));
  }
}

uint16_t m3r_oracle_f2h(float f) { return f2h_sw(f); }
float m3r_oracle_h2f(uint16_t h) { return h2f_sw(h); }

/* ------------------------------------------------------------------------------------ */
/* iter_proj  (matching_kernels.cu:119-275)                                              */
/* ------------------------------------------------------------------------------------ */
static inline void clampf(float *x, const float lo, const float hi) {
  /* matching_kernels.cu:21-23 */
  *x = fminf(fmaxf(*x, lo), hi);
}

/* Rounding.  The reference is compiled by nvcc with its default -fmad=true, so which products
 * are rounded on their own and which are fused into the following add is part of what its
 * binary computes.  The sequence below is the one nvcc 12.9 emits for these expressions
 * (read off the FFMA / FMUL / FADD order of the compiled kernel, oracle/_ref, and of this
 * repo's own kernel, which agree); this file is built with -ffp-contract=off, so every fusion
 * is an explicit fmaf and nothing else is contracted:
 *   a0*b0 + a1*b1 + a2*b2         ->  fma(a2,b2, fma(a0,b0, a1*b1))     (dot3)
 *   w11*x11 + w12*x12 + ...       ->  fma(w22,x22, fma(w21,x21, fma(w11,x11, w12*x12)))
 *   r[j]*inv - pt[j]              ->  fma(r[j], inv, -pt[j])            (r[j] *= inv is not rounded)
 *   A00*A11 - A01*A01             ->  fma(A00,A11, -(A01*A01))
 *   A11*b0 - A01*b1               ->  fma(-A01,b1, A11*b0)
 *   -A01*b0 + A00*b1              ->  fma(-A01,b0, A00*b1)
 *   u + det_inv*(...)             ->  fma(..., det_inv, u)
 * The double-promoted pieces (`1.0 - du`, `1.0 / x`, `lambda *= 0.1`) are kept as written;
 * du is a multiple of 2^-23, so (1.0 - du)*dv rounds once either way.  On smooth scenes the
 * contraction is invisible (match_a/b.npz were reproduced bit for bit without it); on chaotic
 * inputs (tests/golden/mast3r_rand.npz: 4 % of the iterations converge) an uncontracted build
 * leaves the reference's orbit on 0.4 % of the pixels, this one on none
 * (tests/test_matching_gpu.py::test_oracle_twin_of_reference_cuda). */
static inline float dot3c(const float *a, const float *b) {
  return fmaf(a[2], b[2], fmaf(a[0], b[0], a[1] * b[1]));
}

static inline float bilerp(float w11, float w12, float w21, float w22, float x11, float x12,
                           float x21, float x22) {
  return fmaf(w22, x22, fmaf(w21, x21, fmaf(w11, x11, w12 * x12)));
}

static void iter_proj_point(const float *img /* [h,w,9] of batch b */, const float *pt,
                            const float *pinit, float *pout, uint8_t *conv, int h, int w,
                            int max_iter, float lambda_init, float cost_thresh) {
  const int c = 9;
  float u = pinit[0];
  float v = pinit[1];
  clampf(&u, 1, w - 2); /* :143-144 */
  clampf(&v, 1, h - 2);

  float r[3], gx[3], gy[3], err[3];
  uint8_t cv = 0; /* converged is torch::zeros-initialised (:299-301) */
  float lambda = lambda_init;
  for (int i = 0; i < max_iter; i++) {
    /* :155-164 */
    int u11 = (int)floorf(u);
    int v11 = (int)floorf(v);
    float du = u - (float)u11;
    float dv = v - (float)v11;
    float w11 = du * dv;
    float w12 = (1.0 - du) * dv;
    float w21 = du * (1.0 - dv);
    float w22 = (1.0 - du) * (1.0 - dv);
    /* :167-170 */
    const float *r11 = img + ((size_t)(v11 + 1) * w + (u11 + 1)) * c;
    const float *r12 = img + ((size_t)(v11 + 1) * w + u11) * c;
    const float *r21 = img + ((size_t)v11 * w + (u11 + 1)) * c;
    const float *r22 = img + ((size_t)v11 * w + u11) * c;
    for (int j = 0; j < 3; j++) r[j] = bilerp(w11, w12, w21, w22, r11[j], r12[j], r21[j], r22[j]);
    for (int j = 3; j < 6; j++) gx[j - 3] = bilerp(w11, w12, w21, w22, r11[j], r12[j], r21[j], r22[j]);
    for (int j = 6; j < 9; j++) gy[j - 6] = bilerp(w11, w12, w21, w22, r11[j], r12[j], r21[j], r22[j]);
    /* :186-191 */
    float r_norm = sqrtf(dot3c(r, r));
    float r_norm_inv = 1.0 / r_norm;
    /* :195-198 */
    for (int j = 0; j < 3; j++) err[j] = fmaf(r[j], r_norm_inv, -pt[j]);
    float cost = dot3c(err, err);
    /* :202-210 */
    float A00 = dot3c(gx, gx);
    float A01 = dot3c(gx, gy);
    float A11 = dot3c(gy, gy);
    float b0 = -dot3c(err, gx);
    float b1 = -dot3c(err, gy);
    A00 += lambda;
    A11 += lambda;
    /* :213-215 */
    float det_inv = 1.0 / fmaf(A00, A11, -(A01 * A01));
    float num_u = fmaf(-A01, b1, A11 * b0);
    float num_v = fmaf(-A01, b0, A00 * b1);
    /* :218-221 */
    float u_new = fmaf(num_u, det_inv, u);
    float v_new = fmaf(num_v, det_inv, v);
    clampf(&u_new, 1, w - 2);
    clampf(&v_new, 1, h - 2);
    /* :225-233 */
    u11 = (int)floorf(u_new);
    v11 = (int)floorf(v_new);
    du = u_new - u11;
    dv = v_new - v11;
    w11 = du * dv;
    w12 = (1.0 - du) * dv;
    w21 = du * (1.0 - dv);
    w22 = (1.0 - du) * (1.0 - dv);
    r11 = img + ((size_t)(v11 + 1) * w + (u11 + 1)) * c;
    r12 = img + ((size_t)(v11 + 1) * w + u11) * c;
    r21 = img + ((size_t)v11 * w + (u11 + 1)) * c;
    r22 = img + ((size_t)v11 * w + u11) * c;
    /* :241-256 */
    for (int j = 0; j < 3; j++) r[j] = bilerp(w11, w12, w21, w22, r11[j], r12[j], r21[j], r22[j]);
    r_norm = sqrtf(dot3c(r, r));
    r_norm_inv = 1.0 / r_norm;
    for (int j = 0; j < 3; j++) err[j] = fmaf(r[j], r_norm_inv, -pt[j]);
    float new_cost = dot3c(err, err);
    /* :259-268 */
    if (new_cost < cost) {
      u = u_new;
      v = v_new;
      lambda *= 0.1;
      cv = new_cost < cost_thresh;
    } else {
      lambda *= 10.0;
      cv = cost < cost_thresh;
    }
  }
  pout[0] = u; /* :272-273 */
  pout[1] = v;
  *conv = cv;
}

void m3r_oracle_iter_proj(const float *rays, const float *pts, const float *p_init,
                          float *p_new, uint8_t *converged, int b, int h, int w, int n,
                          int max_iter, float lambda_init, float cost_thresh) {
  const long total = (long)b * n;
#pragma omp parallel for schedule(static)
  for (long t = 0; t < total; t++) {
    const int bi = (int)(t / n);
    iter_proj_point(rays + (size_t)bi * h * w * 9, pts + (size_t)t * 3, p_init + (size_t)t * 2,
                    p_new + (size_t)t * 2, converged + t, h, w, max_iter, lambda_init,
                    cost_thresh);
  }
}

/* ------------------------------------------------------------------------------------ */
/* refine_matches  (matching_kernels.cu:25-81)                                           */
/* ------------------------------------------------------------------------------------ */
static inline int inside_image(long u, long v, int W, int H) {
  return v >= 0 && v < H && u >= 0 && u < W; /* :17-19 */
}

/* Scan order, recentring and strict `>` follow :51-77.  SCORE_T / MUL_ADD differ per dtype. */
#define REFINE_BODY(LOAD_INIT, SCORE_DECL, SCORE_ZERO, SCORE_ACC, SCORE_GT, SCORE_SET)      \
  for (long t = 0; t < total; t++) {                                                        \
    const int bi = (int)(t / n);                                                            \
    long u0 = p1[t * 2 + 0], v0 = p1[t * 2 + 1];                                            \
    long u_new = u0, v_new = v0;                                                            \
    LOAD_INIT;                                                                              \
    for (int d = dilation_max; d > 0; d--) {                                                \
      const int rd = radius * d;                                                            \
      const int diam = 2 * rd + 1;                                                          \
      for (int i = 0; i < diam; i += d) {                                                   \
        for (int j = 0; j < diam; j += d) {                                                 \
          const long u = u0 - rd + i;                                                       \
          const long v = v0 - rd + j;                                                       \
          if (inside_image(u, v, w, h)) {                                                   \
            const size_t row = (((size_t)bi * h + (size_t)v) * w + (size_t)u) * fdim;       \
            SCORE_DECL;                                                                     \
            SCORE_ZERO;                                                                     \
            for (int k = 0; k < fdim; k++) { SCORE_ACC; }                                   \
            if (SCORE_GT) {                                                                 \
              SCORE_SET;                                                                    \
              u_new = u;                                                                    \
              v_new = v;                                                                    \
            }                                                                               \
          }                                                                                 \
        }                                                                                   \
      }                                                                                     \
      u0 = u_new;                                                                           \
      v0 = v_new;                                                                           \
    }                                                                                       \
    p1_new[t * 2 + 0] = u_new;                                                              \
    p1_new[t * 2 + 1] = v_new;                                                              \
  }

/* half: score is a c10::Half; `D21*D11` rounds to half, `score +=` rounds to half, both
 * through float (Half.h operator*, operator+).  max_score starts at
 * ::cuda::std::numeric_limits<c10::Half>::min() (:47): libcu++ has no specialisation for
 * c10::Half, so the primary template returns c10::Half() = +0. */
static void refine_f16_sw(const uint16_t *D11, const uint16_t *D21, const int64_t *p1,
                          int64_t *p1_new, int b, int h, int w, int n, int fdim, int radius,
                          int dilation_max) {
  const long total = (long)b * n;
#pragma omp parallel for schedule(static)
  REFINE_BODY(float max_score = 0.0f; const uint16_t *q = D21 + (size_t)t * fdim, float score,
              score = 0.0f,
              {
                float prod = h2f_sw(f2h_sw(h2f_sw(q[k]) * h2f_sw(D11[row + k])));
                score = h2f_sw(f2h_sw(score + prod));
              },
              score > max_score, max_score = score)
}

#if defined(__x86_64__)
__attribute__((target("f16c"))) static inline float rh_hw(float f) {
  return _cvtsh_ss(_cvtss_sh(f, _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC));
}
__attribute__((target("f16c"))) static void refine_f16_hw(const uint16_t *D11,
                                                          const uint16_t *D21,
                                                          const int64_t *p1, int64_t *p1_new,
                                                          int b, int h, int w, int n, int fdim,
                                                          int radius, int dilation_max) {
  const long total = (long)b * n;
#pragma omp parallel for schedule(static)
  REFINE_BODY(float max_score = 0.0f; const uint16_t *q = D21 + (size_t)t * fdim, float score,
              score = 0.0f,
              {
                float prod = rh_hw(_cvtsh_ss(q[k]) * _cvtsh_ss(D11[row + k]));
                score = rh_hw(score + prod);
              },
              score > max_score, max_score = score)
}
#endif

/* float / double: plain accumulation in scalar_t; max_score = numeric_limits<T>::min()
 * = smallest positive normal (FLT_MIN / DBL_MIN). */
static void refine_f32(const float *D11, const float *D21, const int64_t *p1, int64_t *p1_new,
                       int b, int h, int w, int n, int fdim, int radius, int dilation_max) {
  const long total = (long)b * n;
#pragma omp parallel for schedule(static)
  REFINE_BODY(float max_score = 1.17549435e-38f; const float *q = D21 + (size_t)t * fdim,
              float score, score = 0.0f, score += q[k] * D11[row + k], score > max_score,
              max_score = score)
}
static void refine_f64(const double *D11, const double *D21, const int64_t *p1,
                       int64_t *p1_new, int b, int h, int w, int n, int fdim, int radius,
                       int dilation_max) {
  const long total = (long)b * n;
#pragma omp parallel for schedule(static)
  REFINE_BODY(double max_score = 2.2250738585072014e-308;
              const double *q = D21 + (size_t)t * fdim, double score, score = 0.0,
              score += q[k] * D11[row + k], score > max_score, max_score = score)
}

static int g_force_sw_half = 0;
void m3r_oracle_force_sw_half(int on) { g_force_sw_half = on; }

void m3r_oracle_refine_matches(const void *D11, const void *D21, const int64_t *p1,
                               int64_t *p1_new, int b, int h, int w, int n, int fdim,
                               int dtype, int radius, int dilation_max) {
  if (dtype == 0) {
#if defined(__x86_64__)
    if (!g_force_sw_half && __builtin_cpu_supports("f16c")) {
      refine_f16_hw((const uint16_t *)D11, (const uint16_t *)D21, p1, p1_new, b, h, w, n, fdim,
                    radius, dilation_max);
      return;
    }
#endif
    refine_f16_sw((const uint16_t *)D11, (const uint16_t *)D21, p1, p1_new, b, h, w, n, fdim,
                  radius, dilation_max);
  } else if (dtype == 1) {
    refine_f32((const float *)D11, (const float *)D21, p1, p1_new, b, h, w, n, fdim, radius,
               dilation_max);
  } else {
    refine_f64((const double *)D11, (const double *)D21, p1, p1_new, b, h, w, n, fdim, radius,
               dilation_max);
  }
}

/* ------------------------------------------------------------------------------------ */
/* Sim(3) helpers  (gn_kernels.cu:172-413)                                               */
/* ------------------------------------------------------------------------------------ */
#define EPS 1e-6 /* :34 (a double literal in the reference too) */

static inline float huber(float r) { /* :172-175 */
  const float r_abs = fabsf(r);
  return r_abs < 1.345 ? 1.0 : 1.345 / r_abs;
}

static inline void quat_comp(const float *qi, const float *qj, float *out) { /* :178-184 */
  out[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  out[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  out[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  out[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}

static inline void quat_inv(const float *q, float *out) { /* :187-193 */
  out[0] = -q[0];
  out[1] = -q[1];
  out[2] = -q[2];
  out[3] = q[3];
}

static inline void actSO3(const float *q, const float *X, float *Y) { /* :195-205 */
  float uv[3];
  uv[0] = 2.0 * (q[1] * X[2] - q[2] * X[1]);
  uv[1] = 2.0 * (q[2] * X[0] - q[0] * X[2]);
  uv[2] = 2.0 * (q[0] * X[1] - q[1] * X[0]);
  /* Y may alias X (relSim3 calls actSO3(qi_inv, tij, tij)); each Y[k] reads only X[k] */
  Y[0] = X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  Y[1] = X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  Y[2] = X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
}

static inline void actSim3(const float *t, const float *q, const float *s, const float *X,
                           float *Y) { /* :207-219 */
  actSO3(q, X, Y);
  Y[0] *= s[0];
  Y[1] *= s[0];
  Y[2] *= s[0];
  Y[0] += t[0];
  Y[1] += t[1];
  Y[2] += t[2];
}

static inline void scale_vec3_inplace(float *t, float s) { /* :222-227 */
  t[0] *= s;
  t[1] *= s;
  t[2] *= s;
}

static inline void crossInplace(const float *a, float *b) { /* :229-240 */
  float x[3] = {
      a[1] * b[2] - a[2] * b[1],
      a[2] * b[0] - a[0] * b[2],
      a[0] * b[1] - a[1] * b[0],
  };
  b[0] = x[0];
  b[1] = x[1];
  b[2] = x[2];
}

static inline float dot3(const float *t, const float *s) { /* :242-245 */
  return t[0] * s[0] + t[1] * s[1] + t[2] * s[2];
}

static inline float squared_norm3(const float *v) { /* :247-250 */
  return v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
}

static void relSim3(const float *ti, const float *qi, const float *si, const float *tj,
                    const float *qj, const float *sj, float *tij, float *qij,
                    float *sij) { /* :252-272 */
  float si_inv = 1.0 / si[0];
  sij[0] = si_inv * sj[0];
  float qi_inv[4];
  quat_inv(qi, qi_inv);
  quat_comp(qi_inv, qj, qij);
  tij[0] = tj[0] - ti[0];
  tij[1] = tj[1] - ti[1];
  tij[2] = tj[2] - ti[2];
  actSO3(qi_inv, tij, tij);
  scale_vec3_inplace(tij, si_inv);
}

static inline void apply_Sim3_adj_inv(const float *t, const float *q, const float *s,
                                      const float *X, float *Y) { /* :277-297 */
  const float s_inv = 1.0 / s[0];
  float Ra[3];
  actSO3(q, &X[0], Ra);
  Y[0] = s_inv * Ra[0];
  Y[1] = s_inv * Ra[1];
  Y[2] = s_inv * Ra[2];
  actSO3(q, &X[3], &Y[3]);
  Y[3] += s_inv * (t[1] * Ra[2] - t[2] * Ra[1]);
  Y[4] += s_inv * (t[2] * Ra[0] - t[0] * Ra[2]);
  Y[5] += s_inv * (t[0] * Ra[1] - t[1] * Ra[0]);
  Y[6] = X[6] + (s_inv * dot3(t, Ra));
}

static void expSO3(const float *phi, float *q) { /* :299-321 */
  float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float imag, real;
  if (theta_sq < EPS) {
    float theta_p4 = theta_sq * theta_sq;
    imag = 0.5 - (1.0 / 48.0) * theta_sq + (1.0 / 3840.0) * theta_p4;
    real = 1.0 - (1.0 / 8.0) * theta_sq + (1.0 / 384.0) * theta_p4;
  } else {
    float theta = sqrtf(theta_sq);
    imag = sinf(0.5 * theta) / theta;
    real = cosf(0.5 * theta);
  }
  q[0] = imag * phi[0];
  q[1] = imag * phi[1];
  q[2] = imag * phi[2];
  q[3] = real;
}

static void expSim3(const float *xi, float *t, float *q, float *s) { /* :323-390 */
  float tau[3] = {xi[0], xi[1], xi[2]};
  float phi[3] = {xi[3], xi[4], xi[5]};
  float sigma = xi[6];
  float scale = expf(sigma);
  expSO3(phi, q);
  s[0] = scale;
  float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float theta = sqrtf(theta_sq);
  float A, B, C;
  const float one = 1.0;
  const float half = 0.5;
  if (fabsf(sigma) < EPS) {
    C = one;
    if (fabsf(theta) < EPS) {
      A = half;
      B = 1.0 / 6.0;
    } else {
      A = (one - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - one) / sigma;
    if (fabsf(theta) < EPS) {
      float sigma_sq = sigma * sigma;
      A = ((sigma - one) * scale + one) / sigma_sq;
      B = (scale * half * sigma_sq + scale - one - sigma * scale) / (sigma_sq * sigma);
    } else {
      float a = scale * sinf(theta);
      float b = scale * cosf(theta);
      float c = theta_sq + sigma * sigma;
      A = (a * sigma + (one - b) * theta) / (theta * c);
      B = (C - ((b - one) * sigma + a * theta) / (c)) / (theta_sq);
    }
  }
  t[0] = C * tau[0];
  t[1] = C * tau[1];
  t[2] = C * tau[2];
  crossInplace(phi, tau);
  t[0] += A * tau[0];
  t[1] += A * tau[1];
  t[2] += A * tau[2];
  crossInplace(phi, tau);
  t[0] += B * tau[0];
  t[1] += B * tau[1];
  t[2] += B * tau[2];
}

static void retrSim3(const float *xi, const float *t, const float *q, const float *s, float *t1,
                     float *q1, float *s1) { /* :392-413 */
  float dt[3] = {0, 0, 0};
  float dq[4] = {0, 0, 0, 1};
  float ds[1] = {0};
  expSim3(xi, dt, dq, ds);
  quat_comp(dq, q, q1);
  actSO3(dq, t, t1);
  scale_vec3_inplace(t1, ds[0]);
  t1[0] += dt[0];
  t1[1] += dt[1];
  t1[2] += dt[2];
  s1[0] = ds[0] * s[0];
}

void m3r_oracle_pose_retr(float *Twc, const float *dx, int num_poses, int num_fix) {
  /* pose_retr_kernel :415-453 */
  for (int k = num_fix; k < num_poses; k++) {
    float xi[7], q[4], q1[4], t[3], t1[3], s[1], s1[1];
    float *P = Twc + (size_t)k * 8;
    t[0] = P[0], t[1] = P[1], t[2] = P[2];
    q[0] = P[3], q[1] = P[4], q[2] = P[5], q[3] = P[6];
    s[0] = P[7];
    for (int n = 0; n < 7; n++) xi[n] = dx[(size_t)(k - num_fix) * 7 + n];
    retrSim3(xi, t, q, s, t1, q1, s1);
    P[0] = t1[0], P[1] = t1[1], P[2] = t1[2];
    P[3] = q1[0], P[4] = q1[1], P[5] = q1[2], P[6] = q1[3];
    P[7] = s1[0];
  }
}

/* ------------------------------------------------------------------------------------ */
/* index remap  (gn_kernels.cu:161-170)                                                  */
/* ------------------------------------------------------------------------------------ */
static int cmp_i64(const void *a, const void *b) {
  int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
  return (x > y) - (x < y);
}

int m3r_oracle_edge_indices(const int64_t *ii, const int64_t *jj, int E, int32_t *ii_edge,
                            int32_t *jj_edge) {
  int64_t *all = (int64_t *)malloc(sizeof(int64_t) * 2 * (size_t)(E > 0 ? E : 1));
  for (int e = 0; e < E; e++) {
    all[e] = ii[e];
    all[E + e] = jj[e];
  }
  qsort(all, 2 * (size_t)E, sizeof(int64_t), cmp_i64);
  int nu = 0;
  for (int k = 0; k < 2 * E; k++)
    if (k == 0 || all[k] != all[k - 1]) all[nu++] = all[k];
  for (int e = 0; e < E; e++) {
    /* searchsorted(left) on exact members = position in the unique list */
    int64_t *pi = (int64_t *)bsearch(&ii[e], all, nu, sizeof(int64_t), cmp_i64);
    int64_t *pj = (int64_t *)bsearch(&jj[e], all, nu, sizeof(int64_t), cmp_i64);
    ii_edge[e] = (int32_t)(pi - all);
    jj_edge[e] = (int32_t)(pj - all);
  }
  free(all);
  return nu;
}

/* ------------------------------------------------------------------------------------ */
/* accumulation kernels                                                                  */
/* ------------------------------------------------------------------------------------ */
#define REF_THREADS 256 /* gn_kernels.cu:28 */

/* One residual row: Jj = apply_Sim3_adj_inv(T_i, Ji_local); Ji = -Jj (:999-1000 etc.);
 * hij[l] += w*Jx[n]*Jx[m]; vi += w*err*Ji; vj += w*err*Jj (:1002-1013 etc.).
 * Because Ji = -Jj exactly, (w*Jx[n])*Jx[m] over the 14x14 packed triangle is, bit for bit,
 * +/-G[a][c] with G[a][c] = sum (w*Jj[a])*Jj[c] over ALL 49 (a,c): the ii and jj blocks use
 * G[a][c] for a >= c, the ij/ji blocks use -G[a][c] for every (a,c).  So 49 + 7 accumulators
 * per virtual thread reproduce the reference's 105 + 14 exactly. */
static inline void accum_row(const float *ti, const float *qi, const float *si,
                             const float *Jloc, float w, float err, float *G, float *vj) {
  float Jj[7];
  apply_Sim3_adj_inv(ti, qi, si, Jloc, Jj);
  for (int a = 0; a < 7; a++) {
    const float wa = w * Jj[a];
    for (int c = 0; c < 7; c++) G[a * 7 + c] += wa * Jj[c];
  }
  const float we = w * err;
  for (int a = 0; a < 7; a++) vj[a] += we * Jj[a];
}

static void edge_accumulate(const m3r_oracle_gn_params *p, const float *Twc, const float *Xs,
                            const float *Cs, int ix, int jx, const int64_t *idx,
                            const uint8_t *vm, const float *Qe, int n, float *Hs_ii,
                            float *Hs_ij, float *Hs_ji, float *Hs_jj, float *gs_i, float *gs_j,
                            float *acc /* scratch [256*56] */) {
  float ti[3], tj[3], tij[3], qi[4], qj[4], qij[4], si[1], sj[1], sij[1];
  for (int k = 0; k < 3; k++) ti[k] = Twc[ix * 8 + k], tj[k] = Twc[jx * 8 + k];
  for (int k = 0; k < 4; k++) qi[k] = Twc[ix * 8 + 3 + k], qj[k] = Twc[jx * 8 + 3 + k];
  si[0] = Twc[ix * 8 + 7], sj[0] = Twc[jx * 8 + 7];
  relSim3(ti, qi, si, tj, qj, sj, tij, qij, sij);

  memset(acc, 0, sizeof(float) * REF_THREADS * 56);
  const float *Xi_base = Xs + (size_t)ix * n * 3;
  const float *Xj_base = Xs + (size_t)jx * n * 3;
  const float *Ci_base = Cs + (size_t)ix * n;
  const float *Cj_base = Cs + (size_t)jx * n;

  const float sigma_ray_inv = 1.0 / p->sigma_ray;
  const float sigma_dist_inv = 1.0 / p->sigma_dist;
  const float sigma_pixel_inv = 1.0 / p->sigma_pixel;
  const float sigma_depth_inv = 1.0 / p->sigma_depth;
  const float sigma_point_inv = 1.0 / p->sigma_point;
  const float fx = p->fx, fy = p->fy, cx = p->cx, cy = p->cy;
  const int width = p->width, height = p->height, pixel_border = p->pixel_border;

  for (int k = 0; k < n; k++) {
    float *G = acc + (size_t)(k % REF_THREADS) * 56; /* virtual thread of GPU_1D_KERNEL_LOOP */
    float *vj = G + 49;
    float Xi[3], Xj[3], Xj_Ci[3], Ji[7];

    const int valid_match_ind = vm[k] != 0;
    const int64_t ind_Xi = valid_match_ind ? idx[k] : 0;
    Xi[0] = Xi_base[ind_Xi * 3 + 0], Xi[1] = Xi_base[ind_Xi * 3 + 1],
    Xi[2] = Xi_base[ind_Xi * 3 + 2];
    Xj[0] = Xj_base[(size_t)k * 3 + 0], Xj[1] = Xj_base[(size_t)k * 3 + 1],
    Xj[2] = Xj_base[(size_t)k * 3 + 2];
    const float q = Qe[k];
    const float ci = Ci_base[ind_Xi];
    const float cj = Cj_base[k];

    if (p->kind == 0) { /* ray_align_kernel :910-1090 */
      float err[4], w[4];
      const float norm2_i = squared_norm3(Xi);
      const float norm1_i = sqrtf(norm2_i);
      const float norm1_i_inv = 1.0 / norm1_i;
      float ri[3];
      for (int i = 0; i < 3; i++) ri[i] = norm1_i_inv * Xi[i];
      actSim3(tij, qij, sij, Xj, Xj_Ci);
      const float norm2_j = squared_norm3(Xj_Ci);
      const float norm1_j = sqrtf(norm2_j);
      const float norm1_j_inv = 1.0 / norm1_j;
      float rj_Ci[3];
      for (int i = 0; i < 3; i++) rj_Ci[i] = norm1_j_inv * Xj_Ci[i];
      err[0] = rj_Ci[0] - ri[0];
      err[1] = rj_Ci[1] - ri[1];
      err[2] = rj_Ci[2] - ri[2];
      err[3] = norm1_j - norm1_i;
      const int valid =
          valid_match_ind & (q > p->Q_thresh) & (ci > p->C_thresh) & (cj > p->C_thresh);
      const float conf_weight = q;
      const float sqrt_w_ray = valid ? sigma_ray_inv * sqrtf(conf_weight) : 0;
      const float sqrt_w_dist = valid ? sigma_dist_inv * sqrtf(conf_weight) : 0;
      w[0] = huber(sqrt_w_ray * err[0]);
      w[1] = huber(sqrt_w_ray * err[1]);
      w[2] = huber(sqrt_w_ray * err[2]);
      w[3] = huber(sqrt_w_dist * err[3]);
      const float w_const_ray = sqrt_w_ray * sqrt_w_ray;
      const float w_const_dist = sqrt_w_dist * sqrt_w_dist;
      w[0] *= w_const_ray;
      w[1] *= w_const_ray;
      w[2] *= w_const_ray;
      w[3] *= w_const_dist;
      const float norm3_j_inv = norm1_j_inv / norm2_j;
      const float drx_dPx = norm1_j_inv - Xj_Ci[0] * Xj_Ci[0] * norm3_j_inv;
      const float dry_dPy = norm1_j_inv - Xj_Ci[1] * Xj_Ci[1] * norm3_j_inv;
      const float drz_dPz = norm1_j_inv - Xj_Ci[2] * Xj_Ci[2] * norm3_j_inv;
      const float drx_dPy = -Xj_Ci[0] * Xj_Ci[1] * norm3_j_inv;
      const float drx_dPz = -Xj_Ci[0] * Xj_Ci[2] * norm3_j_inv;
      const float dry_dPz = -Xj_Ci[1] * Xj_Ci[2] * norm3_j_inv;
      /* rx :991-997 */
      Ji[0] = drx_dPx, Ji[1] = drx_dPy, Ji[2] = drx_dPz, Ji[3] = 0.0, Ji[4] = rj_Ci[2],
      Ji[5] = -rj_Ci[1], Ji[6] = 0.0;
      accum_row(ti, qi, si, Ji, w[0], err[0], G, vj);
      /* ry :1016-1022 */
      Ji[0] = drx_dPy, Ji[1] = dry_dPy, Ji[2] = dry_dPz, Ji[3] = -rj_Ci[2], Ji[4] = 0.0,
      Ji[5] = rj_Ci[0], Ji[6] = 0.0;
      accum_row(ti, qi, si, Ji, w[1], err[1], G, vj);
      /* rz :1041-1047 */
      Ji[0] = drx_dPz, Ji[1] = dry_dPz, Ji[2] = drz_dPz, Ji[3] = rj_Ci[1], Ji[4] = -rj_Ci[0],
      Ji[5] = 0.0, Ji[6] = 0.0;
      accum_row(ti, qi, si, Ji, w[2], err[2], G, vj);
      /* dist :1067-1073 */
      Ji[0] = rj_Ci[0], Ji[1] = rj_Ci[1], Ji[2] = rj_Ci[2], Ji[3] = 0.0, Ji[4] = 0.0,
      Ji[5] = 0.0, Ji[6] = norm1_j;
      accum_row(ti, qi, si, Ji, w[3], err[3], G, vj);
    } else if (p->kind == 1) { /* calib_proj_kernel :1346-1497 */
      float err[3], w[3];
      const int u_target = ind_Xi % width;
      const int v_target = ind_Xi / width;
      actSim3(tij, qij, sij, Xj, Xj_Ci);
      const int valid_z = ((Xj_Ci[2] > p->z_eps) && (Xi[2] > p->z_eps));
      const float zj_inv = valid_z ? 1.0 / Xj_Ci[2] : 0.0;
      const float zj_log = valid_z ? logf(Xj_Ci[2]) : 0.0;
      const float zi_log = valid_z ? logf(Xi[2]) : 0.0;
      const float x_div_z = Xj_Ci[0] * zj_inv;
      const float y_div_z = Xj_Ci[1] * zj_inv;
      const float u = fx * x_div_z + cx;
      const float v = fy * y_div_z + cy;
      const int valid_u = ((u > pixel_border) && (u < width - 1 - pixel_border));
      const int valid_v = ((v > pixel_border) && (v < height - 1 - pixel_border));
      err[0] = u - u_target;
      err[1] = v - v_target;
      err[2] = zj_log - zi_log;
      const int valid = valid_match_ind & (q > p->Q_thresh) & (ci > p->C_thresh) &
                        (cj > p->C_thresh) & valid_u & valid_v & valid_z;
      const float conf_weight = q;
      const float sqrt_w_pixel = valid ? sigma_pixel_inv * sqrtf(conf_weight) : 0;
      const float sqrt_w_depth = valid ? sigma_depth_inv * sqrtf(conf_weight) : 0;
      w[0] = huber(sqrt_w_pixel * err[0]);
      w[1] = huber(sqrt_w_pixel * err[1]);
      w[2] = huber(sqrt_w_depth * err[2]);
      const float w_const_pixel = sqrt_w_pixel * sqrt_w_pixel;
      const float w_const_depth = sqrt_w_depth * sqrt_w_depth;
      w[0] *= w_const_pixel;
      w[1] *= w_const_pixel;
      w[2] *= w_const_depth;
      /* x :1422-1428 */
      Ji[0] = fx * zj_inv, Ji[1] = 0.0, Ji[2] = -fx * x_div_z * zj_inv,
      Ji[3] = -fx * x_div_z * y_div_z, Ji[4] = fx * (1 + x_div_z * x_div_z),
      Ji[5] = -fx * y_div_z, Ji[6] = 0.0;
      accum_row(ti, qi, si, Ji, w[0], err[0], G, vj);
      /* y :1448-1454 */
      Ji[0] = 0.0, Ji[1] = fy * zj_inv, Ji[2] = -fy * y_div_z * zj_inv,
      Ji[3] = -fy * (1 + y_div_z * y_div_z), Ji[4] = fy * x_div_z * y_div_z,
      Ji[5] = fy * x_div_z, Ji[6] = 0.0;
      accum_row(ti, qi, si, Ji, w[1], err[1], G, vj);
      /* z :1473-1479 */
      Ji[0] = 0.0, Ji[1] = 0.0, Ji[2] = zj_inv, Ji[3] = y_div_z, Ji[4] = -x_div_z, Ji[5] = 0.0,
      Ji[6] = 1.0;
      accum_row(ti, qi, si, Ji, w[2], err[2], G, vj);
    } else { /* point_align_kernel :550-677 */
      float err[3], w[3];
      actSim3(tij, qij, sij, Xj, Xj_Ci);
      err[0] = Xj_Ci[0] - Xi[0];
      err[1] = Xj_Ci[1] - Xi[1];
      err[2] = Xj_Ci[2] - Xi[2];
      const int valid =
          valid_match_ind & (q > p->Q_thresh) & (ci > p->C_thresh) & (cj > p->C_thresh);
      const float conf_weight = q;
      const float sqrt_w_point = valid ? sigma_point_inv * sqrtf(conf_weight) : 0;
      w[0] = huber(sqrt_w_point * err[0]);
      w[1] = huber(sqrt_w_point * err[1]);
      w[2] = huber(sqrt_w_point * err[2]);
      const float w_const_point = sqrt_w_point * sqrt_w_point;
      w[0] *= w_const_point;
      w[1] *= w_const_point;
      w[2] *= w_const_point;
      /* x :602-608 */
      Ji[0] = 1.0, Ji[1] = 0.0, Ji[2] = 0.0, Ji[3] = 0.0, Ji[4] = Xj_Ci[2], Ji[5] = -Xj_Ci[1],
      Ji[6] = Xj_Ci[0];
      accum_row(ti, qi, si, Ji, w[0], err[0], G, vj);
      /* y :627-633 */
      Ji[0] = 0.0, Ji[1] = 1.0, Ji[2] = 0.0, Ji[3] = -Xj_Ci[2], Ji[4] = 0, Ji[5] = Xj_Ci[0],
      Ji[6] = Xj_Ci[1];
      accum_row(ti, qi, si, Ji, w[1], err[1], G, vj);
      /* z :652-658 */
      Ji[0] = 0.0, Ji[1] = 0.0, Ji[2] = 1.0, Ji[3] = Xj_Ci[1], Ji[4] = -Xj_Ci[0], Ji[5] = 0,
      Ji[6] = Xj_Ci[2];
      accum_row(ti, qi, si, Ji, w[2], err[2], G, vj);
    }
  }

  /* blockReduce (:45-55) + warpReduce (:36-43): s[t] += s[t+128]; s[t] += s[t+64]; then
   * offsets 32,16,8,4,2,1 on the first warp.  Only s[0] is read afterwards, and at each step
   * it depends on s[0..off) += s[off..2*off), which is what this loop computes. */
  float red[56];
  for (int c = 0; c < 56; c++) {
    float s[REF_THREADS];
    for (int t = 0; t < REF_THREADS; t++) s[t] = acc[(size_t)t * 56 + c];
    for (int off = REF_THREADS / 2; off >= 1; off >>= 1)
      for (int t = 0; t < off; t++) s[t] += s[t + off];
    red[c] = s[0];
  }
  /* stores :1097-1137 */
  for (int a = 0; a < 7; a++) {
    gs_j[a] = red[49 + a];
    gs_i[a] = -red[49 + a];
    for (int c = 0; c < 7; c++) {
      const float Gac = red[a * 7 + c];
      if (a >= c) {
        Hs_jj[a * 7 + c] = Gac;
        Hs_jj[c * 7 + a] = Gac;
        Hs_ii[a * 7 + c] = Gac;
        Hs_ii[c * 7 + a] = Gac;
      }
      Hs_ij[c * 7 + a] = -Gac; /* Hs[1][m][n-7], m=c (Ji index), n-7=a (Jj index) */
      Hs_ji[a * 7 + c] = -Gac; /* Hs[2][n-7][m] */
    }
  }
}

void m3r_oracle_gn_accumulate(const m3r_oracle_gn_params *p, const float *Twc, const float *Xs,
                              const float *Cs, const int32_t *ii_edge, const int32_t *jj_edge,
                              const int64_t *idx_ii2jj, const uint8_t *valid_match,
                              const float *Q, int num_poses, int num_points, int num_edges,
                              float *Hs, float *gs) {
  (void)num_poses;
  const int E = num_edges, n = num_points;
#pragma omp parallel
  {
    float *acc = (float *)malloc(sizeof(float) * REF_THREADS * 56);
#pragma omp for schedule(dynamic, 1)
    for (int e = 0; e < E; e++) {
      edge_accumulate(p, Twc, Xs, Cs, ii_edge[e], jj_edge[e], idx_ii2jj + (size_t)e * n,
                      valid_match + (size_t)e * n, Q + (size_t)e * n, n,
                      Hs + ((size_t)0 * E + e) * 49, Hs + ((size_t)1 * E + e) * 49,
                      Hs + ((size_t)2 * E + e) * 49, Hs + ((size_t)3 * E + e) * 49,
                      gs + ((size_t)0 * E + e) * 7, gs + ((size_t)1 * E + e) * 7, acc);
    }
    free(acc);
  }
}

/* ------------------------------------------------------------------------------------ */
/* assembly + solve  (SparseBlock, gn_kernels.cu:57-159)                                 */
/* ------------------------------------------------------------------------------------ */
int m3r_oracle_gn_solve(const float *Hs, const float *gs, const int32_t *ii_edge,
                        const int32_t *jj_edge, int num_poses, int num_edges, double *x) {
  const int M = 7, num_fix = 1;
  const int NB = num_poses - num_fix;
  const int D = NB * M;
  const int E = num_edges;
  if (D <= 0) return 1;
  double *A = (double *)calloc((size_t)D * D, sizeof(double));
  double *bv = (double *)calloc((size_t)D, sizeof(double));
  /* update_lhs(Hs.reshape(-1,7,7), cat(ii,ii,jj,jj), cat(ii,jj,ii,jj)) :1201-1203;
   * blocks with a negative (fixed) index are dropped :86; duplicates are summed :95 */
  for (int blk = 0; blk < 4; blk++) {
    for (int e = 0; e < E; e++) {
      const int i = ((blk < 2) ? ii_edge[e] : jj_edge[e]) - num_fix;
      const int j = ((blk % 2 == 0) ? ii_edge[e] : jj_edge[e]) - num_fix;
      if (i >= 0 && j >= 0) {
        const float *B = Hs + ((size_t)blk * E + e) * 49;
        for (int k = 0; k < M; k++)
          for (int l = 0; l < M; l++) A[(size_t)(M * i + k) * D + (M * j + l)] += (double)B[k * 7 + l];
      }
    }
  }
  /* update_rhs(gs.reshape(-1,7), cat(ii,jj)) :1205-1206, :105-112 */
  for (int blk = 0; blk < 2; blk++) {
    for (int e = 0; e < E; e++) {
      const int i = ((blk == 0) ? ii_edge[e] : jj_edge[e]) - num_fix;
      if (i >= 0)
        for (int j = 0; j < M; j++) bv[i * M + j] += (double)gs[((size_t)blk * E + e) * 7 + j];
    }
  }
  /* solve(): L.diagonal() += ep + lm*diag with lm = ep = 0 (:132-137); LLT on the lower
   * triangle (Eigen's default UpLo); info() != Success when a pivot is not positive. */
  int fail = 0;
  for (int j = 0; j < D && !fail; j++) {
    double *Lj = A + (size_t)j * D;
    double d = Lj[j];
    for (int k = 0; k < j; k++) d -= Lj[k] * Lj[k];
    if (!(d > 0.0)) {
      fail = 1;
      break;
    }
    const double ljj = sqrt(d);
    Lj[j] = ljj;
    const double inv = 1.0 / ljj;
#pragma omp parallel for schedule(static) if (D - j > 256)
    for (int i = j + 1; i < D; i++) {
      double *Li = A + (size_t)i * D;
      double s = Li[j];
      for (int k = 0; k < j; k++) s -= Li[k] * Lj[k];
      Li[j] = s * inv;
    }
  }
  if (!fail) {
    for (int i = 0; i < D; i++) { /* L y = b */
      double s = bv[i];
      const double *Li = A + (size_t)i * D;
      for (int k = 0; k < i; k++) s -= Li[k] * x[k];
      x[i] = s / Li[i];
    }
    for (int i = D - 1; i >= 0; i--) { /* L^T x = y */
      double s = x[i];
      for (int k = i + 1; k < D; k++) s -= A[(size_t)k * D + i] * x[k];
      x[i] = s / A[(size_t)i * D + i];
    }
    for (int i = 0; i < D; i++)
      if (!isfinite(x[i])) fail = 1;
  }
  free(A);
  free(bv);
  return fail;
}

/* ------------------------------------------------------------------------------------ */
/* driver loop  (gn_kernels.cu:1140-1228 and twins)                                      */
/* ------------------------------------------------------------------------------------ */
int m3r_oracle_gauss_newton(const m3r_oracle_gn_params *p, float *Twc, const float *Xs,
                            const float *Cs, const int64_t *ii, const int64_t *jj,
                            const int64_t *idx_ii2jj, const uint8_t *valid_match,
                            const float *Q, int num_poses, int num_points, int num_edges,
                            int max_iter, float delta_thresh, float *dx) {
  const int E = num_edges, num_fix = 1;
  const int D = (num_poses - num_fix) * 7;
  int32_t *ie = (int32_t *)malloc(sizeof(int32_t) * (size_t)(E > 0 ? E : 1));
  int32_t *je = (int32_t *)malloc(sizeof(int32_t) * (size_t)(E > 0 ? E : 1));
  m3r_oracle_edge_indices(ii, jj, E, ie, je);
  float *Hs = (float *)calloc((size_t)4 * E * 49 + 1, sizeof(float));
  float *gs = (float *)calloc((size_t)2 * E * 7 + 1, sizeof(float));
  double *x = (double *)calloc((size_t)(D > 0 ? D : 1), sizeof(double));
  int itr = 0;
  for (itr = 0; itr < max_iter; itr++) {
    m3r_oracle_gn_accumulate(p, Twc, Xs, Cs, ie, je, idx_ii2jj, valid_match, Q, num_poses,
                             num_points, E, Hs, gs);
    const int fail = m3r_oracle_gn_solve(Hs, gs, ie, je, num_poses, E, x);
    for (int k = 0; k < D; k++) dx[k] = fail ? -0.0f : (float)(-x[k]); /* dx = -A.solve() */
    m3r_oracle_pose_retr(Twc, dx, num_poses, num_fix);
    double nrm = 0.0;
    for (int k = 0; k < D; k++) nrm += (double)dx[k] * (double)dx[k];
    if ((float)sqrt(nrm) < delta_thresh) { /* :1219-1222 */
      itr++;
      break;
    }
  }
  free(ie), free(je), free(Hs), free(gs), free(x);
  return itr;
}
