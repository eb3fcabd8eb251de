// Iterative projective matching and descriptor refinement for sm_100a.
//
// Replaces the reference's matching_kernels.cu (iter_proj_kernel :119-275,
// refine_matches_kernel :25-81) behind the C ABI of include/m3r_backends.h.  Written from
// the behaviour of those kernels, not from their code: iter_proj memoises its pure
// evaluations (one gather per LM iteration at most, none once a point has converged); the
// descriptor search runs two candidates per half2 lane with per-operation rounding.
#include <cuda_fp16.h>

#include <climits>
#include <mutex>

#include "m3r_common.cuh"

namespace m3r {

// =========================================================================================
// iter_proj
// =========================================================================================
//
// Arithmetic contract (what makes the integer match indices agree with the reference):
//  * bilinear weights: the reference evaluates (1.0-du)*dv etc. in double and rounds to
//    float (matching_kernels.cu:162-164).  du = u - floor(u) with u >= 1 is a multiple of
//    2^-23, so 1-du is exact in float and the float product rounds the same exact value
//    once: identical bits, no double needed.
//  * 1.0/r_norm and 1.0/det (:187,:213,:246): double division rounded to float equals the
//    correctly rounded float reciprocal (double rounding is innocuous for p=24 -> p=53),
//    i.e. __frcp_rn.
//  * lambda *= 0.1 (:262) is a double multiply by the double 0.1: reproduced in double.
//    lambda *= 10.0 (:266) is exact in double, hence equal to the float multiply.
//  * sqrtf is IEEE (default -prec-sqrt=true), fmin/fmax clamp as :21-23.
//  * multiply-add chains are written in the reference's association order so that nvcc's
//    FMA contraction produces the same instruction pattern.

struct Texel9 {
  float c[9];
};

__device__ __forceinline__ float clampf(float x, float lo, float hi) {
  return fminf(fmaxf(x, lo), hi);
}

// One LM evaluation at (u,v): interpolated ray (normalised), gradients, cost.
template <bool kWithGrad, typename Fetch>
__device__ __forceinline__ void eval_at(const Fetch &fetch, float u, float v, const float pt[3],
                                        float gx[3], float gy[3], float err[3], float &cost) {
  const int u11 = static_cast<int>(floorf(u));
  const int v11 = static_cast<int>(floorf(v));
  const float du = u - static_cast<float>(u11);
  const float dv = v - static_cast<float>(v11);
  const float w11 = du * dv;
  const float w12 = (1.0f - du) * dv;
  const float w21 = du * (1.0f - dv);
  const float w22 = (1.0f - du) * (1.0f - dv);
  constexpr int NC = kWithGrad ? 9 : 3;
  float r11[NC], r12[NC], r21[NC], r22[NC];
  // r11 = texel (u11+1, v11+1), weighted by w11 (:167); r12 = (u11, v11+1); r21 = (u11+1, v11); r22 = (u11, v11)
  fetch.template quad<NC>(u11, v11, r11, r12, r21, r22);
  float r[3];
#pragma unroll
  for (int j = 0; j < 3; j++) r[j] = w11 * r11[j] + w12 * r12[j] + w21 * r21[j] + w22 * r22[j];
  if (kWithGrad) {
#pragma unroll
    for (int j = 0; j < 3; j++) {
      gx[j] = w11 * r11[3 + j] + w12 * r12[3 + j] + w21 * r21[3 + j] + w22 * r22[3 + j];
      gy[j] = w11 * r11[6 + j] + w12 * r12[6 + j] + w21 * r21[6 + j] + w22 * r22[6 + j];
    }
  }
  const float r_norm = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  const float r_norm_inv = __frcp_rn(r_norm);
#pragma unroll
  for (int j = 0; j < 3; j++) {
    r[j] *= r_norm_inv;
    err[j] = r[j] - pt[j];
  }
  cost = err[0] * err[0] + err[1] * err[1] + err[2] * err[2];
}

// Gather of the four bilinear corners straight from the [h,w,9] image the API fixes
// (36-byte texels, only 4-byte aligned, so no vector loads).  The whole image (7 MB at
// 512x384) lives in L2; what the kernel has to economise is the number of gathers.
struct GlobalFetch {
  const float *img;
  int w;
  template <int NC>
  __device__ __forceinline__ void operator()(int uu, int vv, float *out) const {
    const float *s = img + (static_cast<size_t>(vv) * w + uu) * 9;
#pragma unroll
    for (int c = 0; c < NC; c++) out[c] = __ldg(s + c);
  }
  template <int NC>
  __device__ __forceinline__ void quad(int u11, int v11, float *r11, float *r12, float *r21, float *r22) const {
    operator()<NC>(u11 + 1, v11 + 1, r11);
    operator()<NC>(u11, v11 + 1, r12);
    operator()<NC>(u11 + 1, v11, r21);
    operator()<NC>(u11, v11, r22);
  }
};

// The same [h,w,9] API image read with 128-bit loads.  The two texels (u, u+1) a bilinear footprint
// takes from one image row are 18 consecutive floats starting at word 9t (t = texel index in the
// whole tensor).  9t mod 4 = t mod 4 =: s, so they lie inside the 16-byte-aligned words
// [9t - s, 9t - s + 21): five aligned float4 plus, when s = 3, one more word.  A footprint is then
// 10 LDG.128 (+ 2 predicated LDG.32) instead of 36 LDG.32 -- and, what the kernel is really bound
// by, a third of the L1 wavefronts: lanes of a warp sit 36 bytes apart, so EVERY load instruction
// of the warp touches 9-10 cache lines whatever its width.  The 18 values are moved into place with
// two rounds of selects on the bits of s (no local memory, no divergence).  Values are the same
// bits as the scalar path.
// Preconditions (checked on the host, else the scalar path runs): the tensor is 16-byte aligned and
// holds a multiple of 4 floats, so an aligned chunk that starts inside it ends inside it.
struct VecFetch {
  const float4 *img4;  // the whole [b,h,w,9] tensor
  size_t base_texel;   // b * h * w
  int w;
  __device__ __forceinline__ void row(int u, int v, float *L, float *R) const {
    const size_t t = base_texel + static_cast<size_t>(v) * w + u;
    const unsigned s = static_cast<unsigned>(t) & 3u;
    const float4 *p = img4 + ((9 * t - s) >> 2);
    const float4 c0 = __ldg(p), c1 = __ldg(p + 1), c2 = __ldg(p + 2), c3 = __ldg(p + 3), c4 = __ldg(p + 4);
    float x20 = 0.0f;
    if (s == 3u) x20 = __ldg(reinterpret_cast<const float *>(p + 5));  // word 9t + 17: last one of texel u+1
    const float W[21] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w, c2.x, c2.y, c2.z,
                         c2.w, c3.x, c3.y, c3.z, c3.w, c4.x, c4.y, c4.z, c4.w, x20};
    const bool s1 = (s & 1u) != 0, s2 = (s & 2u) != 0;
    float A[20];
#pragma unroll
    for (int i = 0; i < 20; i++) A[i] = s1 ? W[i + 1] : W[i];
#pragma unroll
    for (int i = 0; i < 9; i++) {
      L[i] = s2 ? A[i + 2] : A[i];
      R[i] = s2 ? A[i + 11] : A[i + 9];
    }
  }
  template <int NC>
  __device__ __forceinline__ void quad(int u11, int v11, float *r11, float *r12, float *r21, float *r22) const {
    static_assert(NC == 9, "the vector path fetches whole texels");
    row(u11, v11, r22, r21);
    row(u11, v11 + 1, r12, r11);
  }
};

// The same gather from the padded image the fused front end writes for itself (48-byte texels:
// ray, d/du, d/dv, 3 floats of padding): three 128-bit loads per texel instead of nine 32-bit
// ones (the `lg_throttle` stall of the 9-float path: 36 load instructions per evaluation).
struct PaddedFetch {
  const float4 *img;
  int w;
  template <int NC>
  __device__ __forceinline__ void operator()(int uu, int vv, float *out) const {
    const float4 *s = img + (static_cast<size_t>(vv) * w + uu) * 3;
    const float4 a = __ldg(s), b = __ldg(s + 1);
    out[0] = a.x, out[1] = a.y, out[2] = a.z;
    if (NC > 3) {
      const float4 c = __ldg(s + 2);
      out[3] = a.w, out[4] = b.x, out[5] = b.y, out[6] = b.z, out[7] = b.w, out[8] = c.x;
    }
  }
  template <int NC>
  __device__ __forceinline__ void quad(int u11, int v11, float *r11, float *r12, float *r21, float *r22) const {
    operator()<NC>(u11 + 1, v11 + 1, r11);
    operator()<NC>(u11, v11 + 1, r12);
    operator()<NC>(u11 + 1, v11, r21);
    operator()<NC>(u11, v11, r22);
  }
};

// 2x2 normal equations with the LM diagonal and the clamped trial pixel (:202-222)
__device__ __forceinline__ void lm_step(const float gx[3], const float gy[3], const float err[3],
                                        float lambda, float u, float v, float u_hi, float v_hi,
                                        float &u_new, float &v_new) {
  float A00 = gx[0] * gx[0] + gx[1] * gx[1] + gx[2] * gx[2];
  const float A01 = gx[0] * gy[0] + gx[1] * gy[1] + gx[2] * gy[2];
  float A11 = gy[0] * gy[0] + gy[1] * gy[1] + gy[2] * gy[2];
  const float b0 = -(err[0] * gx[0] + err[1] * gx[1] + err[2] * gx[2]);
  const float b1 = -(err[0] * gy[0] + err[1] * gy[1] + err[2] * gy[2]);
  A00 += lambda;
  A11 += lambda;
  const float det_inv = __frcp_rn(A00 * A11 - A01 * A01);
  const float delta_u = det_inv * (A11 * b0 - A01 * b1);
  const float delta_v = det_inv * (-A01 * b0 + A00 * b1);
  u_new = clampf(u + delta_u, 1.0f, u_hi);
  v_new = clampf(v + delta_v, 1.0f, v_hi);
}

// The memoised LM loop (MODE 0 of iter_proj_kernel, see below), shared with the fused match
// kernel.
#ifdef M3R_IP_COUNT  // experiments: number of fetched evaluations and the last iteration that fetched
__device__ int g_ip_evals, g_ip_last;
#define IP_COUNT(it_) (n_evals++, last_fetch = (it_))
#else
#define IP_COUNT(it_) ((void)0)
#endif
template <typename Fetch>
__device__ __forceinline__ void iter_proj_memo(const Fetch &fetch, const float pt[3], float &u,
                                               float &v, float u_hi, float v_hi, int max_iter,
                                               float lambda_init, float cost_thresh, bool &conv,
                                               int *stats = nullptr) {
#ifdef M3R_IP_COUNT
  int n_evals = 1, last_fetch = -1;
#endif
  float lambda = lambda_init;
  conv = false;
  if (max_iter <= 0) return;
  float gx[3], gy[3], err[3], cost;
  eval_at<true>(fetch, u, v, pt, gx, gy, err, cost);
  float ut = 0.f, vt = 0.f, cost_t = 0.f;  // last rejected trial since the pixel last moved
  bool have_trial = false;
  for (int it = 0; it < max_iter; it++) {
    float u_new, v_new;
    lm_step(gx, gy, err, lambda, u, v, u_hi, v_hi, u_new, v_new);
    float ngx[3], ngy[3], nerr[3], new_cost;
    bool fresh = false;
    if (u_new == u && v_new == v) {
      new_cost = cost;
    } else if (have_trial && u_new == ut && v_new == vt) {
      new_cost = cost_t;
    } else {
      eval_at<true>(fetch, u_new, v_new, pt, ngx, ngy, nerr, new_cost);
      fresh = true;
      IP_COUNT(it);
    }
    // a memoised cost is never below `cost` (it was rejected against the same `cost`)
    if (fresh && new_cost < cost) {
      u = u_new;
      v = v_new;
      cost = new_cost;
#pragma unroll
      for (int j = 0; j < 3; j++) gx[j] = ngx[j], gy[j] = ngy[j], err[j] = nerr[j];
      have_trial = false;
      lambda = static_cast<float>(static_cast<double>(lambda) * 0.1);
      conv = new_cost < cost_thresh;
    } else {
      ut = u_new, vt = v_new, cost_t = new_cost, have_trial = true;
      lambda = lambda * 10.0f;
      conv = cost < cost_thresh;
    }
  }
#ifdef M3R_IP_COUNT
  if (stats) stats[0] = n_evals, stats[1] = last_fetch;
#endif
}

#ifndef M3R_IP_THREADS
#define M3R_IP_THREADS 128
#endif
#ifndef M3R_IP_MINCTAS
#define M3R_IP_MINCTAS 1
#endif
constexpr int kIpThreads = M3R_IP_THREADS;
constexpr int kIpMinCtas = M3R_IP_MINCTAS;  // register cap of iter_proj_kernel through __launch_bounds__

// MODE 1: every iteration evaluates the current pixel (9 channels) and the trial pixel
// (3 channels), as the reference kernel does: two gathers per iteration.
//
// MODE 0 (default): the same arithmetic with the evaluations memoised.  An evaluation is a
// pure function of (u,v), so
//   * the trial evaluation fetches all 9 channels; if the step is accepted it IS the next
//     iteration's evaluation of the current pixel (bitwise), if rejected the current one is kept;
//   * a trial pixel equal to the current pixel, or to the last rejected trial while the current
//     pixel has not moved, has a known cost and is not fetched again.
// After convergence (2-4 iterations on smooth maps) the LM step is below half an ulp of the
// coordinate and the remaining iterations touch no memory.  At most one gather per iteration,
// typically 3-4 per point instead of 20; results are bit-identical to MODE 1.
template <int MODE>
__global__ void __launch_bounds__(kIpThreads, kIpMinCtas)
    iter_proj_kernel(const float *__restrict__ rays, const float *__restrict__ pts,
                     const float *__restrict__ p_init, float *__restrict__ p_new,
                     uint8_t *__restrict__ converged, int h, int w, int n, int max_iter,
                     float lambda_init, float cost_thresh) {
  const int b = blockIdx.y;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const size_t gi = static_cast<size_t>(b) * n + idx;
  const GlobalFetch fetch{rays + static_cast<size_t>(b) * h * w * 9, w};

  const float2 p0 = __ldg(reinterpret_cast<const float2 *>(p_init) + gi);
  const float pt[3] = {__ldg(pts + gi * 3 + 0), __ldg(pts + gi * 3 + 1), __ldg(pts + gi * 3 + 2)};
  const float u_hi = static_cast<float>(w - 2), v_hi = static_cast<float>(h - 2);
  float u = clampf(p0.x, 1.0f, u_hi);
  float v = clampf(p0.y, 1.0f, v_hi);

  float lambda = lambda_init;
  bool conv = false;
  if (MODE == 1) {
    for (int it = 0; it < max_iter; it++) {
      float gx[3], gy[3], err[3], cost;
      eval_at<true>(fetch, u, v, pt, gx, gy, err, cost);
      float u_new, v_new;
      lm_step(gx, gy, err, lambda, u, v, u_hi, v_hi, u_new, v_new);
      float err2[3], new_cost;  // cost at the trial pixel, ray channels only (:225-256)
      eval_at<false>(fetch, u_new, v_new, pt, nullptr, nullptr, err2, new_cost);
      if (new_cost < cost) {  // accept (:259-264)
        u = u_new;
        v = v_new;
        lambda = static_cast<float>(static_cast<double>(lambda) * 0.1);
        conv = new_cost < cost_thresh;
      } else {  // reject (:265-268)
        lambda = lambda * 10.0f;
        conv = cost < cost_thresh;
      }
    }
  } else if (MODE == 2) {  // memoised, 128-bit loads (host checked the alignment preconditions)
    const VecFetch vfetch{reinterpret_cast<const float4 *>(rays), static_cast<size_t>(b) * h * w, w};
#ifdef M3R_IP_COUNT
    int st[2] = {0, 0};
    iter_proj_memo(vfetch, pt, u, v, u_hi, v_hi, max_iter, lambda_init, cost_thresh, conv, st);
    reinterpret_cast<float2 *>(p_new)[gi] = make_float2(u, v);
    converged[gi] = static_cast<uint8_t>((conv ? 1 : 0) | (st[0] << 1) | ((st[1] + 1) << 5));  // evals <= 11, last+1 <= 10
    return;
#else
    iter_proj_memo(vfetch, pt, u, v, u_hi, v_hi, max_iter, lambda_init, cost_thresh, conv);
#endif
  } else {
    iter_proj_memo(fetch, pt, u, v, u_hi, v_hi, max_iter, lambda_init, cost_thresh, conv);
  }
  reinterpret_cast<float2 *>(p_new)[gi] = make_float2(u, v);
  converged[gi] = conv ? 1 : 0;
}

// =========================================================================================
// fused matching front end (callers: mast3r_slam/matching.py)
// =========================================================================================
//
// prep_rays: `prep_for_iter_proj` + `img_gradient` (matching.py:25-36, image.py:5-38) in one
// pass: rays = X / max(|X|, 1e-12) (F.normalize), gx/gy = the reference's 3x3 taps
// (1/32)[-3 0 3; -10 0 10; -3 0 3] and its transpose on the reflect-padded ray image, written
// straight in the [b,h,w,9] layout iter_proj reads.  The reference runs ~8 torch kernels and
// moves about 10x the bytes of the result for this.
constexpr int kPrTx = 32, kPrTy = 8;

__device__ __forceinline__ int reflect_idx(int i, int n) {  // F.pad(mode="reflect"), pad 1
  return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i);
}

template <bool PADDED>
__global__ void __launch_bounds__(kPrTx * kPrTy)
    prep_rays_kernel(const float *__restrict__ X11, float *__restrict__ rays, int h, int w) {
  __shared__ float tile[kPrTy + 2][kPrTx + 2][3];
  const int b = blockIdx.z;
  const int u0 = blockIdx.x * kPrTx, v0 = blockIdx.y * kPrTy;
  const float *img = X11 + static_cast<size_t>(b) * h * w * 3;
  for (int x = threadIdx.x; x < (kPrTy + 2) * (kPrTx + 2); x += kPrTx * kPrTy) {
    const int ty = x / (kPrTx + 2), tx = x - ty * (kPrTx + 2);
    const int vv = reflect_idx(min(v0 + ty - 1, h), h), uu = reflect_idx(min(u0 + tx - 1, w), w);
    const float *p = img + (static_cast<size_t>(min(vv, h - 1)) * w + min(uu, w - 1)) * 3;
    const float X = __ldg(p), Y = __ldg(p + 1), Z = __ldg(p + 2);
    const float d = fmaxf(sqrtf(X * X + Y * Y + Z * Z), 1e-12f);
    tile[ty][tx][0] = X / d, tile[ty][tx][1] = Y / d, tile[ty][tx][2] = Z / d;
  }
  __syncthreads();
  const int tx = threadIdx.x % kPrTx, ty = threadIdx.x / kPrTx;
  const int u = u0 + tx, v = v0 + ty;
  if (u >= w || v >= h) return;
  float out[9];
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float a00 = tile[ty][tx][c], a01 = tile[ty][tx + 1][c], a02 = tile[ty][tx + 2][c];
    const float a10 = tile[ty + 1][tx][c], a11 = tile[ty + 1][tx + 1][c], a12 = tile[ty + 1][tx + 2][c];
    const float a20 = tile[ty + 2][tx][c], a21 = tile[ty + 2][tx + 1][c], a22 = tile[ty + 2][tx + 2][c];
    out[c] = a11;
    out[3 + c] = 0.09375f * (a02 - a00) + 0.3125f * (a12 - a10) + 0.09375f * (a22 - a20);
    out[6 + c] = 0.09375f * (a20 - a00) + 0.3125f * (a21 - a01) + 0.09375f * (a22 - a02);
  }
  if (PADDED) {
    float4 *dst = reinterpret_cast<float4 *>(rays) + ((static_cast<size_t>(b) * h + v) * w + u) * 3;
    dst[0] = make_float4(out[0], out[1], out[2], out[3]);
    dst[1] = make_float4(out[4], out[5], out[6], out[7]);
    dst[2] = make_float4(out[8], 0.0f, 0.0f, 0.0f);
  } else {
    float *dst = rays + ((static_cast<size_t>(b) * h + v) * w + u) * 9;
#pragma unroll
    for (int c = 0; c < 9; c++) dst[c] = out[c];
  }
}

// match_proj: iter_proj with its torch prologue and epilogue folded in (matching.py:38-76):
// the query ray is normalised here (F.normalize of X21), the initial pixel comes from the
// previous match index (lin_to_pixel) or the identity, and the result leaves as the truncated
// integer pixel (`p1.long()`) together with valid = converged & |X11[p1] - X21| < dist_thresh.
__global__ void __launch_bounds__(kIpThreads)
    match_proj_kernel(const float *__restrict__ rays, const float *__restrict__ X11,
                      const float *__restrict__ X21, const long long *__restrict__ idx_init,
                      long long *__restrict__ p1, uint8_t *__restrict__ valid, int h, int w, int n,
                      int max_iter, float lambda_init, float cost_thresh, float dist_thresh) {
  const int b = blockIdx.y;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const size_t gi = static_cast<size_t>(b) * n + idx;
  const PaddedFetch fetch{reinterpret_cast<const float4 *>(rays) + static_cast<size_t>(b) * h * w * 3, w};
  const float X = __ldg(X21 + gi * 3), Y = __ldg(X21 + gi * 3 + 1), Z = __ldg(X21 + gi * 3 + 2);
  const float d = fmaxf(sqrtf(X * X + Y * Y + Z * Z), 1e-12f);
  const float pt[3] = {X / d, Y / d, Z / d};
  const long long lin = idx_init ? idx_init[gi] : static_cast<long long>(idx);
  const float u_hi = static_cast<float>(w - 2), v_hi = static_cast<float>(h - 2);
  float u = clampf(static_cast<float>(lin % w), 1.0f, u_hi);
  float v = clampf(static_cast<float>(lin / w), 1.0f, v_hi);
  bool conv;
  iter_proj_memo(fetch, pt, u, v, u_hi, v_hi, max_iter, lambda_init, cost_thresh, conv);
  const int u1 = static_cast<int>(u), v1 = static_cast<int>(v);  // .long(): truncation, values >= 1
  const float *q = X11 + ((static_cast<size_t>(b) * h + v1) * w + u1) * 3;
  const float dx = __ldg(q) - X, dy = __ldg(q + 1) - Y, dz = __ldg(q + 2) - Z;
  const float dist = sqrtf(dx * dx + dy * dy + dz * dz);
  reinterpret_cast<longlong2 *>(p1)[gi] = make_longlong2(u1, v1);
  valid[gi] = (conv && dist < dist_thresh) ? 1 : 0;
}

__global__ void pixel_to_lin_kernel(const long long *__restrict__ p1, long long *__restrict__ lin,
                                    long long total, int w) {
  const long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= total) return;
  const longlong2 p = reinterpret_cast<const longlong2 *>(p1)[k];
  lin[k] = p.x + static_cast<long long>(w) * p.y;
}


// match_confidence: the gathers and reductions that follow every match:
// tracker.py:41,62-67,104-110 and global_opt.py:55-68.  One pass over idx produces
//   Q      = sqrt(Q_src[idx] * Q_dst)                       (bit-identical to torch's mul + sqrt)
//   valid_q   = valid_match & (Q > Q_conf)                  ("valid_kf" / "valid_j")
//   valid_opt = valid_q & (C_src[idx] > C_conf) & (C_dst > C_conf)      (when C given)
// and the three integer counts the callers turn into fractions: sum(valid_q), sum(valid_opt)
// and the number of distinct idx among valid_match (torch.unique(...).shape[0], counted with a
// first-setter-wins bitmap instead of a sort).
constexpr int kMcThreads = 256;

__global__ void __launch_bounds__(kMcThreads)
    match_confidence_kernel(const long long *__restrict__ idx, const float *__restrict__ Q_src,
                            const float *__restrict__ Q_dst, const uint8_t *__restrict__ valid_match,
                            const float *__restrict__ C_src, const float *__restrict__ C_dst,
                            float Q_conf, float C_conf, float *__restrict__ Q_out,
                            uint8_t *__restrict__ valid_q, uint8_t *__restrict__ valid_opt,
                            int *__restrict__ counts, unsigned *__restrict__ bitmap, int n, int n_src,
                            int bitmap_words) {
  const int b = blockIdx.y;
  const size_t row = static_cast<size_t>(b) * n, srow = static_cast<size_t>(b) * n_src;
  int c_q = 0, c_opt = 0, c_uni = 0;
  for (int k = blockIdx.x * kMcThreads + threadIdx.x; k < n; k += gridDim.x * kMcThreads) {
    long long j = idx[row + k];
    const bool vm = valid_match[row + k] != 0;
    const bool inb = j >= 0 && j < n_src;
    if (!inb) j = 0;
    const float q = sqrtf(__ldg(Q_src + srow + j) * __ldg(Q_dst + row + k));
    const bool vq = vm && (q > Q_conf);
    Q_out[row + k] = q;
    valid_q[row + k] = vq ? 1 : 0;
    c_q += vq;
    if (valid_opt) {
      const bool vo = vq && (__ldg(C_src + srow + j) > C_conf) && (__ldg(C_dst + row + k) > C_conf);
      valid_opt[row + k] = vo ? 1 : 0;
      c_opt += vo;
    }
    // distinct-index count: lanes of a warp that hit the same bitmap word (the common case: a
    // smooth match field maps 32 neighbouring points into one or two words) merge their bits
    // and issue ONE atomicOr; the bits that were clear before are the first occurrences
    const bool mark = bitmap && vm && inb;
    const unsigned marking = __ballot_sync(__activemask(), mark);
    if (mark) {
      const int word = static_cast<int>(j >> 5);
      const unsigned peers = __match_any_sync(marking, word);
      const unsigned bits = __reduce_or_sync(peers, 1u << (j & 31));
      if ((threadIdx.x & 31) == __ffs(peers) - 1) {
        const unsigned old = atomicOr(bitmap + static_cast<size_t>(b) * bitmap_words + word, bits);
        c_uni += __popc(bits & ~old);
      }
    }
  }
  __shared__ int red[3][kMcThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    c_q += __shfl_xor_sync(0xffffffffu, c_q, o);
    c_opt += __shfl_xor_sync(0xffffffffu, c_opt, o);
    c_uni += __shfl_xor_sync(0xffffffffu, c_uni, o);
  }
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = c_q, red[1][threadIdx.x >> 5] = c_opt, red[2][threadIdx.x >> 5] = c_uni;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    int v = 0;
#pragma unroll
    for (int w = 0; w < kMcThreads / 32; w++) v += red[threadIdx.x][w];
    if (v) atomicAdd(counts + b * 3 + threadIdx.x, v);  // integers: order does not matter
  }
}

// =========================================================================================
// refine_matches
// =========================================================================================
//
// Bit-exactness contract: the reference accumulates `score += D21[k]*D11[k]` in the tensor
// dtype (matching_kernels.cu:60-63).  For c10::Half both the product and the sum are formed
// in float and rounded to binary16 (torch/headeronly/util/Half.h operator*, operator+);
// since float carries 24 >= 2*11+2 bits that equals one correctly rounded half multiply
// followed by one correctly rounded half add: __hmul2_rn / __hadd2_rn (the _rn forms also
// forbid ptxas from fusing them into HFMA2, which rounds once and would differ).  The k loop
// is sequential, so the two half2 lanes carry two different CANDIDATES, never two k's.
// Ties and the starting score follow :47-69: candidates are visited i-outer/j-inner, a
// candidate wins only with a strictly greater score, and max_score starts at
// numeric_limits<c10::Half>::min(), which libcu++ leaves unspecialised, i.e. Half() = +0.

constexpr int kRefineThreads = 128;

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  return __byte_perm(a, b, sel);
}
__device__ __forceinline__ __half2 as_h2(uint32_t x) { return *reinterpret_cast<__half2 *>(&x); }

template <int FDIM>  // multiple of 8 (uint4 rows)
__global__ void __launch_bounds__(kRefineThreads)
    refine_f16_kernel(const __half *__restrict__ D11, const __half *__restrict__ D21,
                      const long long *__restrict__ p1, long long *__restrict__ p1_new, int h,
                      int w, int n, int radius, int dilation_max) {
  constexpr int NW = FDIM / 2;  // 32-bit words per descriptor row
  constexpr int NV = FDIM / 8;  // uint4 per row
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const int b = blockIdx.y;
  const size_t gi = static_cast<size_t>(b) * n + idx;

  const longlong2 p = reinterpret_cast<const longlong2 *>(p1)[gi];
  // Candidates further than radius*dilation_max from the image can never be inside it.
  const long long reach = static_cast<long long>(radius) * dilation_max;
  if (p.x < -reach || p.x >= w + reach || p.y < -reach || p.y >= h + reach) {
    reinterpret_cast<longlong2 *>(p1_new)[gi] = p;
    return;
  }
  int u0 = static_cast<int>(p.x), v0 = static_cast<int>(p.y);
  int u_new = u0, v_new = v0;

  // query descriptor, kept in registers with each element broadcast to both half2 lanes
  uint32_t qlo[NW], qhi[NW];
  {
    const uint4 *qrow = reinterpret_cast<const uint4 *>(D21 + gi * FDIM);
#pragma unroll
    for (int i = 0; i < NV; i++) {
      const uint4 t = __ldg(qrow + i);
      const uint32_t wds[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
      for (int j = 0; j < 4; j++) {
        qlo[i * 4 + j] = prmt(wds[j], wds[j], 0x1010);
        qhi[i * 4 + j] = prmt(wds[j], wds[j], 0x3232);
      }
    }
  }
  const uint4 *img = reinterpret_cast<const uint4 *>(D11 + static_cast<size_t>(b) * h * w * FDIM);

  __half max_score = __ushort_as_half(0);  // +0
  const int nside = 2 * radius + 1;
  const int ncand = nside * nside;

  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    for (int c = 0; c < ncand; c += 2) {
      // candidate pair (c, c+1) in the reference's scan order: i outer (u), j inner (v)
      const int ia = c / nside, ja = c - ia * nside;
      int ib = ia, jb = ja + 1;
      if (jb == nside) {
        jb = 0;
        ib = ia + 1;
      }
      const int ua = u0 - rd + ia * d, va = v0 - rd + ja * d;
      const int ub = u0 - rd + ib * d, vb = v0 - rd + jb * d;
      const bool ina = (va >= 0) & (va < h) & (ua >= 0) & (ua < w);
      const bool inb = (c + 1 < ncand) & (vb >= 0) & (vb < h) & (ub >= 0) & (ub < w);
      if (!(ina | inb)) continue;
      // out-of-image lanes read a clamped (valid) row and are masked afterwards
      const int uac = min(max(ua, 0), w - 1), vac = min(max(va, 0), h - 1);
      const int ubc = min(max(ub, 0), w - 1), vbc = min(max(vb, 0), h - 1);
      const uint4 *ra = img + (static_cast<size_t>(vac) * w + uac) * NV;
      const uint4 *rb = img + (static_cast<size_t>(vbc) * w + ubc) * NV;
      uint32_t wa[NW], wb[NW];
#pragma unroll
      for (int i = 0; i < NV; i++) {
        const uint4 ta = __ldg(ra + i);
        const uint4 tb = __ldg(rb + i);
        wa[i * 4 + 0] = ta.x, wa[i * 4 + 1] = ta.y, wa[i * 4 + 2] = ta.z, wa[i * 4 + 3] = ta.w;
        wb[i * 4 + 0] = tb.x, wb[i * 4 + 1] = tb.y, wb[i * 4 + 2] = tb.z, wb[i * 4 + 3] = tb.w;
      }
      __half2 acc = as_h2(0u);
#pragma unroll
      for (int m = 0; m < NW; m++) {
        const __half2 even = as_h2(prmt(wa[m], wb[m], 0x5410));  // (a[2m],   b[2m])
        const __half2 odd = as_h2(prmt(wa[m], wb[m], 0x7632));   // (a[2m+1], b[2m+1])
        acc = __hadd2_rn(acc, __hmul2_rn(as_h2(qlo[m]), even));
        acc = __hadd2_rn(acc, __hmul2_rn(as_h2(qhi[m]), odd));
      }
      const __half sa = __low2half(acc), sb = __high2half(acc);
      if (ina && __hgt(sa, max_score)) {
        max_score = sa;
        u_new = ua;
        v_new = va;
      }
      if (inb && __hgt(sb, max_score)) {
        max_score = sb;
        u_new = ub;
        v_new = vb;
      }
    }
    u0 = u_new;  // recentre for the next dilation level (:75-76)
    v0 = v_new;
  }
  reinterpret_cast<longlong2 *>(p1_new)[gi] =
      make_longlong2(static_cast<long long>(u_new), static_cast<long long>(v_new));
}

// -----------------------------------------------------------------------------------------
// Tiled kernel (default for f16, fdim 24, radius 3, one query per pixel of an h x w grid).
//
// A CTA owns a 16x16 tile of query pixels.  While the match field is smooth the 245 candidate
// windows of those 256 queries all fall inside one 48x48-texel box of D11 (110.6 KB): the CTA
// fetches that box ONCE with TMA bulk copies (cp.async.bulk global->shared, one per tile row,
// completion on an mbarrier) and every candidate row is then three conflict-free LDS.128
// (48-byte rows: 8 consecutive lanes cover all 32 banks).  The direct kernel above re-reads
// every candidate row through L1 (53 % hit rate, the rest from L2).
//
// Per dilation level a warp votes: if every lane's 7x7 window lies inside the staged box the
// level runs fully unrolled from shared memory (out-of-image candidates are masked, the box's
// out-of-image rows are simply never written); otherwise that warp runs the level with the
// checked per-candidate path from global memory.  Same candidates, same order, same
// arithmetic: results are identical either way.
#ifndef M3R_RT_QY
#define M3R_RT_QY 16  // measured at 512x384: 16x8 153.5, 16x12 148.9, 16x16 145.4 us (profiles/r1_b_matching_ncu_full.md)
#endif
constexpr int kRtQ = 16;                    // query tile width
constexpr int kRtQy = M3R_RT_QY;            // query tile height
constexpr int kRtT = 48;                    // staged box width (texels): 16 + 2*15 + slack
constexpr int kRtTy = kRtQy + 32;           // staged box height
constexpr int kRtThreads = kRtQ * kRtQy;
constexpr int kRtRowU4 = 3;                 // uint4 per 24-half descriptor row
constexpr int kRtTileBytes = kRtT * kRtTy * kRtRowU4 * 16;
constexpr int kRtSmemBytes = kRtTileBytes + 64;

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// scores of two candidate rows against the query: lanes of the result = (row a, row b)
__device__ __forceinline__ __half2 score_pair(const uint4 *ra, const uint4 *rb,
                                              const uint32_t q[12]) {
  uint32_t wa[12], wb[12];
#pragma unroll
  for (int i = 0; i < kRtRowU4; i++) {
    const uint4 ta = ra[i];
    const uint4 tb = rb[i];
    wa[i * 4 + 0] = ta.x, wa[i * 4 + 1] = ta.y, wa[i * 4 + 2] = ta.z, wa[i * 4 + 3] = ta.w;
    wb[i * 4 + 0] = tb.x, wb[i * 4 + 1] = tb.y, wb[i * 4 + 2] = tb.z, wb[i * 4 + 3] = tb.w;
  }
  __half2 acc = as_h2(0u);
#pragma unroll
  for (int m = 0; m < 12; m++) {
    const __half2 even = as_h2(prmt(wa[m], wb[m], 0x5410));  // (a[2m],   b[2m])
    const __half2 odd = as_h2(prmt(wa[m], wb[m], 0x7632));   // (a[2m+1], b[2m+1])
    acc = __hadd2_rn(acc, __hmul2_rn(__low2half2(as_h2(q[m])), even));
    acc = __hadd2_rn(acc, __hmul2_rn(__high2half2(as_h2(q[m])), odd));
  }
  return acc;
}

#ifdef M3R_RT_PROF  // experiments only (tools/refine_prof.py): per-CTA timeline of this kernel
__device__ unsigned long long g_rt_prof[4096 * 16];
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#endif

// Plan pass (one warp per tile, ~4 us at 512x384): the bounding box of each tile's matches and
// a "heavy" flag.  The timeline of the tiled kernel (profiles/r1_n_refine_timeline.md) shows
// that a launch ends when its slowest tiles end: tiles whose matches collapsed onto the image
// border (view-2 pixels that see nothing of view 1, clamped by iter_proj) live 2.5-3x as long as
// the others, because their searches fan out of any box, and one of them starting in the last
// round holds an SM for 100 us while the rest of the machine idles.  The planned kernel
// therefore takes the heavy tiles first (longest-processing-time-first list scheduling, the
// order computed by every CTA from the flag array), and reads its box from the plan instead of
// reducing it, so the TMA copies start at the first instruction instead of after a dependent
// p1 load and two barriers.
constexpr int kRtMaxPlanned = 16 * kRtThreads;  // tiles the flag scan of the main kernel covers (16 per thread)
// The order only pays while a launch is a few rounds of 2 x 148 CTAs long.  Measured at 512x384
// (768 tiles per image), planned against unplanned: b = 1: 135 / 147 us, b = 2: 242 / 242, b = 3:
// 354 / 369, b = 4: 455 / 474, b = 5: 569 / 553, b = 8 (with a wider scan): 902 / 859: with many
// rounds the heavy tiles are spread by the index order anyway, and taking them first puts two of
// them on the same SM.  Above this many tiles the unplanned grid of tiles is launched.
constexpr int kRtPlanUpTo = kRtMaxPlanned < 3200 ? kRtMaxPlanned : 3200;
static_assert(kRtPlanUpTo <= 16 * kRtThreads, "the flag scan covers 16 tiles per thread of the main kernel");
__global__ void __launch_bounds__(kRtThreads)
    refine_plan_kernel(const long long *__restrict__ p1, int4 *__restrict__ plan,
                       unsigned char *__restrict__ heavy, int h, int w, int dilation_max, int tiles_x,
                       int tiles_y) {
  // one CTA per tile, one query per thread: every load of the pass is in flight at once
  // (the main kernel is launched with programmatic stream serialisation: its CTAs may be set up
  // while this grid runs; they wait for this grid's results at their griddepcontrol.wait)
  asm volatile("griddepcontrol.launch_dependents;");
  __shared__ int red[5];
  const int t = blockIdx.x, tid = threadIdx.x;
  const int per_img = tiles_x * tiles_y;
  const int b = t / per_img, r = t - b * per_img;
  const int ty = r / tiles_x, tx = r - ty * tiles_x;
  const long long reach = 3LL * dilation_max;
  const int qx = tx * kRtQ + (tid & (kRtQ - 1)), qy = ty * kRtQy + (tid >> 4);
  if (tid == 0) red[0] = INT_MAX, red[1] = INT_MIN, red[2] = INT_MAX, red[3] = INT_MIN, red[4] = 0;
  bool live = false;
  int u = 0, v = 0;
  if (qx < w && qy < h) {
    const longlong2 p = reinterpret_cast<const longlong2 *>(
        p1)[static_cast<size_t>(b) * h * w + static_cast<size_t>(qy) * w + qx];
    live = !(p.x < -reach || p.x >= w + reach || p.y < -reach || p.y >= h + reach);
    u = static_cast<int>(p.x), v = static_cast<int>(p.y);
  }
  __syncthreads();
  const int umin = __reduce_min_sync(0xffffffffu, live ? u : INT_MAX);
  const int umax = __reduce_max_sync(0xffffffffu, live ? u : INT_MIN);
  const int vmin = __reduce_min_sync(0xffffffffu, live ? v : INT_MAX);
  const int vmax = __reduce_max_sync(0xffffffffu, live ? v : INT_MIN);
  const int cnt = __popc(__ballot_sync(0xffffffffu, live));
  if ((tid & 31) == 0 && cnt) {
    atomicMin(&red[0], umin), atomicMax(&red[1], umax), atomicMin(&red[2], vmin), atomicMax(&red[3], vmax);
    atomicAdd(&red[4], cnt);
  }
  __syncthreads();
  if (tid == 0) {
    plan[t] = make_int4(red[0], red[1], red[2], red[3]);
    // collapsed: many matches on a span far below the tile's own 16 pixels
    heavy[t] = (red[4] >= kRtThreads / 4 && (red[1] - red[0] < kRtQ / 2 || red[3] - red[2] < kRtQy / 2)) ? 1 : 0;
  }
}

template <bool PLANNED>
__global__ void __launch_bounds__(kRtThreads, 2)
    refine_f16_tiled_kernel(const __half *__restrict__ D11, const __half *__restrict__ D21,
                            const long long *__restrict__ p1, long long *__restrict__ p1_new,
                            long long *__restrict__ lin_out, int h, int w, int dilation_max,
                            const int4 *__restrict__ plan, const unsigned char *__restrict__ heavy,
                            int tiles_x, int tiles_y, int ntiles) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint4 *tile = reinterpret_cast<uint4 *>(smem_raw);
  unsigned long long *mbar = reinterpret_cast<unsigned long long *>(smem_raw + kRtTileBytes);
  int *red = reinterpret_cast<int *>(smem_raw + kRtTileBytes + 16);  // umin, umax, vmin, vmax

  constexpr int radius = 3, nside = 7;
  const int tid = threadIdx.x, lane = tid & 31;
  int tile_x = blockIdx.x, tile_y = blockIdx.y, b = blockIdx.z;
  int4 box = make_int4(INT_MAX, INT_MIN, INT_MAX, INT_MIN);
  if (PLANNED) {
    // this CTA's tile: the blockIdx.x-th entry of (heavy tiles in index order, then the others).
    // Every thread owns 16 consecutive flags; a block-wide prefix sum of the heavy counts finds
    // the owner of the wanted rank.
    asm volatile("griddepcontrol.wait;" ::: "memory");  // the plan pass has finished and is visible
    int *scan = red;  // [0..7] warp totals, [8] the tile
    const int first_id = tid * 16;
    const int nvalid = min(max(ntiles - first_id, 0), 16);
    uint4 f = make_uint4(0u, 0u, 0u, 0u);
    if (nvalid > 0) f = reinterpret_cast<const uint4 *>(heavy)[tid];
    const uint32_t fw[4] = {f.x, f.y, f.z, f.w};
    unsigned bits = 0;  // bit k = flag of tile first_id + k
#pragma unroll
    for (int k = 0; k < 16; k++) {
      if (k < nvalid && ((fw[k >> 2] >> (8 * (k & 3))) & 0xffu) != 0) bits |= 1u << k;
    }
    const int hv = __popc(bits);
    int incl = hv;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) scan[tid >> 5] = incl;
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int k = 0; k < kRtThreads / 32; k++) {
      const int v = scan[k];
      before += (k < (tid >> 5)) ? v : 0;
      total += v;
    }
    const int hp = before + incl - hv;              // heavy tiles before my 16
    const int lp = min(first_id, ntiles) - hp;      // light tiles before my 16
    const int c = blockIdx.x;
    unsigned pick = 0;
    int want = -1;
    if (c < total) {
      if (c >= hp && c < hp + hv) pick = bits, want = c - hp;
    } else {
      const int cl = c - total, nl = nvalid - hv;
      if (cl >= lp && cl < lp + nl) pick = ~bits & ((1u << nvalid) - 1u), want = cl - lp;
    }
    if (want >= 0) {
      for (int k = 0; k < want; k++) pick &= pick - 1;  // drop the `want` lowest set bits
      scan[8] = first_id + __ffs(pick) - 1;
    }
    __syncthreads();
    const int t = scan[8];
    const int per_img = tiles_x * tiles_y;
    b = t / per_img;
    const int r = t - b * per_img;
    tile_y = r / tiles_x, tile_x = r - tile_y * tiles_x;
    box = plan[t];
    __syncthreads();  // scan[] is red[]: re-initialised below
  }
  const int qx = tile_x * kRtQ + (tid & (kRtQ - 1));
  const int qy = tile_y * kRtQy + (tid >> 4);
  const bool active = (qx < w) && (qy < h);
  const size_t n = static_cast<size_t>(h) * w;
  const size_t gi = static_cast<size_t>(b) * n + static_cast<size_t>(qy) * w + qx;

#ifdef M3R_RT_PROF
  unsigned long long *prof = g_rt_prof + ((b * tiles_y + tile_y) * tiles_x + tile_x) % 4096 * 16;
  int n_fallback = 0;
  if (tid == 0) {
    unsigned smid;
    asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
    prof[0] = smid, prof[1] = gtime();
  }
#endif
  if (tid == 0) {
    red[0] = INT_MAX, red[1] = INT_MIN, red[2] = INT_MAX, red[3] = INT_MIN;
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(mbar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  longlong2 p = make_longlong2(0, 0);
  if (active) p = reinterpret_cast<const longlong2 *>(p1)[gi];
  // candidates further than radius*(sum of dilations) ... the first level already decides:
  // beyond radius*dilation_max of the image no candidate is ever inside it, and the centre
  // never moves (same early-out as the direct kernel)
  const long long reach = static_cast<long long>(radius) * dilation_max;
  const bool live = active && !(p.x < -reach || p.x >= w + reach || p.y < -reach || p.y >= h + reach);
  int u0 = live ? static_cast<int>(p.x) : 0, v0 = live ? static_cast<int>(p.y) : 0;
  __syncthreads();  // mbarrier initialised
  if (!PLANNED) {
    const int a0 = __reduce_min_sync(0xffffffffu, live ? u0 : INT_MAX);
    const int a1 = __reduce_max_sync(0xffffffffu, live ? u0 : INT_MIN);
    const int a2 = __reduce_min_sync(0xffffffffu, live ? v0 : INT_MAX);
    const int a3 = __reduce_max_sync(0xffffffffu, live ? v0 : INT_MIN);
    if (lane == 0 && a0 != INT_MAX) {
      atomicMin(&red[0], a0), atomicMax(&red[1], a1), atomicMin(&red[2], a2), atomicMax(&red[3], a3);
    }
    __syncthreads();
    box = make_int4(red[0], red[1], red[2], red[3]);
  }
  const bool any_live = box.x != INT_MAX;
  // staged box [ou, ou+T) x [ov, ov+T), centred on the bounding box of the CTA's matches
  const int ou = any_live ? ((box.x + box.y + 1 - kRtT) >> 1) : 0;
  const int ov = any_live ? ((box.z + box.w + 1 - kRtTy) >> 1) : 0;
  const int c0 = max(ou, 0), c1 = min(ou + kRtT, w);  // columns of the box inside the image
  const int r0 = max(ov, 0), r1 = min(ov + kRtTy, h);
  const bool have_box = any_live && c1 > c0 && r1 > r0;
  const __half *img_h = D11 + static_cast<size_t>(b) * n * 24;
  if (tid < 32 && have_box) {
    const uint32_t row_bytes = static_cast<uint32_t>(c1 - c0) * 48u;
    if (lane == 0) {
      const uint32_t total = row_bytes * static_cast<uint32_t>(r1 - r0);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(mbar)),
                   "r"(total)
                   : "memory");
    }
    __syncwarp();
    for (int gv = r0 + lane; gv < r1; gv += 32) {
      const __half *src = img_h + (static_cast<size_t>(gv) * w + c0) * 24;
      const uint4 *dst = tile + (static_cast<size_t>(gv - ov) * kRtT + (c0 - ou)) * kRtRowU4;
      asm volatile(
          "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
              "r"(smem_u32(dst)),
          "l"(src), "r"(row_bytes), "r"(smem_u32(mbar))
          : "memory");
    }
  }
  // the query descriptor streams in while the box lands
  uint32_t q[12];
  if (active) {
    const uint4 *qrow = reinterpret_cast<const uint4 *>(D21 + gi * 24);
#pragma unroll
    for (int i = 0; i < kRtRowU4; i++) {
      const uint4 t = __ldg(qrow + i);
      q[i * 4 + 0] = t.x, q[i * 4 + 1] = t.y, q[i * 4 + 2] = t.z, q[i * 4 + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 12; i++) q[i] = 0u;
  }
  if (have_box) {
    uint32_t done = 0;
    while (!done) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(smem_u32(mbar)), "r"(0u)
          : "memory");
    }
  }

#ifdef M3R_RT_PROF
  if (tid == 0) prof[2] = gtime();
#endif
  const uint4 *img = reinterpret_cast<const uint4 *>(img_h);
  __half max_score = __ushort_as_half(0);  // +0 (see the contract above)
  int u_new = u0, v_new = v0;
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    const bool inside = have_box && (u0 - rd >= ou) && (u0 + rd < ou + kRtT) && (v0 - rd >= ov) &&
                        (v0 + rd < ov + kRtTy) && (6 * d < kRtTy);
    if (!__any_sync(0xffffffffu, live)) break;  // nothing to do in this warp (no barriers below)
    if (__all_sync(0xffffffffu, inside || !live)) {
      // ---- whole warp inside the staged box: unrolled, unchecked loads from shared memory
      // (at least one live lane is inside, hence 6*d < T and dead lanes stay addressable too)
      unsigned umask = 0, vmask = 0;  // which of the 7 columns / rows are inside the image
      if (live) {
#pragma unroll
        for (int i = 0; i < nside; i++) {
          const int uu = u0 - rd + i * d, vv = v0 - rd + i * d;
          umask |= (uu >= 0 && uu < w) ? (1u << i) : 0u;
          vmask |= (vv >= 0 && vv < h) ? (1u << i) : 0u;
        }
      }
      // dead lanes read the box corner (always addressable: 6*d < T) and are masked
      const int bu = live ? (u0 - rd - ou) : 0, bv = live ? (v0 - rd - ov) : 0;
      const uint4 *base = tile + (bv * kRtT + bu) * kRtRowU4;
      const int du4 = d * kRtRowU4, dv4 = d * kRtT * kRtRowU4;
      int best = -1;
#pragma unroll
      for (int c = 0; c < nside * nside; c += 2) {
        const int ia = c / nside, ja = c % nside;
        const int cb = (c + 1 < nside * nside) ? c + 1 : c;  // the last pair repeats the row
        const int ib = cb / nside, jb = cb % nside;
        const __half2 acc = score_pair(base + ia * du4 + ja * dv4, base + ib * du4 + jb * dv4, q);
        const bool oka = ((umask >> ia) & (vmask >> ja) & 1u) != 0;
        const bool okb = (c + 1 < nside * nside) && (((umask >> ib) & (vmask >> jb) & 1u) != 0);
        const __half sa = __low2half(acc), sb = __high2half(acc);
        if (oka && __hgt(sa, max_score)) max_score = sa, best = c;
        if (okb && __hgt(sb, max_score)) max_score = sb, best = c + 1;
      }
      if (best >= 0) {
        const int bi = best / nside, bj = best - bi * nside;
        u_new = u0 - rd + bi * d;
        v_new = v0 - rd + bj * d;
      }
    } else if (live) {
#ifdef M3R_RT_PROF
      n_fallback++;
#endif
      // ---- checked path from global memory (a window leaves the box)
      for (int c = 0; c < nside * nside; c += 2) {
        const int ia = c / nside, ja = c - ia * nside;
        int ib = ia, jb = ja + 1;
        if (jb == nside) jb = 0, ib = ia + 1;
        const int ua = u0 - rd + ia * d, va = v0 - rd + ja * d;
        const int ub = u0 - rd + ib * d, vb = v0 - rd + jb * d;
        const bool ina = (va >= 0) & (va < h) & (ua >= 0) & (ua < w);
        const bool inb = (c + 1 < nside * nside) & (vb >= 0) & (vb < h) & (ub >= 0) & (ub < w);
        if (!(ina | inb)) continue;
        const int uac = min(max(ua, 0), w - 1), vac = min(max(va, 0), h - 1);
        const int ubc = min(max(ub, 0), w - 1), vbc = min(max(vb, 0), h - 1);
        const __half2 acc = score_pair(img + (static_cast<size_t>(vac) * w + uac) * kRtRowU4,
                                       img + (static_cast<size_t>(vbc) * w + ubc) * kRtRowU4, q);
        const __half sa = __low2half(acc), sb = __high2half(acc);
        if (ina && __hgt(sa, max_score)) max_score = sa, u_new = ua, v_new = va;
        if (inb && __hgt(sb, max_score)) max_score = sb, u_new = ub, v_new = vb;
      }
    }
    u0 = u_new;  // recentre for the next dilation level (:75-76)
    v0 = v_new;
  }
  if (active) {
    const longlong2 res =
        live ? make_longlong2(static_cast<long long>(u_new), static_cast<long long>(v_new)) : p;
    if (p1_new) reinterpret_cast<longlong2 *>(p1_new)[gi] = res;
    if (lin_out) lin_out[gi] = res.x + static_cast<long long>(w) * res.y;  // pixel_to_lin (matching.py:88)
  }
#ifdef M3R_RT_PROF
  __syncwarp();
  if (lane == 0) prof[3 + (tid >> 5)] = gtime();
  const int nf = __reduce_max_sync(0xffffffffu, n_fallback);  // levels this warp ran from global memory
  if (lane == 0 && nf) atomicAdd(prof + 15, static_cast<unsigned long long>(nf));
#endif
}

// Generic path: any fdim, any of the three dtypes; accumulation in T as the reference does.
template <typename T>
struct RefineOps;
template <>
struct RefineOps<__half> {
  using acc_t = __half;
  static __device__ __forceinline__ acc_t zero() { return __ushort_as_half(0); }
  static __device__ __forceinline__ acc_t init_max() { return __ushort_as_half(0); }
  static __device__ __forceinline__ acc_t mac(acc_t s, __half a, __half b) {
    return __hadd_rn(s, __hmul_rn(a, b));
  }
  static __device__ __forceinline__ bool gt(acc_t a, acc_t b) { return __hgt(a, b); }
};
template <>
struct RefineOps<float> {
  using acc_t = float;
  static __device__ __forceinline__ acc_t zero() { return 0.0f; }
  static __device__ __forceinline__ acc_t init_max() { return 1.17549435e-38f; }  // FLT_MIN
  static __device__ __forceinline__ acc_t mac(acc_t s, float a, float b) {
    s += a * b;  // contracted to FFMA by nvcc exactly as in the reference build
    return s;
  }
  static __device__ __forceinline__ bool gt(acc_t a, acc_t b) { return a > b; }
};
template <>
struct RefineOps<double> {
  using acc_t = double;
  static __device__ __forceinline__ acc_t zero() { return 0.0; }
  static __device__ __forceinline__ acc_t init_max() { return 2.2250738585072014e-308; }
  static __device__ __forceinline__ acc_t mac(acc_t s, double a, double b) {
    s += a * b;
    return s;
  }
  static __device__ __forceinline__ bool gt(acc_t a, acc_t b) { return a > b; }
};

template <typename T>
__global__ void __launch_bounds__(kRefineThreads)
    refine_generic_kernel(const T *__restrict__ D11, const T *__restrict__ D21,
                          const long long *__restrict__ p1, long long *__restrict__ p1_new,
                          int h, int w, int n, int fdim, int radius, int dilation_max) {
  using Ops = RefineOps<T>;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const int b = blockIdx.y;
  const size_t gi = static_cast<size_t>(b) * n + idx;
  long long u0 = p1[gi * 2 + 0], v0 = p1[gi * 2 + 1];
  long long u_new = u0, v_new = v0;
  const T *q = D21 + gi * fdim;
  const T *img = D11 + static_cast<size_t>(b) * h * w * fdim;
  typename Ops::acc_t max_score = Ops::init_max();
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    const int diam = 2 * rd + 1;
    for (int i = 0; i < diam; i += d) {
      for (int j = 0; j < diam; j += d) {
        const long long u = u0 - rd + i;
        const long long v = v0 - rd + j;
        if (v >= 0 && v < h && u >= 0 && u < w) {
          const T *row = img + (static_cast<size_t>(v) * w + static_cast<size_t>(u)) * fdim;
          typename Ops::acc_t score = Ops::zero();
          for (int k = 0; k < fdim; k++) score = Ops::mac(score, q[k], row[k]);
          if (Ops::gt(score, max_score)) {
            max_score = score;
            u_new = u;
            v_new = v;
          }
        }
      }
    }
    u0 = u_new;
    v0 = v_new;
  }
  p1_new[gi * 2 + 0] = u_new;
  p1_new[gi * 2 + 1] = v_new;
}

}  // namespace m3r

// =========================================================================================
// C ABI
// =========================================================================================
using namespace m3r;

static int launch_refine_tiled(const __half *D11, const __half *D21, const long long *p1,
                               long long *p1_new, long long *lin_out, int b, int h, int w,
                               int dilation_max, cudaStream_t s, bool planned = true) {
  static bool attr_set_dev[64] = {false};  // per device; benign race: idempotent
  int dev = 0;
  M3R_CUDA_TRY(cudaGetDevice(&dev));
  bool &attr_set = attr_set_dev[dev & 63];
  if (!attr_set) {
    M3R_CUDA_TRY(cudaFuncSetAttribute(refine_f16_tiled_kernel<false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, kRtSmemBytes));
    M3R_CUDA_TRY(cudaFuncSetAttribute(refine_f16_tiled_kernel<true>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, kRtSmemBytes));
    attr_set = true;
  }
  const int tiles_x = (w + kRtQ - 1) / kRtQ, tiles_y = (h + kRtQy - 1) / kRtQy;
  const long long ntiles = static_cast<long long>(b) * tiles_x * tiles_y;
  M3R_REQUIRE(tiles_y <= 65535, "refine_matches: image too tall");
  if (planned && ntiles <= kRtPlanUpTo) {
    // stream-ordered scratch: one int4 box per tile, then the flag bytes padded to 16 per thread
    // (from a private pool that keeps its memory across host synchronisations: the default
    // pool's release threshold of 0 hands the pages back at every sync and the next allocation
    // then costs hundreds of microseconds)
    static cudaMemPool_t pools[64] = {nullptr};
    static std::mutex pool_mu;
    cudaMemPool_t pool;
    {
      std::lock_guard<std::mutex> lk(pool_mu);
      if (!pools[dev & 63]) {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        M3R_CUDA_TRY(cudaMemPoolCreate(&pools[dev & 63], &props));
        unsigned long long keep = ~0ull;
        M3R_CUDA_TRY(cudaMemPoolSetAttribute(pools[dev & 63], cudaMemPoolAttrReleaseThreshold, &keep));
      }
      pool = pools[dev & 63];
    }
    const size_t plan_bytes = static_cast<size_t>(ntiles) * sizeof(int4);
    void *scratch = nullptr;
    // flag bytes: every thread of the main kernel loads 16 of them as one uint4, whatever ntiles is
    M3R_CUDA_TRY(cudaMallocFromPoolAsync(&scratch, plan_bytes + kRtMaxPlanned, pool, s));
    int4 *plan = static_cast<int4 *>(scratch);
    unsigned char *heavy = static_cast<unsigned char *>(scratch) + plan_bytes;
    const int nt = static_cast<int>(ntiles);
    refine_plan_kernel<<<nt, kRtThreads, 0, s>>>(p1, plan, heavy, h, w, dilation_max, tiles_x, tiles_y);
    g_launch_count.fetch_add(1, std::memory_order_relaxed);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nt), cfg.blockDim = dim3(kRtThreads), cfg.dynamicSmemBytes = kRtSmemBytes, cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr, cfg.numAttrs = 1;
    const int4 *plan_c = plan;
    const unsigned char *heavy_c = heavy;
    cudaError_t e = cudaLaunchKernelEx(&cfg, refine_f16_tiled_kernel<true>, D11, D21, p1, p1_new, lin_out, h, w,
                                       dilation_max, plan_c, heavy_c, tiles_x, tiles_y, nt);
    if (e == cudaSuccess) e = cudaGetLastError();
    const cudaError_t ef = cudaFreeAsync(scratch, s);
    g_launch_count.fetch_add(1, std::memory_order_relaxed);
    if (e != cudaSuccess) return cuda_fail(e, "refine_f16_tiled_kernel<planned>");
    if (ef != cudaSuccess) return cuda_fail(ef, "cudaFreeAsync");
    return M3R_OK;
  }
  const dim3 tgrid(tiles_x, tiles_y, b);
  refine_f16_tiled_kernel<false><<<tgrid, kRtThreads, kRtSmemBytes, s>>>(
      D11, D21, p1, p1_new, lin_out, h, w, dilation_max, nullptr, nullptr, tiles_x, tiles_y, 0);
  M3R_LAUNCH_CHECK("refine_f16_tiled_kernel");
  return M3R_OK;
}

#ifdef M3R_RT_PROF
extern "C" int m3r_debug_refine_profile(unsigned long long *out, int n_entries) {
  M3R_CUDA_TRY(cudaDeviceSynchronize());
  M3R_CUDA_TRY(cudaMemcpyFromSymbol(out, g_rt_prof, sizeof(unsigned long long) * n_entries));
  static unsigned long long zero[4096 * 16];
  M3R_CUDA_TRY(cudaMemcpyToSymbol(g_rt_prof, zero, sizeof(zero)));
  return M3R_OK;
}
#endif

extern "C" int m3r_iter_proj_v(const float *rays, const float *pts, const float *p_init,
                               float *p_new, uint8_t *converged, int b, int h, int w, int n,
                               int max_iter, float lambda_init, float cost_thresh, int variant,
                               void *stream) {
  M3R_REQUIRE(b >= 0 && n >= 0, "iter_proj: negative batch or point count");
  if (b == 0 || n == 0) return M3R_OK;
  M3R_REQUIRE(rays && pts && p_init && p_new && converged, "iter_proj: null pointer");
  M3R_REQUIRE(h >= 3 && w >= 3, "iter_proj: image must be at least 3x3 (got %dx%d)", h, w);
  M3R_REQUIRE(b <= 65535, "iter_proj: batch %d exceeds grid.y limit", b);
  M3R_REQUIRE((reinterpret_cast<uintptr_t>(p_new) & 7u) == 0, "iter_proj: p_new not 8-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const dim3 block(kIpThreads);
  const dim3 grid((n + block.x - 1) / block.x, b);
  if (variant == 1) {
    iter_proj_kernel<1><<<grid, block, 0, s>>>(rays, pts, p_init, p_new, converged, h, w, n,
                                               max_iter, lambda_init, cost_thresh);
  } else if (variant == 0 || variant == 2) {
    // variant 0: memoised LM with 128-bit loads when the image allows it (16-byte aligned, a
    // multiple of 4 floats: true for every contiguous torch tensor whose b*h*w is a multiple of 4);
    // variant 2: the same with scalar loads (what variant 0 falls back to)
    const bool vec = variant == 0 && aligned16(rays) && ((static_cast<long long>(b) * h * w) & 3LL) == 0;
    if (vec)
      iter_proj_kernel<2><<<grid, block, 0, s>>>(rays, pts, p_init, p_new, converged, h, w, n,
                                                 max_iter, lambda_init, cost_thresh);
    else
      iter_proj_kernel<0><<<grid, block, 0, s>>>(rays, pts, p_init, p_new, converged, h, w, n,
                                                 max_iter, lambda_init, cost_thresh);
  } else {
    set_error("iter_proj: unknown variant %d", variant);
    return M3R_ERR_INVALID_ARGUMENT;
  }
  M3R_LAUNCH_CHECK("iter_proj_kernel");
  return M3R_OK;
}

extern "C" int m3r_iter_proj(const float *rays, const float *pts, const float *p_init,
                             float *p_new, uint8_t *converged, int b, int h, int w, int n,
                             int max_iter, float lambda_init, float cost_thresh, void *stream) {
  return m3r_iter_proj_v(rays, pts, p_init, p_new, converged, b, h, w, n, max_iter, lambda_init,
                         cost_thresh, 0, stream);
}

extern "C" int m3r_refine_matches_v(const void *D11, const void *D21, const int64_t *p1,
                                    int64_t *p1_new, int b, int h, int w, int n, int fdim,
                                    int dtype, int radius, int dilation_max, int variant,
                                    void *stream) {
  M3R_REQUIRE(b >= 0 && n >= 0, "refine_matches: negative batch or point count");
  if (b == 0 || n == 0) return M3R_OK;
  M3R_REQUIRE(D11 && D21 && p1 && p1_new, "refine_matches: null pointer");
  M3R_REQUIRE(h > 0 && w > 0 && fdim >= 0, "refine_matches: bad shape");
  M3R_REQUIRE(b <= 65535, "refine_matches: batch %d exceeds grid.y limit", b);
  M3R_REQUIRE(radius >= 0 && radius <= 1024 && dilation_max <= 1024,
              "refine_matches: radius/dilation out of range");
  M3R_REQUIRE((reinterpret_cast<uintptr_t>(p1) & 15u) == 0 &&
                  (reinterpret_cast<uintptr_t>(p1_new) & 15u) == 0,
              "refine_matches: p1/p1_new must be 16-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const dim3 block(kRefineThreads);
  const dim3 grid((n + block.x - 1) / block.x, b);
  const long long *p1l = reinterpret_cast<const long long *>(p1);
  long long *p1nl = reinterpret_cast<long long *>(p1_new);
  if (dtype == M3R_DTYPE_F16) {
    const bool fast = (fdim == 24) && aligned16(D11) && aligned16(D21) && variant != 1;
    const bool tiled = fast && (variant == 0 || variant == 3) && radius == 3 && dilation_max >= 1 &&
                       static_cast<long long>(n) == static_cast<long long>(h) * w;
    if (tiled) {  // variant 3: the same kernel without the plan pass (grid of tiles in index order)
      const int rc = launch_refine_tiled(static_cast<const __half *>(D11), static_cast<const __half *>(D21),
                                         p1l, p1nl, nullptr, b, h, w, dilation_max, s, variant == 0);
      if (rc) return rc;
      return M3R_OK;
    } else if (fast) {
      refine_f16_kernel<24><<<grid, block, 0, s>>>(static_cast<const __half *>(D11),
                                                   static_cast<const __half *>(D21), p1l, p1nl,
                                                   h, w, n, radius, dilation_max);
    } else {
      refine_generic_kernel<__half><<<grid, block, 0, s>>>(
          static_cast<const __half *>(D11), static_cast<const __half *>(D21), p1l, p1nl, h, w, n,
          fdim, radius, dilation_max);
    }
  } else if (dtype == M3R_DTYPE_F32) {
    refine_generic_kernel<float><<<grid, block, 0, s>>>(static_cast<const float *>(D11),
                                                        static_cast<const float *>(D21), p1l,
                                                        p1nl, h, w, n, fdim, radius,
                                                        dilation_max);
  } else if (dtype == M3R_DTYPE_F64) {
    refine_generic_kernel<double><<<grid, block, 0, s>>>(static_cast<const double *>(D11),
                                                         static_cast<const double *>(D21), p1l,
                                                         p1nl, h, w, n, fdim, radius,
                                                         dilation_max);
  } else {
    set_error("refine_matches: unknown dtype code %d", dtype);
    return M3R_ERR_INVALID_ARGUMENT;
  }
  M3R_LAUNCH_CHECK("refine_matches_kernel");
  return M3R_OK;
}

// -----------------------------------------------------------------------------------------
// Allocation-free variant of the planned, tiled refine_matches (no counterpart in the
// reference; added for callers that cannot allocate:
// CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// The planned kernel needs scratch: one int4 box per query tile and one flag byte per tile
// (padded to the 16-per-thread flag scan).  m3r_refine_matches takes it from a stream-ordered
// pool; here it lives in ONE file-scope device array of fixed size, so the call allocates
// nothing (no cudaMalloc / cudaMallocAsync / cudaFree, no torch allocation).  Consequences:
//   * every GPU that loads the library holds the array (kRtStaticScratchBytes = 55 296 B)
//     until the process exits;
//   * at most kRtPlanUpTo = 3200 query tiles of 16x16 per call, i.e. b <= 4 frames at 512x384;
//     larger problems are rejected with an error (use m3r_refine_matches);
//   * all calls on a device share the array: calls that overlap on the GPU (different streams,
//     or graph branches) race on it.  Serialise them.  Nothing in it has to start at zero: the
//     plan pass writes boxes and flags of tiles [0, ntiles) before the main kernel reads them,
//     and flag bytes beyond ntiles are masked by the scan, so the last call's contents are
//     harmless.
// In host code `g_rt_static_scratch` is only a stand-in for the symbol: the device address
// comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr size_t kRtStaticPlanBytes = static_cast<size_t>(kRtPlanUpTo) * sizeof(int4);
constexpr size_t kRtStaticScratchBytes = kRtStaticPlanBytes + kRtMaxPlanned;
static __device__ __align__(256) unsigned char g_rt_static_scratch[kRtStaticScratchBytes];

static int rt_static_scratch_address(int dev, unsigned char **out) {
  static unsigned char *addr[64] = {nullptr};
  static std::mutex mu;
  std::lock_guard<std::mutex> lk(mu);
  M3R_REQUIRE(dev >= 0 && dev < 64, "refine_matches_static: device index %d out of range", dev);
  if (!addr[dev]) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_rt_static_scratch);  // `dev` is current (checked by the caller)
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_rt_static_scratch)");
    addr[dev] = static_cast<unsigned char *>(p);
  }
  *out = addr[dev];
  return M3R_OK;
}

extern "C" int64_t m3r_refine_matches_static_max_tiles(void) { return kRtPlanUpTo; }

extern "C" int m3r_refine_matches_static(const void *D11, const void *D21, const int64_t *p1,
                                         int64_t *p1_new, int b, int h, int w, int n, int fdim,
                                         int dtype, int radius, int dilation_max, void *stream) {
  M3R_REQUIRE(b >= 0 && n >= 0, "refine_matches_static: negative batch or point count");
  if (b == 0 || n == 0) return M3R_OK;
  M3R_REQUIRE(D11 && D21 && p1 && p1_new, "refine_matches_static: null pointer");
  M3R_REQUIRE(h > 0 && w > 0, "refine_matches_static: bad shape");
  M3R_REQUIRE(dtype == M3R_DTYPE_F16 && fdim == 24 && radius == 3 && dilation_max >= 1 && dilation_max <= 1024 &&
                  static_cast<long long>(n) == static_cast<long long>(h) * w,
              "refine_matches_static: only the tiled configuration is available without allocation "
              "(f16 descriptors, fdim 24, radius 3, one query per pixel); use m3r_refine_matches");
  M3R_REQUIRE(aligned16(D11) && aligned16(D21) && aligned16(p1) && aligned16(p1_new),
              "refine_matches_static: D11, D21, p1, p1_new must be 16-byte aligned");
  const int tiles_x = (w + kRtQ - 1) / kRtQ, tiles_y = (h + kRtQy - 1) / kRtQy;
  const long long ntiles = static_cast<long long>(b) * tiles_x * tiles_y;
  M3R_REQUIRE(ntiles <= kRtPlanUpTo,
              "refine_matches_static: %lld query tiles of %dx%d exceed the static scratch (max %d tiles, "
              "%zu bytes); use m3r_refine_matches",
              ntiles, kRtQ, kRtQy, kRtPlanUpTo, kRtStaticScratchBytes);
  int dev = 0;
  M3R_CUDA_TRY(cudaGetDevice(&dev));
  unsigned char *scratch = nullptr;
  int rc = rt_static_scratch_address(dev, &scratch);
  if (rc) return rc;
  static bool attr_set_dev[64] = {false};  // per device; benign race: idempotent (no allocation)
  if (!attr_set_dev[dev & 63]) {
    M3R_CUDA_TRY(cudaFuncSetAttribute(refine_f16_tiled_kernel<true>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, kRtSmemBytes));
    attr_set_dev[dev & 63] = true;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int4 *plan = reinterpret_cast<int4 *>(scratch);
  unsigned char *heavy = scratch + kRtStaticPlanBytes;
  const int nt = static_cast<int>(ntiles);
  const long long *p1l = reinterpret_cast<const long long *>(p1);
  refine_plan_kernel<<<nt, kRtThreads, 0, s>>>(p1l, plan, heavy, h, w, dilation_max, tiles_x, tiles_y);
  M3R_LAUNCH_CHECK("refine_plan_kernel");
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nt), cfg.blockDim = dim3(kRtThreads), cfg.dynamicSmemBytes = kRtSmemBytes, cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr, cfg.numAttrs = 1;
  const int4 *plan_c = plan;
  const unsigned char *heavy_c = heavy;
  long long *lin_none = nullptr;
  cudaError_t e = cudaLaunchKernelEx(&cfg, refine_f16_tiled_kernel<true>, static_cast<const __half *>(D11),
                                     static_cast<const __half *>(D21), p1l, reinterpret_cast<long long *>(p1_new),
                                     lin_none, h, w, dilation_max, plan_c, heavy_c, tiles_x, tiles_y, nt);
  if (e == cudaSuccess) e = cudaGetLastError();
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  if (e != cudaSuccess) return cuda_fail(e, "refine_f16_tiled_kernel<planned, static scratch>");
  return M3R_OK;
}

extern "C" int m3r_refine_matches(const void *D11, const void *D21, const int64_t *p1,
                                  int64_t *p1_new, int b, int h, int w, int n, int fdim,
                                  int dtype, int radius, int dilation_max, void *stream) {
  return m3r_refine_matches_v(D11, D21, p1, p1_new, b, h, w, n, fdim, dtype, radius,
                              dilation_max, 0, stream);
}

// -----------------------------------------------------------------------------------------
// fused matching front end
// -----------------------------------------------------------------------------------------
extern "C" int m3r_prep_rays(const float *X11, float *rays_img_with_grad, int b, int h, int w,
                             void *stream) {
  M3R_REQUIRE(b >= 0, "prep_rays: negative batch");
  if (b == 0) return M3R_OK;
  M3R_REQUIRE(X11 && rays_img_with_grad, "prep_rays: null pointer");
  M3R_REQUIRE(h >= 2 && w >= 2, "prep_rays: reflect padding needs at least 2x2 (got %dx%d)", h, w);
  M3R_REQUIRE(b <= 65535 && (h + kPrTy - 1) / kPrTy <= 65535, "prep_rays: image or batch too large");
  const dim3 grid((w + kPrTx - 1) / kPrTx, (h + kPrTy - 1) / kPrTy, b);
  prep_rays_kernel<false><<<grid, kPrTx * kPrTy, 0, static_cast<cudaStream_t>(stream)>>>(
      X11, rays_img_with_grad, h, w);
  M3R_LAUNCH_CHECK("prep_rays_kernel");
  return M3R_OK;
}

extern "C" int m3r_match(const float *X11, const float *X21, const void *D11, const void *D21,
                         int fdim, int dtype, const int64_t *idx_init, int64_t *idx_1_to_2,
                         uint8_t *valid, float *rays_ws, int64_t *p1_ws, int b, int h, int w,
                         int max_iter, float lambda_init, float cost_thresh, float dist_thresh,
                         int radius, int dilation_max, void *stream) {
  M3R_REQUIRE(b >= 0, "match: negative batch");
  if (b == 0) return M3R_OK;
  M3R_REQUIRE(X11 && X21 && idx_1_to_2 && valid && rays_ws && p1_ws, "match: null pointer");
  M3R_REQUIRE(h >= 3 && w >= 3, "match: image must be at least 3x3 (got %dx%d)", h, w);
  M3R_REQUIRE(b <= 65535, "match: batch %d exceeds grid.y limit", b);
  M3R_REQUIRE(static_cast<long long>(h) * w < (1LL << 31), "match: image too large");
  M3R_REQUIRE((reinterpret_cast<uintptr_t>(p1_ws) & 15u) == 0, "match: p1_ws must be 16-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  M3R_REQUIRE(aligned16(rays_ws), "match: rays_ws must be 16-byte aligned");
  const int n = h * w;
  int rc = M3R_OK;
  {
    M3R_REQUIRE((h + kPrTy - 1) / kPrTy <= 65535, "match: image too tall");
    const dim3 pgrid((w + kPrTx - 1) / kPrTx, (h + kPrTy - 1) / kPrTy, b);
    prep_rays_kernel<true><<<pgrid, kPrTx * kPrTy, 0, s>>>(X11, rays_ws, h, w);  // 12-float texels
    M3R_LAUNCH_CHECK("prep_rays_kernel");
  }
  long long *p1 = reinterpret_cast<long long *>(p1_ws);
  long long *p1n = p1 + static_cast<size_t>(b) * n * 2;
  const dim3 grid((n + kIpThreads - 1) / kIpThreads, b);
  match_proj_kernel<<<grid, kIpThreads, 0, s>>>(rays_ws, X11, X21,
                                                reinterpret_cast<const long long *>(idx_init), p1, valid,
                                                h, w, n, max_iter, lambda_init, cost_thresh, dist_thresh);
  M3R_LAUNCH_CHECK("match_proj_kernel");
  const long long total = static_cast<long long>(b) * n;
  if (radius > 0) {  // matching.py:78
    M3R_REQUIRE(D11 && D21, "match: null descriptor pointer");
    const bool tiled = dtype == M3R_DTYPE_F16 && fdim == 24 && radius == 3 && dilation_max >= 1 &&
                       aligned16(D11) && aligned16(D21);
    if (tiled) {
      rc = launch_refine_tiled(static_cast<const __half *>(D11), static_cast<const __half *>(D21), p1,
                               nullptr, reinterpret_cast<long long *>(idx_1_to_2), b, h, w, dilation_max, s);
      if (rc) return rc;
      return M3R_OK;
    }
    rc = m3r_refine_matches(D11, D21, p1_ws, reinterpret_cast<int64_t *>(p1n), b, h, w, n, fdim, dtype,
                            radius, dilation_max, stream);
    if (rc) return rc;
    p1 = p1n;
  }
  pixel_to_lin_kernel<<<static_cast<int>((total + 255) / 256), 256, 0, s>>>(
      p1, reinterpret_cast<long long *>(idx_1_to_2), total, w);
  M3R_LAUNCH_CHECK("pixel_to_lin_kernel");
  return M3R_OK;
}

extern "C" int64_t m3r_match_confidence_scratch_bytes(int b, int n_src) {
  return static_cast<int64_t>(b) * ((n_src + 31) / 32) * 4;
}

extern "C" int m3r_match_confidence(const int64_t *idx, const float *Q_src, const float *Q_dst,
                                    const uint8_t *valid_match, const float *C_src, const float *C_dst,
                                    float Q_conf, float C_conf, float *Q_out, uint8_t *valid_q,
                                    uint8_t *valid_opt, int32_t *counts, void *scratch, int b, int n,
                                    int n_src, void *stream) {
  M3R_REQUIRE(b >= 0 && n >= 0 && n_src >= 0, "match_confidence: negative size");
  M3R_REQUIRE(counts != nullptr, "match_confidence: null counts");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (b == 0) return M3R_OK;
  M3R_CUDA_TRY(cudaMemsetAsync(counts, 0, sizeof(int32_t) * 3 * b, s));
  if (n == 0) return M3R_OK;
  M3R_REQUIRE(idx && Q_src && Q_dst && valid_match && Q_out && valid_q, "match_confidence: null pointer");
  M3R_REQUIRE(n_src >= 1, "match_confidence: empty source view");
  M3R_REQUIRE((valid_opt == nullptr) == (C_src == nullptr) && (C_src == nullptr) == (C_dst == nullptr),
              "match_confidence: C_src, C_dst and valid_opt go together");
  M3R_REQUIRE(b <= 65535, "match_confidence: batch %d exceeds grid.y limit", b);
  const int words = (n_src + 31) / 32;
  if (scratch) M3R_CUDA_TRY(cudaMemsetAsync(scratch, 0, static_cast<size_t>(b) * words * 4, s));
  int gx = (n + kMcThreads - 1) / kMcThreads;
  if (gx > kNumSMs * 8) gx = kNumSMs * 8;
  match_confidence_kernel<<<dim3(gx, b), kMcThreads, 0, s>>>(
      reinterpret_cast<const long long *>(idx), Q_src, Q_dst, valid_match, C_src, C_dst, Q_conf, C_conf,
      Q_out, valid_q, valid_opt, counts, static_cast<unsigned *>(scratch), n, n_src, words);
  M3R_LAUNCH_CHECK("match_confidence_kernel");
  return M3R_OK;
}
