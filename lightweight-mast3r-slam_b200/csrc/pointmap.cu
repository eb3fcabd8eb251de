// Pointmap producers of the Gauss-Newton inputs: the running fusion of
// a keyframe's canonical pointmap, Frame.update_pointmap (mast3r_slam/frame.py:41-102), the
// Sim(3) transport in front of it (tracker.py:97-98: `Xkk = T_CkCf.act(Xkf)`), and
// constrain_points_to_ray (mast3r_slam/geometry.py:37-42, 107-115).
//
// All of it is elementwise over n points: 32 B read + 16 B written per point for the fusion,
// HBM-bound.  The reference spends 6-20 torch kernels per update (each a full pass over the
// pointmap); here every update is one pass, optionally reading the pose straight from the
// device struct the tracking solve left behind, so the frame loop needs no host sync between
// the solve and the keyframe update.
//
// Rounding: torch evaluates `((C*Xc) + (Cn*X)) / (C + Cn)` as separate correctly rounded
// float kernels; the _rn intrinsics below reproduce that bit for bit (and keep ptxas from
// contracting the products into FMAs).
#include "m3r_common.cuh"
#include "sim3.cuh"

namespace m3r {

constexpr int kPmThreads = 256;

__device__ __forceinline__ void to_spherical(const float P[3], float S[3]) {
  // frame.py:79-85: r = |P|, phi = atan2(y, x), theta = acos(z / r)
  const float r = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(P[0], P[0]), __fmul_rn(P[1], P[1])), __fmul_rn(P[2], P[2])));
  S[0] = r, S[1] = atan2f(P[1], P[0]), S[2] = acosf(__fdiv_rn(P[2], r));
}

template <int MODE, bool TRANSFORM>
__global__ void __launch_bounds__(kPmThreads)
    update_pointmap_kernel(float *__restrict__ Xc, float *__restrict__ Cc, const float *__restrict__ X,
                           const float *__restrict__ C, const float *__restrict__ T, long long n) {
  Pose P{};
  if (TRANSFORM) P = load_pose(T, 0);
  for (long long k = static_cast<long long>(blockIdx.x) * kPmThreads + threadIdx.x; k < n;
       k += static_cast<long long>(gridDim.x) * kPmThreads) {
    float x[3] = {__ldg(X + 3 * k), __ldg(X + 3 * k + 1), __ldg(X + 3 * k + 2)};
    if (TRANSFORM) act_sim3(P, x, x);
    const float cn = __ldg(C + k);
    const float co = Cc[k];
    float xo[3] = {Xc[3 * k], Xc[3 * k + 1], Xc[3 * k + 2]};
    if (MODE == M3R_FILTER_WEIGHTED_POINTMAP) {
      const float cs = __fadd_rn(co, cn);
#pragma unroll
      for (int c = 0; c < 3; c++)
        Xc[3 * k + c] = __fdiv_rn(__fadd_rn(__fmul_rn(co, xo[c]), __fmul_rn(cn, x[c])), cs);
      Cc[k] = cs;
    } else if (MODE == M3R_FILTER_WEIGHTED_SPHERICAL) {
      float s1[3], s2[3], s[3];
      to_spherical(xo, s1);
      to_spherical(x, s2);
      const float cs = __fadd_rn(co, cn);
#pragma unroll
      for (int c = 0; c < 3; c++) s[c] = __fdiv_rn(__fadd_rn(__fmul_rn(co, s1[c]), __fmul_rn(cn, s2[c])), cs);
      // frame.py:87-93
      const float st = sinf(s[2]);
      Xc[3 * k] = __fmul_rn(__fmul_rn(s[0], st), cosf(s[1]));
      Xc[3 * k + 1] = __fmul_rn(__fmul_rn(s[0], st), sinf(s[1]));
      Xc[3 * k + 2] = __fmul_rn(s[0], cosf(s[2]));
      Cc[k] = cs;
    } else {  // M3R_FILTER_INDEP_CONF, frame.py:69-73
      if (cn > co) {
        Xc[3 * k] = x[0], Xc[3 * k + 1] = x[1], Xc[3 * k + 2] = x[2];
        Cc[k] = cn;
      }
    }
  }
}

// constrain_points_to_ray: X <- z * ((u - cx)/fx, (v - cy)/fy, 1), (u, v) the pixel of the point
__global__ void __launch_bounds__(kPmThreads)
    constrain_to_ray_kernel(const float *__restrict__ X, float *__restrict__ out,
                            const float *__restrict__ K, int w, long long hw, long long total) {
  const float fx = __ldg(K), cx = __ldg(K + 2), fy = __ldg(K + 4), cy = __ldg(K + 5);
  for (long long k = static_cast<long long>(blockIdx.x) * kPmThreads + threadIdx.x; k < total;
       k += static_cast<long long>(gridDim.x) * kPmThreads) {
    const long long p = k % hw;
    const float u = static_cast<float>(p % w), v = static_cast<float>(p / w);
    const float z = __ldg(X + 3 * k + 2);
    // geometry.py:108-114: tmp = (p - c) / f; P = z * (tmp1, tmp2, 1)
    out[3 * k] = __fmul_rn(z, __fdiv_rn(__fsub_rn(u, cx), fx));
    out[3 * k + 1] = __fmul_rn(z, __fdiv_rn(__fsub_rn(v, cy), fy));
    out[3 * k + 2] = z;
  }
}

static int grid_for(long long n) {
  long long g = (n + kPmThreads - 1) / kPmThreads;
  const long long cap = static_cast<long long>(kNumSMs) * 16;
  return static_cast<int>(g < cap ? (g > 0 ? g : 1) : cap);
}

template <int MODE>
static int launch_update(float *Xc, float *Cc, const float *X, const float *C, const float *T, long long n,
                         cudaStream_t s) {
  if (T)
    update_pointmap_kernel<MODE, true><<<grid_for(n), kPmThreads, 0, s>>>(Xc, Cc, X, C, T, n);
  else
    update_pointmap_kernel<MODE, false><<<grid_for(n), kPmThreads, 0, s>>>(Xc, Cc, X, C, T, n);
  M3R_LAUNCH_CHECK("update_pointmap_kernel");
  return M3R_OK;
}

}  // namespace m3r

using namespace m3r;

extern "C" int m3r_update_pointmap(int mode, float *X_canon, float *C_canon, const float *X, const float *C,
                                   const float *T, int64_t n, void *stream) {
  M3R_REQUIRE(n >= 0, "update_pointmap: negative point count");
  if (n == 0) return M3R_OK;
  M3R_REQUIRE(X_canon && C_canon && X && C, "update_pointmap: null pointer");
  M3R_REQUIRE(T == nullptr || (reinterpret_cast<uintptr_t>(T) & 15u) == 0,
              "update_pointmap: the pose must be 16-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (mode) {
    case M3R_FILTER_WEIGHTED_POINTMAP:
      return launch_update<M3R_FILTER_WEIGHTED_POINTMAP>(X_canon, C_canon, X, C, T, n, s);
    case M3R_FILTER_WEIGHTED_SPHERICAL:
      return launch_update<M3R_FILTER_WEIGHTED_SPHERICAL>(X_canon, C_canon, X, C, T, n, s);
    case M3R_FILTER_INDEP_CONF:
      return launch_update<M3R_FILTER_INDEP_CONF>(X_canon, C_canon, X, C, T, n, s);
    default:
      set_error("update_pointmap: unknown filtering mode %d", mode);
      return M3R_ERR_INVALID_ARGUMENT;
  }
}

extern "C" int m3r_constrain_points_to_ray(const float *Xs, float *out, const float *K, int b, int h, int w,
                                           void *stream) {
  M3R_REQUIRE(b >= 0 && h >= 0 && w >= 0, "constrain_points_to_ray: negative size");
  const long long hw = static_cast<long long>(h) * w, total = hw * b;
  if (total == 0) return M3R_OK;
  M3R_REQUIRE(Xs && out && K, "constrain_points_to_ray: null pointer");
  constrain_to_ray_kernel<<<grid_for(total), kPmThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      Xs, out, K, w, hw, total);
  M3R_LAUNCH_CHECK("constrain_to_ray_kernel");
  return M3R_OK;
}
