// Tracking pose solve for sm_100a: the per-frame Gauss-Newton of
// FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 (mast3r_slam/tracker.py:157-266)
// as ONE persistent cooperative kernel.
//
// The reference runs, per iteration, ~40 torch kernels that materialise [n,3,7], [n,4,3] and
// [n,4,7] Jacobians in HBM, a [4n,7]^T [4n,7] GEMM, a 7x7 Cholesky, and two host syncs
// (`.item()` for the cost and the convergence test), for up to 50 iterations per frame.
//
// Here (not a translation):
//  * one edge, one free pose, n points: the whole problem (32 B per point, 6.3 MB at 512x384)
//    is staged ONCE into the shared memory of the 148 resident CTAs (42 KB each) -- including
//    the gather Xf = X_canon_f[idx_f2k] (tracker.py:155) and the keyframe's ray/distance
//    (tracker.py:183) -- and every iteration then runs out of shared memory;
//  * an iteration is: per-point residual + 4x7 (3x7) Jacobian in registers -> 28+7+1 sums per
//    thread -> warp shuffles -> one fp64 partial per CTA -> grid barrier -> every CTA adds the
//    partials in the same fixed order (bitwise identical on all CTAs, so no broadcast and no
//    second barrier), factors the 7x7 system in fp64, applies the left retraction and the
//    reference's convergence test.  One grid barrier per iteration, no host involvement;
//  * the result (T_WCf = T_WCk * T_CkCf, T_CkCf, status, iteration count, cost) is written to
//    one small device struct.
#include "m3r_common.cuh"

#include "sim3.cuh"

namespace m3r {

constexpr int kTrThreads = 512;
constexpr int kTrWarps = kTrThreads / 32;
constexpr int kTrAcc = 36;        // 28 (upper triangle of H) + 7 (J^T W r) + 1 (cost)

struct TrackArgs {
  const float *Xf;            // [n_f,3] frame pointmap (gathered through idx when non-null)
  const long long *idx;       // [n] or null
  const float *Xk;            // [n,3]
  const float *Qk;            // [n]
  const uint8_t *valid;       // [n]
  const float *meas;          // [n,3] calib: (u, v, log z) of the keyframe points
  const uint8_t *valid_meas;  // [n]   calib
  const float *T_WCf, *T_WCk; // [8]
  m3r_track_result *result;
  double *partial;            // [2][grid][kTrAcc]
  unsigned *barrier;          // [0] arrival counter (monotonic), [1] its value at kernel start
  const float *Kp;            // [9] calib intrinsics (device)
  int n;
  int n_f;                    // rows of Xf (bounds the gather)
  int max_iters;
  float inv_sigma_a, inv_sigma_b, huber_k, rel_error, delta_norm;
  float u_lo, u_hi, v_lo, v_hi, z_eps;
};

struct TrackSmem {
  double warp_part[kTrWarps][kTrAcc];
  double all_part[kNumSMs * kTrAcc];
  double sys[kTrAcc];
  Pose T;
  int stop;    // 0 continue, 1 converged, 2 factorisation failed
  double cost;
};
constexpr int kTrSmemMax = 227 * 1024;
constexpr int kTrRecBase = (static_cast<int>(sizeof(TrackSmem)) + 15) & ~15;
// staged points per CTA: 8 floats each in what the fixed part leaves of the 227 KB
constexpr int kTrCap = ((kTrSmemMax - kTrRecBase) / 32) & ~31;

// 7x7 SPD solve H x = g in fp64 (H given by its upper triangle, row-major packed).  Fully
// unrolled: the factor lives in registers, one rsqrt per pivot and no divisions.  Returns
// false when a pivot is not positive (torch.linalg.cholesky would raise).
__device__ __forceinline__ bool solve7(const double *__restrict__ sys, double x[7]) {
  double L[7][7], ri[7];
  {
    int k = 0;
#pragma unroll
    for (int i = 0; i < 7; i++)
#pragma unroll
      for (int j = i; j < 7; j++) L[j][i] = sys[k++];  // lower triangle
  }
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 7; j++) {
    double d = L[j][j];
#pragma unroll
    for (int p = 0; p < j; p++) d -= L[j][p] * L[j][p];
    ok = ok && (d > 0.0) && isfinite(d);
    ri[j] = rsqrt(d);
#pragma unroll
    for (int i = j + 1; i < 7; i++) {
      double v = L[i][j];
#pragma unroll
      for (int p = 0; p < j; p++) v -= L[i][p] * L[j][p];
      L[i][j] = v * ri[j];
    }
  }
  double y[7];
#pragma unroll
  for (int i = 0; i < 7; i++) {
    double v = sys[28 + i];
#pragma unroll
    for (int p = 0; p < i; p++) v -= L[i][p] * y[p];
    y[i] = v * ri[i];
  }
#pragma unroll
  for (int i = 6; i >= 0; i--) {
    double v = y[i];
#pragma unroll
    for (int p = i + 1; p < 7; p++) v -= L[p][i] * x[p];
    x[i] = v * ri[i];
  }
  return ok;
}

// Grid barrier on a monotonically increasing arrival counter (scratch word 0); `target` is
// the value the counter reaches when every CTA has arrived.  One release add and an acquire
// poll by thread 0: about half the cost of cooperative_groups' grid.sync(), whose L1
// invalidate this kernel does not need (the partials are read with ld.global.cg).
__device__ __forceinline__ void grid_barrier(unsigned *counter, unsigned target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
    unsigned v;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
    } while (static_cast<int>(v - target) < 0);
  }
  __syncthreads();
}

// Sum each of v[0..31] over the 32 lanes with 31 shuffles (instead of 160): every step
// halves the number of live values by exchanging one half with the partner lane.  On return
// v[0] of lane l holds the warp total of value l.
__device__ __forceinline__ void warp_transpose_sum32(float v[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; i++) {
      const float send = upper ? v[i] : v[i + off];
      const float keep = upper ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
}

// acc += w * (J^T J | -J^T e | 0.5 e^2) for one residual row
__device__ __forceinline__ void add_row(float acc[kTrAcc], const float J[7], float e, float w) {
  int k = 0;
#pragma unroll
  for (int a = 0; a < 7; a++) {
    const float wa = w * J[a];
#pragma unroll
    for (int b = a; b < 7; b++) acc[k++] += wa * J[b];
    acc[28 + a] -= wa * e;
  }
  acc[35] += 0.5f * w * e * e;
}

// J = -Jp @ [I, -skew(P), P]  (tracker.py:193 / :247 with geometry.act_Sim3, geometry.py:45-53)
__device__ __forceinline__ void chain_row(const float a[3], const float P[3], float J[7]) {
  J[0] = -a[0], J[1] = -a[1], J[2] = -a[2];
  J[3] = a[1] * P[2] - a[2] * P[1];  // (a x P)
  J[4] = a[2] * P[0] - a[0] * P[2];
  J[5] = a[0] * P[1] - a[1] * P[0];
  J[6] = -(a[0] * P[0] + a[1] * P[1] + a[2] * P[2]);
}

// whitened Huber weight: sqrt_info^2 * huber(sqrt_info * r)  (tracker.py:158-161,
// nonlinear_optimizer.py:29-35); si == 0 for masked points
__device__ __forceinline__ float robust_weight(float si, float e, float k) {
  const float wh = fabsf(si * e);
  const float hub = wh < k ? 1.0f : k / wh;
  return si * si * hub;
}

template <int KIND>
__device__ __forceinline__ void point_rows(const TrackArgs &A, const float K[9], const Pose &T, const float X[3],
                                           const float m[4], float s, float acc[kTrAcc]) {
  if (s == 0.0f) return;  // masked point: contributes exact zeros in the reference
  float P[3];
  act_sim3(T, X, P);
  float J[7];
  if (KIND == M3R_GN_RAYS) {
    // geometry.point_to_ray_dist (geometry.py:12-33)
    const float d = sqrtf(P[0] * P[0] + P[1] * P[1] + P[2] * P[2]);
    const float d_inv = 1.0f / d;
    const float r[3] = {d_inv * P[0], d_inv * P[1], d_inv * P[2]};
    const float d_inv2 = d_inv * d_inv;
    const float sa = s * A.inv_sigma_a, sb = s * A.inv_sigma_b;
#pragma unroll
    for (int c = 0; c < 3; c++) {
      float a[3];
#pragma unroll
      for (int j = 0; j < 3; j++) a[j] = d_inv * ((c == j ? 1.0f : 0.0f) - d_inv2 * (P[c] * P[j]));
      chain_row(a, P, J);
      const float e = m[c] - r[c];
      add_row(acc, J, e, robust_weight(sa, e, A.huber_k));
    }
    chain_row(r, P, J);
    const float e = m[3] - d;
    add_row(acc, J, e, robust_weight(sb, e, A.huber_k));
  } else {
    // geometry.project_calib (geometry.py:64-107)
    const float x = P[0], y = P[1], z = P[2];
    const float p0 = K[0] * x + K[1] * y + K[2] * z;
    const float p1 = K[3] * x + K[4] * y + K[5] * z;
    const float p2 = K[6] * x + K[7] * y + K[8] * z;
    const float u = p0 / p2, v = p1 / p2;
    const bool ok = (u > A.u_lo) && (u < A.u_hi) && (v > A.v_lo) && (v < A.v_hi) && (z > A.z_eps);
    if (!ok) return;
    const float z_inv = 1.0f / z;
    const float fx = K[0], fy = K[4];
    const float sa = s * A.inv_sigma_a, sb = s * A.inv_sigma_b;
    const float a0[3] = {fx * z_inv, 0.0f, -fx * x * z_inv * z_inv};
    const float a1[3] = {0.0f, fy * z_inv, -fy * y * z_inv * z_inv};
    const float a2[3] = {0.0f, 0.0f, z_inv};
    float e = m[0] - u;
    chain_row(a0, P, J);
    add_row(acc, J, e, robust_weight(sa, e, A.huber_k));
    e = m[1] - v;
    chain_row(a1, P, J);
    add_row(acc, J, e, robust_weight(sa, e, A.huber_k));
    e = m[2] - logf(z);
    chain_row(a2, P, J);
    add_row(acc, J, e, robust_weight(sb, e, A.huber_k));
  }
}

// what one point contributes to the staged record: (X[3], m[4], s)
template <int KIND>
__device__ __forceinline__ void load_point(const TrackArgs &A, int i, float X[3], float m[4], float &s) {
  long long src = i;
  bool ok = A.valid[i] != 0;
  if (A.idx) {
    src = A.idx[i];
    ok = ok && src >= 0 && src < A.n_f;
    if (!ok) src = 0;
  }
  const float *xf = A.Xf + src * 3;
  X[0] = __ldg(xf), X[1] = __ldg(xf + 1), X[2] = __ldg(xf + 2);
  const float q = __ldg(A.Qk + i);
  if (KIND == M3R_GN_RAYS) {
    const float *xk = A.Xk + static_cast<size_t>(i) * 3;
    const float a = __ldg(xk), b = __ldg(xk + 1), c = __ldg(xk + 2);
    const float d = sqrtf(a * a + b * b + c * c);  // tracker.py:183
    m[0] = a / d, m[1] = b / d, m[2] = c / d, m[3] = d;
  } else {
    const float *mk = A.meas + static_cast<size_t>(i) * 3;
    m[0] = __ldg(mk), m[1] = __ldg(mk + 1), m[2] = __ldg(mk + 2), m[3] = 0.0f;
    ok = ok && A.valid_meas[i] != 0;
  }
  s = ok ? sqrtf(q) : 0.0f;  // sqrt_info = valid * sqrt(Qk) / sigma  (tracker.py:175-176)
}

template <int KIND, bool STAGED>
__global__ void __launch_bounds__(kTrThreads, 1) track_pose_kernel(const TrackArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  TrackSmem &S = *reinterpret_cast<TrackSmem *>(smem_raw);
  float *rec = reinterpret_cast<float *>(smem_raw + kTrRecBase);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = gridDim.x;
  const int chunk = (A.n + G - 1) / G;
  const int c0 = min(blockIdx.x * chunk, A.n), c1 = min(c0 + chunk, A.n);
  const int cnt = c1 - c0;
  // counter value when this launch started: written by the previous launch's last action
  const unsigned base = __ldcg(A.barrier + 1);
  float K[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (KIND == M3R_GN_CALIB) {
#pragma unroll
    for (int k = 0; k < 9; k++) K[k] = __ldg(A.Kp + k);
  }

  if (STAGED) {
    // structure of arrays with stride `chunk`: conflict-free in the iteration loop
    for (int k = tid; k < cnt; k += kTrThreads) {
      float X[3], m[4], s;
      load_point<KIND>(A, c0 + k, X, m, s);
      rec[0 * chunk + k] = X[0], rec[1 * chunk + k] = X[1], rec[2 * chunk + k] = X[2];
      rec[3 * chunk + k] = m[0], rec[4 * chunk + k] = m[1], rec[5 * chunk + k] = m[2];
      rec[6 * chunk + k] = m[3], rec[7 * chunk + k] = s;
    }
  }
  if (tid == 0) {
    const Pose Twf = load_pose(A.T_WCf, 0), Twk = load_pose(A.T_WCk, 0);
    S.T = rel_sim3(Twk, Twf);  // T_CkCf = T_WCk^-1 * T_WCf  (tracker.py:180)
    S.stop = 0;
    S.cost = 0.0;
  }
  __syncthreads();

  double old_cost = INFINITY;
  int it = 0;
  for (; it < A.max_iters; it++) {
    const Pose T = S.T;
    float acc[kTrAcc];
#pragma unroll
    for (int k = 0; k < kTrAcc; k++) acc[k] = 0.0f;
    for (int k = tid; k < cnt; k += kTrThreads) {
      float X[3], m[4], s;
      if (STAGED) {
        X[0] = rec[0 * chunk + k], X[1] = rec[1 * chunk + k], X[2] = rec[2 * chunk + k];
        m[0] = rec[3 * chunk + k], m[1] = rec[4 * chunk + k], m[2] = rec[5 * chunk + k];
        m[3] = rec[6 * chunk + k], s = rec[7 * chunk + k];
      } else {
        load_point<KIND>(A, c0 + k, X, m, s);
      }
      point_rows<KIND>(A, K, T, X, m, s, acc);
    }
    // warp: fp32 shuffles; CTA and grid: fp64 in a fixed order
    {
      float tail[4] = {acc[32], acc[33], acc[34], acc[35]};
#pragma unroll
      for (int k = 0; k < 4; k++)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) tail[k] += __shfl_xor_sync(0xffffffffu, tail[k], o);
      warp_transpose_sum32(acc, lane);
      S.warp_part[warp][lane] = static_cast<double>(acc[0]);
      if (lane < 4) S.warp_part[warp][32 + lane] = static_cast<double>(tail[lane]);
    }
    __syncthreads();
    double *part = A.partial + static_cast<size_t>(it & 1) * G * kTrAcc;
    if (tid < kTrAcc) {
      double v = 0.0;
#pragma unroll
      for (int w = 0; w < kTrWarps; w++) v += S.warp_part[w][tid];
      part[static_cast<size_t>(blockIdx.x) * kTrAcc + tid] = v;
    }
    grid_barrier(A.barrier, base + static_cast<unsigned>(it + 1) * G);
    // every CTA: pull all G partials into shared memory (independent coalesced L2 loads), then
    // the same ordered sum on every CTA (8 interleaved slices per value)
    for (int k = tid; k < G * kTrAcc; k += kTrThreads) S.all_part[k] = __ldcg(part + k);
    __syncthreads();
    if (tid < kTrAcc * 8) {
      const int v = tid >> 3, sl = tid & 7;
      double s = 0.0;
      for (int c = sl; c < G; c += 8) s += S.all_part[c * kTrAcc + v];
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      if (sl == 0) S.sys[v] = s;
    }
    __syncthreads();
    if (tid == 0) {
      double x[7];
      const double cost = S.sys[35];
      S.cost = cost;
      if (!solve7(S.sys, x)) {
        S.stop = 2;
      } else {
        float tau[7];
        double nrm = 0.0;
#pragma unroll
        for (int k = 0; k < 7; k++) {
          tau[k] = static_cast<float>(x[k]);
          nrm += static_cast<double>(tau[k]) * tau[k];
        }
        Pose Tn = S.T;
        retr_sim3(tau, Tn);  // T_CkCf.retr(tau)  (tracker.py:196)
        S.T = Tn;
        // nonlinear_optimizer.check_convergence (:5-26); NaN on the first pass compares false
        const double rel_dec = fabs((old_cost - cost) / old_cost);
        if (rel_dec < static_cast<double>(A.rel_error) || sqrt(nrm) < static_cast<double>(A.delta_norm))
          S.stop = 1;
      }
    }
    __syncthreads();
    old_cost = S.cost;
    if (S.stop) {
      it++;
      break;
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
    m3r_track_result *R = A.result;
    const Pose Twk = load_pose(A.T_WCk, 0);
    const Pose Tc = S.T;
    const Pose Tw = mul_sim3(Twk, Tc);  // T_WCf = T_WCk * T_CkCf  (tracker.py:213)
    R->T_WCf[0] = Tw.t[0], R->T_WCf[1] = Tw.t[1], R->T_WCf[2] = Tw.t[2];
    R->T_WCf[3] = Tw.q[0], R->T_WCf[4] = Tw.q[1], R->T_WCf[5] = Tw.q[2], R->T_WCf[6] = Tw.q[3];
    R->T_WCf[7] = Tw.s;
    R->T_CkCf[0] = Tc.t[0], R->T_CkCf[1] = Tc.t[1], R->T_CkCf[2] = Tc.t[2];
    R->T_CkCf[3] = Tc.q[0], R->T_CkCf[4] = Tc.q[1], R->T_CkCf[5] = Tc.q[2], R->T_CkCf[6] = Tc.q[3];
    R->T_CkCf[7] = Tc.s;
    // every CTA has passed `it` barriers and read `base` long ago: safe to publish the new base
    A.barrier[1] = base + static_cast<unsigned>(it) * G;
    R->status = S.stop == 2 ? M3R_TRACK_CHOLESKY_FAILED : M3R_TRACK_OK;
    R->iterations = it;
    R->cost = S.cost;
  }
}

template <int KIND, bool STAGED>
static int launch_track(const TrackArgs &A, int grid, size_t smem, cudaStream_t s) {
  static bool attr_set_dev[64] = {false};  // per device; benign race: idempotent
  int dev = 0;
  M3R_CUDA_TRY(cudaGetDevice(&dev));
  bool &attr_set = attr_set_dev[dev & 63];
  if (!attr_set) {
    M3R_CUDA_TRY(cudaFuncSetAttribute(track_pose_kernel<KIND, STAGED>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, kTrSmemMax));
    attr_set = true;
  }
  TrackArgs args = A;
  void *params[] = {&args};
  M3R_CUDA_TRY(cudaLaunchCooperativeKernel(reinterpret_cast<void *>(track_pose_kernel<KIND, STAGED>),
                                           dim3(grid), dim3(kTrThreads), params, smem, s));
  M3R_LAUNCH_CHECK("track_pose_kernel");
  return M3R_OK;
}

}  // namespace m3r

using namespace m3r;

extern "C" int64_t m3r_track_scratch_doubles(void) { return 2 + 2LL * kNumSMs * kTrAcc; }

extern "C" int m3r_track_pose(const m3r_track_params *prm, const float *Xf, int64_t n_f,
                              const int64_t *idx_f2k, const float *Xk, const float *Qk,
                              const uint8_t *valid, const float *meas_k, const uint8_t *valid_meas_k,
                              const float *K, int64_t n, const float *T_WCf, const float *T_WCk,
                              m3r_track_result *result, double *scratch, void *stream) {
  M3R_REQUIRE(prm && result && scratch && T_WCf && T_WCk, "track_pose: null pointer");
  M3R_REQUIRE(prm->kind == M3R_GN_RAYS || prm->kind == M3R_GN_CALIB,
              "track_pose: kind must be M3R_GN_RAYS or M3R_GN_CALIB (got %d)", prm->kind);
  M3R_REQUIRE(n >= 1 && n < (1LL << 31) && n_f >= 1 && n_f < (1LL << 31), "track_pose: bad point count");
  M3R_REQUIRE(Xf && Qk && valid, "track_pose: null pointer");
  M3R_REQUIRE(idx_f2k || n_f == n, "track_pose: without idx_f2k, Xf must have n rows");
  M3R_REQUIRE(prm->max_iters >= 1, "track_pose: max_iters must be >= 1");
  M3R_REQUIRE(prm->sigma_a > 0.0f && prm->sigma_b > 0.0f, "track_pose: sigmas must be positive");
  if (prm->kind == M3R_GN_RAYS) M3R_REQUIRE(Xk != nullptr, "track_pose: rays need Xk");
  else M3R_REQUIRE(meas_k && valid_meas_k && K, "track_pose: calib needs meas_k, valid_meas_k and K");
  M3R_REQUIRE((reinterpret_cast<uintptr_t>(T_WCf) & 15u) == 0 && (reinterpret_cast<uintptr_t>(T_WCk) & 15u) == 0,
              "track_pose: poses must be 16-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  TrackArgs A{};
  A.Xf = Xf, A.idx = reinterpret_cast<const long long *>(idx_f2k), A.Xk = Xk, A.Qk = Qk, A.valid = valid;
  A.meas = meas_k, A.valid_meas = valid_meas_k, A.T_WCf = T_WCf, A.T_WCk = T_WCk;
  A.result = result, A.partial = scratch + 2, A.barrier = reinterpret_cast<unsigned *>(scratch), A.Kp = K;
  A.n = static_cast<int>(n), A.n_f = static_cast<int>(n_f), A.max_iters = prm->max_iters;
  A.inv_sigma_a = 1.0f / prm->sigma_a, A.inv_sigma_b = 1.0f / prm->sigma_b;
  A.huber_k = prm->huber_k, A.rel_error = prm->rel_error, A.delta_norm = prm->delta_norm;
  // geometry.py:79-80: u > border && u < w - 1 - border (and v with h)
  A.u_lo = prm->pixel_border, A.u_hi = static_cast<float>(prm->width - 1) - prm->pixel_border;
  A.v_lo = prm->pixel_border, A.v_hi = static_cast<float>(prm->height - 1) - prm->pixel_border;
  A.z_eps = prm->z_eps;

  int grid = static_cast<int>((n + kTrThreads - 1) / kTrThreads);
  if (grid > kNumSMs) grid = kNumSMs;
  const int chunk = static_cast<int>((n + grid - 1) / grid);
  const bool staged = chunk <= kTrCap;
  const size_t smem = static_cast<size_t>(kTrRecBase) + (staged ? static_cast<size_t>(chunk) * 8 * sizeof(float) : 0);
  if (prm->kind == M3R_GN_RAYS)
    return staged ? launch_track<M3R_GN_RAYS, true>(A, grid, smem, s)
                  : launch_track<M3R_GN_RAYS, false>(A, grid, smem, s);
  return staged ? launch_track<M3R_GN_CALIB, true>(A, grid, smem, s)
                : launch_track<M3R_GN_CALIB, false>(A, grid, smem, s);
}
