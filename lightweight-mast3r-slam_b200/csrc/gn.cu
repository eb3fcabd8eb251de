// Sim(3) keyframe-graph Gauss-Newton for sm_100a: accumulation, deterministic assembly,
// on-device fp64 Cholesky solve, left retraction and the stop test, with no host round trip.
//
// Replaces the reference's gn_kernels.cu (ray_align_kernel :813-1138, calib_proj_kernel
// :1231-1543, point_align_kernel :455-723, pose_retr_kernel :415-453, the Eigen SparseBlock
// host solve :57-159 and the three driver loops) behind include/m3r_backends.h.
//
// Design (not a translation of the reference):
//  * The reference carries the full 14x14 packed Hessian of (pose i, pose j) per thread and
//    maps every residual row into the world frame per point.  Its own Ji = -Jj
//    (gn_kernels.cu:1000) means the 14x14 block is [[H,-H],[-H,H]] with H = A S A^T, where S
//    is the 7x7 sum of w * J_loc^T J_loc in the LOCAL frame of camera i and A is the 7x7
//    matrix of apply_Sim3_adj_inv(T_i) (:277-297), constant over the edge.  So the streaming
//    kernel accumulates only S (28 values, sparsity of J_loc resolved at compile time) and
//    s = sum w*e*J_loc (7 values); A is applied once per edge in fp64.
//  * Work is split into (edge, chunk) items over a persistent grid; every item writes its
//    partial sums to its own slot, and the chunk partials are combined in a fixed order in
//    fp64: bitwise reproducible, no atomics.
//  * Assembly into the dense lower-triangular system is a gather per 7x7 block pair over the
//    edge list in edge order (deterministic).  The system is factored by a blocked fp64
//    Cholesky on the device; failure (non-positive pivot) yields dx = 0 as in :147-150.
//  * Retraction, ||dx|| and the stop flag live in the same final kernel; once the flag is set
//    the remaining enqueued kernels exit at their first instruction.
#include "m3r_common.cuh"

#include <cooperative_groups.h>

#include <cstdlib>
#include <cstring>
#include <vector>

#include "sim3.cuh"

namespace m3r {

// ----------------------------------------------------------------------------------------
// small math
// ----------------------------------------------------------------------------------------
// single-MUFU approximations (<= 2 ulp): the streaming kernel is instruction-issue bound and
// the 1e-4 pose tolerance leaves five orders of magnitude of head-room
__device__ __forceinline__ float rsqrt_fast(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sqrt_fast(float x) {
  float y;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_fast(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float huber_w(float r) {
  // gn_kernels.cu:172-175 (k = 1.345): 1 if |r| < k else k/|r|  ==  min(1, k/|r|)
  return fminf(1.0f, 1.345f * rcp_fast(fabsf(r)));
}

// ----------------------------------------------------------------------------------------
// local-frame accumulators: S (packed lower 7x7, index a*(a+1)/2 + c, a >= c) and s (7)
// ----------------------------------------------------------------------------------------
constexpr int kNS = 28;       // packed S
constexpr int kNAcc = 35;     // S + s

// Adds one residual row with weight w, error e and local Jacobian J.  MASK has bit a set when
// J[a] can be non-zero; everything else is removed at compile time (adding an exact zero
// product never changes a sum, so this is value-identical to the dense update).
template <unsigned MASK>
__device__ __forceinline__ void add_row(float w, float e, const float J[7], float acc[kNAcc]) {
  const float we = w * e;
#pragma unroll
  for (int a = 0; a < 7; a++) {
    if (!((MASK >> a) & 1u)) continue;
    const float wa = w * J[a];
#pragma unroll
    for (int c = 0; c <= a; c++) {
      if (!((MASK >> c) & 1u)) continue;
      acc[a * (a + 1) / 2 + c] = fmaf(wa, J[c], acc[a * (a + 1) / 2 + c]);
    }
    acc[kNS + a] = fmaf(we, J[a], acc[kNS + a]);
  }
}

struct GnParams {
  float sigma_a_inv;  // rays: 1/sigma_ray   calib: 1/sigma_pixel  points: 1/sigma_point
  float sigma_b_inv;  // rays: 1/sigma_dist  calib: 1/sigma_depth
  float C_thresh, Q_thresh;
  float fx, fy, cx, cy;  // calib (read on device from K)
  int width, height;
  float u_lo, u_hi, v_lo, v_hi;  // calib: pixel_border bounds as floats
  float z_eps;
};

template <int KIND>
struct Residual;

// ray_align_kernel :910-1090 -- residual (unit-ray difference, distance difference)
template <>
struct Residual<M3R_GN_RAYS> {
  static __device__ __forceinline__ void point(const GnParams &P, const Pose &Tij,
                                               const float Xi[3], const float Xj[3], float q,
                                               int, float acc[kNAcc]) {
    const float n2i = Xi[0] * Xi[0] + Xi[1] * Xi[1] + Xi[2] * Xi[2];
    const float n1i_inv = rsqrt_fast(n2i);
    const float n1i = n2i * n1i_inv;
    float Pj[3];
    act_sim3(Tij, Xj, Pj);
    const float n2j = Pj[0] * Pj[0] + Pj[1] * Pj[1] + Pj[2] * Pj[2];
    const float n1j_inv = rsqrt_fast(n2j);
    const float n1j = n2j * n1j_inv;
    const float r[3] = {n1j_inv * Pj[0], n1j_inv * Pj[1], n1j_inv * Pj[2]};
    const float e0 = r[0] - n1i_inv * Xi[0];
    const float e1 = r[1] - n1i_inv * Xi[1];
    const float e2 = r[2] - n1i_inv * Xi[2];
    const float e3 = n1j - n1i;
    const float sq = sqrt_fast(q);
    const float swr = P.sigma_a_inv * sq;
    const float swd = P.sigma_b_inv * sq;
    const float wr = swr * swr, wd = swd * swd;
    const float w0 = huber_w(swr * e0) * wr;
    const float w1 = huber_w(swr * e1) * wr;
    const float w2 = huber_w(swr * e2) * wr;
    const float w3 = huber_w(swd * e3) * wd;
    const float n3 = n1j_inv * n1j_inv * n1j_inv;  // 1/|P|^3 (:982)
    const float axx = n1j_inv - Pj[0] * Pj[0] * n3;
    const float ayy = n1j_inv - Pj[1] * Pj[1] * n3;
    const float azz = n1j_inv - Pj[2] * Pj[2] * n3;
    const float axy = -Pj[0] * Pj[1] * n3;
    const float axz = -Pj[0] * Pj[2] * n3;
    const float ayz = -Pj[1] * Pj[2] * n3;
    {
      const float J[7] = {axx, axy, axz, 0.f, r[2], -r[1], 0.f};  // :991-997
      add_row<0b0110111>(w0, e0, J, acc);
    }
    {
      const float J[7] = {axy, ayy, ayz, -r[2], 0.f, r[0], 0.f};  // :1016-1022
      add_row<0b0101111>(w1, e1, J, acc);
    }
    {
      const float J[7] = {axz, ayz, azz, r[1], -r[0], 0.f, 0.f};  // :1041-1047
      add_row<0b0011111>(w2, e2, J, acc);
    }
    {
      const float J[7] = {r[0], r[1], r[2], 0.f, 0.f, 0.f, n1j};  // :1067-1073
      add_row<0b1000111>(w3, e3, J, acc);
    }
  }
  static __device__ __forceinline__ bool extra_valid(const GnParams &, const Pose &,
                                                     const float *, const float *) {
    return true;
  }
};

// point_align_kernel :550-677 -- residual X_j->i - X_i
template <>
struct Residual<M3R_GN_POINTS> {
  static __device__ __forceinline__ void point(const GnParams &P, const Pose &Tij,
                                               const float Xi[3], const float Xj[3], float q,
                                               int, float acc[kNAcc]) {
    float Pj[3];
    act_sim3(Tij, Xj, Pj);
    const float e0 = Pj[0] - Xi[0], e1 = Pj[1] - Xi[1], e2 = Pj[2] - Xi[2];
    const float sw = P.sigma_a_inv * sqrt_fast(q);
    const float wc = sw * sw;
    const float w0 = huber_w(sw * e0) * wc;
    const float w1 = huber_w(sw * e1) * wc;
    const float w2 = huber_w(sw * e2) * wc;
    {
      const float J[7] = {1.f, 0.f, 0.f, 0.f, Pj[2], -Pj[1], Pj[0]};  // :602-608
      add_row<0b1110001>(w0, e0, J, acc);
    }
    {
      const float J[7] = {0.f, 1.f, 0.f, -Pj[2], 0.f, Pj[0], Pj[1]};  // :627-633
      add_row<0b1101010>(w1, e1, J, acc);
    }
    {
      const float J[7] = {0.f, 0.f, 1.f, Pj[1], -Pj[0], 0.f, Pj[2]};  // :652-658
      add_row<0b1011100>(w2, e2, J, acc);
    }
  }
};

// calib_proj_kernel :1346-1497 -- residual (pixel u, pixel v, log depth)
template <>
struct Residual<M3R_GN_CALIB> {
  static __device__ __forceinline__ void point(const GnParams &P, const Pose &Tij,
                                               const float Xi[3], const float Xj[3], float q,
                                               int ind_Xi, float acc[kNAcc]) {
    float Pj[3];
    act_sim3(Tij, Xj, Pj);
    const bool valid_z = (Pj[2] > P.z_eps) && (Xi[2] > P.z_eps);  // :1368
    if (!valid_z) return;
    const float iz = rcp_fast(Pj[2]);
    const float x = Pj[0] * iz, y = Pj[1] * iz;
    const float u = P.fx * x + P.cx;
    const float v = P.fy * y + P.cy;
    const bool valid_uv = (u > P.u_lo) && (u < P.u_hi) && (v > P.v_lo) && (v < P.v_hi);  // :1382-1383
    if (!valid_uv) return;
    const int v_t = ind_Xi / P.width;  // :1361-1362
    const int u_t = ind_Xi - v_t * P.width;
    const float e0 = u - static_cast<float>(u_t);
    const float e1 = v - static_cast<float>(v_t);
    const float e2 = __logf(Pj[2] * rcp_fast(Xi[2]));  // log z_j->i - log z_i
    const float sq = sqrt_fast(q);
    const float swp = P.sigma_a_inv * sq;
    const float swd = P.sigma_b_inv * sq;
    const float wp = swp * swp, wd = swd * swd;
    const float w0 = huber_w(swp * e0) * wp;
    const float w1 = huber_w(swp * e1) * wp;
    const float w2 = huber_w(swd * e2) * wd;
    {
      const float J[7] = {P.fx * iz,       0.f, -P.fx * x * iz, -P.fx * x * y,
                          P.fx * (1.f + x * x), -P.fx * y,      0.f};  // :1422-1428
      add_row<0b0111101>(w0, e0, J, acc);
    }
    {
      const float J[7] = {0.f,          P.fy * iz, -P.fy * y * iz, -P.fy * (1.f + y * y),
                          P.fy * x * y, P.fy * x,  0.f};  // :1448-1454
      add_row<0b0111110>(w1, e1, J, acc);
    }
    {
      const float J[7] = {0.f, 0.f, iz, y, -x, 0.f, 1.f};  // :1473-1479
      add_row<0b1011100>(w2, e2, J, acc);
    }
  }
};

// ----------------------------------------------------------------------------------------
// per-call preparation (m3r_gn_begin): packed keyframes and compacted match lists
// ----------------------------------------------------------------------------------------
// Xs, Cs, idx_ii2jj, valid_match and Q do not change over the iterations of one call, and
// neither does the part of the validity test that depends only on them
// (valid_match & q > Q_thresh & c_i > C_thresh & c_j > C_thresh, gn_kernels.cu:914-951).
// So the call first
//   * packs every keyframe point as one 16-byte (x, y, z, c) record: the per-iteration gathers
//     become single aligned 128-bit loads instead of 3+1 scalar ones;
//   * compacts, per (edge, 1024-point segment), the points that pass the test into a dense
//     list of (k, idx, q): the per-iteration kernel then runs with all 32 lanes of a warp on
//     contributing points (a quarter of the lanes idled on invalid points before), reads 12
//     bytes per surviving point instead of 13 per point, and carries no validity logic.
// Order within a segment is preserved and segments are summed in a fixed order, so results
// stay bitwise reproducible.
#ifndef M3R_ACC_THREADS
#define M3R_ACC_THREADS 256
#endif
#ifndef M3R_ACC_CTAS
#define M3R_ACC_CTAS 2
#endif
constexpr int kAccThreads = M3R_ACC_THREADS;  // accumulation kernel: threads per CTA ...
constexpr int kAccCtasPerSM = M3R_ACC_CTAS;    // ... and resident CTAs per SM (register budget)
constexpr int kAccWarps = kAccThreads / 32;
constexpr int kSeg = 2048;                  // points per warp segment
constexpr int kChunk = kSeg * kAccWarps;    // points per work item (one CTA pass)

__global__ void __launch_bounds__(256)
    gn_pack_kernel(const float *__restrict__ Xs, const float *__restrict__ Cs,
                   float4 *__restrict__ P, float *__restrict__ cs_copy, size_t total, int n,
                   const unsigned char *__restrict__ kf_needed) {
  // cs_copy: local copy of Cs for the compaction's gathers when Xs / Cs live on another device
  // (single-process multi-GPU, gn_multi.cuh); null otherwise.  kf_needed: keyframes that an edge
  // of this shard touches; the others are never gathered and are skipped (on 8 ranks a shard of a
  // 64-keyframe graph touches about a third of them: less HBM traffic, and in the single-process
  // mode less NVLink traffic out of the caller's device, which is what bounds its scatter).
  // grid: (chunks of a keyframe's points, keyframes)
  const int kf = blockIdx.y;
  if (!kf_needed[kf]) return;
  const size_t base = static_cast<size_t>(kf) * n;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
    const size_t k = base + p;
    const float c = __ldcs(Cs + k);
    P[k] = make_float4(__ldcs(Xs + 3 * k), __ldcs(Xs + 3 * k + 1), __ldcs(Xs + 3 * k + 2), c);
    if (cs_copy) cs_copy[k] = c;
  }
}

struct CompactArgs {
  const float *Cs;             // [N, n]
  const long long *idx;        // shard rows [ES, n]
  const unsigned char *valid;  // shard rows
  const float *Q;              // shard rows
  const int *pose_i, *pose_j;  // [E] whole graph
  int *ck, *cidx;              // [ES * chunks * kChunk]
  float *cq;
  int *counts;                 // [ES * chunks * kAccWarps]
  int n, edge_begin, num_shard_edges, chunks_per_edge;
  float C_thresh, Q_thresh;
};

__global__ void __launch_bounds__(kAccThreads) gn_compact_kernel(const CompactArgs A) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n = A.n;
  const int items = A.num_shard_edges * A.chunks_per_edge;
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int le = item / A.chunks_per_edge;
    const int ch = item - le * A.chunks_per_edge;
    const int e = A.edge_begin + le;
    const float *Ci = A.Cs + static_cast<size_t>(A.pose_i[e]) * n;
    const float *Cj = A.Cs + static_cast<size_t>(A.pose_j[e]) * n;
    const long long *idx_e = A.idx + static_cast<size_t>(le) * n;
    const unsigned char *val_e = A.valid + static_cast<size_t>(le) * n;
    const float *Q_e = A.Q + static_cast<size_t>(le) * n;
    const size_t slot = (static_cast<size_t>(item) * kAccWarps + warp) * kSeg;
    const int k_begin = ch * kChunk + warp * kSeg;
    int fill = 0;
#ifndef M3R_COMPACT_U
#define M3R_COMPACT_U 4
#endif
    constexpr int U = M3R_COMPACT_U;  // rounds in flight: the kernel is a latency-bound stream + gather
    for (int r0 = 0; r0 < kSeg / 32; r0 += U) {
      float q[U], cj[U];
      long long raw[U];
      unsigned char vm[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int k = k_begin + (r0 + u) * 32 + lane;
        const bool in = k < n;
        q[u] = in ? __ldcs(Q_e + k) : 0.f;
        raw[u] = in ? __ldcs(idx_e + k) : 0LL;
        vm[u] = in ? __ldcs(val_e + k) : static_cast<unsigned char>(0);
        cj[u] = in ? __ldg(Cj + k) : 0.f;
      }
      bool ok[U];
      int id[U];
      float ci[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        ok[u] = (vm[u] != 0) && (q[u] > A.Q_thresh) && (cj[u] > A.C_thresh);
        // clamp keeps a corrupt index from reading outside keyframe i's pointmap
        id[u] = ok[u] ? static_cast<int>(min(max(raw[u], 0LL), static_cast<long long>(n - 1))) : 0;
        ci[u] = ok[u] ? __ldg(Ci + id[u]) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < U; u++) {
        const bool keep = ok[u] && (ci[u] > A.C_thresh);
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (keep) {
          const size_t pos = slot + fill + __popc(m & ((1u << lane) - 1u));
          A.ck[pos] = k_begin + (r0 + u) * 32 + lane;
          A.cidx[pos] = id[u];
          A.cq[pos] = q[u];
        }
        fill += __popc(m);
      }
    }
    if (lane == 0) A.counts[item * kAccWarps + warp] = fill;
  }
}

// ----------------------------------------------------------------------------------------
// accumulation kernel (every iteration)
// ----------------------------------------------------------------------------------------
struct AccArgs {
  const float *Twc;
  const float4 *P;
  const int *ck, *cidx;
  const float *cq;
  const int *counts;
  const int *pose_i;           // [E] whole graph, packed pose index
  const int *pose_j;
  const float *K;              // calib
  double *partials;            // [items, 35]
  const int *flags;
  int n, edge_begin, num_shard_edges, chunks_per_edge;
  GnParams P_;
};

// Latency is hidden with per-warp cp.async rings in shared memory rather than with registers
// (the 35 accumulators leave no room for a deep register pipeline, and a one-round register
// pipeline left the warps waiting on the gathers: ncu source view, profiles/).  Per round of 32
// list entries a warp commits one group holding
//   * the (k, idx, q) entries of round r+6  (three 128-byte rows, 24 lanes x 16 B), and
//   * the two 16-byte point records of round r+3 for every lane (gathers),
// and waits until at most two groups are pending, i.e. until the group of round r-3 has landed:
// the records of round r and the entries needed to issue the gathers of round r+3.
constexpr int kEntSlots = 8;  // entry ring (rounds r .. r+6 are live)
constexpr int kRecSlots = 4;  // record ring (rounds r .. r+3)
constexpr int kGatherLead = 3, kEntryLead = 6;
struct __align__(16) AccWarpSmem {
  int ent[kEntSlots][3][32];     // [slot][k | idx | q bits][lane]
  float4 rec[kRecSlots][2][32];  // [slot][Xj | Xi][lane]
};
constexpr int kAccSmemBytes = kAccWarps * static_cast<int>(sizeof(AccWarpSmem));

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(
                   static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))),
               "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void cp_async16_ca(void *smem_dst, const void *gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(
                   static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))),
               "l"(gsrc)
               : "memory");
}

template <int KIND>
__global__ void __launch_bounds__(kAccThreads, kAccCtasPerSM) gn_accumulate_kernel(const AccArgs A) {
  if (A.flags[0]) return;  // converged earlier: nothing to do
  extern __shared__ __align__(16) unsigned char acc_smem_raw[];
  __shared__ double red[kAccWarps][kNAcc];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  AccWarpSmem &W = reinterpret_cast<AccWarpSmem *>(acc_smem_raw)[warp];
  const int n = A.n;
  const int items = A.num_shard_edges * A.chunks_per_edge;

  GnParams P = A.P_;
  if (KIND == M3R_GN_CALIB) {  // intrinsics live on the device (gn_kernels.cu:1274-1279)
    P.fx = __ldg(A.K + 0), P.fy = __ldg(A.K + 4), P.cx = __ldg(A.K + 2), P.cy = __ldg(A.K + 5);
  }

  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int le = item / A.chunks_per_edge;  // edge within the shard
    const int e = A.edge_begin + le;
    const int ix = A.pose_i[e], jx = A.pose_j[e];
    const Pose Tij = rel_sim3(load_pose(A.Twc, ix), load_pose(A.Twc, jx));
    const float4 *Pi = A.P + static_cast<size_t>(ix) * n;
    const float4 *Pj = A.P + static_cast<size_t>(jx) * n;
    const size_t slot = (static_cast<size_t>(item) * kAccWarps + warp) * kSeg;
    const int cnt = A.counts[item * kAccWarps + warp];
    const int nr = (cnt + 31) >> 5;
    // lanes 0-7 copy the k row of a round, 8-15 the idx row, 16-23 the q row (16 B each)
    const int erow = lane >> 3;
    const int *esrc = (erow == 0 ? A.ck : erow == 1 ? A.cidx : reinterpret_cast<const int *>(A.cq)) +
                      slot + (lane & 7) * 4;

    float acc[kNAcc];
#pragma unroll
    for (int k = 0; k < kNAcc; k++) acc[k] = 0.f;

    for (int t = -kEntryLead; t < nr; t++) {
      asm volatile("cp.async.wait_group %0;" ::"n"(kGatherLead - 1) : "memory");
      __syncwarp();
      if (t + kEntryLead < nr && lane < 24) {
        cp_async16(&W.ent[(t + kEntryLead) & (kEntSlots - 1)][erow][(lane & 7) * 4],
                   esrc + (t + kEntryLead) * 32);
      }
      const int tg = t + kGatherLead;
      if (tg >= 0 && tg < nr) {
        const bool in = tg * 32 + lane < cnt;
        const int kk = in ? W.ent[tg & (kEntSlots - 1)][0][lane] : 0;
        const int ii = in ? W.ent[tg & (kEntSlots - 1)][1][lane] : 0;
        cp_async16_ca(&W.rec[tg & (kRecSlots - 1)][0][lane], Pj + kk);
        cp_async16_ca(&W.rec[tg & (kRecSlots - 1)][1][lane], Pi + ii);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      if (t >= 0 && t * 32 + lane < cnt) {
        const float4 xj = W.rec[t & (kRecSlots - 1)][0][lane];
        const float4 xi = W.rec[t & (kRecSlots - 1)][1][lane];
        const int i1 = W.ent[t & (kEntSlots - 1)][1][lane];
        const float q1 = __int_as_float(W.ent[t & (kEntSlots - 1)][2][lane]);
        const float Xi[3] = {xi.x, xi.y, xi.z}, Xj[3] = {xj.x, xj.y, xj.z};
        Residual<KIND>::point(P, Tij, Xi, Xj, q1, i1, acc);
      }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");

    // warp tree in fp32, then fp64 across warps in a fixed order
#pragma unroll
    for (int k = 0; k < kNAcc; k++) {
      float v = acc[k];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
      if (lane == 0) red[warp][k] = static_cast<double>(v);
    }
    __syncthreads();
    if (threadIdx.x < kNAcc) {
      double s = 0.0;
#pragma unroll
      for (int wq = 0; wq < kAccWarps; wq++) s += red[wq][threadIdx.x];
      A.partials[static_cast<size_t>(item) * kNAcc + threadIdx.x] = s;
    }
    __syncthreads();
  }
}

// ----------------------------------------------------------------------------------------
// accumulation kernel, two list entries per lane in packed fp32 (gn_packed.cuh): the one that runs
// ----------------------------------------------------------------------------------------
#include "gn_packed.cuh"

#ifndef M3R_ACC_PACKED
#define M3R_ACC_PACKED 1
#endif
#ifndef M3R_ACC2_CTAS
#define M3R_ACC2_CTAS 2
#endif
constexpr int kAcc2CtasPerSM = M3R_ACC2_CTAS;
// Same rings as gn_accumulate_kernel, but a round is 64 list entries: lane l takes entries l and l + 32
// of the round as the two halves of its register pairs.
// A round being twice as long, half the ring depth gives the same look-ahead in list entries, and
// two CTAs (16 warps) fit an SM again: with one the kernel was latency-bound and slower than scalar.
#ifndef M3R_ACC2_GATHER_LEAD
#define M3R_ACC2_GATHER_LEAD 2
#endif
constexpr int kGather2Lead = M3R_ACC2_GATHER_LEAD;
// the entries of round t + lead are read when its gathers are issued, so they must have landed by then:
// the wait at the top of round t completes the groups up to round t - lead, hence entry lead = 2 x lead
constexpr int kEntry2Lead = 2 * kGather2Lead;
constexpr int kEnt2Slots = kEntry2Lead + 1, kRec2Slots = kGather2Lead + 1;  // live rounds
struct __align__(16) AccWarpSmem2 {
  int ent[kEnt2Slots][3][64];     // [slot][k | idx | q bits][entry of the round]
  float4 rec[kRec2Slots][2][64];  // [slot][Xj | Xi][entry of the round]
};
constexpr int kAcc2SmemBytes = kAccWarps * static_cast<int>(sizeof(AccWarpSmem2));
static_assert(kSeg % 64 == 0, "a warp segment is a whole number of 64-entry rounds");

template <int KIND>
__global__ void __launch_bounds__(kAccThreads, kAcc2CtasPerSM) gn_accumulate2_kernel(const AccArgs A) {
  if (A.flags[0]) return;  // converged earlier: nothing to do
  extern __shared__ __align__(16) unsigned char acc_smem_raw[];
  __shared__ double red[kAccWarps][kNAcc];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  AccWarpSmem2 &W = reinterpret_cast<AccWarpSmem2 *>(acc_smem_raw)[warp];
  const int n = A.n;
  const int items = A.num_shard_edges * A.chunks_per_edge;

  GnParams P = A.P_;
  if (KIND == M3R_GN_CALIB) {  // intrinsics live on the device (gn_kernels.cu:1274-1279)
    P.fx = __ldg(A.K + 0), P.fy = __ldg(A.K + 4), P.cx = __ldg(A.K + 2), P.cy = __ldg(A.K + 5);
  }

  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int le = item / A.chunks_per_edge;  // edge within the shard
    const int e = A.edge_begin + le;
    const int ix = A.pose_i[e], jx = A.pose_j[e];
    const Pose2 Tij(rel_sim3(load_pose(A.Twc, ix), load_pose(A.Twc, jx)));
    const float4 *Pi = A.P + static_cast<size_t>(ix) * n;
    const float4 *Pj = A.P + static_cast<size_t>(jx) * n;
    const size_t slot = (static_cast<size_t>(item) * kAccWarps + warp) * kSeg;
    const int cnt = A.counts[item * kAccWarps + warp];
    const int nr = (cnt + 63) >> 6;
    // a round's entries are three 256-byte rows: lanes 0-15 copy the k row, 16-31 the idx row, and
    // lanes 0-15 the q row with a second instruction (16 B each)
    const int c4 = (lane & 15) * 4;
    const int *esrc01 = (lane < 16 ? A.ck : A.cidx) + slot + c4;
    const int *esrc2 = reinterpret_cast<const int *>(A.cq) + slot + c4;

    F2 acc[kNAcc];
#pragma unroll
    for (int k = 0; k < kNAcc; k++) acc[k] = F2(0.f);

    for (int t = -kEntry2Lead; t < nr; t++) {
      asm volatile("cp.async.wait_group %0;" ::"n"(kGather2Lead - 1) : "memory");
      __syncwarp();
      if (t + kEntry2Lead < nr) {
        const int sl = (t + kEntry2Lead) % kEnt2Slots;
        cp_async16(&W.ent[sl][lane >> 4][c4], esrc01 + (t + kEntry2Lead) * 64);
        if (lane < 16) cp_async16(&W.ent[sl][2][c4], esrc2 + (t + kEntry2Lead) * 64);
      }
      const int tg = t + kGather2Lead;
      if (tg >= 0 && tg < nr) {
        const int sl = tg % kEnt2Slots, rs = tg % kRec2Slots;
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int x = h * 32 + lane;
          const bool in = tg * 64 + x < cnt;
          const int kk = in ? W.ent[sl][0][x] : 0;
          const int ii = in ? W.ent[sl][1][x] : 0;
          cp_async16_ca(&W.rec[rs][0][x], Pj + kk);
          cp_async16_ca(&W.rec[rs][1][x], Pi + ii);
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      if (t >= 0 && t * 64 + lane < cnt) {  // the first half exists; the second may not (last round)
        const int sl = t % kEnt2Slots, rs = t % kRec2Slots;
        const bool in1 = t * 64 + 32 + lane < cnt;
        const float4 xj0 = W.rec[rs][0][lane], xi0 = W.rec[rs][1][lane];
        float4 xj1 = W.rec[rs][0][32 + lane], xi1 = W.rec[rs][1][32 + lane];
        float q1 = __int_as_float(W.ent[sl][2][32 + lane]);
        if (!in1) {  // harmless inputs and a zero confidence: all weights of that half are zero
          xj1 = make_float4(0.f, 0.f, 1.f, 0.f), xi1 = xj1, q1 = 0.f;
        }
        const F2 Xi[3] = {F2(xi0.x, xi1.x), F2(xi0.y, xi1.y), F2(xi0.z, xi1.z)};
        const F2 Xj[3] = {F2(xj0.x, xj1.x), F2(xj0.y, xj1.y), F2(xj0.z, xj1.z)};
        const F2 q(__int_as_float(W.ent[sl][2][lane]), q1);
        Residual2<KIND>::point(P, Tij, Xi, Xj, q, W.ent[sl][1][lane], in1 ? W.ent[sl][1][32 + lane] : 0, acc);
      }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");

    // the two halves meet, then the warp tree in fp32, then fp64 across warps in a fixed order
#pragma unroll
    for (int k = 0; k < kNAcc; k++) {
      float v = acc[k].v.x + acc[k].v.y;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
      if (lane == 0) red[warp][k] = static_cast<double>(v);
    }
    __syncthreads();
    if (threadIdx.x < kNAcc) {
      double s = 0.0;
#pragma unroll
      for (int wq = 0; wq < kAccWarps; wq++) s += red[wq][threadIdx.x];
      A.partials[static_cast<size_t>(item) * kNAcc + threadIdx.x] = s;
    }
    __syncthreads();
  }
}

// ----------------------------------------------------------------------------------------
// per-edge finalisation: S = sum of chunk partials (fixed order), H_jj = A S A^T, g_j = A s
// ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64)
    gn_edge_finalize_kernel(const double *__restrict__ partials, const float *__restrict__ Twc,
                            const int *__restrict__ pose_i, int edge_begin, int chunks_per_edge,
                            double *__restrict__ edge_H, double *__restrict__ edge_g,
                            const int *__restrict__ flags) {
  if (flags[0]) return;
  __shared__ double S[7][7], sv[7], Am[7][7], AS[7][7];
  const int le = blockIdx.x;
  const int t = threadIdx.x;
  if (t < kNAcc) {
    double s = 0.0;
    for (int c = 0; c < chunks_per_edge; c++)
      s += partials[(static_cast<size_t>(le) * chunks_per_edge + c) * kNAcc + t];
    if (t < kNS) {
      int a = 0;
      while ((a + 1) * (a + 2) / 2 <= t) a++;
      const int c = t - a * (a + 1) / 2;
      S[a][c] = s;
      S[c][a] = s;
    } else {
      sv[t - kNS] = s;
    }
  }
  if (t == 63) {
    // A = matrix of apply_Sim3_adj_inv(T_i) (gn_kernels.cu:277-297):
    //   Y[0:3] = R a / s,  Y[3:6] = R b + t x (R a) / s,  Y[6] = c + t . (R a) / s
    const int ix = pose_i[edge_begin + le];
    const float *Pp = Twc + ix * 8;
    const double tt[3] = {Pp[0], Pp[1], Pp[2]};
    const double q[4] = {Pp[3], Pp[4], Pp[5], Pp[6]};
    const double s_inv = 1.0 / static_cast<double>(Pp[7]);
    for (int r = 0; r < 7; r++)
      for (int c = 0; c < 7; c++) Am[r][c] = 0.0;
    for (int c = 0; c < 3; c++) {
      double ex[3] = {0, 0, 0}, Rc[3];
      ex[c] = 1.0;
      act_so3<double>(q, ex, Rc);  // column c of R
      for (int r = 0; r < 3; r++) {
        Am[r][c] = s_inv * Rc[r];
        Am[3 + r][3 + c] = Rc[r];
      }
      Am[3][c] = s_inv * (tt[1] * Rc[2] - tt[2] * Rc[1]);
      Am[4][c] = s_inv * (tt[2] * Rc[0] - tt[0] * Rc[2]);
      Am[5][c] = s_inv * (tt[0] * Rc[1] - tt[1] * Rc[0]);
      Am[6][c] = s_inv * (tt[0] * Rc[0] + tt[1] * Rc[1] + tt[2] * Rc[2]);
    }
    Am[6][6] = 1.0;
  }
  __syncthreads();
  if (t < 49) {
    const int r = t / 7, c = t % 7;
    double v = 0.0;
    for (int k = 0; k < 7; k++) v += Am[r][k] * S[k][c];
    AS[r][c] = v;
  }
  __syncthreads();
  if (t < 49) {
    const int r = t / 7, c = t % 7;
    if (r >= c) {  // compute the lower triangle, mirror: exactly symmetric blocks
      double v = 0.0;
      for (int k = 0; k < 7; k++) v += AS[r][k] * Am[c][k];
      edge_H[static_cast<size_t>(le) * 49 + r * 7 + c] = v;
      edge_H[static_cast<size_t>(le) * 49 + c * 7 + r] = v;
    }
  } else if (t < 56) {
    const int r = t - 49;
    double v = 0.0;
    for (int k = 0; k < 7; k++) v += Am[r][k] * sv[k];
    edge_g[static_cast<size_t>(le) * 7 + r] = v;
  }
}

// ----------------------------------------------------------------------------------------
// deterministic assembly of the dense system  (SparseBlock::update_lhs/update_rhs, :71-113)
//   block (a,b), a >= b, of H (D x D row-major, D = 7*(N-1)); pose p -> block p-1, pose 0 fixed
//   H[a][a] += H_e for every edge touching a;  H[a][b] -= H_e for every edge joining a and b
//   rhs[a]  += g_e (a = j side) or -g_e (a = i side)
// ----------------------------------------------------------------------------------------
constexpr int kAsmMaxStagedEdges = 4096;  // (io, jo) pairs staged in shared memory: 32 KB

// Which 7x7 blocks of the lower triangle receive anything from this shard's edges does not change
// over the iterations of a call: m3r_gn_begin marks them (gn_mark_blocks_kernel) and lists them in
// row-major order (gn_list_blocks_kernel).  Per iteration the system is zeroed with one memset and
// only the listed blocks get a CTA.  (One CTA per block of the whole NB x NB grid, each staging all
// edge pairs, was 24 us per iteration at 64 keyframes and 65 025 CTAs x 16 KB of L2 reads at 256;
// a 64-keyframe graph touches 309 of its 3 969 blocks.)
__global__ void gn_mark_blocks_kernel(const int *__restrict__ pose_i, const int *__restrict__ pose_j,
                                      int edge_begin, int num_shard_edges, int NB,
                                      unsigned char *__restrict__ blk_flag,
                                      unsigned char *__restrict__ kf_needed) {
  const int le = blockIdx.x * blockDim.x + threadIdx.x;
  if (le >= num_shard_edges) return;
  kf_needed[pose_i[edge_begin + le]] = 1;  // keyframes whose points this shard gathers (gn_pack_kernel)
  kf_needed[pose_j[edge_begin + le]] = 1;
  const int io = pose_i[edge_begin + le] - 1, jo = pose_j[edge_begin + le] - 1;  // -1: the fixed pose
  if (io >= 0) blk_flag[static_cast<size_t>(io) * NB + io] = 1;
  if (jo >= 0) blk_flag[static_cast<size_t>(jo) * NB + jo] = 1;
  if (io >= 0 && jo >= 0 && io != jo) blk_flag[static_cast<size_t>(max(io, jo)) * NB + min(io, jo)] = 1;
}

constexpr int kListThreads = 1024;
__global__ void __launch_bounds__(kListThreads)
    gn_list_blocks_kernel(const unsigned char *__restrict__ blk_flag, int NB, int2 *__restrict__ blk_list,
                          int *__restrict__ blk_count, int capacity) {
  // ordered compaction of the NB * NB flags by one CTA: thread t owns a contiguous range
  __shared__ int warp_tot[kListThreads / 32];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int total = NB * NB;
  const int per = (total + kListThreads - 1) / kListThreads;
  const int k0 = min(t * per, total), k1 = min(k0 + per, total);
  int mine = 0;
  for (int k = k0; k < k1; k++) mine += blk_flag[k] != 0;
  int incl = mine;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += o;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  if (t < 32) {
    int w = warp_tot[t];
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, w, off);
      if (t >= off) w += o;
    }
    warp_tot[t] = w;
  }
  __syncthreads();
  int pos = incl - mine + (warp > 0 ? warp_tot[warp - 1] : 0);
  for (int k = k0; k < k1; k++) {
    if (blk_flag[k] != 0) {
      if (pos < capacity) blk_list[pos] = make_int2(k / NB, k % NB);  // (block row a, block col b), a >= b
      pos++;
    }
  }
  if (t == kListThreads - 1) *blk_count = min(pos, capacity);
}

template <bool STAGED>
__global__ void __launch_bounds__(64)
    gn_assemble_kernel(const double *__restrict__ edge_H, const double *__restrict__ edge_g,
                       const int *__restrict__ pose_i, const int *__restrict__ pose_j,
                       int edge_begin, int num_shard_edges, int NB, double *__restrict__ sys,
                       const int *__restrict__ flags, const int2 *__restrict__ blk_list,
                       const int *__restrict__ blk_count) {
  if (flags[0]) return;
  if (static_cast<int>(blockIdx.x) >= *blk_count) return;  // the grid is an upper bound (edges + poses)
  const int2 ab = blk_list[blockIdx.x];
  const int a = ab.x, b = ab.y;  // block row, block col (a >= b); everything else was zeroed by the memset
  const int t = threadIdx.x;
  const int D = NB * 7;
  const int r = t / 7, c = t % 7;
  // the shard's (pose_i, pose_j) pairs: one coalesced pass into shared memory, so that the scan
  // below costs shared-memory latency per 64-edge tile instead of an L2 round trip
  extern __shared__ int2 ep[];
  if (STAGED) {
    for (int e = t; e < num_shard_edges; e += 64)
      ep[e] = make_int2(pose_i[edge_begin + e] - 1, pose_j[edge_begin + e] - 1);
    __syncthreads();
  }
  // Two passes: the scan over the shard's edges only appends the edges that touch this block to a
  // list in shared memory (edge order preserved); their 7x7 blocks are fetched afterwards, eight per
  // round with the loads of a round issued together and the adds in list order.  (Fetching per
  // 64-edge tile made every tile with a hit wait for its own L2 round trip: 22 us per launch at 64
  // keyframes for whichever block had the longest chain.)
  constexpr int kHitCap = 1024;
  __shared__ int hit[kHitCap];
  __shared__ unsigned masks[2][2];  // double-buffered by tile parity: one barrier per tile
  double accH = 0.0, accg = 0.0;
  auto flush = [&](int nh) {
    __syncthreads();  // the list is complete
    constexpr int kHitUnroll = 8;
    for (int k0 = 0; k0 < nh; k0 += kHitUnroll) {
      double v[kHitUnroll];
#pragma unroll
      for (int u = 0; u < kHitUnroll; u++) {
        v[u] = 0.0;
        if (k0 + u < nh) {
          const int hle = hit[k0 + u] >> 2, hc = hit[k0 + u] & 3;
          if (t < 49) {
            const double x = edge_H[static_cast<size_t>(hle) * 49 + t];
            v[u] = (hc == 3) ? -x : x;
          } else if (t < 56 && a == b) {
            const double g = edge_g[static_cast<size_t>(hle) * 7 + (t - 49)];
            v[u] = (hc == 1) ? -g : g;  // g_i = -g_j
          }
        }
      }
#pragma unroll
      for (int u = 0; u < kHitUnroll; u++) {
        if (k0 + u < nh) {
          if (t < 49) accH += v[u];
          else accg += v[u];
        }
      }
    }
    __syncthreads();  // the list may be overwritten
  };
  int nh = 0;  // uniform over the CTA
  for (int e0 = 0, tile = 0; e0 < num_shard_edges; e0 += 64, tile++) {
    const int le = e0 + t;
    int code = 0;  // 1: +H (diag, i side)  2: +H (diag, j side)  3: -H (off-diagonal)
    if (le < num_shard_edges) {
      int io, jo;
      if (STAGED) io = ep[le].x, jo = ep[le].y;
      else io = pose_i[edge_begin + le] - 1, jo = pose_j[edge_begin + le] - 1;
      if (a == b) {
        if (io == a && jo == a) code = 0;  // self edge: the four blocks cancel exactly
        else if (io == a) code = 1;
        else if (jo == a) code = 2;
      } else if ((io == a && jo == b) || (io == b && jo == a)) {
        code = 3;
      }
    }
    const unsigned m0 = __ballot_sync(0xffffffffu, code != 0);
    if ((t & 31) == 0) masks[tile & 1][t >> 5] = m0;
    __syncthreads();
    const unsigned lo = masks[tile & 1][0], hi = masks[tile & 1][1];
    const int cnt = __popc(lo) + __popc(hi);
    if (cnt == 0) continue;  // no edge of this tile touches the block (the common case)
    if (nh + cnt > kHitCap) {  // uniform: the list is full (a pose with > 1000 edges in this shard)
      flush(nh);
      nh = 0;
    }
    if (code != 0) {
      const int pos = (t < 32) ? __popc(lo & ((1u << t) - 1u))
                               : __popc(lo) + __popc(hi & ((1u << (t - 32)) - 1u));
      hit[nh + pos] = (le << 2) | code;
    }
    nh += cnt;
  }
  flush(nh);
  if (t < 49) sys[static_cast<size_t>(a * 7 + r) * D + (b * 7 + c)] = accH;
  else if (t < 56 && a == b) sys[static_cast<size_t>(D) * D + a * 7 + (t - 49)] = accg;
}

// ----------------------------------------------------------------------------------------
// blocked Cholesky (fp64, lower, in place, row-major with leading dimension D)
// ----------------------------------------------------------------------------------------
constexpr int kPB = 32;  // panel width

#ifdef M3R_TEST_VARIANTS  // the multi-launch solver: a test-only variant (tests/_variants build)
// Factor the kPB x kPB diagonal block at k0 in shared memory (every CTA does it redundantly,
// it is 11k flops), then solve this CTA's 64 rows of the panel: L21 = A21 L11^-T.
// CTA 0 also stores L11.
// L11 goes to `ldiag` (panel-major [D/kPB][kPB][kPB]) rather than back into the matrix: the
// other CTAs of this launch are still reading the unfactored block from there.
__global__ void __launch_bounds__(64)
    chol_panel_kernel(double *__restrict__ Amat, double *__restrict__ ldiag, int D, int k0,
                      int *__restrict__ flags) {
  if (flags[0]) return;
  __shared__ double L[kPB][kPB + 1];
  const int t = threadIdx.x;
  const int pb = min(kPB, D - k0);
  for (int x = t; x < kPB * kPB; x += 64) {
    const int r = x / kPB, c = x % kPB;
    L[r][c] = (r < pb && c <= r) ? Amat[static_cast<size_t>(k0 + r) * D + (k0 + c)] : 0.0;
  }
  __syncthreads();
  __shared__ int bad;
  if (t == 0) bad = 0;
  for (int j = 0; j < pb; j++) {
    __syncthreads();
    if (t == 0) {
      const double d = L[j][j];
      if (!(d > 0.0)) {
        bad = 1;
        L[j][j] = 1.0;  // keep going with finite numbers; the result is discarded
      } else {
        L[j][j] = sqrt(d);
      }
    }
    __syncthreads();
    const double inv = 1.0 / L[j][j];
    for (int r = j + 1 + t; r < pb; r += 64) L[r][j] *= inv;
    __syncthreads();
    // trailing update of the small block: rows r > j, cols j < c <= r
    for (int x = t; x < (pb - j - 1) * (pb - j - 1); x += 64) {
      const int r = j + 1 + x / (pb - j - 1), c = j + 1 + x % (pb - j - 1);
      if (c <= r) L[r][c] -= L[r][j] * L[c][j];
    }
  }
  __syncthreads();
  if (blockIdx.x == 0) {
    if (t == 0 && bad) flags[2] = 1;
    double *Ld = ldiag + static_cast<size_t>(k0 / kPB) * kPB * kPB;
    for (int x = t; x < kPB * kPB; x += 64) Ld[x] = L[x / kPB][x % kPB];
  }
  // rows below the diagonal block: one row per thread, forward substitution against L11
  const int row = k0 + pb + blockIdx.x * 64 + t;
  if (row < D) {
    double x[kPB];
    double *Ar = Amat + static_cast<size_t>(row) * D + k0;
#pragma unroll
    for (int c = 0; c < kPB; c++) x[c] = (c < pb) ? Ar[c] : 0.0;
#pragma unroll
    for (int c = 0; c < kPB; c++) {
      if (c < pb) {
        double s = x[c];
#pragma unroll
        for (int k = 0; k < kPB; k++)
          if (k < c) s -= x[k] * L[c][k];
        x[c] = s / L[c][c];
      }
    }
#pragma unroll
    for (int c = 0; c < kPB; c++)
      if (c < pb) Ar[c] = x[c];
  }
}

// A22 -= L21 L21^T on the lower 64x64 tiles of the trailing matrix.
__global__ void __launch_bounds__(256)
    chol_update_kernel(double *__restrict__ Amat, int D, int k0, const int *__restrict__ flags) {
  if (flags[0]) return;
  const int tr = blockIdx.y, tc = blockIdx.x;
  if (tc > tr) return;
  const int base = k0 + kPB;
  __shared__ double Ra[64][kPB + 1], Rb[64][kPB + 1];
  const int t = threadIdx.x;
  for (int x = t; x < 64 * kPB; x += 256) {
    const int r = x / kPB, c = x % kPB;
    const int ga = base + tr * 64 + r, gb = base + tc * 64 + r;
    Ra[r][c] = (ga < D) ? Amat[static_cast<size_t>(ga) * D + k0 + c] : 0.0;
    Rb[r][c] = (gb < D) ? Amat[static_cast<size_t>(gb) * D + k0 + c] : 0.0;
  }
  __syncthreads();
  const int ty = t / 16, tx = t % 16;  // 16x16 threads, 4x4 outputs each
  double acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j] = 0.0;
#pragma unroll 8
  for (int k = 0; k < kPB; k++) {
    double av[4], bv[4];
#pragma unroll
    for (int i = 0; i < 4; i++) av[i] = Ra[ty + 16 * i][k], bv[i] = Rb[tx + 16 * i][k];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) acc[i][j] += av[i] * bv[j];
  }
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int gr = base + tr * 64 + ty + 16 * i, gc = base + tc * 64 + tx + 16 * j;
      if (gr < D && gc <= gr) Amat[static_cast<size_t>(gr) * D + gc] -= acc[i][j];
    }
}

#endif  // M3R_TEST_VARIANTS

// ----------------------------------------------------------------------------------------
// triangular solves, dx = -x, retraction, stop test (one CTA)
// ----------------------------------------------------------------------------------------
// dx = -x, left retraction, stop test: the tail of every solve kernel (NT threads of one CTA)
template <int NT>
__device__ void finish_step(const int t, const bool fail, const int D, const double *xbuf,
                            float *__restrict__ Twc, float *__restrict__ dx, const int num_poses,
                            const float delta_thresh, int *__restrict__ flags) {
  __shared__ double redn[NT / 32];
  // dx = -x (zeros when the factorisation failed or produced non-finite values)
  __shared__ int nonfinite;
  if (t == 0) nonfinite = 0;
  __syncthreads();
  if (!fail) {
    int bad = 0;
    for (int k = t; k < D; k += NT)
      if (!isfinite(xbuf[k])) bad = 1;
    if (bad) nonfinite = 1;
  }
  __syncthreads();
  const bool zero_step = fail || nonfinite;
  double nrm = 0.0;
  for (int k = t; k < D; k += NT) {
    const float d = zero_step ? 0.0f : static_cast<float>(-xbuf[k]);
    dx[k] = d;
    nrm += static_cast<double>(d) * static_cast<double>(d);
  }
  __syncthreads();

  // left retraction of poses 1..N-1 (pose_retr_kernel :415-453, retrSim3 :392-413)
  for (int k = 1 + t; k < num_poses; k += NT) {
    float xi[7];
#pragma unroll
    for (int m = 0; m < 7; m++) xi[m] = dx[(k - 1) * 7 + m];
    float *P = Twc + k * 8;
    const float tt[3] = {P[0], P[1], P[2]};
    const float q[4] = {P[3], P[4], P[5], P[6]};
    const float s = P[7];
    float dt[3], dq[4], ds;
    exp_sim3(xi, dt, dq, ds);
    float q1[4], t1[3];
    quat_mul(dq, q, q1);
    act_so3<float>(dq, tt, t1);
    P[0] = t1[0] * ds + dt[0];
    P[1] = t1[1] * ds + dt[1];
    P[2] = t1[2] * ds + dt[2];
    P[3] = q1[0], P[4] = q1[1], P[5] = q1[2], P[6] = q1[3];
    P[7] = ds * s;
  }

  // stop test (:1219-1222)
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) nrm += __shfl_xor_sync(0xffffffffu, nrm, off);
  if ((t & 31) == 0) redn[t >> 5] = nrm;
  __syncthreads();
  if (t == 0) {
    double s = 0.0;
    for (int k = 0; k < NT / 32; k++) s += redn[k];
    flags[1] += 1;  // iterations executed
    flags[2] = 0;   // clear the failure flag for the next iteration
    if (static_cast<float>(sqrt(s)) < delta_thresh) flags[0] = 1;
  }
}

#ifdef M3R_TEST_VARIANTS
constexpr int kSolveThreads = 512;

__global__ void __launch_bounds__(kSolveThreads)
    gn_solve_retract_kernel(const double *__restrict__ sys, const double *__restrict__ ldiag,
                            int D, double *__restrict__ xbuf,
                            float *__restrict__ Twc, float *__restrict__ dx, int num_poses,
                            float delta_thresh, int *__restrict__ flags) {
  if (flags[0]) return;
  const int t = threadIdx.x;
  const double *L = sys;
  const double *rhs = sys + static_cast<size_t>(D) * D;
  __shared__ double xs[kPB];
  const bool fail = flags[2] != 0;

  if (!fail) {
    for (int k = t; k < D; k += kSolveThreads) xbuf[k] = rhs[k];
    __syncthreads();
    // forward: L y = b, panel by panel
    for (int k0 = 0; k0 < D; k0 += kPB) {
      const int pb = min(kPB, D - k0);
      const double *Ld = ldiag + static_cast<size_t>(k0 / kPB) * kPB * kPB;
      if (t < 32) {  // one warp solves the diagonal block
        double mine = (t < pb) ? xbuf[k0 + t] : 0.0;
        for (int c = 0; c < pb; c++) {
          double yc = 0.0;
          if (t == c) yc = mine / Ld[c * kPB + c];
          yc = __shfl_sync(0xffffffffu, yc, c);
          if (t == c) mine = yc;
          else if (t > c && t < pb) mine -= Ld[t * kPB + c] * yc;
        }
        if (t < pb) {
          xs[t] = mine;
          xbuf[k0 + t] = mine;
        }
      }
      __syncthreads();
      for (int r = k0 + pb + t; r < D; r += kSolveThreads) {
        double s = xbuf[r];
        const double *Lr = L + static_cast<size_t>(r) * D + k0;
        for (int c = 0; c < pb; c++) s -= Lr[c] * xs[c];
        xbuf[r] = s;
      }
      __syncthreads();
    }
    // backward: L^T x = y
    const int last = ((D - 1) / kPB) * kPB;
    for (int k0 = last; k0 >= 0; k0 -= kPB) {
      const int pb = min(kPB, D - k0);
      const double *Ld = ldiag + static_cast<size_t>(k0 / kPB) * kPB * kPB;
      if (t < 32) {
        double mine = (t < pb) ? xbuf[k0 + t] : 0.0;
        for (int c = pb - 1; c >= 0; c--) {
          double xc = 0.0;
          if (t == c) xc = mine / Ld[c * kPB + c];
          xc = __shfl_sync(0xffffffffu, xc, c);
          if (t == c) mine = xc;
          else if (t < c) mine -= Ld[c * kPB + t] * xc;
        }
        if (t < pb) {
          xs[t] = mine;
          xbuf[k0 + t] = mine;
        }
      }
      __syncthreads();
      for (int r = t; r < k0; r += kSolveThreads) {
        double s = xbuf[r];
        for (int c = 0; c < pb; c++) s -= L[static_cast<size_t>(k0 + c) * D + r] * xs[c];
        xbuf[r] = s;
      }
      __syncthreads();
    }
  }

  finish_step<kSolveThreads>(t, fail, D, xbuf, Twc, dx, num_poses, delta_thresh, flags);
}

#endif  // M3R_TEST_VARIANTS

// ----------------------------------------------------------------------------------------
// small systems: the whole solve in the shared memory of one CTA
// ----------------------------------------------------------------------------------------
// Up to kSmallMaxD unknowns (8 keyframes: the graph of a SLAM run's first seconds) the cooperative
// kernel is all fixed cost (launch, one grid barrier per 32-column panel, the substitutions): 21 us
// at D = 21, 46 us at D = 49.  This kernel takes 17 and 36 us there; at D = 63 the two are equal
// (50 / 48 us) and at D = 105 it takes 105 us against 75 (a plain column algorithm: two barriers
// and one shared-memory round trip per column), hence the limit.  Here the lower triangle is
// copied once into shared memory as packed rows with the right-hand side as row D, so that the
// forward substitution is part of the factorisation (row D ends up as y = L^-1 b); a right-looking
// column Cholesky follows (every thread recomputes the pivot's root and reciprocal instead of
// waiting for one thread's: two barriers per column), then the backward substitution on one warp
// and the common finish (dx = -x, retraction, stop test).  fp64 throughout, like the Eigen LLT of
// the reference (gn_kernels.cu:132-150); a non-positive or non-finite pivot gives dx = 0.
constexpr int kSmallThreads = 512;
constexpr int kSmallMaxD = 49;  // 7 * 7: where it was measured to beat the cooperative kernel

__host__ __device__ constexpr size_t small_smem_doubles(int D) {
  return static_cast<size_t>(D + 1) * (D + 2) / 2 + 3 * static_cast<size_t>(D) + 8;
}

__global__ void __launch_bounds__(kSmallThreads)
    gn_chol_solve_small_kernel(const double *__restrict__ sys, int D, float *__restrict__ Twc,
                               float *__restrict__ dx, int num_poses, float delta_thresh,
                               int *__restrict__ flags) {
  if (flags[0]) return;
  extern __shared__ double small_sm[];
  double *A = small_sm;                                          // packed rows 0..D, row i at i(i+1)/2
  double *col = A + static_cast<size_t>(D + 1) * (D + 2) / 2;    // current column, scaled  [D+1]
  double *invd = col + D + 1;                                    // 1 / L_jj                [D]
  double *xv = invd + D;                                         // solution               [D]
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  constexpr int NW = kSmallThreads / 32;
  auto row = [&](int i) { return A + static_cast<size_t>(i) * (i + 1) / 2; };

  // rows of the lower triangle (the assembled system is block-lower: every element with
  // column <= row is meaningful), then the right-hand side as row D
  for (int i = warp; i <= D; i += NW) {
    const double *src = i < D ? sys + static_cast<size_t>(i) * D : sys + static_cast<size_t>(D) * D;
    const int len = i < D ? i + 1 : D;
    double *dst = row(i);
    for (int j = lane; j < len; j += 32) dst[j] = src[j];
  }
  __syncthreads();

  bool fail = flags[2] != 0;
  for (int j = 0; j < D && !fail; j++) {
    const double d = row(j)[j];  // same value in every thread: the decision below is uniform
    if (!(d > 0.0) || !isfinite(d)) {
      fail = true;
      break;
    }
    const double inv = rsqrt(d);
    for (int i = j + 1 + t; i <= D; i += kSmallThreads) {
      const double v = row(i)[j] * inv;
      row(i)[j] = v;
      col[i] = v;
    }
    if (t == 0) invd[j] = inv;
    __syncthreads();
    // trailing update A[r][c] -= L[r][j] L[c][j] for j < c <= min(r, D-1), j < r <= D
    for (int r = j + 1 + warp; r <= D; r += NW) {
      const double lr = col[r];
      double *ar = row(r);
      const int cmax = r < D ? r : D - 1;
      for (int c = j + 1 + lane; c <= cmax; c += 32) ar[c] -= lr * col[c];
    }
    __syncthreads();
  }

  if (!fail && warp == 0) {
    // backward: L^T x = y, y = row D; x_j = y_j / L_jj = y_j * inv_j^... (L_jj = d_j * inv_j, so
    // 1 / L_jj = inv_j of the already updated pivot d_j: invd holds rsqrt(d_j) = 1 / L_jj)
    double *y = row(D);
    for (int j = D - 1; j >= 0; j--) {
      const double xj = y[j] * invd[j];
      const double *lj = row(j);
      for (int i = lane; i < j; i += 32) y[i] -= lj[i] * xj;
      if (lane == 0) xv[j] = xj;
      __syncwarp();
    }
  }
  __syncthreads();
  finish_step<kSmallThreads>(t, fail, D, xv, Twc, dx, num_poses, delta_thresh, flags);
}

#include "gn_chol_coop.cuh"
#ifdef M3R_TEST_VARIANTS
#include "gn_chol_cluster.cuh"
#endif

// ----------------------------------------------------------------------------------------
// keyframe id -> packed pose index (sorted-unique rank), no host sync
// ----------------------------------------------------------------------------------------
// get_unique_kf_idx + create_inds (gn_kernels.cu:161-170: unique(cat(ii, jj)) sorted, then
// searchsorted) without a sort: keyframe ids index the shared keyframe buffer (at most 512 slots,
// frame.py:221), so one CTA marks the ids that occur in a bitmap in shared memory, an exclusive
// prefix sum over the words' popcounts gives the number of distinct ids below each word, and
//   rank(id) = prefix[id / 32] + popc(word[id / 32] & ((1 << id % 32) - 1)).
// O(E + kRemapMaxId / 32) instead of the O(E^2) pair of kernels this replaces (93 us at E = 492).
// Ids outside [0, kRemapMaxId) take the general kernels below.  flags[3] = kGnErrBadPoseCount when
// the graph holds more distinct ids than the caller's num_poses (the reference would index out of
// bounds; here the rank is clamped for memory safety and the error is reported by
// m3r_gn_iterations_done).
constexpr int kGnErrExchangeTimeout = 1;  // flags[3]
constexpr int kGnErrBadPoseCount = 2;
constexpr int kRemapMaxId = 1 << 16;
constexpr int kRemapWords = kRemapMaxId / 32;
constexpr int kRemapThreads = 1024;

__device__ __forceinline__ void gn_reset_flags(int *flags) {
  flags[0] = 0;  // done
  flags[1] = 0;  // iterations executed
  flags[2] = 0;  // factorisation failed
  flags[3] = 0;  // error code
}

__global__ void __launch_bounds__(kRemapThreads)
    gn_remap_kernel(const long long *__restrict__ ii, const long long *__restrict__ jj, int E,
                    int num_poses, int *__restrict__ pose_i, int *__restrict__ pose_j,
                    int *__restrict__ flags, int *__restrict__ fallback) {
  __shared__ unsigned bits[kRemapWords];
  __shared__ int prefix[kRemapWords];
  __shared__ int warp_tot[kRemapThreads / 32];
  __shared__ int out_of_range;
  const int t = threadIdx.x;
  if (t == 0) {
    gn_reset_flags(flags);
    out_of_range = 0;
  }
  for (int k = t; k < kRemapWords; k += kRemapThreads) bits[k] = 0u;
  __syncthreads();
  for (int k = t; k < 2 * E; k += kRemapThreads) {
    const long long v = (k < E) ? ii[k] : jj[k - E];
    if (v < 0 || v >= kRemapMaxId) out_of_range = 1;
    else atomicOr(&bits[v >> 5], 1u << (v & 31));
  }
  __syncthreads();
  if (out_of_range) {  // uniform
    if (t == 0) *fallback = 1;
    return;
  }
  if (t == 0) *fallback = 0;
  // exclusive scan of popc(bits[]) : kRemapWords / kRemapThreads consecutive words per thread
  constexpr int kPer = kRemapWords / kRemapThreads;
  int loc[kPer], sum = 0;
#pragma unroll
  for (int u = 0; u < kPer; u++) {
    loc[u] = sum;
    sum += __popc(bits[t * kPer + u]);
  }
  int incl = sum;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, off);
    if ((t & 31) >= off) incl += o;
  }
  if ((t & 31) == 31) warp_tot[t >> 5] = incl;
  __syncthreads();
  if (t < 32) {
    int w = warp_tot[t];
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, w, off);
      if (t >= off) w += o;
    }
    warp_tot[t] = w;  // inclusive over warps
  }
  __syncthreads();
  const int base = incl - sum + ((t >> 5) > 0 ? warp_tot[(t >> 5) - 1] : 0);
#pragma unroll
  for (int u = 0; u < kPer; u++) prefix[t * kPer + u] = base + loc[u];
  __syncthreads();
  if (t == 0 && warp_tot[kRemapThreads / 32 - 1] > num_poses) flags[3] = kGnErrBadPoseCount;
  for (int k = t; k < 2 * E; k += kRemapThreads) {
    const long long v = (k < E) ? ii[k] : jj[k - E];
    int rank = prefix[v >> 5] + __popc(bits[v >> 5] & ((1u << (v & 31)) - 1u));
    rank = min(rank, num_poses - 1);  // memory safety when the caller's N is inconsistent
    if (k < E) pose_i[k] = rank;
    else pose_j[k - E] = rank;
  }
}

// general ids (any int64): first occurrences, then rank = number of smaller first occurrences.
// O(E^2); only runs when gn_remap_kernel met an id outside [0, kRemapMaxId).
__global__ void gn_first_occurrence_kernel(const long long *__restrict__ ii,
                                           const long long *__restrict__ jj, int E,
                                           unsigned char *__restrict__ first,
                                           const int *__restrict__ fallback) {
  if (*fallback == 0) return;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= 2 * E) return;
  const long long v = (k < E) ? ii[k] : jj[k - E];
  unsigned char f = 1;
  for (int m = 0; m < k; m++) {
    const long long o = (m < E) ? ii[m] : jj[m - E];
    if (o == v) {
      f = 0;
      break;
    }
  }
  first[k] = f;
}

__global__ void gn_rank_kernel(const long long *__restrict__ ii, const long long *__restrict__ jj,
                               int E, const unsigned char *__restrict__ first, int num_poses,
                               int *__restrict__ pose_i, int *__restrict__ pose_j,
                               int *__restrict__ flags, const int *__restrict__ fallback) {
  if (*fallback == 0) return;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= 2 * E) return;
  const long long v = (k < E) ? ii[k] : jj[k - E];
  int rank = 0;
  for (int m = 0; m < 2 * E; m++) {
    const long long o = (m < E) ? ii[m] : jj[m - E];
    rank += (first[m] && o < v) ? 1 : 0;
  }
  if (rank >= num_poses) flags[3] = kGnErrBadPoseCount;
  rank = min(rank, num_poses - 1);
  if (k < E) pose_i[k] = rank;
  else pose_j[k - E] = rank;
}

}  // namespace m3r

// =========================================================================================
// host side
// =========================================================================================
using namespace m3r;

template <typename T>
struct DevBuf {
  T *p = nullptr;
  size_t cap = 0;
  // grow-only; contents are not preserved
  int reserve(size_t want) {
    if (want <= cap) return M3R_OK;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    const size_t alloc = want + want / 4 + 16;
    cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&p), alloc * sizeof(T));
    if (e != cudaSuccess) {
      set_error("cudaMalloc(%zu bytes): %s", alloc * sizeof(T), cudaGetErrorString(e));
      return M3R_ERR_NO_MEMORY;
    }
    cap = alloc;
    return M3R_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

struct m3r_gn_workspace {
  int device = 0;
  DevBuf<int> pose_i, pose_j;     // [E] packed pose index per directed edge
  DevBuf<unsigned char> first;    // [2E] scratch of the id remap
  DevBuf<float4> packed;          // [N, n] (x, y, z, c) records of every keyframe point
  DevBuf<float> cs_copy;          // [N, n] local copy of Cs (copy_cs: the caller's Cs is on a peer device)
  bool copy_cs = false;
  DevBuf<int> ck, cidx;           // [items * kChunk] compacted match lists (point k, match idx)
  DevBuf<float> cq;               //                  and their confidences
  DevBuf<int> counts;             // [items * kAccWarps] entries per warp segment
  DevBuf<double> partials;        // [items, 35]
  DevBuf<double> edge_H, edge_g;  // [E_shard, 49], [E_shard, 7]
  DevBuf<double> system;          // D*D + D (single-GPU entry points)
  DevBuf<unsigned char> kf_needed; // [N] keyframes this shard's edges touch
  DevBuf<unsigned char> blk_flag; // [NB * NB] 7x7 blocks of the lower triangle this shard's edges touch
  DevBuf<int2> blk_list;          // their (row, col), row-major order; the count lives in flags[5]
  DevBuf<double> ldiag;           // factored diagonal blocks of the Cholesky
  DevBuf<double> xbuf;            // [D]
  int *flags = nullptr;           // [0] done, [1] iterations, [2] factorisation failed, [3] error code
                                  // (kGnErr*), [4] id remap took the general path, [5] non-empty blocks
  int begun_edges = -1;
  int begun_shard_begin = -1, begun_shard_end = -1, begun_points = -1;
};

#include "gn_comm.cuh"
#include "gn_multi.cuh"

extern "C" int m3r_gn_workspace_create(m3r_gn_workspace **out) {
  M3R_REQUIRE(out != nullptr, "gn_workspace_create: null out pointer");
  m3r_gn_workspace *ws = new m3r_gn_workspace();
  cudaError_t e = cudaGetDevice(&ws->device);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&ws->flags), 8 * sizeof(int));
  if (e == cudaSuccess) e = cudaMemset(ws->flags, 0, 8 * sizeof(int));
  if (e != cudaSuccess) {
    delete ws;
    return cuda_fail(e, "gn_workspace_create");
  }
  *out = ws;
  return M3R_OK;
}

extern "C" void m3r_gn_workspace_destroy(m3r_gn_workspace *ws) {
  if (!ws) return;
  ws->pose_i.release(), ws->pose_j.release(), ws->first.release(), ws->partials.release();
  ws->cs_copy.release();
  ws->blk_flag.release(), ws->blk_list.release(), ws->kf_needed.release();
  ws->packed.release(), ws->ck.release(), ws->cidx.release(), ws->cq.release(), ws->counts.release();
  ws->edge_H.release(), ws->edge_g.release(), ws->system.release(), ws->ldiag.release();
  ws->xbuf.release();
  cudaFree(ws->flags);
  delete ws;
}

extern "C" int64_t m3r_gn_system_size(int num_poses) {
  const int64_t D = 7 * static_cast<int64_t>(num_poses > 1 ? num_poses - 1 : 0);
  return D * D + D;
}

static int check_problem(const m3r_gn_problem *p) {
  M3R_REQUIRE(p != nullptr, "gn: null problem");
  M3R_REQUIRE(p->kind == M3R_GN_RAYS || p->kind == M3R_GN_CALIB || p->kind == M3R_GN_POINTS,
              "gn: unknown kind %d", p->kind);
  M3R_REQUIRE(p->num_poses >= 2, "gn: need at least 2 poses (got %d)", p->num_poses);
  M3R_REQUIRE(p->num_points >= 1 && p->num_edges >= 0, "gn: bad sizes");
  M3R_REQUIRE(0 <= p->edge_begin && p->edge_begin <= p->edge_end && p->edge_end <= p->num_edges,
              "gn: bad edge shard [%d,%d) of %d", p->edge_begin, p->edge_end, p->num_edges);
  M3R_REQUIRE(p->Twc && p->Xs && p->Cs, "gn: null pose/point pointer");
  M3R_REQUIRE(p->num_edges == 0 || (p->ii && p->jj), "gn: null edge index pointer");
  M3R_REQUIRE(p->edge_end == p->edge_begin || (p->idx_ii2jj && p->valid_match && p->Q),
              "gn: null edge data pointer");
  M3R_REQUIRE(aligned16(p->Twc), "gn: Twc must be 16-byte aligned");
  if (p->kind == M3R_GN_CALIB) {
    M3R_REQUIRE(p->K != nullptr, "gn: calib needs K");
    M3R_REQUIRE(p->width > 0 && p->height > 0, "gn: calib needs image size");
  }
  return M3R_OK;
}

extern "C" int m3r_gn_begin(m3r_gn_workspace *ws, const m3r_gn_problem *prob, void *stream) {
  M3R_REQUIRE(ws != nullptr, "gn_begin: null workspace");
  int rc = check_problem(prob);
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int E = prob->num_edges;
  const size_t want = static_cast<size_t>(E > 0 ? E : 1);
  if ((rc = ws->pose_i.reserve(want))) return rc;
  if ((rc = ws->pose_j.reserve(want))) return rc;
  if ((rc = ws->first.reserve(2 * want))) return rc;
  ws->begun_edges = E;
  const int M = 2 * E;
  const int blocks = (M + 255) / 256;
  const long long *ii_p = reinterpret_cast<const long long *>(prob->ii);
  const long long *jj_p = reinterpret_cast<const long long *>(prob->jj);
  int *fallback = ws->flags + 4;
  gn_remap_kernel<<<1, kRemapThreads, 0, s>>>(ii_p, jj_p, E, prob->num_poses, ws->pose_i.p, ws->pose_j.p,
                                              ws->flags, fallback);
  M3R_LAUNCH_CHECK("gn_remap_kernel");
  if (blocks > 0) {  // both return at once unless an id fell outside the bitmap
    gn_first_occurrence_kernel<<<blocks, 256, 0, s>>>(ii_p, jj_p, E, ws->first.p, fallback);
    M3R_LAUNCH_CHECK("gn_first_occurrence_kernel");
    gn_rank_kernel<<<blocks, 256, 0, s>>>(ii_p, jj_p, E, ws->first.p, prob->num_poses, ws->pose_i.p,
                                          ws->pose_j.p, ws->flags, fallback);
    M3R_LAUNCH_CHECK("gn_rank_kernel");
  }

  const int n = prob->num_points;
  const int ES = prob->edge_end - prob->edge_begin;
  {  // the 7x7 blocks this shard's edges touch (gn_assemble_kernel)
    const int NB = prob->num_poses - 1;
    const size_t nblk = static_cast<size_t>(NB) * NB;
    const int cap = ES + NB;  // an edge adds at most one off-diagonal block; NB diagonal blocks
    if ((rc = ws->blk_flag.reserve(nblk))) return rc;
    if ((rc = ws->blk_list.reserve(static_cast<size_t>(cap)))) return rc;
    if ((rc = ws->kf_needed.reserve(static_cast<size_t>(prob->num_poses)))) return rc;
    M3R_CUDA_TRY(cudaMemsetAsync(ws->blk_flag.p, 0, nblk, s));
    M3R_CUDA_TRY(cudaMemsetAsync(ws->kf_needed.p, 0, static_cast<size_t>(prob->num_poses), s));
    if (ES > 0) {
      gn_mark_blocks_kernel<<<(ES + 255) / 256, 256, 0, s>>>(ws->pose_i.p, ws->pose_j.p, prob->edge_begin, ES, NB,
                                                             ws->blk_flag.p, ws->kf_needed.p);
      M3R_LAUNCH_CHECK("gn_mark_blocks_kernel");
    }
    gn_list_blocks_kernel<<<1, kListThreads, 0, s>>>(ws->blk_flag.p, NB, ws->blk_list.p, ws->flags + 5, cap);
    M3R_LAUNCH_CHECK("gn_list_blocks_kernel");
  }
  // per-call preparation: packed keyframes, compacted match lists of this shard
  const int chunks = (n + kChunk - 1) / kChunk;
  const size_t items = static_cast<size_t>(ES) * chunks;
  M3R_REQUIRE(items * kAccWarps < (1ull << 31), "gn_begin: shard too large (%zu items)", items);
  const size_t total = static_cast<size_t>(prob->num_poses) * n;
  if ((rc = ws->packed.reserve(total))) return rc;
  if (ws->copy_cs && (rc = ws->cs_copy.reserve(total))) return rc;
  if ((rc = ws->ck.reserve(items * kChunk + 1))) return rc;
  if ((rc = ws->cidx.reserve(items * kChunk + 1))) return rc;
  if ((rc = ws->cq.reserve(items * kChunk + 1))) return rc;
  if ((rc = ws->counts.reserve(items * kAccWarps + 1))) return rc;
  // Everything m3r_gn_linearize and m3r_gn_solve_retract use is reserved HERE, before any kernel of
  // the call is enqueued: growing a buffer means cudaFree + cudaMalloc, which synchronise the
  // device, and once the multi-GPU exchange is running a kernel may be spinning on a peer's flag
  // whose signal this host thread has not enqueued yet (ranks that share a process: the single-
  // process multi-GPU driver and the loop-back tests) -- the host would block on the device, the
  // device on the host, until the exchange times out.  After m3r_gn_begin the stages never allocate.
  {
    const int D = 7 * (prob->num_poses - 1);
    if ((rc = ws->partials.reserve(items * kNAcc + 1))) return rc;
    if ((rc = ws->edge_H.reserve(static_cast<size_t>(ES > 0 ? ES : 1) * 49))) return rc;
    if ((rc = ws->edge_g.reserve(static_cast<size_t>(ES > 0 ? ES : 1) * 7))) return rc;
    if ((rc = ws->xbuf.reserve(static_cast<size_t>(D)))) return rc;
    if ((rc = ws->ldiag.reserve(static_cast<size_t>((D + kPB - 1) / kPB) * kLdStride + 2 * kPB * kPB))) return rc;
  }
  {
    // 4 points per thread in flight; keyframes in grid.y (at most 512 keyframe slots, frame.py:221)
    M3R_REQUIRE(prob->num_poses <= 65535, "gn_begin: %d poses exceed grid.y", prob->num_poses);
    const int chunks = (n + 1023) / 1024;
    gn_pack_kernel<<<dim3(chunks, prob->num_poses), 256, 0, s>>>(
        prob->Xs, prob->Cs, ws->packed.p, ws->copy_cs ? ws->cs_copy.p : nullptr, total, n, ws->kf_needed.p);
    M3R_LAUNCH_CHECK("gn_pack_kernel");
  }
  if (items > 0) {
    CompactArgs c{};
    c.Cs = ws->copy_cs ? ws->cs_copy.p : prob->Cs;
    c.idx = reinterpret_cast<const long long *>(prob->idx_ii2jj);
    c.valid = prob->valid_match;
    c.Q = prob->Q;
    c.pose_i = ws->pose_i.p, c.pose_j = ws->pose_j.p;
    c.ck = ws->ck.p, c.cidx = ws->cidx.p, c.cq = ws->cq.p, c.counts = ws->counts.p;
    c.n = n, c.edge_begin = prob->edge_begin, c.num_shard_edges = ES, c.chunks_per_edge = chunks;
    c.C_thresh = prob->C_thresh, c.Q_thresh = prob->Q_thresh;
    const size_t cap = static_cast<size_t>(kNumSMs) * 8;
    gn_compact_kernel<<<static_cast<int>(items < cap ? items : cap), kAccThreads, 0, s>>>(c);
    M3R_LAUNCH_CHECK("gn_compact_kernel");
  }
  ws->begun_shard_begin = prob->edge_begin;
  ws->begun_shard_end = prob->edge_end;
  ws->begun_points = n;
  return M3R_OK;
}

static GnParams make_params(const m3r_gn_problem *p) {
  GnParams P{};
  P.C_thresh = p->C_thresh;
  P.Q_thresh = p->Q_thresh;
  if (p->kind == M3R_GN_RAYS) {
    P.sigma_a_inv = 1.0f / p->sigma_ray;
    P.sigma_b_inv = 1.0f / p->sigma_dist;
  } else if (p->kind == M3R_GN_CALIB) {
    P.sigma_a_inv = 1.0f / p->sigma_pixel;
    P.sigma_b_inv = 1.0f / p->sigma_depth;
    P.width = p->width;
    P.height = p->height;
    // (u > pixel_border) && (u < width - 1 - pixel_border), ints converted to float (:1382-1383)
    P.u_lo = static_cast<float>(p->pixel_border);
    P.u_hi = static_cast<float>(p->width - 1 - p->pixel_border);
    P.v_lo = static_cast<float>(p->pixel_border);
    P.v_hi = static_cast<float>(p->height - 1 - p->pixel_border);
    P.z_eps = p->z_eps;
  } else {
    P.sigma_a_inv = 1.0f / p->sigma_point;
  }
  return P;
}

extern "C" int m3r_gn_linearize(m3r_gn_workspace *ws, const m3r_gn_problem *prob, double *system,
                                void *stream) {
  M3R_REQUIRE(ws != nullptr && system != nullptr, "gn_linearize: null workspace/system");
  int rc = check_problem(prob);
  if (rc) return rc;
  M3R_REQUIRE(prob->num_edges == ws->begun_edges, "gn_linearize: call m3r_gn_begin first");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int n = prob->num_points;
  const int ES = prob->edge_end - prob->edge_begin;
  const int chunks = (n + kChunk - 1) / kChunk;
  const size_t items = static_cast<size_t>(ES) * chunks;
  if ((rc = ws->partials.reserve(items * kNAcc + 1))) return rc;
  if ((rc = ws->edge_H.reserve(static_cast<size_t>(ES > 0 ? ES : 1) * 49))) return rc;
  if ((rc = ws->edge_g.reserve(static_cast<size_t>(ES > 0 ? ES : 1) * 7))) return rc;

  M3R_REQUIRE(prob->edge_begin == ws->begun_shard_begin && prob->edge_end == ws->begun_shard_end &&
                  n == ws->begun_points,
              "gn_linearize: shard differs from the one given to m3r_gn_begin");
  if (ES > 0) {
    AccArgs a{};
    a.Twc = prob->Twc;
    a.P = ws->packed.p;
    a.ck = ws->ck.p, a.cidx = ws->cidx.p, a.cq = ws->cq.p, a.counts = ws->counts.p;
    a.pose_i = ws->pose_i.p, a.pose_j = ws->pose_j.p;
    a.K = prob->K;
    a.partials = ws->partials.p;
    a.flags = ws->flags;
    a.n = n, a.edge_begin = prob->edge_begin, a.num_shard_edges = ES, a.chunks_per_edge = chunks;
    a.P_ = make_params(prob);
#if M3R_ACC_PACKED
    const size_t cap = static_cast<size_t>(kNumSMs) * kAcc2CtasPerSM;
#else
    const size_t cap = static_cast<size_t>(kNumSMs) * kAccCtasPerSM;
#endif
    const int grid = static_cast<int>(items < cap ? items : cap);
    static bool attr_set_dev[64] = {false};  // per device; benign race: idempotent
    bool &attr_set = attr_set_dev[ws->device & 63];
    if (!attr_set) {
      M3R_CUDA_TRY(cudaFuncSetAttribute(gn_accumulate2_kernel<M3R_GN_RAYS>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kAcc2SmemBytes));
      M3R_CUDA_TRY(cudaFuncSetAttribute(gn_accumulate2_kernel<M3R_GN_CALIB>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kAcc2SmemBytes));
      M3R_CUDA_TRY(cudaFuncSetAttribute(gn_accumulate2_kernel<M3R_GN_POINTS>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kAcc2SmemBytes));
      M3R_CUDA_TRY(cudaFuncSetAttribute(gn_accumulate_kernel<M3R_GN_RAYS>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kAccSmemBytes));
      M3R_CUDA_TRY(cudaFuncSetAttribute(gn_accumulate_kernel<M3R_GN_CALIB>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kAccSmemBytes));
      M3R_CUDA_TRY(cudaFuncSetAttribute(gn_accumulate_kernel<M3R_GN_POINTS>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kAccSmemBytes));
      attr_set = true;
    }
#if M3R_ACC_PACKED
    if (prob->kind == M3R_GN_RAYS)
      gn_accumulate2_kernel<M3R_GN_RAYS><<<grid, kAccThreads, kAcc2SmemBytes, s>>>(a);
    else if (prob->kind == M3R_GN_CALIB)
      gn_accumulate2_kernel<M3R_GN_CALIB><<<grid, kAccThreads, kAcc2SmemBytes, s>>>(a);
    else
      gn_accumulate2_kernel<M3R_GN_POINTS><<<grid, kAccThreads, kAcc2SmemBytes, s>>>(a);
#else
    if (prob->kind == M3R_GN_RAYS)
      gn_accumulate_kernel<M3R_GN_RAYS><<<grid, kAccThreads, kAccSmemBytes, s>>>(a);
    else if (prob->kind == M3R_GN_CALIB)
      gn_accumulate_kernel<M3R_GN_CALIB><<<grid, kAccThreads, kAccSmemBytes, s>>>(a);
    else
      gn_accumulate_kernel<M3R_GN_POINTS><<<grid, kAccThreads, kAccSmemBytes, s>>>(a);
#endif
    M3R_LAUNCH_CHECK("gn_accumulate_kernel");

    gn_edge_finalize_kernel<<<ES, 64, 0, s>>>(ws->partials.p, prob->Twc, ws->pose_i.p,
                                              prob->edge_begin, chunks, ws->edge_H.p,
                                              ws->edge_g.p, ws->flags);
    M3R_LAUNCH_CHECK("gn_edge_finalize_kernel");
  }
  const int NB = prob->num_poses - 1;
  // zero the whole system (also what the exchange and the solver expect of blocks nobody writes),
  // then one CTA per block that has edges; ES + NB bounds their number
  M3R_CUDA_TRY(cudaMemsetAsync(system, 0, static_cast<size_t>(m3r_gn_system_size(prob->num_poses)) * sizeof(double), s));
  const int asm_grid = ES + NB;
  if (ES <= kAsmMaxStagedEdges)
    gn_assemble_kernel<true><<<asm_grid, 64, static_cast<size_t>(ES) * sizeof(int2), s>>>(
        ws->edge_H.p, ws->edge_g.p, ws->pose_i.p, ws->pose_j.p, prob->edge_begin, ES, NB, system, ws->flags,
        ws->blk_list.p, ws->flags + 5);
  else
    gn_assemble_kernel<false><<<asm_grid, 64, 0, s>>>(ws->edge_H.p, ws->edge_g.p, ws->pose_i.p, ws->pose_j.p,
                                                     prob->edge_begin, ES, NB, system, ws->flags, ws->blk_list.p,
                                                     ws->flags + 5);
  M3R_LAUNCH_CHECK("gn_assemble_kernel");
  return M3R_OK;
}

extern "C" int m3r_gn_solve_retract(m3r_gn_workspace *ws, const m3r_gn_problem *prob, double *system,
                                    float *dx, float delta_thresh, void *stream) {
  M3R_REQUIRE(ws != nullptr && system != nullptr && dx != nullptr, "gn_solve_retract: null pointer");
  int rc = check_problem(prob);
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int D = 7 * (prob->num_poses - 1);
  if ((rc = ws->xbuf.reserve(static_cast<size_t>(D)))) return rc;
  if ((rc = ws->ldiag.reserve(static_cast<size_t>((D + kPB - 1) / kPB) * kLdStride + 2 * kPB * kPB))) return rc;
#ifdef M3R_TEST_VARIANTS
  // variants build only (tests/_variants): M3R_GN_SOLVER = cluster (D <= 448; measured slower: 430
  // vs 340 us at D = 441, profiles/r1_d_cluster_solver_phases.txt) | multi (one launch per panel
  // phase) | coop (the cooperative kernel at every size).  The product library has no switch.
  static const int solver_choice = [] {
    const char *e = getenv("M3R_GN_SOLVER");
    if (e && strcmp(e, "multi") == 0) return 2;
    if (e && strcmp(e, "cluster") == 0) return 0;
    if (e && strcmp(e, "coop") == 0) return 3;  // the cooperative kernel at every size
    return 1;
  }();
#else
  constexpr int solver_choice = 1;
#endif
  if (solver_choice == 1 && D <= kSmallMaxD) {
    // small system: one CTA, matrix in shared memory
    const size_t smem = small_smem_doubles(D) * sizeof(double);
    static bool small_attr_dev[64] = {false};
    if (!small_attr_dev[ws->device & 63]) {
      M3R_CUDA_TRY(cudaFuncSetAttribute(gn_chol_solve_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        static_cast<int>(small_smem_doubles(kSmallMaxD) * sizeof(double))));
      small_attr_dev[ws->device & 63] = true;
    }
    gn_chol_solve_small_kernel<<<1, kSmallThreads, smem, s>>>(system, D, prob->Twc, dx, prob->num_poses,
                                                              delta_thresh, ws->flags);
    M3R_LAUNCH_CHECK("gn_chol_solve_small_kernel");
    return M3R_OK;
  }
#ifdef M3R_TEST_VARIANTS
  if (solver_choice == 0 && D <= kClMaxD) {
    // one 8-CTA cluster, matrix resident in distributed shared memory
    if ((rc = ws->xbuf.reserve(static_cast<size_t>(kClMaxD)))) return rc;
    if ((rc = ws->ldiag.reserve(static_cast<size_t>(kClMaxD)))) return rc;
    static bool attr_set_dev[64] = {false};
    bool &attr_set = attr_set_dev[ws->device & 63];
    if (!attr_set) {
      M3R_CUDA_TRY(cudaFuncSetAttribute(gn_chol_solve_cluster_kernel,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kClSmemBytes));
      attr_set = true;
    }
    gn_chol_solve_cluster_kernel<<<kClSize, kClThreads, kClSmemBytes, s>>>(
        system, ws->ldiag.p, D, ws->xbuf.p, prob->Twc, dx, prob->num_poses, delta_thresh, ws->flags);
    M3R_LAUNCH_CHECK("gn_chol_solve_cluster_kernel");
    return M3R_OK;
  }
  const bool use_multi = solver_choice == 2;
  if (!use_multi)
#endif
  {

    // one cooperative launch: Cholesky + substitutions + retraction + stop test
    static bool coop_attr_dev[64] = {false};
    if (!coop_attr_dev[ws->device & 63]) {
      M3R_CUDA_TRY(cudaFuncSetAttribute(gn_chol_solve_coop_kernel,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kCoopSmemBytes));
      coop_attr_dev[ws->device & 63] = true;
    }
    static int max_coresident_dev[64] = {0};  // per device (SM counts differ under MIG / mixed parts)
    int &max_coresident = max_coresident_dev[ws->device & 63];
    if (max_coresident == 0) {
      int sms = 0, per_sm = 0;
      const int dev = ws->device;
      M3R_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      M3R_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gn_chol_solve_coop_kernel,
                                                                kCoopThreads, kCoopSmemBytes));
      max_coresident = sms * (per_sm > 0 ? 1 : 0);  // one CTA per SM keeps the barrier cheap
      if (max_coresident <= 0) {
        set_error("gn_solve_retract: cooperative kernel cannot be resident");
        return M3R_ERR_CUDA;
      }
    }
    const int m0 = D - kPB > 0 ? D - kPB : 0;
    const int T0 = (m0 + 31) / 32;  // 32x32 tiles of the first trailing update
    const int row_ctas = (m0 + kTrsmRows - 1) / kTrsmRows;
    int want = T0 * (T0 + 1) / 2 + row_ctas;  // update CTAs + panel CTAs (look-ahead)
    if (want < 1) want = 1;
    const int grid = want < max_coresident ? want : max_coresident;
    double *sys_p = system, *ld_p = ws->ldiag.p, *x_p = ws->xbuf.p;
    float *twc_p = prob->Twc, *dx_p = dx;
    int D_ = D, np = prob->num_poses;
    float dth = delta_thresh;
    int *fl = ws->flags;
    void *args[] = {&sys_p, &ld_p, &D_, &x_p, &twc_p, &dx_p, &np, &dth, &fl};
    M3R_CUDA_TRY(cudaLaunchCooperativeKernel(reinterpret_cast<void *>(gn_chol_solve_coop_kernel),
                                             dim3(grid), dim3(kCoopThreads), args, kCoopSmemBytes, s));
    g_launch_count.fetch_add(1, std::memory_order_relaxed);
    return M3R_OK;
  }
#ifdef M3R_TEST_VARIANTS
  for (int k0 = 0; k0 < D; k0 += kPB) {
    const int rows_below = D - k0 - kPB;
    const int pblocks = rows_below > 0 ? (rows_below + 63) / 64 : 1;
    chol_panel_kernel<<<pblocks, 64, 0, s>>>(system, ws->ldiag.p, D, k0, ws->flags);
    M3R_LAUNCH_CHECK("chol_panel_kernel");
    if (rows_below > 0) {
      const int tiles = (rows_below + 63) / 64;
      chol_update_kernel<<<dim3(tiles, tiles), 256, 0, s>>>(system, D, k0, ws->flags);
      M3R_LAUNCH_CHECK("chol_update_kernel");
    }
  }
  gn_solve_retract_kernel<<<1, kSolveThreads, 0, s>>>(system, ws->ldiag.p, D, ws->xbuf.p,
                                                      prob->Twc, dx, prob->num_poses,
                                                      delta_thresh, ws->flags);
  M3R_LAUNCH_CHECK("gn_solve_retract_kernel");
  return M3R_OK;
#endif
}

#ifdef M3R_COOP_PROF
// experiments (tools/kbench.py): cumulative per-phase cycle counts of the cooperative solver
extern "C" int m3r_debug_coop_profile(long long *out16) {
  long long zero[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  M3R_CUDA_TRY(cudaDeviceSynchronize());
  M3R_CUDA_TRY(cudaMemcpyFromSymbol(out16, g_coop_prof, sizeof(zero)));
  M3R_CUDA_TRY(cudaMemcpyToSymbol(g_coop_prof, zero, sizeof(zero)));
  return M3R_OK;
}
extern "C" int m3r_debug_coop_stamps(long long *out16) {
  long long zero[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  M3R_CUDA_TRY(cudaDeviceSynchronize());
  M3R_CUDA_TRY(cudaMemcpyFromSymbol(out16, g_coop_stamp, sizeof(zero)));
  M3R_CUDA_TRY(cudaMemcpyToSymbol(g_coop_stamp, zero, sizeof(zero)));
  return M3R_OK;
}
#endif
#ifdef M3R_TEST_VARIANTS
extern "C" int m3r_debug_cluster_profile(long long *out9) {
  long long zero[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  M3R_CUDA_TRY(cudaDeviceSynchronize());
  M3R_CUDA_TRY(cudaMemcpyFromSymbol(out9, g_cl_prof, sizeof(zero)));
  M3R_CUDA_TRY(cudaMemcpyToSymbol(g_cl_prof, zero, sizeof(zero)));
  return M3R_OK;
}
#endif

extern "C" int m3r_gn_edge_blocks(m3r_gn_workspace *ws, const double **Hjj, const double **gj) {
  M3R_REQUIRE(ws && Hjj && gj, "gn_edge_blocks: null pointer");
  *Hjj = ws->edge_H.p;
  *gj = ws->edge_g.p;
  return M3R_OK;
}

extern "C" int m3r_gn_iterations_done(m3r_gn_workspace *ws, void *stream, int *iters) {
  M3R_REQUIRE(ws && iters, "gn_iterations_done: null pointer");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int host[4] = {0, 0, 0, 0};
  M3R_CUDA_TRY(cudaMemcpyAsync(host, ws->flags, sizeof(host), cudaMemcpyDeviceToHost, s));
  M3R_CUDA_TRY(cudaStreamSynchronize(s));
  *iters = host[1];
  if (host[3] == kGnErrExchangeTimeout) {
    set_error("gn: a multi-GPU exchange timed out waiting for a peer (the solve was stopped)");
    return M3R_ERR_CUDA;
  }
  if (host[3] == kGnErrBadPoseCount) {
    set_error("gn: ii/jj hold more distinct keyframe ids than num_poses");
    return M3R_ERR_INVALID_ARGUMENT;
  }
  return M3R_OK;
}

static int run_gn(m3r_gn_workspace *ws, m3r_gn_problem *p, float *dx, int max_iter,
                  float delta_thresh, void *stream) {
  M3R_REQUIRE(ws != nullptr, "gauss_newton: null workspace");
  M3R_REQUIRE(dx != nullptr, "gauss_newton: null dx");
  p->edge_begin = 0;
  p->edge_end = p->num_edges;
  int rc = m3r_gn_begin(ws, p, stream);
  if (rc) return rc;
  const size_t need = static_cast<size_t>(m3r_gn_system_size(p->num_poses));
  if ((rc = ws->system.reserve(need))) return rc;
  for (int it = 0; it < max_iter; it++) {
    if ((rc = m3r_gn_linearize(ws, p, ws->system.p, stream))) return rc;
    if ((rc = m3r_gn_solve_retract(ws, p, ws->system.p, dx, delta_thresh, stream))) return rc;
  }
  return M3R_OK;
}

extern "C" int m3r_gauss_newton_rays(m3r_gn_workspace *ws, float *Twc, const float *Xs,
                                     const float *Cs, const int64_t *ii, const int64_t *jj,
                                     const int64_t *idx_ii2jj, const uint8_t *valid_match,
                                     const float *Q, float *dx, int num_poses, int num_points,
                                     int num_edges, float sigma_ray, float sigma_dist,
                                     float C_thresh, float Q_thresh, int max_iter,
                                     float delta_thresh, void *stream) {
  m3r_gn_problem p{};
  p.kind = M3R_GN_RAYS;
  p.num_poses = num_poses, p.num_points = num_points, p.num_edges = num_edges;
  p.Twc = Twc, p.Xs = Xs, p.Cs = Cs, p.ii = ii, p.jj = jj;
  p.idx_ii2jj = idx_ii2jj, p.valid_match = valid_match, p.Q = Q;
  p.sigma_ray = sigma_ray, p.sigma_dist = sigma_dist;
  p.C_thresh = C_thresh, p.Q_thresh = Q_thresh;
  return run_gn(ws, &p, dx, max_iter, delta_thresh, stream);
}

extern "C" int m3r_gauss_newton_calib(m3r_gn_workspace *ws, float *Twc, const float *Xs,
                                      const float *Cs, const float *K, const int64_t *ii,
                                      const int64_t *jj, const int64_t *idx_ii2jj,
                                      const uint8_t *valid_match, const float *Q, float *dx,
                                      int num_poses, int num_points, int num_edges, int height,
                                      int width, int pixel_border, float z_eps, float sigma_pixel,
                                      float sigma_depth, float C_thresh, float Q_thresh,
                                      int max_iter, float delta_thresh, void *stream) {
  m3r_gn_problem p{};
  p.kind = M3R_GN_CALIB;
  p.num_poses = num_poses, p.num_points = num_points, p.num_edges = num_edges;
  p.Twc = Twc, p.Xs = Xs, p.Cs = Cs, p.ii = ii, p.jj = jj;
  p.idx_ii2jj = idx_ii2jj, p.valid_match = valid_match, p.Q = Q;
  p.K = K, p.height = height, p.width = width, p.pixel_border = pixel_border;
  p.z_eps = z_eps, p.sigma_pixel = sigma_pixel, p.sigma_depth = sigma_depth;
  p.C_thresh = C_thresh, p.Q_thresh = Q_thresh;
  return run_gn(ws, &p, dx, max_iter, delta_thresh, stream);
}

extern "C" int m3r_gauss_newton_points(m3r_gn_workspace *ws, float *Twc, const float *Xs,
                                       const float *Cs, const int64_t *ii, const int64_t *jj,
                                       const int64_t *idx_ii2jj, const uint8_t *valid_match,
                                       const float *Q, float *dx, int num_poses, int num_points,
                                       int num_edges, float sigma_point, float C_thresh,
                                       float Q_thresh, int max_iter, float delta_thresh,
                                       void *stream) {
  m3r_gn_problem p{};
  p.kind = M3R_GN_POINTS;
  p.num_poses = num_poses, p.num_points = num_points, p.num_edges = num_edges;
  p.Twc = Twc, p.Xs = Xs, p.Cs = Cs, p.ii = ii, p.jj = jj;
  p.idx_ii2jj = idx_ii2jj, p.valid_match = valid_match, p.Q = Q;
  p.sigma_point = sigma_point;
  p.C_thresh = C_thresh, p.Q_thresh = Q_thresh;
  return run_gn(ws, &p, dx, max_iter, delta_thresh, stream);
}
