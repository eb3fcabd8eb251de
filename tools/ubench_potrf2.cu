// Micro-benchmark 2: why does warp_chol32 take 386 cycles per pivot inside gn_chol_solve_coop_kernel
// when the same loop runs at 176 stand-alone (tools/ubench_potrf.cu, variant E)?  Adds the kernel's
// context one factor at a time:
//   P0  the product's loop with its special-case branches, one warp per CTA
//   P1  P0 as warp 0 of a 256-thread CTA (the other warps wait at a barrier): real WARPSYNC
//   P2  P1 + four follower warps polling the progress counter (wait_pivot, groups of 4)
//   P3  P2 + the followers doing TRSM-like work (LDS of L11 + DFMA) per published pivot
//   Q*  the same four with the branch-free loop (sticky "abnormal" flag instead of branches)
#include <cstdio>
#include <cuda_runtime.h>
constexpr int kPB = 32;
__device__ __forceinline__ double chain_rsqrt(double d) {
  double r = static_cast<double>(rsqrtf(__double2float_rn(d)));
  double e = fma(-d * r, r, 1.0);
  r = fma(0.5 * r, e, r);
  e = fma(-d * r, r, 1.0);
  r = fma(0.5 * r, e, r);
  return r;
}
__device__ __forceinline__ double chain_rcp(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  double e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  return r;
}
__device__ __noinline__ double rsqrt_wide_range(double d) { return rsqrt(d); }

// the product's loop (gn_chol_coop.cuh)
__device__ __forceinline__ void chol_branchy(double a[kPB], int lane, double *col, double *dinv, int *bad,
                                             double (*L11)[kPB + 1], volatile int *prog) {
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    __syncwarp();
    if (j > 0) *prog = j;
    col[lane] = a[j];
    __syncwarp();
    const double d = col[j];
    double r, rc;
    if (!(d > 0.0)) {
      if (lane == 0) *bad = 1;
      r = 1.0, rc = 1.0;
    } else if (d < 1e-30 || d > 1e30) {
      r = rsqrt_wide_range(d);
      rc = r * r;
    } else {
      rc = chain_rcp(d);
      r = chain_rsqrt(d);
    }
    const double tj = a[j] * rc;
#pragma unroll
    for (int c = j + 1; c < kPB; c++) a[c] -= tj * col[c];
    dinv[j] = r;
    a[j] *= r;
    L11[lane][j] = (lane >= j) ? a[j] : 0.0;
  }
  __syncwarp();
  *prog = kPB;
}
// branch-free: an abnormal pivot is replaced by 1 and remembered
__device__ __forceinline__ bool chol_flat(double a[kPB], int lane, double *col, double *dinv,
                                          double (*L11)[kPB + 1], volatile int *prog) {
  bool abn = false;
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    __syncwarp();
    if (j > 0) *prog = j;
    col[lane] = a[j];
    __syncwarp();
    const double d0 = col[j];
    const bool normal = d0 > 1e-30 && d0 < 1e30;
    abn |= !normal;
    const double d = normal ? d0 : 1.0;
    const double rc = chain_rcp(d);
    const double r = chain_rsqrt(d);
    const double tj = a[j] * rc;
#pragma unroll
    for (int c = j + 1; c < kPB; c++) a[c] -= tj * col[c];
    dinv[j] = r;
    a[j] *= r;
    L11[lane][j] = (lane >= j) ? a[j] : 0.0;
  }
  __syncwarp();
  *prog = kPB;
  return abn;
}

// branch-free AND barrier-free: inside this straight-line loop the warp cannot diverge, the LSU
// handles one warp's shared-memory instructions in order, so a compiler fence is all that is needed
// between the column store and the loads (the warp is brought together once, in front of the loop)
#define WARP_FENCE() asm volatile("" ::: "memory")
__device__ __forceinline__ bool chol_flat_nobar(double a[kPB], int lane, volatile double *col, double *dinv,
                                                double (*L11)[kPB + 1], volatile int *prog) {
  bool abn = false;
  __syncwarp();
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    WARP_FENCE();
    if (j > 0) *prog = j;
    col[lane] = a[j];
    WARP_FENCE();
    const double d0 = col[j];
    const bool normal = d0 > 1e-30 && d0 < 1e30;
    abn |= !normal;
    const double d = normal ? d0 : 1.0;
    const double rc = chain_rcp(d);
    const double r = chain_rsqrt(d);
    const double tj = a[j] * rc;
#pragma unroll
    for (int c = j + 1; c < kPB; c++) a[c] -= tj * col[c];
    dinv[j] = r;
    a[j] *= r;
    L11[lane][j] = (lane >= j) ? a[j] : 0.0;
  }
  WARP_FENCE();
  __syncwarp();
  *prog = kPB;
  return abn;
}
template <int FENCE>  // 0: compiler fence only, 1: __threadfence_block (MEMBAR.SC.CTA), 2: fence.acq_rel.cta
__device__ __forceinline__ void wait_pivot(volatile int *prog, int c) {
  while (*prog <= c) __nanosleep(100);
  if (FENCE == 1) __threadfence_block();
  else if (FENCE == 2) asm volatile("fence.acq_rel.cta;" ::: "memory");
  else asm volatile("" ::: "memory");
}

struct Smem {
  double L11[kPB][kPB + 1];
  double col[kPB], dinv[kPB];
  int bad, prog;
};

template <int FLAT, int MODE, int FENCE = 1>  // FLAT 0 product loop, 1 branch-free, 2 branch- and barrier-free; MODE 0: 32 threads; 1: 256 threads, others idle; 2: pollers; 3: pollers working
__global__ void __launch_bounds__(256, 1) k(const double *A, double *out, long long *cyc, int reps) {
  __shared__ Smem S;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  for (int rep = 0; rep < reps; rep++) {
    if (t == 0) S.prog = 0, S.bad = 0;
    __syncthreads();
    if (warp == 0) {
      double a[kPB];
#pragma unroll
      for (int c = 0; c < kPB; c++) a[c] = A[lane * kPB + c];
      const long long t0 = clock64();
      if (FLAT == 2) {
        if (chol_flat_nobar(a, lane, S.col, S.dinv, S.L11, &S.prog)) S.bad = 1;
      } else if (FLAT == 1) {
        if (chol_flat(a, lane, S.col, S.dinv, S.L11, &S.prog)) S.bad = 1;
      } else {
        chol_branchy(a, lane, S.col, S.dinv, &S.bad, S.L11, &S.prog);
      }
      const long long t1 = clock64();
      if (lane == 0) cyc[rep] = t1 - t0, cyc[8] = t1;
#pragma unroll
      for (int c = 0; c < kPB; c++) out[lane * kPB + c] = (c <= lane) ? a[c] : 0.0;
    } else if (MODE >= 2 && (warp == 1 || warp == 2 || warp == 3 || warp == 5)) {
      double x[kPB];
#pragma unroll
      for (int c = 0; c < kPB; c++) x[c] = A[lane * kPB + c] + warp;
#pragma unroll
      for (int c = 0; c < kPB; c++) {
        if ((c % 4) == 0) wait_pivot<FENCE>(&S.prog, c + 3);
        if (MODE == 3) {
          const double xc = x[c] * S.dinv[c];
          x[c] = xc;
#pragma unroll
          for (int q = c + 1; q < kPB; q++) x[q] -= xc * S.L11[q][c];
        }
      }
      if (lane == 0) cyc[8 + warp] = clock64();
      if (MODE == 3) {
        double s = 0;
#pragma unroll
        for (int c = 0; c < kPB; c++) s += x[c];
        out[kPB * kPB + t] = s;
      }
    }
    __syncthreads();
  }
}

template <int FLAT, int MODE>
void run(const char *name, const double *dA, double *dO, long long *dc) {
  long long hc[8];
  k<FLAT, MODE><<<1, MODE == 0 ? 32 : 256>>>(dA, dO, dc, 6);
  cudaMemcpy(hc, dc, 48, cudaMemcpyDeviceToHost);
  printf("%-44s cycles: %lld (cold) %lld %lld %lld %lld  (%.0f per pivot)  %s\n", name, hc[0], hc[2], hc[3], hc[4], hc[5],
         hc[5] / 32.0, cudaGetErrorString(cudaGetLastError()));
}
int main() {
  double h[kPB * kPB];
  for (int r = 0; r < kPB; r++)
    for (int c = 0; c < kPB; c++)
      h[r * kPB + c] = (r == c ? 40.0 : 0.0) + 1.0 / (1 + r + c) + 0.01 * ((r * 7 + c * 3) % 5 + (c * 7 + r * 3) % 5);
  double *dA, *dO;
  long long *dc;
  cudaMalloc(&dA, sizeof(h));
  cudaMalloc(&dO, sizeof(h) + 4096);
  cudaMalloc(&dc, 256);
  cudaMemcpy(dA, h, sizeof(h), cudaMemcpyHostToDevice);
  run<0, 0>("P0 product loop, one warp", dA, dO, dc);
  run<0, 1>("P1 product loop, warp 0 of 256 threads", dA, dO, dc);
  run<0, 2>("P2 + 4 warps polling the counter", dA, dO, dc);
  run<0, 3>("P3 + the pollers doing TRSM-like work", dA, dO, dc);
  run<1, 0>("Q0 branch-free loop, one warp", dA, dO, dc);
  run<1, 1>("Q1 branch-free, warp 0 of 256 threads", dA, dO, dc);
  run<1, 2>("Q2 + 4 warps polling", dA, dO, dc);
  run<1, 3>("Q3 + pollers working", dA, dO, dc);
  run<2, 0>("R0 branch- and barrier-free, one warp", dA, dO, dc);
  run<2, 1>("R1 warp 0 of 256 threads", dA, dO, dc);
  run<2, 2>("R2 + 4 warps polling", dA, dO, dc);
  run<2, 3>("R3 + pollers working", dA, dO, dc);
  {
    long long hc[16];
    k<2, 3, 1><<<1, 256>>>(dA, dO, dc, 6);
    cudaMemcpy(hc, dc, 128, cudaMemcpyDeviceToHost);
    printf("followers behind the factorisation, MEMBAR.SC.CTA per group: %lld %lld %lld %lld cycles (potrf %lld)\n", hc[9] - hc[8], hc[10] - hc[8], hc[11] - hc[8], hc[13] - hc[8], hc[5]);
    k<2, 3, 2><<<1, 256>>>(dA, dO, dc, 6);
    cudaMemcpy(hc, dc, 128, cudaMemcpyDeviceToHost);
    printf("followers behind the factorisation, fence.acq_rel.cta:      %lld %lld %lld %lld cycles (potrf %lld)\n", hc[9] - hc[8], hc[10] - hc[8], hc[11] - hc[8], hc[13] - hc[8], hc[5]);
    k<2, 3, 0><<<1, 256>>>(dA, dO, dc, 6);
    cudaMemcpy(hc, dc, 128, cudaMemcpyDeviceToHost);
    printf("followers behind the factorisation, compiler fence only:    %lld %lld %lld %lld cycles (potrf %lld)\n", hc[9] - hc[8], hc[10] - hc[8], hc[11] - hc[8], hc[13] - hc[8], hc[5]);
  }
  double o[kPB * kPB], ref[kPB * kPB];
  k<0, 1><<<1, 256>>>(dA, dO, dc, 1);
  cudaMemcpy(ref, dO, sizeof(ref), cudaMemcpyDeviceToHost);
  k<2, 3><<<1, 256>>>(dA, dO, dc, 1);
  cudaMemcpy(o, dO, sizeof(o), cudaMemcpyDeviceToHost);
  double err = 0;
  for (int i = 0; i < kPB * kPB; i++) err = fmax(err, fabs(o[i] - ref[i]));
  printf("R3 vs P1 factor: max abs diff %.3e\n", err);
  return 0;
}
