// Micro-benchmark: one warp factors a 32x32 SPD block (fp64).  Variants:
//  A  rows in registers, fully unrolled, rsqrt (2 Newton) on the pivot chain   (current kernel)
//  B  block in shared memory, rolled loops, rsqrt on the chain
//  C  rows in registers, LDL^T chain (reciprocal, 2 Newton), sqrt scaling after the chain
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
  double r = static_cast<double>(__frcp_rn(__double2float_rn(d)));
  double e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  return r;
}
__device__ __forceinline__ void cholA(double a[kPB], int lane, double *dinv) {
  const unsigned full = 0xffffffffu;
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    const double d = __shfl_sync(full, a[j], j);
    const double r = chain_rsqrt(d);
    if (lane == 0) dinv[j] = r;
    const double l = a[j] * r;
    a[j] = l;
#pragma unroll
    for (int c = j + 1; c < kPB; c++) {
      const double lc = __shfl_sync(full, l, c);
      a[c] -= l * lc;
    }
  }
}
__device__ __forceinline__ void cholC(double a[kPB], int lane, double *dinv) {
  const unsigned full = 0xffffffffu;
  double dj[kPB];
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    const double d = __shfl_sync(full, a[j], j);
    dj[j] = d;
    const double r = chain_rcp(d);
    const double t = a[j] * r;  // unit-lower column of LDL^T
#pragma unroll
    for (int c = j + 1; c < kPB; c++) {
      const double ac = __shfl_sync(full, a[j], c);  // unscaled A[c][j]
      a[c] -= t * ac;
    }
  }
  // scale: L[lane][j] = a[j] / sqrt(d_j)
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    const double r = chain_rsqrt(dj[j]);
    a[j] *= r;
    if (lane == 0) dinv[j] = r;
  }
}
__device__ __noinline__ void cholB(double (*S)[kPB + 1], int lane, double *dinv) {
  for (int j = 0; j < kPB; j++) {
    const double d = S[j][j];
    const double r = chain_rsqrt(d);
    const double l = S[lane][j] * r;
    __syncwarp();
    S[lane][j] = l;
    if (lane == 0) dinv[j] = r;
    __syncwarp();
#pragma unroll 4
    for (int c = j + 1; c < kPB; c++) S[lane][c] -= l * S[c][j];
    __syncwarp();
  }
}
// D: rows in registers, column j broadcast through shared memory, one round trip per pivot
__device__ __forceinline__ void cholD(double a[kPB], int lane, double *col, double *dinv) {
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    __syncwarp();
    col[lane] = a[j];
    __syncwarp();
    const double d = col[j];
    const double r = chain_rsqrt(d);
    if (lane == 0) dinv[j] = r;
    const double tj = a[j] * (r * r);
    a[j] *= r;
#pragma unroll
    for (int c = j + 1; c < kPB; c++) a[c] -= tj * col[c];
  }
}
template <int V>
__global__ void k(const double *A, double *out, long long *cyc, int reps) {
  __shared__ double S[kPB][kPB + 1];
  __shared__ double dinv[kPB];
  const int lane = threadIdx.x;
  for (int rep = 0; rep < reps; rep++) {
    double a[kPB];
#pragma unroll
    for (int c = 0; c < kPB; c++) a[c] = A[lane * kPB + c];
    if (V == 1) {
      for (int c = 0; c < kPB; c++) S[lane][c] = a[c];
      __syncwarp();
    }
    const long long t0 = clock64();
    if (V == 0) cholA(a, lane, dinv);
    if (V == 1) cholB(S, lane, dinv);
    if (V == 2) cholC(a, lane, dinv);
    if (V == 3) cholD(a, lane, S[0], dinv);
    const long long t1 = clock64();
    if (lane == 0) cyc[rep] = t1 - t0;
    if (V == 1) {
      __syncwarp();
      for (int c = 0; c < kPB; c++) a[c] = S[lane][c];
    }
#pragma unroll
    for (int c = 0; c < kPB; c++) out[lane * kPB + c] = (c <= lane) ? a[c] : 0.0;
  }
}
int main() {
  double h[kPB * kPB], o[kPB * kPB];
  for (int r = 0; r < kPB; r++)
    for (int c = 0; c < kPB; c++) h[r * kPB + c] = (r == c ? 40.0 : 0.0) + 1.0 / (1 + r + c) + 0.01 * ((r * 7 + c * 3) % 5 + (c * 7 + r * 3) % 5);
  double *dA, *dO; long long *dc, hc[8];
  cudaMalloc(&dA, sizeof(h)); cudaMalloc(&dO, sizeof(h)); cudaMalloc(&dc, 64);
  cudaMemcpy(dA, h, sizeof(h), cudaMemcpyHostToDevice);
  double ref[kPB * kPB];
  const char *names[] = {"A registers+rsqrt chain", "B shared memory rolled", "C registers LDL^T chain",
                         "D registers + smem broadcast"};
  for (int v = 0; v < 4; v++) {
    if (v == 0) k<0><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 1) k<1><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 2) k<2><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 3) k<3><<<1, 32>>>(dA, dO, dc, 4);
    cudaMemcpy(hc, dc, 32, cudaMemcpyDeviceToHost);
    cudaMemcpy(o, dO, sizeof(o), cudaMemcpyDeviceToHost);
    if (v == 0) for (int i = 0; i < kPB * kPB; i++) ref[i] = o[i];
    double err = 0; for (int i = 0; i < kPB * kPB; i++) err = fmax(err, fabs(o[i] - ref[i]));
    printf("%-28s cycles per call: %lld (cold) %lld %lld %lld   max |L - L_A| %.2e  %s\n", names[v], hc[0], hc[1], hc[2], hc[3], err,
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
