// Micro-benchmark: fp64 throughput of one SM on B200, DFMA with fresh operands (as in the solver's update
// loops) against mma.sync.m8n8k4.f64 (DMMA).  One CTA of 256 threads (2 warps per scheduler).
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// 4x4 register tile, two fresh operands per DFMA (the pattern of coop_update_tile / the diagonal update)
__global__ void __launch_bounds__(256, 1) k_dfma_tile(const double *in, double *out, long long *cyc, int n) {
  __shared__ double sa[64][33], sb[64][33];
  const int t = threadIdx.x;
  for (int x = t; x < 64 * 32; x += 256) sa[x / 32][x % 32] = in[x], sb[x / 32][x % 32] = in[x + 2048];
  __syncthreads();
  const int ty = t / 16, tx = t % 16;
  double acc[4][4] = {};
  const long long t0 = clock64();
  for (int rep = 0; rep < n; rep++) {
#pragma unroll 8
    for (int k = 0; k < 32; k++) {
      double av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; i++) av[i] = sa[ty + 16 * i][k], bv[i] = sb[tx + 16 * i][k];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = fma(av[i], bv[j], acc[i][j]);
    }
  }
  const long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) s += acc[i][j];
  out[t] = s;
  if (t == 0) cyc[0] = t1 - t0;
}

// the same 64x64x32 product per repetition with DMMA: warp w owns rows 8w..8w+7 (x 64 columns = 8 tiles of
// 8x8), k in steps of 4.  Fragment layout of m8n8k4: a = A[lane/4][lane%4], b = B[lane%4][lane/4] (col
// operand), c0,c1 = C[lane/4][2*(lane%4) + 0,1].
__global__ void __launch_bounds__(256, 1) k_dmma(const double *in, double *out, long long *cyc, int n) {
  __shared__ double sa[64][33], sb[64][33];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  for (int x = t; x < 64 * 32; x += 256) sa[x / 32][x % 32] = in[x], sb[x / 32][x % 32] = in[x + 2048];
  __syncthreads();
  double c[8][2] = {};
  const int g = lane >> 2, q = lane & 3;
  const long long t0 = clock64();
  for (int rep = 0; rep < n; rep++) {
#pragma unroll
    for (int k = 0; k < 32; k += 4) {
      const double a = sa[8 * warp + g][k + q];
      double b[8];
#pragma unroll
      for (int j = 0; j < 8; j++) b[j] = sb[8 * j + g][k + q];
#pragma unroll
      for (int j = 0; j < 8; j++) dmma(c[j][0], c[j][1], a, b[j]);
    }
  }
  const long long t1 = clock64();
  double s = 0;
  for (int j = 0; j < 8; j++) s += c[j][0] + c[j][1];
  out[t] = s;
  if (t == 0) cyc[0] = t1 - t0;
}

// check the fragment layout: C = A B^T for small integers
__global__ void k_check(double *out) {
  const int lane = threadIdx.x, g = lane >> 2, q = lane & 3;
  double c0 = 0, c1 = 0;
  const double a = 1 + g * 4 + q;          // A[g][q]
  const double b = 100 + g * 10 + q;       // Bt[g][q]  (B[k = q][n = g])
  dmma(c0, c1, a, b);
  out[g * 8 + 2 * q] = c0, out[g * 8 + 2 * q + 1] = c1;
}

int main() {
  double *in, *out;
  long long *cyc, h;
  cudaMalloc(&in, 8192 * 8);
  cudaMalloc(&out, 4096 * 8);
  cudaMalloc(&cyc, 64);
  double hin[4096];
  for (int i = 0; i < 4096; i++) hin[i] = 1.0 / (1 + i % 37);
  cudaMemcpy(in, hin, sizeof(hin), cudaMemcpyHostToDevice);
  const int n = 200;
  const double fma_per_rep = 64.0 * 64 * 32;
  for (int it = 0; it < 2; it++) {
    k_dfma_tile<<<1, 256>>>(in, out, cyc, n);
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("DFMA 4x4 register tiles from shared memory: %.1f cycles per 64x64x32 tile, %.1f FMA/clk/SM\n", double(h) / n,
           fma_per_rep * n / h);
    k_dmma<<<1, 256>>>(in, out, cyc, n);
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("DMMA m8n8k4 fragments from shared memory:   %.1f cycles per 64x64x32 tile, %.1f FMA/clk/SM  %s\n", double(h) / n,
           fma_per_rep * n / h, cudaGetErrorString(cudaGetLastError()));
  }
  k_check<<<1, 32>>>(out);
  double ho[64];
  cudaMemcpy(ho, out, sizeof(ho), cudaMemcpyDeviceToHost);
  double err = 0;
  for (int m = 0; m < 8; m++)
    for (int nn = 0; nn < 8; nn++) {
      double ref = 0;
      for (int k = 0; k < 4; k++) ref += (1 + m * 4 + k) * (100.0 + nn * 10 + k);
      err = fmax(err, fabs(ref - ho[m * 8 + nn]));
    }
  printf("fragment layout check (C[m][n] = sum_k A[m][k] Bt[n][k]): max error %.1f  %s\n", err,
         cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
