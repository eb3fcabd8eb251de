// Micro-benchmark: fp64 FMA latency/throughput and 64-bit shuffle cost on this GPU.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_lat(double *out, long long *cyc, int n) {
  double a = out[0], b = out[1], c = out[2];
  long long t0 = clock64();
  for (int i = 0; i < n; i++) { a = fma(a, b, c); a = fma(a, b, c); a = fma(a, b, c); a = fma(a, b, c); }
  long long t1 = clock64();
  out[3 + threadIdx.x + blockIdx.x * blockDim.x] = a;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_tput(double *out, long long *cyc, int n) {
  double a[8];
  for (int k = 0; k < 8; k++) a[k] = out[k] + threadIdx.x;
  double b = out[1], c = out[2];
  long long t0 = clock64();
  for (int i = 0; i < n; i++) {
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = fma(a[k], b, c);
  }
  long long t1 = clock64();
  double s = 0;
  for (int k = 0; k < 8; k++) s += a[k];
  out[16 + threadIdx.x + blockIdx.x * blockDim.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_shfl(double *out, long long *cyc, int n) {
  double a = out[0] + threadIdx.x;
  long long t0 = clock64();
  for (int i = 0; i < n; i++) {
    a += __shfl_sync(0xffffffffu, a, (i * 7) & 31);
    a += __shfl_sync(0xffffffffu, a, (i * 5) & 31);
  }
  long long t1 = clock64();
  out[16 + threadIdx.x + blockIdx.x * blockDim.x] = a;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_f32lat(float *out, long long *cyc, int n) {
  float a = out[0], b = out[1], c = out[2];
  long long t0 = clock64();
  for (int i = 0; i < n; i++) { a = fmaf(a, b, c); a = fmaf(a, b, c); a = fmaf(a, b, c); a = fmaf(a, b, c); }
  long long t1 = clock64();
  out[3 + threadIdx.x] = a;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
  double *d; long long *c, h; float *f;
  cudaMalloc(&d, 1 << 20); cudaMemset(d, 0, 1 << 20); cudaMalloc(&c, 64); cudaMalloc(&f, 1 << 16); cudaMemset(f, 0, 1 << 16);
  const int n = 4096;
  for (int threads : {32, 128, 256, 1024}) {
    k_lat<<<1, threads>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("DFMA dependent chain, %4d thr/CTA: %.1f cyc per DFMA\n", threads, double(h) / (4.0 * n));
    k_tput<<<1, threads>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("DFMA 8 independent,  %4d thr/CTA: %.2f cyc per warp-DFMA per SMSP (warps/SMSP=%d)\n", threads,
           double(h) / (8.0 * n) / (threads >= 128 ? threads / 128.0 : 1.0), threads >= 128 ? threads / 128 : 1);
  }
  k_shfl<<<1, 32>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
  printf("dependent (shfl64 + DADD): %.1f cyc per pair\n", double(h) / (2.0 * n));
  k_f32lat<<<1, 32>>>(f, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
  printf("FFMA dependent chain: %.1f cyc\n", double(h) / (4.0 * n));
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
