// Issue-rate microbenchmark for the instruction mixes the matching / GN kernels are made of
// (sm_100a).  Prints warp-instructions per clock per SM for each mix, 32 warps per SM, 8
// independent chains per thread.   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench_pipes ubench_pipes.cu
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CHAINS 8
#define ITERS 2048

__device__ __forceinline__ uint32_t hadd2(uint32_t a, uint32_t b) {
  uint32_t d; asm volatile("add.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t hmul2(uint32_t a, uint32_t b) {
  uint32_t d; asm volatile("mul.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t hfma2(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d; asm volatile("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t s) {
  uint32_t d; asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s)); return d; }
__device__ __forceinline__ float ffma(float a, float b, float c) {
  float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ uint32_t lop(uint32_t a, uint32_t b) {
  uint32_t d; asm volatile("xor.b32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t iadd(uint32_t a, uint32_t b) {
  uint32_t d; asm volatile("add.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }

template <int MIX>
__global__ void __launch_bounds__(256) k(uint32_t *out, uint32_t seed, int iters) {
  uint32_t x[CHAINS], y[CHAINS];
  float f[CHAINS];
  unsigned long long g[CHAINS];
  const uint32_t s2 = seed ^ 0x3c003c00u;
  for (int c = 0; c < CHAINS; c++) {
    x[c] = seed + c + threadIdx.x; y[c] = seed * 3 + c; f[c] = (float)(seed + c); g[c] = seed + c;
  }
  const float fa = (float)seed, fb = 1.0f + (float)seed;
  const unsigned long long ga = seed, gb = seed + 1;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < CHAINS; c++) {
      if (MIX == 0) x[c] = hfma2(x[c], s2, y[c]);
      if (MIX == 1) x[c] = hadd2(x[c], y[c]);
      if (MIX == 2) x[c] = hmul2(x[c], s2);
      if (MIX == 3) x[c] = prmt(x[c], y[c], 0x5410);
      if (MIX == 4) f[c] = ffma(f[c], fa, fb);
      if (MIX == 5) g[c] = ffma2(g[c], ga, gb);
      if (MIX == 6) { x[c] = hadd2(x[c], y[c]); y[c] = hmul2(y[c], s2); }                 // hadd2 + hmul2
      if (MIX == 7) { x[c] = hadd2(x[c], y[c]); y[c] = prmt(y[c], x[c], 0x7632); }        // hadd2 + prmt
      if (MIX == 8) { x[c] = hfma2(x[c], s2, y[c]); y[c] = prmt(y[c], x[c], 0x7632); }    // hfma2 + prmt
      if (MIX == 9) { x[c] = hfma2(x[c], s2, y[c]); y[c] = hadd2(y[c], s2); }             // hfma2 + hadd2
      if (MIX == 10) { x[c] = hadd2(x[c], hmul2(y[c], s2)); y[c] = prmt(y[c], x[c], 0x5410); }  // refine inner
      if (MIX == 11) { f[c] = ffma(f[c], fa, fb); x[c] = lop(x[c], y[c]); }               // ffma + lop
      if (MIX == 12) { f[c] = ffma(f[c], fa, fb); x[c] = hfma2(x[c], s2, y[c]); }         // ffma + hfma2
      if (MIX == 13) x[c] = lop(x[c], y[c]);
      if (MIX == 14) x[c] = iadd(x[c], y[c]);
      if (MIX == 15) { g[c] = ffma2(g[c], ga, gb); x[c] = lop(x[c], y[c]); }
      if (MIX == 16) { x[c] = hfma2(x[c], s2, y[c]); y[c] = hmul2(y[c], s2); }            // hfma2 + hmul2
      if (MIX == 17) { x[c] = hfma2(x[c], s2, y[c]); f[c] = __int_as_float(lop(__float_as_int(f[c]), y[c])); }
    }
  }
  uint32_t r = 0;
  for (int c = 0; c < CHAINS; c++) r ^= x[c] ^ y[c] ^ __float_as_int(f[c]) ^ (uint32_t)g[c] ^ (uint32_t)(g[c] >> 32);
  if (r == 0x12345678u) out[threadIdx.x] = r;
}

static const char *names[] = {"hfma2", "hadd2", "hmul2", "prmt", "ffma", "ffma2(x2 flops)", "hadd2+hmul2", "hadd2+prmt",
                              "hfma2+prmt", "hfma2+hadd2", "hadd2+hmul2+prmt", "ffma+lop", "ffma+hfma2", "lop", "iadd",
                              "ffma2+lop", "hfma2+hmul2", "hfma2+lop"};
static const int per_iter[] = {1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 3, 2, 2, 1, 1, 2, 2, 2};

template <int MIX>
void run(uint32_t *d, int sms, double mhz) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  k<MIX><<<sms * 4, 256>>>(d, 1, 16);
  cudaEventRecord(a);
  k<MIX><<<sms * 4, 256>>>(d, 1, ITERS);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double warp_instr = (double)sms * 4 * 8 * ITERS * CHAINS * per_iter[MIX];
  const double clk = ms * 1e-3 * mhz * 1e6;
  printf("%-20s %8.3f ms  %6.3f warp-instr/clk/SM  (%.3f per SMSP)\n", names[MIX], ms, warp_instr / clk / sms,
         warp_instr / clk / sms / 4);
}

int main() {
  int sms, khz; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double mhz = khz / 1e3;
  printf("SMs %d, clock %.0f MHz (rates assume the max clock)\n", sms, mhz);
  uint32_t *d; cudaMalloc(&d, 4096);
  run<0>(d, sms, mhz); run<1>(d, sms, mhz); run<2>(d, sms, mhz); run<3>(d, sms, mhz); run<4>(d, sms, mhz);
  run<5>(d, sms, mhz); run<6>(d, sms, mhz); run<7>(d, sms, mhz); run<8>(d, sms, mhz); run<9>(d, sms, mhz);
  run<10>(d, sms, mhz); run<11>(d, sms, mhz); run<12>(d, sms, mhz); run<13>(d, sms, mhz); run<14>(d, sms, mhz);
  run<15>(d, sms, mhz); run<16>(d, sms, mhz); run<17>(d, sms, mhz);
  return 0;
}
