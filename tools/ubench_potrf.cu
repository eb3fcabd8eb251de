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

// ---- round 2: the product's loop (E) and software-pipelined forms of it.  All publish L11 / dinv
// as the kernel's warp_chol32 does (the followers' progress counter included).
__device__ __forceinline__ double rcp64(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  double e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  return r;
}
// E: as shipped (gn_chol_coop.cuh warp_chol32)
__device__ __forceinline__ void cholE(double a[kPB], int lane, double *col, double *dinv,
                                      double (*L11)[kPB + 1], volatile int *prog) {
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    __syncwarp();
    if (j > 0) *prog = j;
    col[lane] = a[j];
    __syncwarp();
    const double d = col[j];
    const double rc = rcp64(d);
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
}
// F: next column published after ONE update; the rest, the root and the stores behind it
__device__ __forceinline__ void cholF(double a[kPB], int lane, double *col, double *dinv,
                                      double (*L11)[kPB + 1], volatile int *prog) {
  col[lane] = a[0];
  __syncwarp();
  double d = col[0];
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    const double *cb = col + (j % 3) * kPB;
    double *cn = col + ((j + 1) % 3) * kPB;
    const double rc = rcp64(d);
    const double tj = a[j] * rc;
    if (j + 1 < kPB) {
      a[j + 1] -= tj * cb[j + 1];
      cn[lane] = a[j + 1];
    }
    __syncwarp();
    if (j > 0) *prog = j;
    const double dnext = (j + 1 < kPB) ? cn[j + 1] : 1.0;
#pragma unroll
    for (int c = j + 2; c < kPB; c++) a[c] -= tj * cb[c];
    const double r = chain_rsqrt(d);
    dinv[j] = r;
    a[j] *= r;
    L11[lane][j] = (lane >= j) ? a[j] : 0.0;
    d = dnext;
  }
  __syncwarp();
  *prog = kPB;
}
// G: F with the root and the stores of column j delayed into round j + 1, beside its reciprocal
__device__ __forceinline__ void cholG(double a[kPB], int lane, double *col, double *dinv,
                                      double (*L11)[kPB + 1], volatile int *prog) {
  col[lane] = a[0];
  __syncwarp();
  double d = col[0], dprev = 1.0;
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    const double *cb = col + (j % 3) * kPB;
    double *cn = col + ((j + 1) % 3) * kPB;
    const double rc = rcp64(d);
    const double rp = chain_rsqrt(dprev);  // root of the PREVIOUS pivot: independent of rc, same block
    const double tj = a[j] * rc;
    if (j + 1 < kPB) {
      a[j + 1] -= tj * cb[j + 1];
      cn[lane] = a[j + 1];
    }
    __syncwarp();
    if (j > 1) *prog = j - 1;
    const double dnext = (j + 1 < kPB) ? cn[j + 1] : 1.0;
#pragma unroll
    for (int c = j + 2; c < kPB; c++) a[c] -= tj * cb[c];
    if (j > 0) {
      dinv[j - 1] = rp;
      a[j - 1] *= rp;
      L11[lane][j - 1] = (lane >= j - 1) ? a[j - 1] : 0.0;
    }
    dprev = d;
    d = dnext;
  }
  const double rp = chain_rsqrt(dprev);
  dinv[kPB - 1] = rp;
  a[kPB - 1] *= rp;
  L11[lane][kPB - 1] = (lane >= kPB - 1) ? a[kPB - 1] : 0.0;
  __syncwarp();
  *prog = kPB;
}
// I: G, and the next pivot travels by shuffle (two SHFL) instead of STS -> barrier -> LDS; the column
// buffer still carries the off-chain entries
__device__ __forceinline__ void cholI(double a[kPB], int lane, double *col, double *dinv,
                                      double (*L11)[kPB + 1], volatile int *prog) {
  const unsigned full = 0xffffffffu;
  col[lane] = a[0];
  __syncwarp();
  double d = col[0], dprev = 1.0;
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    const double *cb = col + (j % 3) * kPB;
    double *cn = col + ((j + 1) % 3) * kPB;
    const double rc = rcp64(d);
    const double rp = chain_rsqrt(dprev);
    const double tj = a[j] * rc;
    double dnext = 1.0;
    if (j + 1 < kPB) {
      a[j + 1] -= tj * cb[j + 1];
      dnext = __shfl_sync(full, a[j + 1], j + 1);
      cn[lane] = a[j + 1];
    }
    __syncwarp();
    if (j > 1) *prog = j - 1;
#pragma unroll
    for (int c = j + 2; c < kPB; c++) a[c] -= tj * cb[c];
    if (j > 0) {
      dinv[j - 1] = rp;
      a[j - 1] *= rp;
      L11[lane][j - 1] = (lane >= j - 1) ? a[j - 1] : 0.0;
    }
    dprev = d;
    d = dnext;
  }
  const double rp = chain_rsqrt(dprev);
  dinv[kPB - 1] = rp;
  a[kPB - 1] *= rp;
  L11[lane][kPB - 1] = (lane >= kPB - 1) ? a[kPB - 1] : 0.0;
  __syncwarp();
  *prog = kPB;
}

// K: E with column j read into registers in one batch BEFORE the reciprocal chain (the shipped loop's
// LDS each land in the same register pair and are used at once: 8 serialised round trips per pivot)
__device__ __forceinline__ void cholK(double a[kPB], int lane, double *col, double *dinv,
                                      double (*L11)[kPB + 1], volatile int *prog) {
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    __syncwarp();
    if (j > 0) *prog = j;
    col[lane] = a[j];
    __syncwarp();
    double cv[kPB];
#pragma unroll
    for (int c = j; c < kPB; c++) cv[c] = col[c];
    const double d = cv[j];
    const double rc = rcp64(d);
    const double r = chain_rsqrt(d);
    const double tj = a[j] * rc;
#pragma unroll
    for (int c = j + 1; c < kPB; c++) a[c] -= tj * cv[c];
    dinv[j] = r;
    a[j] *= r;
    L11[lane][j] = (lane >= j) ? a[j] : 0.0;
  }
  __syncwarp();
  *prog = kPB;
}
// J: G with column j read into registers at the top of the round, before the reciprocal
__device__ __forceinline__ void cholJ(double a[kPB], int lane, double *col, double *dinv,
                                      double (*L11)[kPB + 1], volatile int *prog) {
  col[lane] = a[0];
  __syncwarp();
  double d = col[0], dprev = 1.0;
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    const double *cb = col + (j % 3) * kPB;
    double *cn = col + ((j + 1) % 3) * kPB;
    double cv[kPB];
#pragma unroll
    for (int c = j + 1; c < kPB; c++) cv[c] = cb[c];
    const double rc = rcp64(d);
    const double rp = chain_rsqrt(dprev);
    const double tj = a[j] * rc;
    if (j + 1 < kPB) {
      a[j + 1] -= tj * cv[j + 1];
      cn[lane] = a[j + 1];
    }
    __syncwarp();
    if (j > 1) *prog = j - 1;
    const double dnext = (j + 1 < kPB) ? cn[j + 1] : 1.0;
#pragma unroll
    for (int c = j + 2; c < kPB; c++) a[c] -= tj * cv[c];
    if (j > 0) {
      dinv[j - 1] = rp;
      a[j - 1] *= rp;
      L11[lane][j - 1] = (lane >= j - 1) ? a[j - 1] : 0.0;
    }
    dprev = d;
    d = dnext;
  }
  const double rp = chain_rsqrt(dprev);
  dinv[kPB - 1] = rp;
  a[kPB - 1] *= rp;
  L11[lane][kPB - 1] = (lane >= kPB - 1) ? a[kPB - 1] : 0.0;
  __syncwarp();
  *prog = kPB;
}

// L: the elimination chain alone.  No square root, no scaled column: column j + 1 (unscaled) goes to
// U[j + 1][lane] as soon as its one update is done; the rest of pivot j's row updates issue beside the
// reciprocal of pivot j + 1 (one basic block between two barriers: no branch on the pivot value, an
// abnormal pivot only raises a sticky flag).  Scaling by 1/sqrt(d_j) is someone else's job.
__device__ __forceinline__ bool cholL(double a[kPB], int lane, double (*U)[kPB], volatile int *prog) {
  U[0][lane] = a[0];
  __syncwarp();
  double d = U[0][0];
  bool abn = false;
  double tj_prev = 0.0;
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    // (a) what is left of pivot j - 1: rows c >= j + 1 (column j was updated before it was published)
    if (j > 0) {
#pragma unroll
      for (int c = j + 1; c < kPB; c++) a[c] -= tj_prev * U[j - 1][c];
    }
    // (b) pivot j
    const bool normal = d > 1e-30 && d < 1e30;
    abn |= !normal;
    const double rc = rcp64(normal ? d : 1.0);
    const double tj = a[j] * rc;
    if (j + 1 < kPB) {
      a[j + 1] -= tj * U[j][j + 1];
      U[j + 1][lane] = a[j + 1];
    }
    __syncwarp();
    *prog = j + 1;  // columns <= j + 1 of U are complete... (uniform store)
    if (j + 1 < kPB) d = U[j + 1][j + 1];
    tj_prev = tj;
  }
  return abn;
}
template <int V>
__global__ void __launch_bounds__(32, 1) k(const double *A, double *out, long long *cyc, int reps) {
  __shared__ double S[kPB][kPB + 1];
  __shared__ double dinv[kPB];
  __shared__ double L11[kPB][kPB + 1];
  __shared__ double col3[3 * kPB];
  __shared__ int prog;
  __shared__ double U[kPB][kPB];
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
    if (V == 4) cholE(a, lane, col3, dinv, L11, &prog);
    if (V == 5) cholF(a, lane, col3, dinv, L11, &prog);
    if (V == 6) cholG(a, lane, col3, dinv, L11, &prog);
    if (V == 7) cholI(a, lane, col3, dinv, L11, &prog);
    if (V == 8) cholK(a, lane, col3, dinv, L11, &prog);
    if (V == 9) cholJ(a, lane, col3, dinv, L11, &prog);
    if (V == 10) cholL(a, lane, U, &prog);
    const long long t1 = clock64();
    if (lane == 0) cyc[rep] = t1 - t0;
    if (V == 10) {
      __syncwarp();
      for (int c = 0; c < kPB; c++) a[c] = U[c][lane] * chain_rsqrt(U[c][c]);
    }
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
                         "D registers + smem broadcast", "E shipped warp_chol32", "F pipelined",
                         "G pipelined, root delayed", "I G + next pivot by shuffle", "K shipped + column preloaded",
                         "J G + column preloaded", "L chain only (no root)"};
  for (int v = 0; v < 11; v++) {
    if (v == 0) k<0><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 1) k<1><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 2) k<2><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 3) k<3><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 4) k<4><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 5) k<5><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 6) k<6><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 7) k<7><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 8) k<8><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 9) k<9><<<1, 32>>>(dA, dO, dc, 4);
    if (v == 10) k<10><<<1, 32>>>(dA, dO, dc, 4);
    cudaMemcpy(hc, dc, 32, cudaMemcpyDeviceToHost);
    cudaMemcpy(o, dO, sizeof(o), cudaMemcpyDeviceToHost);
    if (v == 0) for (int i = 0; i < kPB * kPB; i++) ref[i] = o[i];
    double err = 0; for (int i = 0; i < kPB * kPB; i++) err = fmax(err, fabs(o[i] - ref[i]));
    printf("%-28s cycles per call: %lld (cold) %lld %lld %lld   max |L - L_A| %.2e  %s\n", names[v], hc[0], hc[1], hc[2], hc[3], err,
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
