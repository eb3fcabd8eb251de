// One cooperative kernel per Gauss-Newton iteration for the linear step:
// blocked right-looking fp64 Cholesky of the dense system (resident in L2), forward
// substitution fused into the panel phase, backward substitution, dx = -x, retraction and the
// stop test.  Replaces 2*ceil(D/32)+1 dependent launches by one launch with two grid-wide
// barriers per 32-column panel.  (Included by gn.cu inside namespace m3r.)
//
// Per panel k (columns k0..k0+pb):
//   phase A  every CTA that owns rows of the panel factors the pb x pb diagonal block
//            redundantly (one warp, rows in registers, shuffles; no barrier between POTRF and
//            TRSM), solves y_k = L11^-1 b_k, then its rows: L21 = A21 L11^-T, b_r -= L21 y_k.
//            CTA 0 stores L11 to `ldiag` (not in place: other CTAs are still reading A11).
//   grid.sync
//   phase B  A22 -= L21 L21^T on the lower tiles (32x32 or 64x64), strided over the CTAs.
//   grid.sync
// After the last panel CTA 0 runs the backward substitution L^T x = y (column oriented: every
// step is one round of independent, coalesced loads) and finish_step().
#pragma once

namespace cg = ::cooperative_groups;  // <cooperative_groups.h> is included by gn.cu at global scope

// cycles CTA 0 spent in: [0] phase A, [1] barrier 1, [2] phase B, [3] barrier 2, [4] backward
// substitution, [5] finish_step; [6] launches.  Read by m3r_debug_coop_profile (tools/).
__device__ long long g_coop_prof[8];
#define COOP_TICK(slot)                                   \
  do {                                                    \
    if (blockIdx.x == 0 && threadIdx.x == 0) {            \
      const long long now_ = clock64();                   \
      prof_acc[slot] += now_ - prof_t;                    \
      prof_t = now_;                                      \
    }                                                     \
  } while (0)

constexpr int kCoopThreads = 256;
constexpr int kLdStride = kPB * (kPB + 1);  // per panel in `ldiag`: L11 (kPB x kPB) + 1/diag (kPB)
constexpr int kTrsmRows = 128;  // panel rows per CTA in phase A (warps 1..4, one row per thread)

// 1/sqrt(d) to fp64 accuracy from the fp32 MUFU.RSQ seed and two Newton steps: this sits on
// the pivot chain of the factorisation (D sequential pivots), so its latency matters.
__device__ __forceinline__ double chain_rsqrt(double d) {
  if (d < 1e-30 || d > 1e30) return rsqrt(d);
  double r = static_cast<double>(rsqrtf(__double2float_rn(d)));
  double e = fma(-d * r, r, 1.0);
  r = fma(0.5 * r, e, r);
  e = fma(-d * r, r, 1.0);
  r = fma(0.5 * r, e, r);
  return r;
}

// 32x32 lower Cholesky by one warp: lane r holds row r in registers (a[c], c <= r meaningful).
// Returns the reciprocal pivots through dinv[] (shared) and sets *bad on a non-positive pivot.
__device__ __forceinline__ void warp_chol32(double a[kPB], int lane, double *dinv, int *bad) {
  const unsigned full = 0xffffffffu;
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    const double d = __shfl_sync(full, a[j], j);
    double r;
    if (!(d > 0.0)) {
      if (lane == 0) *bad = 1;
      r = 1.0;  // keep finite; the step is discarded
    } else {
      r = chain_rsqrt(d);
    }
    if (lane == 0) dinv[j] = r;
    const double l = a[j] * r;  // lane j: d * rsqrt(d) = sqrt(d); lanes > j: L[lane][j]
    a[j] = l;
#pragma unroll
    for (int c = j + 1; c < kPB; c++) {
      const double lc = __shfl_sync(full, l, c);  // L[c][j]
      a[c] -= l * lc;                             // meaningful for lanes >= c
    }
  }
}

template <int TILE>  // 32 or 64: A22 tile (tr, tc) -= L21[tr] L21[tc]^T
__device__ __forceinline__ void coop_update_tile(double *__restrict__ A, int D, int k0, int pb,
                                                 int base, int tr, int tc, double (*Ra)[kPB + 1],
                                                 double (*Rb)[kPB + 1], int t) {
  constexpr int R = TILE / 16;  // outputs per thread per dimension
  const int ty = t / 16, tx = t % 16;
  double acc[R][R];
#pragma unroll
  for (int i = 0; i < R; i++)
#pragma unroll
    for (int j = 0; j < R; j++) {
      const int gr = base + tr * TILE + ty + 16 * i, gc = base + tc * TILE + tx + 16 * j;
      acc[i][j] = (gr < D && gc <= gr) ? A[static_cast<size_t>(gr) * D + gc] : 0.0;
    }
  __syncthreads();
  for (int x = t; x < TILE * kPB; x += kCoopThreads) {
    const int r = x / kPB, c = x % kPB;
    const int ga = base + tr * TILE + r, gb = base + tc * TILE + r;
    Ra[r][c] = (ga < D && c < pb) ? A[static_cast<size_t>(ga) * D + k0 + c] : 0.0;
    Rb[r][c] = (gb < D && c < pb) ? A[static_cast<size_t>(gb) * D + k0 + c] : 0.0;
  }
  __syncthreads();
#pragma unroll 8
  for (int k = 0; k < kPB; k++) {
    double av[R], bv[R];
#pragma unroll
    for (int i = 0; i < R; i++) av[i] = Ra[ty + 16 * i][k], bv[i] = Rb[tx + 16 * i][k];
#pragma unroll
    for (int i = 0; i < R; i++)
#pragma unroll
      for (int j = 0; j < R; j++) acc[i][j] -= av[i] * bv[j];
  }
#pragma unroll
  for (int i = 0; i < R; i++)
#pragma unroll
    for (int j = 0; j < R; j++) {
      const int gr = base + tr * TILE + ty + 16 * i, gc = base + tc * TILE + tx + 16 * j;
      if (gr < D && gc <= gr) A[static_cast<size_t>(gr) * D + gc] = acc[i][j];
    }
}

__device__ __forceinline__ void tile_from_index(int tile, int &tr, int &tc) {
  tr = static_cast<int>((sqrt(8.0 * tile + 1.0) - 1.0) * 0.5);
  while ((tr + 1) * (tr + 2) / 2 <= tile) tr++;
  while (tr * (tr + 1) / 2 > tile) tr--;
  tc = tile - tr * (tr + 1) / 2;
}

__global__ void __launch_bounds__(kCoopThreads)
    gn_chol_solve_coop_kernel(double *__restrict__ sys, double *__restrict__ ldiag, int D,
                              double *__restrict__ xbuf, float *__restrict__ Twc,
                              float *__restrict__ dx, int num_poses, float delta_thresh,
                              int *__restrict__ flags) {
  cg::grid_group grid = cg::this_grid();
  if (flags[0]) return;  // uniform over the grid (the flag only changes after the last barrier)
  __shared__ double L11[kPB][kPB + 1];
  __shared__ double dinv[kPB];
  __shared__ double yk[kPB];
  __shared__ double Ra[64][kPB + 1], Rb[64][kPB + 1];
  __shared__ int bad;
  const int t = threadIdx.x;
  const int lane = t & 31, warp = t >> 5;
  double *A = sys;
  double *rhs = sys + static_cast<size_t>(D) * D;
  if (t == 0) bad = 0;
  long long prof_acc[6] = {0, 0, 0, 0, 0, 0};
  long long prof_t = clock64();
  __syncthreads();

  for (int k0 = 0; k0 < D; k0 += kPB) {
    const int pb = min(kPB, D - k0);
    const int m = D - k0 - pb;  // rows below the panel
    const int row_ctas = max(1, (m + kTrsmRows - 1) / kTrsmRows);
    // ---------------- phase A
    for (int rc = blockIdx.x; rc < row_ctas; rc += gridDim.x) {
      // panel rows of this CTA start loading while warp 0 factors the diagonal block
      const int row = k0 + pb + rc * kTrsmRows + (t - 32);
      const bool has_row = (warp >= 1 && warp <= kTrsmRows / 32) && row < D;
      double x[kPB];
      double *Ar = A + static_cast<size_t>(has_row ? row : 0) * D + k0;
      if (has_row) {
#pragma unroll
        for (int c = 0; c < kPB; c++) x[c] = (c < pb) ? Ar[c] : 0.0;
      }
      if (warp == 0) {
        double a[kPB];
#pragma unroll
        for (int c = 0; c < kPB; c++)
          a[c] = (lane < pb && c <= lane) ? A[static_cast<size_t>(k0 + lane) * D + (k0 + c)]
                                          : ((lane >= pb && c == lane) ? 1.0 : 0.0);
        const double b_mine = (lane < pb) ? rhs[k0 + lane] : 0.0;  // in flight during the factor
        warp_chol32(a, lane, dinv, &bad);
#pragma unroll
        for (int c = 0; c < kPB; c++) L11[lane][c] = (c <= lane) ? a[c] : 0.0;
        __syncwarp();
        // release the TRSM warps; the y_k chain below overlaps with their work
        asm volatile("bar.arrive 1, %0;" ::"r"(kCoopThreads));
        // y_k = L11^-1 b_k (serial chain over the pivots)
        double mine = b_mine;
#pragma unroll
        for (int c = 0; c < kPB; c++) {
          double yc = (lane == c) ? mine * dinv[c] : 0.0;
          yc = __shfl_sync(0xffffffffu, yc, c);
          if (lane == c) mine = yc;
          else if (lane > c) mine -= a[c] * yc;
        }
        yk[lane] = (lane < pb) ? mine : 0.0;
        if (rc == 0) {
          if (lane < pb) xbuf[k0 + lane] = mine;
          double *Ld = ldiag + static_cast<size_t>(k0 / kPB) * kLdStride;
#pragma unroll
          for (int c = 0; c < kPB; c++) Ld[lane * kPB + c] = (c <= lane) ? a[c] : 0.0;
          Ld[kPB * kPB + lane] = dinv[lane];
        }
      } else {
        asm volatile("bar.sync 1, %0;" ::"r"(kCoopThreads));  // L11, dinv ready
      }
      if (has_row) {  // right-looking TRSM: independent updates per column -> ILP
#pragma unroll
        for (int c = 0; c < kPB; c++) {
          const double xc = x[c] * dinv[c];
          x[c] = xc;
#pragma unroll
          for (int k = c + 1; k < kPB; k++) x[k] -= xc * L11[k][c];
        }
#pragma unroll
        for (int c = 0; c < kPB; c++)
          if (c < pb) Ar[c] = x[c];
      }
      __syncthreads();  // y_k ready
      if (has_row) {
        double dot = 0.0;
#pragma unroll
        for (int c = 0; c < kPB; c++) dot += x[c] * yk[c];
        rhs[row] -= dot;
      }
      __syncthreads();
    }
    COOP_TICK(0);
    grid.sync();
    COOP_TICK(1);
    // ---------------- phase B
    if (m > 0) {
      const int base = k0 + pb;
      const int T64 = (m + 63) / 64;
      if (T64 * (T64 + 1) / 2 >= 2 * static_cast<int>(gridDim.x)) {
        const int ntiles = T64 * (T64 + 1) / 2;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
          int tr, tc;
          tile_from_index(tile, tr, tc);
          coop_update_tile<64>(A, D, k0, pb, base, tr, tc, Ra, Rb, t);
        }
      } else {  // few tiles: 32x32 tiles spread the update over more SMs
        const int T32 = (m + 31) / 32;
        const int ntiles = T32 * (T32 + 1) / 2;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
          int tr, tc;
          tile_from_index(tile, tr, tc);
          coop_update_tile<32>(A, D, k0, pb, base, tr, tc, Ra, Rb, t);
        }
      }
      COOP_TICK(2);
      grid.sync();
      COOP_TICK(3);
    }
  }
  if (blockIdx.x != 0) return;

  // ---------------- backward substitution L^T x = y (CTA 0), x overwrites xbuf
  __syncthreads();
  const bool fail = bad != 0;
  if (!fail) {
    const int last = ((D - 1) / kPB) * kPB;
    for (int k0 = last; k0 >= 0; k0 -= kPB) {
      const int pb = min(kPB, D - k0);
      if (warp == 0) {  // x_k = L11^-T y_k
        const double *Ld = ldiag + static_cast<size_t>(k0 / kPB) * kLdStride;
        double col[kPB];  // column `lane` of L11: col[r] = L11[r][lane]
#pragma unroll
        for (int r = 0; r < kPB; r++) col[r] = Ld[r * kPB + lane];
        const double my_dinv = Ld[kPB * kPB + lane];
        double mine = (lane < pb) ? xbuf[k0 + lane] : 0.0;
#pragma unroll
        for (int cc = kPB - 1; cc >= 0; cc--) {
          double xc = (lane == cc) ? mine * my_dinv : 0.0;
          xc = __shfl_sync(0xffffffffu, xc, cc);
          if (cc < pb) {
            if (lane == cc) mine = xc;
            else if (lane < cc) mine -= col[cc] * xc;
          }
        }
        yk[lane] = (lane < pb) ? mine : 0.0;
        if (lane < pb) xbuf[k0 + lane] = mine;
      }
      __syncthreads();
      // y_j -= sum_c L[k0+c][j] x_k[c] for all j < k0: independent, coalesced rows of L
      for (int j = t; j < k0; j += kCoopThreads) {
        double lv[kPB];
#pragma unroll
        for (int c = 0; c < kPB; c++)
          lv[c] = (c < pb) ? A[static_cast<size_t>(k0 + c) * D + j] : 0.0;
        double s = xbuf[j];
#pragma unroll
        for (int c = 0; c < kPB; c++) s -= lv[c] * yk[c];
        xbuf[j] = s;
      }
      __syncthreads();
    }
  }
  COOP_TICK(4);
  finish_step<kCoopThreads>(t, fail, D, xbuf, Twc, dx, num_poses, delta_thresh, flags);
  COOP_TICK(5);
  if (t == 0) {
    for (int k = 0; k < 6; k++) g_coop_prof[k] += prof_acc[k];
    g_coop_prof[6] += 1;
  }
}
