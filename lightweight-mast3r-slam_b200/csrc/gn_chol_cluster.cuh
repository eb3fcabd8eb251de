// Linear step of one Gauss-Newton iteration on ONE thread-block cluster (8 CTAs), with the
// matrix resident in distributed shared memory: blocked right-looking fp64 Cholesky, fused
// forward substitution, backward substitution, dx = -x, retraction and the stop test.
// Used when D = 7(N-1) <= 448 (64 keyframes: D = 441); larger systems use the grid-wide
// cooperative kernel of gn_chol_coop.cuh.  (Included by gn.cu inside namespace m3r.)
//
// Why: at D = 441 the factorisation is a chain of 441 dependent pivots, and the grid-wide kernel
// spends its time between the links of that chain: two grid barriers and three L2 round trips
// per 32-column panel (profiles/r1_a_coop_solver_phases.txt: 24 us per panel, 340 us per solve).
// Here the lower triangle (0.86 MB padded) lives in the shared memory of 8 SMs, panel data
// moves between SMs through DSMEM, and the barriers are cluster barriers.
//
// Layout: 32-row blocks; CTA c owns row blocks c and c+8.  Row block b stores its 32 rows of
// the block-lower triangle, 32(b+1) columns, row-major with an odd leading dimension.
// Per panel k (owner = k mod 8):
//   1. owner, warp 0: POTRF of the diagonal block in registers (lane = row, shuffles), y_k.
//   2. cluster barrier; every CTA copies L11, 1/diag and y_k from the owner (DSMEM).
//   3. TRSM of the owned rows below the panel (one row per thread) and b_r -= L21 y_k.
//   4. cluster barrier.
//   5. SYRK: every CTA updates its own row blocks, tile by tile; the other operand (rows of
//      block bc) is fetched from its owner through DSMEM, double-buffered in registers.
//   6. cluster barrier.
// Then CTA 0 runs the backward substitution reading L straight from the owners' shared memory.
#pragma once

__device__ long long g_cl_prof[9];  // read by m3r_debug_cluster_profile (tools/)
constexpr int kClSize = 8;
constexpr int kClThreads = 256;
constexpr int kClMaxNb = 14;                    // row blocks (D <= 448)
constexpr int kClMaxD = kClMaxNb * kPB;
__host__ __device__ constexpr int cl_ld(int b) { return kPB * (b + 1) + 1; }
// doubles of block storage for CTA c (blocks c and c+8), worst case over c with nb = kClMaxNb
constexpr int kClBlkDoubles = kPB * (cl_ld(5) + cl_ld(13));
struct ClSmem {
  double L11[kPB][kPB + 1];
  double stage[2][kPB][kPB + 1];
  double dinv[kPB];
  double yk[kPB];
  double bl[2][kPB];   // right-hand side entries of the owned rows
  int bad;
  int prog;  // written by warp_chol32, unused here
  int pad_;
  double blk[kClBlkDoubles];
};
constexpr int kClSmemBytes = static_cast<int>(sizeof(ClSmem));

__global__ void __cluster_dims__(kClSize, 1, 1) __launch_bounds__(kClThreads, 1)
    gn_chol_solve_cluster_kernel(const double *__restrict__ sys, double *__restrict__ dinvg, int D,
                                 double *__restrict__ xbuf, float *__restrict__ Twc,
                                 float *__restrict__ dx, int num_poses, float delta_thresh,
                                 int *__restrict__ flags) {
  cg::cluster_group cl = cg::this_cluster();
  if (flags[0]) return;  // uniform over the cluster
  extern __shared__ __align__(16) unsigned char cl_smem_raw[];
  ClSmem &S = *reinterpret_cast<ClSmem *>(cl_smem_raw);
  const int c = static_cast<int>(cl.block_rank());
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int nb = (D + kPB - 1) / kPB;
  const double *rhs = sys + static_cast<size_t>(D) * D;
  // local block lb <-> global row block c + 8*lb, storage base and leading dimension
  const int ld0 = cl_ld(c), ld1 = cl_ld(c + kClSize);
  const int base1 = kPB * ld0;

  // cycles of CTA 0: [0] load, [1] POTRF (incl. waiting for the owner), [2] TRSM, [3] barrier,
  // [4] SYRK, [5] barrier, [6] backward, [7] finish; g_cl_prof[8] = launches
  long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long prof_t = clock64();
#define CL_TICK(slot)                          \
  do {                                         \
    if (c == 0 && t == 0) {                    \
      const long long now_ = clock64();        \
      prof_acc[slot] += now_ - prof_t;         \
      prof_t = now_;                           \
    }                                          \
  } while (0)
  // ---------------- load the owned row blocks (identity on the padding rows)
  if (t == 0) S.bad = 0;
  for (int lb = 0; lb < 2; lb++) {
    const int b = c + kClSize * lb;
    if (b >= nb) break;
    const int ld = lb ? ld1 : ld0, width = kPB * (b + 1);
    double *dst = S.blk + (lb ? base1 : 0);
    for (int x = t; x < kPB * width; x += kClThreads) {
      const int r = x / width, col = x - r * width;
      const int gr = b * kPB + r;
      double v = (gr == col) ? 1.0 : 0.0;
      if (gr < D && col <= gr) v = sys[static_cast<size_t>(gr) * D + col];
      dst[r * ld + col] = v;
    }
    if (t < kPB) S.bl[lb][t] = (b * kPB + t < D) ? rhs[b * kPB + t] : 0.0;
  }
  cl.sync();
  CL_TICK(0);

  for (int k = 0; k < nb; k++) {
    const int owner = k & (kClSize - 1), lbk = k / kClSize;
    // ---------------- 1. POTRF of the diagonal block + y_k (owner, warp 0)
    if (c == owner && warp == 0) {
      double *blkk = S.blk + (lbk ? base1 : 0);
      const int ld = lbk ? ld1 : ld0;
      double a[kPB];
#pragma unroll
      for (int cc = 0; cc < kPB; cc++) a[cc] = (cc <= lane) ? blkk[lane * ld + k * kPB + cc] : 0.0;
      warp_chol32(a, lane, S.stage[0][0], S.dinv, &S.bad, S.L11, &S.prog);  // (this variant ignores the per-pivot counter)
#pragma unroll
      for (int cc = 0; cc < kPB; cc++) {
        S.L11[lane][cc] = (cc <= lane) ? a[cc] : 0.0;
        if (cc <= lane) blkk[lane * ld + k * kPB + cc] = a[cc];
      }
      __syncwarp();
      double mine = S.bl[lbk][lane];
#pragma unroll
      for (int cc = 0; cc < kPB; cc++) {
        double yc = (lane == cc) ? mine * S.dinv[cc] : 0.0;
        yc = __shfl_sync(0xffffffffu, yc, cc);
        if (lane == cc) mine = yc;
        else if (lane > cc) mine -= a[cc] * yc;
      }
      S.yk[lane] = mine;
      xbuf[k * kPB + lane] = mine;            // y (xbuf holds kClMaxD entries)
      dinvg[k * kPB + lane] = S.dinv[lane];
      if (lane == 0 && S.bad) flags[2] = 1;
    }
    cl.sync();
    CL_TICK(1);
    // ---------------- 2./3. TRSM of the owned rows below the panel
    const bool has0 = (c > k) && (c < nb), has1 = (c + kClSize > k) && (c + kClSize < nb);
    if (has0 || has1) {
      if (c != owner) {
        const ClSmem *R = cl.map_shared_rank(&S, owner);
        for (int x = t; x < kPB * kPB; x += kClThreads) S.L11[x / kPB][x % kPB] = R->L11[x / kPB][x % kPB];
        if (t < kPB) S.dinv[t] = R->dinv[t], S.yk[t] = R->yk[t];
      }
      __syncthreads();
      const int lb = t >> 5;
      if (t < 64 && (lb ? has1 : has0)) {
        double *row = S.blk + (lb ? base1 : 0) + lane * (lb ? ld1 : ld0) + k * kPB;
        double x[kPB];
#pragma unroll
        for (int cc = 0; cc < kPB; cc++) x[cc] = row[cc];
#pragma unroll
        for (int cc = 0; cc < kPB; cc++) {
          const double xc = x[cc] * S.dinv[cc];
          x[cc] = xc;
#pragma unroll
          for (int kk = cc + 1; kk < kPB; kk++) x[kk] -= xc * S.L11[kk][cc];
        }
        double dot = 0.0;
#pragma unroll
        for (int cc = 0; cc < kPB; cc++) row[cc] = x[cc], dot += x[cc] * S.yk[cc];
        S.bl[lb][lane] -= dot;
      }
    }
    CL_TICK(2);
    cl.sync();
    CL_TICK(3);
    // ---------------- 5. SYRK on the owned row blocks: A[br][bc] -= L21[br] L21[bc]^T
    {
      const int n0 = has0 ? (c - k) : 0;                 // tiles of local block 0: bc = k+1..c
      const int n1 = has1 ? (c + kClSize - k) : 0;       // tiles of local block 1
      const int ntiles = n0 + n1;
      const int ty = t >> 4, tx = t & 15;
      // the remote operand of tile i: rows of block bc, columns of panel k (4 doubles per thread)
      auto fetch = [&](int i, double v[4]) {
        const int bc = k + 1 + (i < n0 ? i : i - n0);
        const ClSmem *R = cl.map_shared_rank(&S, bc & (kClSize - 1));
        const int lbc = bc / kClSize, cr = bc & (kClSize - 1);
        const double *src = R->blk + (lbc ? kPB * cl_ld(cr) : 0);
        const int ldc = cl_ld(bc);
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int x = t + j * kClThreads;  // 0..1023
          v[j] = src[(x >> 5) * ldc + k * kPB + (x & 31)];
        }
      };
      double pre[4];
      if (ntiles > 0) fetch(0, pre);
      for (int i = 0; i < ntiles; i++) {
        double(*B)[kPB + 1] = S.stage[i & 1];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int x = t + j * kClThreads;
          B[x >> 5][x & 31] = pre[j];
        }
        __syncthreads();
        if (i + 1 < ntiles) fetch(i + 1, pre);
        const int lb = (i < n0) ? 0 : 1;
        const int bc = k + 1 + (i < n0 ? i : i - n0);
        double *Ablk = S.blk + (lb ? base1 : 0);
        const int ld = lb ? ld1 : ld0;
        const double *A0 = Ablk + ty * ld + k * kPB, *A1 = Ablk + (ty + 16) * ld + k * kPB;
        double c00 = 0, c01 = 0, c10 = 0, c11 = 0;
#pragma unroll 8
        for (int kk = 0; kk < kPB; kk++) {
          const double a0 = A0[kk], a1 = A1[kk], b0 = B[tx][kk], b1 = B[tx + 16][kk];
          c00 += a0 * b0, c01 += a0 * b1, c10 += a1 * b0, c11 += a1 * b1;
        }
        double *C0 = Ablk + ty * ld + bc * kPB, *C1 = Ablk + (ty + 16) * ld + bc * kPB;
        C0[tx] -= c00, C0[tx + 16] -= c01, C1[tx] -= c10, C1[tx + 16] -= c11;
        // stage[(i+1)&1] was last read two tiles ago; the barrier at the top of the next
        // iteration orders this tile's reads of stage[i&1] before tile i+2 overwrites it
      }
    }
    CL_TICK(4);
    cl.sync();
    CL_TICK(5);
  }

  // ---------------- backward substitution L^T x = y (CTA 0), x overwrites xbuf
  if (c == 0) {
    const bool fail = flags[2] != 0;
    if (!fail) {
      for (int k = nb - 1; k >= 0; k--) {
        const ClSmem *R = cl.map_shared_rank(&S, k & (kClSize - 1));
        const double *Lk = R->blk + ((k / kClSize) ? kPB * cl_ld(k & (kClSize - 1)) : 0);
        const int ld = cl_ld(k);
        if (warp == 0) {  // x_k = L11^-T y_k
          double col[kPB];  // column `lane` of L11
#pragma unroll
          for (int r = 0; r < kPB; r++) col[r] = (r >= lane) ? Lk[r * ld + k * kPB + lane] : 0.0;
          const double my_dinv = dinvg[k * kPB + lane];
          double mine = xbuf[k * kPB + lane];
#pragma unroll
          for (int cc = kPB - 1; cc >= 0; cc--) {
            double xc = (lane == cc) ? mine * my_dinv : 0.0;
            xc = __shfl_sync(0xffffffffu, xc, cc);
            if (lane == cc) mine = xc;
            else if (lane < cc) mine -= col[cc] * xc;
          }
          S.yk[lane] = mine;
          xbuf[k * kPB + lane] = mine;
        }
        __syncthreads();
        // y_j -= sum_r L[32k + r][j] x_k[r] for every column j left of the panel
        for (int j = t; j < k * kPB; j += kClThreads) {
          double lv[kPB];
#pragma unroll
          for (int r = 0; r < kPB; r++) lv[r] = Lk[r * ld + j];
          double s = xbuf[j];
#pragma unroll
          for (int r = 0; r < kPB; r++) s -= lv[r] * S.yk[r];
          xbuf[j] = s;
        }
        __syncthreads();
      }
    }
  }
  CL_TICK(6);
  cl.sync();  // the other CTAs' shared memory stays alive until CTA 0 has read it
  if (c != 0) return;
  finish_step<kClThreads>(t, flags[2] != 0, D, xbuf, Twc, dx, num_poses, delta_thresh, flags);
  CL_TICK(7);
  if (t == 0) {
    for (int q = 0; q < 8; q++) g_cl_prof[q] += prof_acc[q];
    g_cl_prof[8] += 1;
  }
#undef CL_TICK
}
