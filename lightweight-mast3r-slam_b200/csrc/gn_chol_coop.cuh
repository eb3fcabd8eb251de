// One cooperative kernel per Gauss-Newton iteration for the linear step:
// blocked right-looking fp64 Cholesky of the dense system (resident in L2), forward
// substitution fused into the panel phase, backward substitution, dx = -x, retraction and the
// stop test.  Replaces 2*ceil(D/32)+1 dependent launches by one launch with ONE grid-wide
// barrier per 32-column panel.  (Included by gn.cu inside namespace m3r.)
//
// Per panel k (columns k0..k0+pb), look-ahead by one panel:
//   phase A  (panel CTAs: one per 64 rows below the panel; the warps of a CTA have roles, see the
//            kernel)  loads the diagonal block minus T (the previous panel's update of it, prepared at
//            the end of the previous step) and its rows, folds the previous panel's trailing update
//            into those rows, factors the pb x pb diagonal block redundantly -- one warp runs the
//            elimination chain, another takes the square roots and publishes the scaled columns --
//            while the TRSM rows (L21 = A21 L11^-T), the forward chain y_k = L11^-1 b_k and the inverse
//            of the block follow the published columns; b_r -= L21 y_k.  CTA 0 stores L11^-1 to `ldiag`
//            and computes T for the next panel from the rows it has just solved.
//   phase B  (all other CTAs, at the same time)  A22 -= L21 L21^T of the PREVIOUS panel on the
//            rows/columns behind this panel, lower tiles 32x32 or 64x64, pipelined over the tiles.
//   grid.sync
// Then the backward substitution L^T x = y: CTA 0 alone for D <= 512 (x in shared memory, the next
// block's rows in flight), all CTAs in super-blocks of 8 panels above that; finish_step() on CTA 0.
// What was measured on the way, and what was tried and dropped: profiles/r2_i_potrf_chain.md.
#pragma once

namespace cg = ::cooperative_groups;  // <cooperative_groups.h> is included by gn.cu at global scope

// cycles CTA 0 spent in: [0] phase A, [1] barrier 1, [2] phase B, [3] barrier 2, [4] backward
// substitution, [5] finish_step; [6] launches.  Read by m3r_debug_coop_profile (tools/).
// The clock64() reads are compiled in only with -DM3R_COOP_PROF (tools/, experiments): the product
// kernel carries none of them.
#ifdef M3R_COOP_PROF
__device__ long long g_coop_prof[16];  // [8..13]: phase A split of CTA 0, see COOP_ATICK
#define COOP_TICK(slot)                                   \
  do {                                                    \
    if (blockIdx.x == 0 && threadIdx.x == 0) {            \
      const long long now_ = clock64();                   \
      prof_acc[slot] += now_ - prof_t;                    \
      prof_t = now_;                                      \
    }                                                     \
  } while (0)

// phase A of CTA 0, thread 0 (warp 0): [8] stage previous L21, [9] loads + look-ahead update,
// [10] POTRF, [11] y_k chain + stores, [12] waiting for the TRSM warps, [13] rhs update
// phase A as the first TRSM warp of CTA 0 sees it: [0] row loads + look-ahead fold, [1] waiting
// for L11, [2] TRSM + row stores, [3] waiting for y_k + rhs update  -> g_coop_prof[11,12,14,15]
#define COOP_TTICK(slot)                                  \
  do {                                                    \
    if (blockIdx.x == 0 && threadIdx.x == 32) {           \
      const long long now_ = clock64();                   \
      tprof[slot] += now_ - tprof_t;                      \
      tprof_t = now_;                                     \
    }                                                     \
  } while (0)

#define COOP_ATICK(slot)                                  \
  do {                                                    \
    if (blockIdx.x == 0 && threadIdx.x == 0) {            \
      const long long now_ = clock64();                   \
      aprof[slot - 8] += now_ - aprof_t;                  \
      aprof_t = now_;                                     \
    }                                                     \
  } while (0)
// when each role of CTA 0 reaches a point of the panel step, in cycles since the step began (summed
// over the panels): read by m3r_debug_coop_stamps
__device__ long long g_coop_stamp[16];
#define COOP_STAMP(slot)                                                   \
  do {                                                                     \
    if (blockIdx.x == 0 && lane == 0) S.stamp[slot] = clock64() - S.t0;     \
  } while (0)  /* to shared memory: a read-modify-write of a global counter stalls the warp 0.7 us */
#else
#define COOP_STAMP(slot) do { } while (0)
#define COOP_TICK(slot) do { } while (0)
#define COOP_TTICK(slot) do { } while (0)
#define COOP_ATICK(slot) do { } while (0)
#endif

constexpr int kCoopThreads = 256;
constexpr int kLdStride = kPB * (kPB + 1);  // per panel in `ldiag`: L11 (kPB x kPB) + 1/diag (kPB)
// Panel rows per CTA in phase A, one row per thread of the TRSM warps.  The rows carry almost all
// of the fp64 work of a panel step (look-ahead fold 1024 DFMA per row, TRSM 528), and fp64 issues
// at ~2.2 cycles per warp instruction per scheduler on B200: with 128 rows per CTA (4 warps, two
// of them on one scheduler) the row path took 12.5 us per panel and the POTRF warp waited for
// it.  64 rows on two warps with a scheduler each halve that and double the panel CTAs.
constexpr int kTrsmWarps = 2;
constexpr int kTrsmRows = 32 * kTrsmWarps;

// 1/sqrt(d) to fp64 accuracy from the fp32 MUFU.RSQ seed and two Newton steps: this sits on
// the pivot chain of the factorisation (D sequential pivots), so its latency matters.
__device__ __forceinline__ double chain_rsqrt(double d) {
  double r = static_cast<double>(rsqrtf(__double2float_rn(d)));
  double e = fma(-d * r, r, 1.0);
  r = fma(0.5 * r, e, r);
  e = fma(-d * r, r, 1.0);
  r = fma(0.5 * r, e, r);
  return r;
}

// 1/d to fp64 accuracy: MUFU.RCP64H seed (~20 bits) and two Newton steps of two dependent FMAs
// each.  The pivot chain only needs 1/d (to eliminate below the pivot); the square root that
// scales the column is computed beside it, off the chain.
__device__ __forceinline__ double chain_rcp(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  double e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  return r;
}

// 1/sqrt(d) from the fp64 MUFU.RSQ64H seed (2^-22.9) and two Newton steps: no conversion to fp32 and
// back on the way, and good for any normal double (the fp32-seeded chain_rsqrt needs 1e-30 < d < 1e30).
__device__ __forceinline__ double chain_rsqrt64(double d) {
  double r;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  double e = fma(-d * r, r, 1.0);
  r = fma(0.5 * r, e, r);
  e = fma(-d * r, r, 1.0);
  r = fma(0.5 * r, e, r);
  return r;
}

// A pivot the factorisation can use: positive, finite, 2^-997 <= d < 2^996 (so that the MUFU seeds of
// 1/d and 1/sqrt(d) and their Newton steps stay in range).  One integer compare on the high word
// instead of two DSETP: false for zero, negatives, denormals, infinities and NaN.
__device__ __forceinline__ bool pivot_usable(double d) {
  return static_cast<unsigned>(__double2hiint(d)) - 0x01a00000u < 0x7c900000u;
}

// ---- factorisation of the 32x32 diagonal block: one warp runs the elimination chain, another scales
//
// Measured (tools/ubench_potrf.cu, ubench_potrf2.cu; profiles/r2_i_potrf_chain.md): the 32-pivot loop
// is executed by ONE in-order warp, so every dependent instruction it carries is latency the next
// pivot waits for, whether or not it is on the mathematical critical path.  The loop that did
// everything (reciprocal, square root, column scaling, the d <= 0 / wide-range branches) took 404
// cycles per pivot alone on an SM, 316 without the branches, and 155 when it carries the chain only.
//
// warp_chol32_chain (warp 0): lane r holds row r of the block in registers.  Pivot j needs ONE update
// of pivot j - 1 (column j of every row); that update is done first, the unscaled column j goes to
// U[j][lane] at once, and the other row updates of pivot j - 1 issue beside the reciprocal of pivot j.
// Straight-line code: no branch on the pivot value (an unusable pivot is replaced by 1 and
// remembered), and no __syncwarp inside the loop -- here it would be a BRA.DIV, which ends a basic
// block and with it everything ptxas can overlap.  The warp is brought together once in front of
// the loop; straight-line code cannot diverge, and the LSU handles one warp's shared-memory
// instructions in order, so a store by lane a is visible to the load by lane b that follows it in
// program order.  WARP_FENCE keeps the compiler from reordering them.
// *prog_u = number of complete columns of U (U[c][c] is the pivot d_c).
// back-off of the warps that poll a progress counter in shared memory (build parameters for tools/)
#ifndef M3R_POLL_NS
#define M3R_POLL_NS 100
#endif
#ifndef M3R_POLL_U_NS
#define M3R_POLL_U_NS 32
#endif
#if M3R_POLL_NS > 0
#define M3R_POLL_SLEEP() __nanosleep(M3R_POLL_NS)
#else
#define M3R_POLL_SLEEP() do { } while (0)
#endif
#if M3R_POLL_U_NS > 0
#define M3R_POLL_U_SLEEP() __nanosleep(M3R_POLL_U_NS)
#else
#define M3R_POLL_U_SLEEP() do { } while (0)
#endif
#define M3R_WARP_FENCE() asm volatile("" ::: "memory")
__device__ __forceinline__ bool warp_chol32_chain(double a[kPB], int lane, double (*U)[kPB],
                                                  volatile int *prog_u) {
  __syncwarp();
  U[0][lane] = a[0];
  M3R_WARP_FENCE();
  *prog_u = 1;
  double d = U[0][0];
  bool abn = false;
  double tj_prev = 0.0;
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    if (j > 0) {  // what is left of pivot j - 1 (its update of column j was done before column j went out)
#pragma unroll
      for (int c = j + 1; c < kPB; c++) a[c] -= tj_prev * U[j - 1][c];  // meaningful for lanes >= c
    }
    // the reciprocal (MUFU.RCP64H seed) is good for any normal double; only the square root, which is
    // not computed here, needs the narrower range
    const bool usable = pivot_usable(d);  // false for d <= 0 and NaN
    abn |= !usable;
    const double rc = chain_rcp(usable ? d : 1.0);
    const double tj = a[j] * rc;  // A[lane][j] / d
    if (j + 1 < kPB) {
      a[j + 1] -= tj * U[j][j + 1];
      U[j + 1][lane] = a[j + 1];
      M3R_WARP_FENCE();
      *prog_u = j + 2;  // every lane stores the same value
      d = U[j + 1][j + 1];
    }
    tj_prev = tj;
  }
  M3R_WARP_FENCE();
  __syncwarp();
  return abn;
}

// follower side of prog_u
__device__ __forceinline__ void wait_columns(volatile int *prog_u, int need) {
  while (*prog_u < need) M3R_POLL_U_SLEEP();
  M3R_WARP_FENCE();  // see wait_pivot
}

// warp_chol32_scale (a warp that is idle during the factorisation): column c of L is column c of U times
// 1/sqrt(d_c).  Publishes L11[.][c] (zeros above the diagonal) and dinv[c] in groups of kFollowGroup
// columns, then raises *prog: the TRSM rows, the y_k chain and the inverse of the block follow it.
// An unusable pivot (d <= 0, NaN, outside 2^-997 .. 2^996) was replaced by 1 in the chain, which has
// flagged the step as failed; the factor is kept finite here the same way.
__device__ __forceinline__ void warp_chol32_scale(double (*U)[kPB], int lane, double *dinv,
                                                  double (*L11)[kPB + 1], volatile int *prog_u,
                                                  volatile int *prog);
#ifdef M3R_TEST_VARIANTS  // the one-warp-does-everything loop: the cluster solver of the variants build still uses it
// out-of-line so that the pivot chain's code stays small (never taken for sane systems)
__device__ __noinline__ double rsqrt_wide_range(double d) { return rsqrt(d); }
// 32x32 lower Cholesky by one warp: lane r holds row r in registers (a[c], c <= r meaningful).
// Returns the reciprocal pivots through dinv[] (shared) and sets *bad on a non-positive pivot.
// Column j travels between the lanes through `col` (32 doubles of shared memory), one round trip
// per pivot: every lane publishes its entry of column j, reads the pivot, and updates its row
// with a[c] -= (a[j]/d) * A[c][j].  (A 64-bit __shfl_sync costs ~12 SASS instructions; the
// shuffle version of this routine was 8k instructions and the kernel 300 KB, far beyond the
// instruction cache, which made every panel run at cold-cache speed.)
// After pivot j the warp publishes column j of L (L11[.][j]; zeros above the diagonal) and 1/L_jj,
// then raises *prog to j + 1: the TRSM warps, the y_k chain and (one panel later) the inverse of the
// block follow the factorisation one pivot behind instead of starting when all 32 pivots are done
// (the TRSM alone was 3.6 us per panel behind the POTRF, profiles/r1_i_coop_solver_phases.txt).
__device__ __forceinline__ void warp_chol32(double a[kPB], int lane, double *col, double *dinv,
                                            int *bad, double (*L11)[kPB + 1], volatile int *prog) {
#pragma unroll
  for (int j = 0; j < kPB; j++) {
    // one barrier does two jobs: the previous readers of col are done, and every lane's part of
    // column j-1 (written at the end of the previous round) precedes the counter that lets the
    // followers read it.  All lanes store the same value: no divergent branch on the chain.
    __syncwarp();
    if (j > 0) *prog = j;
    col[lane] = a[j];
    __syncwarp();
    const double d = col[j];
    double r, rc;
    if (!(d > 0.0)) {
      if (lane == 0) *bad = 1;
      r = 1.0, rc = 1.0;  // keep finite; the step is discarded
    } else if (d < 1e-30 || d > 1e30) {
      r = rsqrt_wide_range(d);
      rc = r * r;
    } else {
      rc = chain_rcp(d);    // on the chain: the next pivot waits for tj
      r = chain_rsqrt(d);   // beside it: only the stored factor needs the root
    }
    const double tj = a[j] * rc;  // A[lane][j] / d
#pragma unroll
    for (int c = j + 1; c < kPB; c++) a[c] -= tj * col[c];  // meaningful for lanes >= c
    dinv[j] = r;  // every lane holds the same r
    a[j] *= r;    // lane j: sqrt(d); lanes > j: L[lane][j]
    L11[lane][j] = (lane >= j) ? a[j] : 0.0;
  }
  __syncwarp();
  *prog = kPB;
}

#endif

// follower side: block until pivot c of this panel's POTRF is published
// (Called once per group of pivots, with a back-off: every poll is a shared-memory load in the
// pipe the POTRF chain's own stores and loads go through; polling every pivot from four warps cost
// the chain 15 %.)
__device__ __forceinline__ void wait_pivot(volatile int *prog, int c) {
  while (*prog <= c) M3R_POLL_SLEEP();
  // No MEMBAR: the producer warp stores the columns and then the counter, this warp loads the counter
  // and then the columns, and the SM's shared-memory pipe performs one warp's instructions in
  // order, so the compiler is the only thing that could reorder them.  __threadfence_block() here is a
  // MEMBAR.SC.CTA per group of pivots: measured, it keeps a follower 1050 cycles behind the
  // factorisation instead of 530 (tools/ubench_potrf2.cu), twice in series (scaling warp, then TRSM).
  M3R_WARP_FENCE();
}
constexpr int kFollowGroup = 4;  // TRSM rows and y_k follow the POTRF in groups of this many pivots
constexpr int kInvGroup = 8;     // the block inverse (off the critical path) in groups of this many

__device__ __forceinline__ void warp_chol32_scale(double (*U)[kPB], int lane, double *dinv,
                                                  double (*L11)[kPB + 1], volatile int *prog_u,
                                                  volatile int *prog) {
#pragma unroll 1
  for (int c0 = 0; c0 < kPB; c0 += kFollowGroup) {
    wait_columns(prog_u, c0 + kFollowGroup);
    double r[kFollowGroup], u[kFollowGroup];
#pragma unroll
    for (int i = 0; i < kFollowGroup; i++) {
      const double d = U[c0 + i][c0 + i];
      u[i] = U[c0 + i][lane];
      r[i] = pivot_usable(d) ? d : 1.0;  // an unusable pivot: the chain has flagged the step as failed
    }
#pragma unroll
    for (int i = 0; i < kFollowGroup; i++) r[i] = chain_rsqrt64(r[i]);  // four independent chains
#pragma unroll
    for (int i = 0; i < kFollowGroup; i++) {
      const int c = c0 + i;
      dinv[c] = r[i];  // every lane holds the same value
      L11[lane][c] = (lane >= c) ? u[i] * r[i] : 0.0;
    }
    __syncwarp();
    *prog = c0 + kFollowGroup;
  }
}

__device__ __forceinline__ void tile_from_index(int tile, int &tr, int &tc) {
  tr = static_cast<int>((sqrt(8.0 * tile + 1.0) - 1.0) * 0.5);
  while ((tr + 1) * (tr + 2) / 2 <= tile) tr++;
  while (tr * (tr + 1) / 2 > tile) tr--;
  tc = tile - tr * (tr + 1) / 2;
}

// Trailing update A22 -= L21 L21^T of one panel, the CTA's share of the lower TILE x TILE tiles (TILE = 32
// or 64; tiles first, first + stride, ...), software-pipelined: while tile i is multiplied out of shared
// memory, the operand panels L21[tr], L21[tc] and the accumulators of tile i + 1 are already in flight
// (registers), and the operand panels alternate between two shared-memory buffers so that one CTA
// barrier per tile is enough.  The first version took each tile on its own: accumulators from L2, barrier,
// operands from L2, barrier, 1.7 us of DFMA, store -- two serialised round trips per tile, and at
// D = 1785 CTA 0 waited 250 us per solve for this phase (profiles/r2_i_potrf_chain.md).
template <int TILE>
struct TileRegs {
  static constexpr int R = TILE / 16;                      // outputs per thread per dimension
  static constexpr int NL = TILE * kPB / kCoopThreads;     // operand elements per thread and panel
  double ra[NL], rb[NL], acc[R][R];
  int tr, tc;
};

template <int TILE>
__device__ __forceinline__ void tile_issue_loads(const double *__restrict__ A, int D, int kp, int base,
                                                 int tile, int t, TileRegs<TILE> &q) {
  constexpr int R = TileRegs<TILE>::R, NL = TileRegs<TILE>::NL;
  tile_from_index(tile, q.tr, q.tc);
  const int ty = t / 16, tx = t % 16;
#pragma unroll
  for (int i = 0; i < NL; i++) {
    const int x = t + i * kCoopThreads;
    const int r = x / kPB, c = x % kPB;
    const int ga = base + q.tr * TILE + r, gb = base + q.tc * TILE + r;
    q.ra[i] = (ga < D) ? A[static_cast<size_t>(ga) * D + kp + c] : 0.0;
    q.rb[i] = (gb < D) ? A[static_cast<size_t>(gb) * D + kp + c] : 0.0;
  }
#pragma unroll
  for (int i = 0; i < R; i++)
#pragma unroll
    for (int j = 0; j < R; j++) {
      const int gr = base + q.tr * TILE + ty + 16 * i, gc = base + q.tc * TILE + tx + 16 * j;
      q.acc[i][j] = (gr < D && gc <= gr) ? A[static_cast<size_t>(gr) * D + gc] : 0.0;
    }
}

template <int TILE>
__device__ __noinline__ void coop_update_tiles(double *__restrict__ A, int D, int kp, int base, int ntiles,
                                                  int first, int stride, double (*Ra)[64][kPB + 1],
                                                  double (*Rb)[64][kPB + 1], int t) {
  constexpr int R = TileRegs<TILE>::R, NL = TileRegs<TILE>::NL;
  if (first >= ntiles) return;  // uniform over the CTA
  const int ty = t / 16, tx = t % 16;
  TileRegs<TILE> cur;
  tile_issue_loads<TILE>(A, D, kp, base, first, t, cur);
  int b = 0;
  for (int tile = first;; tile += stride, b ^= 1) {
#pragma unroll
    for (int i = 0; i < NL; i++) {
      const int x = t + i * kCoopThreads;
      Ra[b][x / kPB][x % kPB] = cur.ra[i];
      Rb[b][x / kPB][x % kPB] = cur.rb[i];
    }
    __syncthreads();  // buffer b is staged; every thread is done with buffer b ^ 1 (the tile before)
    const bool more = tile + stride < ntiles;
    TileRegs<TILE> nxt;
    if (more) tile_issue_loads<TILE>(A, D, kp, base, tile + stride, t, nxt);
    double acc[R][R];
#pragma unroll
    for (int i = 0; i < R; i++)
#pragma unroll
      for (int j = 0; j < R; j++) acc[i][j] = cur.acc[i][j];
#pragma unroll 8
    for (int k = 0; k < kPB; k++) {
      double av[R], bv[R];
#pragma unroll
      for (int i = 0; i < R; i++) av[i] = Ra[b][ty + 16 * i][k], bv[i] = Rb[b][tx + 16 * i][k];
#pragma unroll
      for (int i = 0; i < R; i++)
#pragma unroll
        for (int j = 0; j < R; j++) acc[i][j] -= av[i] * bv[j];
    }
#pragma unroll
    for (int i = 0; i < R; i++)
#pragma unroll
      for (int j = 0; j < R; j++) {
        const int gr = base + cur.tr * TILE + ty + 16 * i, gc = base + cur.tc * TILE + tx + 16 * j;
        if (gr < D && gc <= gr) A[static_cast<size_t>(gr) * D + gc] = acc[i][j];
      }
    if (!more) break;
    cur = nxt;
  }
}

// M = L^-1 of one factored 32x32 diagonal block (by one warp, lane = column of M), stored row-major
// with the reciprocal pivots behind it: the layout the backward phase reads from `ldiag`.
__device__ __forceinline__ void coop_invert_block(double (*L)[kPB + 1], const double *dinv_blk,
                                                  double *__restrict__ Ld, int lane, volatile int *prog) {
  // lane c solves L m = e_c, right-looking: once m_k is known every later row takes its update,
  // so a pivot costs two dependent operations (the chain of a left-looking dot product grows with
  // the row: 9 us per block), and step k needs column k of L only -- the warp follows the POTRF of
  // the same panel one pivot behind, like the TRSM rows
  double acc[kPB];
#pragma unroll
  for (int r = 0; r < kPB; r++) acc[r] = (r == lane) ? 1.0 : 0.0;
#pragma unroll
  for (int k = 0; k < kPB; k++) {
    if ((k % kInvGroup) == 0) wait_pivot(prog, k + kInvGroup - 1);
    const double mk = acc[k] * dinv_blk[k];
    acc[k] = mk;
#pragma unroll
    for (int r = k + 1; r < kPB; r++) acc[r] -= L[r][k] * mk;
  }
#pragma unroll
  for (int r = 0; r < kPB; r++) Ld[r * kPB + lane] = acc[r];
  Ld[kPB * kPB + lane] = dinv_blk[lane];
}

// half of the CTA's staged panel rows (L21 after the TRSM) to global memory, one row per instruction.
// Full tiles (all but the last CTA's rows and the last panel) take a branch-free path: eight
// shared-memory loads, then eight stores at 32-bit row offsets.  (A guarded load-store pair per row
// compiles to a branch and a dependent chain of ten integer instructions: 70 cycles per row on an
// in-order warp, 1.4 us per panel behind everything else.)
__device__ __forceinline__ void coop_store_rows(double *__restrict__ A, int D, int k0, int pb, int row0,
                                                int nrows, double (*Xr)[kPB + 1], int half, int lane) {
  constexpr int kRowsPerWarp = kTrsmRows / 2;
  const int n = nrows - half * kRowsPerWarp;  // rows of this half that exist
  const double *src = &Xr[half * kRowsPerWarp][lane];
  double *dst = A + static_cast<size_t>(row0 + half * kRowsPerWarp) * D + k0 + lane;
  if (n >= kRowsPerWarp && pb == kPB) {
#pragma unroll
    for (int i = 0; i < kRowsPerWarp; i += 8) {
      double v[8];
#pragma unroll
      for (int q = 0; q < 8; q++) v[q] = src[(i + q) * (kPB + 1)];
#pragma unroll
      for (int q = 0; q < 8; q++) dst[(i + q) * D] = v[q];  // the system has fewer than 2^31 entries
    }
  } else if (lane < pb) {
#pragma unroll 1
    for (int i = 0; i < n; i++) dst[i * D] = src[i * (kPB + 1)];
  }
}

constexpr int kBackMultiMin = 512;  // systems above this size run the backward substitution on all CTAs

constexpr int kLs = kPB + 2;  // even leading dimension: 16-byte aligned pairs for LDS.128
struct CoopSmem {
  double L11[kPB][kPB + 1];
  double Ad[kPB][kPB + 1];   // diagonal block with the previous panel's update folded in
  double Ls[kPB][kLs];       // previous panel's L21, rows of the diagonal block
  double U[kPB][kPB];        // unscaled columns of the factor, U[c][row] (warp_chol32_chain)
  double dinv[kPB], yk[kPB], bk[kPB];
  union {
    struct {
      double Xr[kTrsmRows][kPB + 1];  // this CTA's rows of the panel
      double Lp[kTrsmRows][kLs];      // their L21 of the previous panel
    } a;
    struct {
      double Ra[2][64][kPB + 1], Rb[2][64][kPB + 1];  // phase B operand panels, double-buffered
    } b;
  } u;
  int bad;
  int prog;    // scaled columns of L11 (and dinv) that are published (warp_chol32_scale)
  int prog_u;  // complete columns of U (warp_chol32_chain)
#ifdef M3R_COOP_PROF
  long long t0;
  long long stamp[16];
#endif
};
constexpr int kCoopSmemBytes = static_cast<int>(sizeof(CoopSmem));

// T = X X^T for the first 32 rows X of the L21 this CTA has just computed: the update the NEXT panel's
// diagonal block needs.  Run by warps 4, 5, 6, 7 (one per scheduler) of the CTA of the first rows when their own work of
// the step is over (slot = 0..3), out of rows 0..31 of S.u.a.Lp (the TRSM warp of those rows stages them there); 64 register
// tiles of 4x4 outputs (rows tr + 8i, columns tc + 8k: consecutive threads read consecutive rows, whose
// 272-byte stride makes the 128-bit loads conflict-free) x 2 halves of the inner dimension, two
// accumulators per output.
__device__ __forceinline__ void coop_next_diag_update(CoopSmem &S, double *__restrict__ dT, int slot, int lane) {
  const int u = slot * 32 + lane;
  const int tile = u & 63, half = u >> 6;
  const int tr = tile >> 3, tc = tile & 7;
  double acc[4][4], acc1[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int k = 0; k < 4; k++) acc[i][k] = 0.0, acc1[i][k] = 0.0;
  asm volatile("bar.sync 6, 160;");  // the rows are staged
#pragma unroll 4
  for (int j = half * (kPB / 2); j < (half + 1) * (kPB / 2); j += 2) {
    double2 av[4], bv[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      av[i] = *reinterpret_cast<const double2 *>(&S.u.a.Lp[tr + 8 * i][j]);
      bv[i] = *reinterpret_cast<const double2 *>(&S.u.a.Lp[tc + 8 * i][j]);
    }
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int k = 0; k < 4; k++) {
        acc[i][k] = fma(av[i].x, bv[k].x, acc[i][k]);
        acc1[i][k] = fma(av[i].y, bv[k].y, acc1[i][k]);
      }
  }
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int k = 0; k < 4; k++) acc[i][k] += acc1[i][k];
  // The two halves of the inner dimension meet in global memory: fire-and-forget fp64 adds (RED) into a
  // zeroed T.  Two addends per entry: the sum does not depend on their order.  (Combining them through
  // shared memory, two named barriers and a staged copy was ~100 dependent instructions: 0.8 us at the
  // very end of the step.)
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int r = tr + 8 * i, c = tc + 8 * k;
      if (c <= r) atomicAdd(&dT[r * kPB + c], acc[i][k]);
    }
}

// Backward substitution L^T x = y for large systems, all CTAs (see the call site).  Out of line: it runs
// once per solve and must not take part in the register allocation of the panel loop.
__device__ __noinline__ void coop_backward_multi(cg::grid_group &grid, CoopSmem &S, double *__restrict__ A,
                                                 const double *__restrict__ ldiag, int D,
                                                 double *__restrict__ xbuf) {
  const int t = threadIdx.x;
  const int lane = t & 31, warp = t >> 5;
  double *yk = S.yk;
  constexpr int kSuper = 8 * kPB;
  constexpr int kLdBufM = kPB * kPB + kPB;
  double *xs = &S.u.a.Lp[0][0];   // the super-block's part of the vector (CTA 0)
  double *ldb = &S.u.a.Xr[0][0];  // 2 x kLdBufM doubles: M = L11^-1 of the current / next panel (CTA 0)
  double *red = &S.u.a.Xr[0][0];  // 8 x 32 partial sums (the other CTAs, and CTA 0 in the fold stage)
  const int nsuper = (D + kSuper - 1) / kSuper;
  for (int sp = nsuper - 1; sp >= 0; sp--) {
    const int c0 = sp * kSuper, c1 = min(D, c0 + kSuper);
    if (blockIdx.x == 0) {
      const int lastp = ((c1 - 1) / kPB) * kPB;
      for (int i = t; i < c1 - c0; i += kCoopThreads) xs[i] = __ldcg(&xbuf[c0 + i]);
      for (int x = t; x < kLdBufM; x += kCoopThreads)
        ldb[x] = ldiag[static_cast<size_t>(lastp / kPB) * kLdStride + x];
      __syncthreads();
      int buf = 0;
      const int jl = t - 32;  // update thread -> column c0 + jl of the super-block (224 >= 256 - 32 columns)
      for (int k0 = lastp; k0 >= c0; k0 -= kPB, buf ^= 1) {
        const int pb = min(kPB, D - k0);
        double lv[kPB];
        if (warp == 0) {  // x_k = M^T y_k
          const double *Md = ldb + buf * kLdBufM;
          const double *ys = xs + (k0 - c0);
          double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
          for (int r = 0; r < kPB; r += 4) {
            s0 += Md[(r + 0) * kPB + lane] * ((r + 0 < pb) ? ys[r + 0] : 0.0);
            s1 += Md[(r + 1) * kPB + lane] * ((r + 1 < pb) ? ys[r + 1] : 0.0);
            s2 += Md[(r + 2) * kPB + lane] * ((r + 2 < pb) ? ys[r + 2] : 0.0);
            s3 += Md[(r + 3) * kPB + lane] * ((r + 3 < pb) ? ys[r + 3] : 0.0);
          }
          const double mine = (s0 + s1) + (s2 + s3);
          __syncwarp();
          yk[lane] = (lane < pb) ? mine : 0.0;
          if (lane < pb) xs[k0 - c0 + lane] = mine;
        } else {
          if (c0 + jl < k0) {
#pragma unroll
            for (int c = 0; c < kPB; c++)
              lv[c] = (c < pb) ? A[static_cast<size_t>(k0 + c) * D + c0 + jl] : 0.0;
          }
          if (k0 > c0) {  // M of the next panel of this super-block
            const double *src = ldiag + static_cast<size_t>(k0 / kPB - 1) * kLdStride;
            for (int x = jl; x < kLdBufM; x += kCoopThreads - 32) ldb[(buf ^ 1) * kLdBufM + x] = src[x];
          }
        }
        __syncthreads();
        if (warp != 0 && c0 + jl < k0) {
          double sacc = xs[jl];
#pragma unroll
          for (int c = 0; c < kPB; c++) sacc -= lv[c] * yk[c];
          xs[jl] = sacc;
        }
        __syncthreads();
      }
      for (int i = t; i < c1 - c0; i += kCoopThreads) xbuf[c0 + i] = xs[i];
    }
    if (c0 == 0) break;  // (uniform) nothing in front of the first super-block
    __threadfence();
    grid.sync();  // x of the super-block is in xbuf
    // fold: y_j -= sum_{r in [c0, c1)} L[r][j] x_r for j < c0; a CTA takes 32 columns per pass, warp w
    // the rows c0 + 32 w ..., lane = column: every load instruction is one 256-byte row segment
    for (int jc = blockIdx.x * 32; jc < c0; jc += gridDim.x * 32) {
      const int j = jc + lane;
      const int r0 = c0 + warp * kPB;
      double xv[kPB], part = 0.0;
#pragma unroll
      for (int c = 0; c < kPB; c++) xv[c] = (r0 + c < c1) ? __ldcg(&xbuf[r0 + c]) : 0.0;
      if (j < c0) {
        double lvf[kPB];
#pragma unroll
        for (int c = 0; c < kPB; c++)
          lvf[c] = (r0 + c < c1) ? A[static_cast<size_t>(r0 + c) * D + j] : 0.0;
        double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
#pragma unroll
        for (int c = 0; c < kPB; c += 4) {
          p0 = fma(lvf[c], xv[c], p0), p1 = fma(lvf[c + 1], xv[c + 1], p1);
          p2 = fma(lvf[c + 2], xv[c + 2], p2), p3 = fma(lvf[c + 3], xv[c + 3], p3);
        }
        part = (p0 + p1) + (p2 + p3);
      }
      __syncthreads();  // red is free (previous pass, or CTA 0's use of the buffer as ldb)
      red[warp * 32 + lane] = part;
      __syncthreads();
      if (warp == 0 && j < c0) {
        double sum = 0.0;
#pragma unroll
        for (int w = 0; w < kCoopThreads / 32; w++) sum += red[w * 32 + lane];
        xbuf[j] = __ldcg(&xbuf[j]) - sum;
      }
    }
    __threadfence();
    grid.sync();  // y in front of the super-block is up to date
  }
}

__global__ void __launch_bounds__(kCoopThreads)
    gn_chol_solve_coop_kernel(double *__restrict__ sys, double *__restrict__ ldiag, int D,
                              double *__restrict__ xbuf, float *__restrict__ Twc,
                              float *__restrict__ dx, int num_poses, float delta_thresh,
                              int *__restrict__ flags) {
  cg::grid_group grid = cg::this_grid();
  if (flags[0]) return;  // uniform over the grid (the flag only changes after the last barrier)
  extern __shared__ __align__(16) unsigned char coop_smem_raw[];
  CoopSmem &S = *reinterpret_cast<CoopSmem *>(coop_smem_raw);
  double (*Ra)[64][kPB + 1] = S.u.b.Ra;
  double (*Rb)[64][kPB + 1] = S.u.b.Rb;
  double *yk = S.yk;
  int &bad = S.bad;
  const int t = threadIdx.x;
  const int lane = t & 31, warp = t >> 5;
  double *A = sys;
  double *rhs = sys + static_cast<size_t>(D) * D;
  // two kPB x kPB scratch blocks behind the diagonal blocks: panel p reads T from block p & 1 and CTA 0
  // builds the next one in block (p + 1) & 1, which it zeroes at the start of the step (its last
  // readers were the loads of panel p - 1, a grid barrier ago)
  double *dT2 = ldiag + static_cast<size_t>((D + kPB - 1) / kPB) * kLdStride;
  if (t == 0) bad = 0;
#ifdef M3R_COOP_PROF
  long long prof_acc[6] = {0, 0, 0, 0, 0, 0};
  long long prof_t = clock64();
  long long aprof[6] = {0, 0, 0, 0, 0, 0};
  long long tprof[4] = {0, 0, 0, 0};  // thread 32 (first TRSM warp) of CTA 0, see COOP_TTICK
  long long tprof_t = 0;
  long long aprof_t = 0;
#endif
  __syncthreads();
  // warp roles in phase A (scheduler = warp % 4): 0 = elimination chain of the diagonal block, 4 = square
  // roots and scaled columns behind it (same scheduler as the chain, whose instructions are sparse; on
  // scheduler 3 it queued behind the DFMAs of the inverse and needed 700 cycles per group of four
  // columns), 1, 2 = the CTA's 64 panel rows (look-ahead update, TRSM, rhs), 3 = inverse of the block
  // (CTA of the first rows only), 5 = forward-substitution chain y_k; 0 and 3 then store the rows of L21;
  // 4, 5, 6, 7 of the CTA of the first rows (one per scheduler): diagonal update of the NEXT panel once
  // their own work is done
  const bool trsm_warp = warp >= 1 && warp <= kTrsmWarps;  // warps 1, 2: schedulers 1, 2
  const int trsm_slot = warp - 1;

  for (int k0 = 0; k0 < D; k0 += kPB) {
    const int pb = min(kPB, D - k0);
    const int m = D - k0 - pb;  // rows below the panel
    const int row_ctas = max(1, (m + kTrsmRows - 1) / kTrsmRows);
    // Look-ahead: the trailing update of the PREVIOUS panel (columns kp..kp+31) has not been
    // applied to this panel's columns yet; the panel CTAs fold it into their loads, while the
    // other CTAs apply it to the rest of the trailing matrix (phase B below) at the same time.
    const bool la = k0 > 0;
    const int kp = k0 - kPB;
    // ---------------- phase A (panel CTAs)
    for (int rc = blockIdx.x; rc < row_ctas; rc += gridDim.x) {
#ifdef M3R_COOP_PROF
      if (blockIdx.x == 0 && t == 0) aprof_t = clock64(), S.t0 = aprof_t;
      __syncthreads();
#endif
      const int row0 = k0 + pb + rc * kTrsmRows;
      const int nrows = max(0, min(kTrsmRows, D - row0));
      // the CTA of the first rows prepares the next panel's diagonal update (those rows ARE its Ls)
      const bool need_T = rc == 0 && k0 + pb < D;
      const double *dT = dT2 + ((k0 / kPB) & 1) * (kPB * kPB);
      double *dTn = dT2 + (((k0 / kPB) + 1) & 1) * (kPB * kPB);
      if (need_T) {
#pragma unroll
        for (int i = 0; i < kPB * kPB / kCoopThreads; i++) dTn[t + i * kCoopThreads] = 0.0;
      }
      double (*L11)[kPB + 1] = S.L11;
      double *dinv = S.dinv;
      volatile int *prog = &S.prog;  // pivots of this panel's POTRF published so far
      // 1. everything phase A reads from L2, in ONE round trip by all threads: the diagonal block, the
      //    previous panel's rows of it, the rhs, and the CTA's panel rows with their L21 of the
      //    previous panel.  (Round 1 staged the rows after this barrier, beside the POTRF, which was
      //    the critical path then; since the factorisation warp carries the chain only, the row path
      //    -- loads, look-ahead fold, TRSM -- is, and a second round trip in front of it costs 1 us per panel.)
      {
        double vd[kPB * kPB / kCoopThreads], vs[kPB * kPB / kCoopThreads], vt[kPB * kPB / kCoopThreads];
        double vx[kTrsmRows * kPB / kCoopThreads], vl[kTrsmRows * kPB / kCoopThreads];
#pragma unroll
        for (int i = 0; i < kPB * kPB / kCoopThreads; i++) {
          const int x = t + i * kCoopThreads;
          const int r = x / kPB, c = x % kPB;
          const size_t g = static_cast<size_t>(k0 + r) * D;
          vd[i] = (r >= pb && r == c) ? 1.0 : 0.0;  // identity on the padding of the last panel
          if (r < pb && c <= r) vd[i] = A[g + k0 + c];
          vt[i] = (la && r < pb && c <= r) ? dT[x] : 0.0;  // subtracted below: no arithmetic between the loads
          vs[i] = (la && r < pb) ? A[g + kp + c] : 0.0;
        }
#pragma unroll
        for (int i = 0; i < kTrsmRows * kPB / kCoopThreads; i++) {
          const int x = t + i * kCoopThreads;
          const int r = x / kPB, c = x % kPB;
          const size_t g = static_cast<size_t>(row0 + r) * D;
          vx[i] = (r < nrows && c < pb) ? A[g + k0 + c] : 0.0;
          vl[i] = (la && r < nrows) ? A[g + kp + c] : 0.0;
        }
        if (t < kPB) S.bk[t] = (t < pb) ? rhs[k0 + t] : 0.0;
        if (t == 0) S.prog = 0, S.prog_u = 0;  // nothing of this panel is published yet
#pragma unroll
        for (int i = 0; i < kPB * kPB / kCoopThreads; i++) {
          const int x = t + i * kCoopThreads;
          S.Ad[x / kPB][x % kPB] = vd[i] - vt[i];
          S.Ls[x / kPB][x % kPB] = vs[i];
        }
#pragma unroll
        for (int i = 0; i < kTrsmRows * kPB / kCoopThreads; i++) {
          const int x = t + i * kCoopThreads;
          S.u.a.Xr[x / kPB][x % kPB] = vx[i];
          S.u.a.Lp[x / kPB][x % kPB] = vl[i];
        }
      }
      __syncthreads();
      COOP_ATICK(8);
#ifdef M3R_COOP_PROF
      if (blockIdx.x == 0 && t == 32) tprof_t = clock64();
#endif
      // 2. (the update of the diagonal block by the previous panel, 1.5 us in front of the chain, is no
      //    longer done here: CTA 0 computed T = L21_top L21_top^T at the end of the previous step, when
      //    four of its warps were idle, and the loads above subtracted it -- coop_next_diag_update)
      COOP_ATICK(9);
      if (warp == 0) {
        // 3. POTRF, elimination chain (rows in registers); warp 4 scales the columns behind it
        double a[kPB];
#pragma unroll
        for (int c = 0; c < kPB; c++) a[c] = (c <= lane) ? S.Ad[lane][c] : 0.0;
        COOP_STAMP(0);
        if (warp_chol32_chain(a, lane, S.U, &S.prog_u) && lane == 0) bad = 1;  // step discarded
        COOP_STAMP(1);
        asm volatile("bar.sync 5, 128;");  // the TRSM warps have staged their rows of L21
        coop_store_rows(A, D, k0, pb, row0, nrows, S.u.a.Xr, 1, lane);
        COOP_ATICK(10);
      } else if (warp == 4) {
        COOP_STAMP(2);
        warp_chol32_scale(S.U, lane, dinv, L11, &S.prog_u, prog);  // publishes L11 / dinv, four columns at a time
        COOP_STAMP(3);
        if (need_T) coop_next_diag_update(S, dTn, 0, lane);
      } else if (trsm_warp) {
        // 4. this thread's row: fold in the previous panel's update while warp 0 factors
        const int r = trsm_slot * 32 + lane;
        const int row = row0 + r;
        const bool has_row = r < nrows;
        double x[kPB];
#pragma unroll
        for (int c = 0; c < kPB; c++) x[c] = S.u.a.Xr[r][c];
        if (la) {
#pragma unroll 2
          for (int j = 0; j < kPB; j += 2) {
            const double2 lp = *reinterpret_cast<const double2 *>(&S.u.a.Lp[r][j]);
#pragma unroll
            for (int c = 0; c < kPB; c++) {
              const double2 b2 = *reinterpret_cast<const double2 *>(&S.Ls[c][j]);
              x[c] = fma(-lp.y, b2.y, fma(-lp.x, b2.x, x[c]));  // two DFMA (a sum of products costs three)
            }
          }
        }
        COOP_TTICK(0);
        if (warp == 1) COOP_STAMP(4);
        const double rhs_old = has_row ? rhs[row] : 0.0;  // in flight during the TRSM
        COOP_TTICK(1);
        // right-looking TRSM: independent updates per column -> ILP.  Column c of L11 is final
        // once pivot c is published, so the row follows the POTRF one pivot behind.
#pragma unroll
        for (int c = 0; c < kPB; c++) {
          if ((c % kFollowGroup) == 0) wait_pivot(prog, c + kFollowGroup - 1);
          const double xc = x[c] * dinv[c];
          x[c] = xc;
#pragma unroll
          for (int k = c + 1; k < kPB; k++) x[k] -= xc * L11[k][c];
        }
        if (warp == 1) COOP_STAMP(5);
        // the rows go back through shared memory: a thread storing its own row issues 32 STG that
        // each touch 32 different lines (1 us of LSU time per panel, behind the factorisation);
        // staged, the two spare warps store whole rows, one coalesced instruction per row
#pragma unroll
        for (int c = 0; c < kPB; c++) S.u.a.Xr[r][c] = x[c];  // this thread's own slots
        asm volatile("bar.arrive 5, 128;");  // warps 0 and 3 store them; this warp goes on to the rhs
        if (need_T && trsm_slot == 0) {
          // rows 0..31 are the next panel's Ls: staged for coop_next_diag_update where only this warp
          // has been reading (its own rows of Lp, in the fold; S.Ls may still be read by the other TRSM warp)
#pragma unroll
          for (int c = 0; c < kPB; c++) S.u.a.Lp[r][c] = x[c];
          asm volatile("bar.arrive 6, 160;");
        }
        COOP_TTICK(2);
        if (warp == 1) COOP_STAMP(6);
        asm volatile("bar.sync 2, %0;" ::"n"(kTrsmRows + 32));  // y_k ready
        if (has_row) {
          double d0 = 0.0, d1 = 0.0, d2 = 0.0, d3 = 0.0;
#pragma unroll
          for (int c = 0; c < kPB; c += 4) {
            d0 += x[c] * yk[c], d1 += x[c + 1] * yk[c + 1];
            d2 += x[c + 2] * yk[c + 2], d3 += x[c + 3] * yk[c + 3];
          }
          rhs[row] = rhs_old - ((d0 + d1) + (d2 + d3));
        }
        COOP_TTICK(3);
        if (warp == 1) COOP_STAMP(12);
      } else if (warp == 3) {
        // 6. M = L11^-1 for the backward phase (x_k = M^T y_k is then 32 independent dot products
        //    instead of a 32-pivot chain), on a warp that has nothing else to do
        if (rc == 0) {
          coop_invert_block(L11, dinv, ldiag + static_cast<size_t>(k0 / kPB) * kLdStride, lane, prog);
          COOP_STAMP(8);
        }
        asm volatile("bar.sync 5, 128;");  // the TRSM warps have staged their rows of L21
        COOP_STAMP(10);
        coop_store_rows(A, D, k0, pb, row0, nrows, S.u.a.Xr, 0, lane);
        COOP_STAMP(11);
      } else if (warp == 6 || warp == 7) {
        if (need_T) coop_next_diag_update(S, dTn, warp == 6 ? 2 : 3, lane);
      } else if (warp == 5) {
        // 5. y_k = L11^-1 b_k (serial chain over the pivots), off the POTRF warp and one pivot
        //    behind it: step c needs L11[.][c] and dinv[c] only
        double mine = S.bk[lane];
#pragma unroll
        for (int c = 0; c < kPB; c++) {  // y_c travels through shared memory (see warp_chol32)
          if ((c % kFollowGroup) == 0) wait_pivot(prog, c + kFollowGroup - 1);
          if (lane == c) mine *= dinv[c], yk[c] = mine;
          __syncwarp();
          if (lane > c) mine -= L11[lane][c] * yk[c];
        }
        __syncwarp();
        if (lane >= pb) yk[lane] = 0.0;
        COOP_STAMP(7);
        asm volatile("bar.sync 2, %0;" ::"n"(kTrsmRows + 32));
        if (rc == 0 && lane < pb) xbuf[k0 + lane] = mine;
        if (need_T) coop_next_diag_update(S, dTn, 1, lane);
      }
      __syncthreads();
      if (warp == 0) COOP_STAMP(9);
#ifdef M3R_COOP_PROF
      __syncthreads();
      if (blockIdx.x == 0 && t < 16) g_coop_stamp[t] += S.stamp[t];
#endif
      COOP_ATICK(13);
    }
    COOP_TICK(0);
    // ---------------- phase B: A22 -= L21 L21^T of the previous panel on rows/cols >= k0 + pb,
    // by the CTAs that are not on the panel's critical path (all of them if the grid is small)
    if (la && m > 0) {
      const int base = k0 + pb;
      const int nu = static_cast<int>(gridDim.x) > row_ctas ? static_cast<int>(gridDim.x) - row_ctas
                                                            : static_cast<int>(gridDim.x);
      const int u = static_cast<int>(gridDim.x) > row_ctas ? static_cast<int>(blockIdx.x) - row_ctas
                                                           : static_cast<int>(blockIdx.x);
      if (u >= 0) {
        const int T64 = (m + 63) / 64;
        if (T64 * (T64 + 1) / 2 >= 2 * nu) {
          coop_update_tiles<64>(A, D, kp, base, T64 * (T64 + 1) / 2, u, nu, Ra, Rb, t);
        } else {  // few tiles: 32x32 tiles spread the update over more SMs
          const int T32 = (m + 31) / 32;
          coop_update_tiles<32>(A, D, kp, base, T32 * (T32 + 1) / 2, u, nu, Ra, Rb, t);
        }
      }
    }
    COOP_TICK(2);
    grid.sync();
    COOP_TICK(1);
  }
  // ---------------- backward substitution L^T x = y, x overwrites xbuf
  // Large systems (D > kBackMultiMin), all CTAs: one CTA streaming the whole of L (12.7 MB at D = 1785)
  // through its SM's L2 port took 213 us of a 0.78 ms solve, a quarter of it.  The columns are taken in
  // super-blocks of 8 panels: CTA 0 solves inside the super-block (the triangle it reads is 256 columns
  // wide at most: one load per thread and step), then every CTA folds the super-block's x into a slice
  // of the columns in front of it (32 columns x 256 rows per CTA and pass, summed over the 8 warps),
  // with a grid barrier on either side.  y and x live in xbuf (global) between the stages.
  if (D > kBackMultiMin) {
    coop_backward_multi(grid, S, A, ldiag, D, xbuf);
    if (blockIdx.x != 0) return;
    __syncthreads();
    const bool failm = bad != 0;
    COOP_TICK(4);
    finish_step<kCoopThreads>(t, failm, D, xbuf, Twc, dx, num_poses, delta_thresh, flags);
    COOP_TICK(5);
#ifdef M3R_COOP_PROF
    if (t == 0) {
      for (int k = 0; k < 6; k++) g_coop_prof[k] += prof_acc[k];
      for (int k = 0; k < 6; k++) g_coop_prof[8 + k] += aprof[k];
      g_coop_prof[6] += 1;
    }
#endif
    return;
  }
  if (blockIdx.x != 0) return;

  // ---------------- small and medium systems: CTA 0 alone
  // The vector lives in shared memory for the whole phase, the diagonal block of the NEXT step
  // is prefetched by the idle warps while warp 0 walks the current 32-pivot chain, and the
  // update threads issue their loads of L before the barrier that publishes x_k: per block the
  // critical path is one chain plus 32 FMAs instead of chain + three L2 round trips.
  __syncthreads();
  const bool fail = bad != 0;
  constexpr int kLdBuf = kPB * kPB + kPB;
  static_assert(kBackMultiMin <= kTrsmRows * kLs, "the vector has to fit S.u.a.Lp");
  if (!fail) {
    double *xs = &S.u.a.Lp[0][0];
    double *ldb = &S.u.a.Xr[0][0];  // 2 x kLdBuf doubles (Xr holds 128 x 33)
    const int last = ((D - 1) / kPB) * kPB;
    for (int j = t; j < D; j += kCoopThreads) xs[j] = xbuf[j];
    for (int x = t; x < kLdBuf; x += kCoopThreads)
      ldb[x] = ldiag[static_cast<size_t>(last / kPB) * kLdStride + x];
    __syncthreads();
    int buf = 0;
    constexpr int kUpd = kCoopThreads - 32;  // update threads (warps 1..7)
    // Update thread j0 loads the step's 32 rows of L at BOTH of its first two columns (j0 and j0 + 224:
    // D <= 480 needs no third) before the barrier, beside warp 0's multiply: one L2 round trip per step
    // where k0 > 224 used to take two.  (Keeping the first column one step ahead in registers as well
    // was tried: 0.1575 ms at D = 441 but spills, and slower at D = 105 and D = 1785.)
    const int j0 = t - 32, j1 = j0 + kUpd;
    for (int k0 = last; k0 >= 0; k0 -= kPB, buf ^= 1) {
      const int pb = min(kPB, D - k0);
      double lvA[kPB], lvB[kPB];
      if (warp == 0) {  // x_k = M^T y_k, M = L11^-1 (row-major in the buffer): one dot product per lane
        const double *Md = ldb + buf * kLdBuf;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
        for (int r = 0; r < kPB; r += 4) {
          s0 += Md[(r + 0) * kPB + lane] * ((r + 0 < pb) ? xs[k0 + r + 0] : 0.0);
          s1 += Md[(r + 1) * kPB + lane] * ((r + 1 < pb) ? xs[k0 + r + 1] : 0.0);
          s2 += Md[(r + 2) * kPB + lane] * ((r + 2 < pb) ? xs[k0 + r + 2] : 0.0);
          s3 += Md[(r + 3) * kPB + lane] * ((r + 3 < pb) ? xs[k0 + r + 3] : 0.0);
        }
        const double mine = (s0 + s1) + (s2 + s3);
        __syncwarp();  // every lane has read y_k before it is overwritten by x_k
        yk[lane] = (lane < pb) ? mine : 0.0;
        if (lane < pb) xs[k0 + lane] = mine;
      } else {
        // loads first (they only land in registers), then the prefetch that has to wait for
        // its data before it can store to shared memory
        if (j0 < k0) {
#pragma unroll
          for (int c = 0; c < kPB; c++)
            lvA[c] = (c < pb) ? A[static_cast<size_t>(k0 + c) * D + j0] : 0.0;
        }
        if (j1 < k0) {
#pragma unroll
          for (int c = 0; c < kPB; c++)
            lvB[c] = (c < pb) ? A[static_cast<size_t>(k0 + c) * D + j1] : 0.0;
        }
        if (k0 > 0) {
          const double *src = ldiag + static_cast<size_t>(k0 / kPB - 1) * kLdStride;
          double pf[(kLdBuf + kUpd - 1) / kUpd];
#pragma unroll
          for (int i = 0; i < (kLdBuf + kUpd - 1) / kUpd; i++)
            pf[i] = (j0 + i * kUpd < kLdBuf) ? src[j0 + i * kUpd] : 0.0;
#pragma unroll
          for (int i = 0; i < (kLdBuf + kUpd - 1) / kUpd; i++)
            if (j0 + i * kUpd < kLdBuf) ldb[(buf ^ 1) * kLdBuf + j0 + i * kUpd] = pf[i];
        }
      }
      __syncthreads();
      // y_j -= sum_c L[k0+c][j] x_k[c] for all j < k0: independent, coalesced rows of L
      if (warp != 0) {
        if (j0 < k0) {
          double sacc = xs[j0];
#pragma unroll
          for (int c = 0; c < kPB; c++) sacc -= lvA[c] * yk[c];
          xs[j0] = sacc;
        }
        if (j1 < k0) {
          double sacc = xs[j1];
#pragma unroll
          for (int c = 0; c < kPB; c++) sacc -= lvB[c] * yk[c];
          xs[j1] = sacc;
        }
        for (int j = j1 + kUpd; j < k0; j += kUpd) {
#pragma unroll
          for (int c = 0; c < kPB; c++)
            lvB[c] = (c < pb) ? A[static_cast<size_t>(k0 + c) * D + j] : 0.0;
          double sacc = xs[j];
#pragma unroll
          for (int c = 0; c < kPB; c++) sacc -= lvB[c] * yk[c];
          xs[j] = sacc;
        }
      }
      __syncthreads();
    }
    for (int j = t; j < D; j += kCoopThreads) xbuf[j] = xs[j];
    __syncthreads();
  }
  COOP_TICK(4);
  finish_step<kCoopThreads>(t, fail, D, xbuf, Twc, dx, num_poses, delta_thresh, flags);
  COOP_TICK(5);
#ifdef M3R_COOP_PROF
  if (t == 0) {
    for (int k = 0; k < 6; k++) g_coop_prof[k] += prof_acc[k];
    for (int k = 0; k < 6; k++) g_coop_prof[8 + k] += aprof[k];
    g_coop_prof[6] += 1;
  }
  if (t == 32) {
    g_coop_prof[11] += tprof[0], g_coop_prof[12] += tprof[1];
    g_coop_prof[14] += tprof[2], g_coop_prof[15] += tprof[3];
  }
#endif
}
