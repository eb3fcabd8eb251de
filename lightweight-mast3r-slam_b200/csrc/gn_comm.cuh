// Multi-GPU exchange of the assembled Gauss-Newton systems over NVLink peer memory
// (include/m3r_backends.h, "Multi-GPU exchange").  Included by gn.cu after the workspace.
//
// One-shot all-reduce, written for this payload (0.8 MB of fp64 at 64 keyframes, 12.8 MB at
// 256): every rank reads the W-1 remote buffers directly through NVLink and adds them in rank
// order, so all ranks obtain bitwise identical sums and the replicated Cholesky solves stay in
// lockstep.  One arrival flag per (reader, writer) pair carries the round number; buffers are
// double-buffered by round parity, which makes a second barrier unnecessary (header comment).
//
// Large systems (256 keyframes: 25.5 MB) on 4+ ranks use a two-phase variant instead, because
// the one-shot form makes every rank pull W-1 whole buffers (89 MB inbound at 8 ranks, 0.22 ms):
// reduce-scatter -- rank p sums rows [b_p, b_p+1) of all buffers in rank order into its `red`
// buffer -- then all-gather -- every rank copies each peer's reduced slice.  Inbound traffic per
// rank drops to 2 (W-1)/W of ONE buffer; every element is still summed by exactly one rank in
// rank order, so the replicas stay bitwise identical.
#pragma once

namespace m3r {

constexpr int kCommMaxWorld = 16;
constexpr int kCommFlagBytes = 1024;  // flag block at the start of the exchange area

__device__ __forceinline__ int ld_acquire_sys(const int *p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(int *p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ double2 ld_relaxed_sys_f64x2(const double *p) {
  double2 v;
  asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double ld_relaxed_sys_f64(const double *p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}

struct CommTable {
  int *flags[kCommMaxWorld];            // flags[p]: flag block of rank p (peer-mapped)
  const double *buf[2][kCommMaxWorld];  // buf[parity][p]: system buffer of rank p
  double *red[2][kCommMaxWorld];        // red[parity][p]: reduced slice of rank p (two-phase)
};

struct CommSlices {
  int begin[kCommMaxWorld + 1];  // rank p reduces rows [begin[p], begin[p+1]) of the D+1 rows
};

// "my buffer of round `round` is complete": one release store per peer.  The buffer was written
// by earlier kernels on this stream (visible at the device's L2, which is where peer reads land).
// The handshake (this signal and the wait in the reduce kernels) runs in EVERY round, also after
// the solve has converged and the data movement is skipped: the wait of round r+1 is what proves
// that every peer has finished reading this rank's buf[r & 1] before round r+2 overwrites it, and
// the host-side round counter advances whether or not the device still iterates.  A call that
// stops early therefore leaves the flags in step with the counter, whatever the parity of the
// rounds it skipped, and the next call on the same communicator starts from a clean handshake.
__global__ void comm_signal_kernel(const CommTable T, int rank, int world, int round, int flag_set) {
  const int p = threadIdx.x;
  if (p < world) {
    __threadfence_system();
    st_release_sys(T.flags[p] + flag_set * kCommMaxWorld + rank, round);
  }
}

// wall clock in ns, the same on every SM (clock64() is a per-SM cycle counter: it is neither
// comparable between SMs nor tied to wall time while the clocks ramp)
__device__ __forceinline__ unsigned long long comm_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
constexpr unsigned long long kCommTimeoutNs = 4000000000ull;  // 4 s: a peer died; do not hang the GPU

// true when every rank's flag of `flag_set` has reached `round` (false: timed out)
__device__ __forceinline__ bool comm_wait_all(const CommTable &T, int rank, int world, int round,
                                              int flag_set, int *timed_out) {
  if (threadIdx.x == 0) *timed_out = 0;
  __syncthreads();
  if (threadIdx.x < world) {
    const int *f = T.flags[rank] + flag_set * kCommMaxWorld + threadIdx.x;
    const unsigned long long t0 = comm_now_ns();
    while (ld_acquire_sys(f) < round) {
      if (comm_now_ns() - t0 > kCommTimeoutNs) {
        *timed_out = 1;
        break;
      }
      __nanosleep(100);
    }
  }
  __syncthreads();
  return *timed_out == 0;
}

__device__ __forceinline__ int comm_row_len(int r, int D) { return (r < D) ? 7 * (r / 7 + 1) : D; }

// A peer did not arrive: report it (flags[3], read by m3r_gn_iterations_done) and stop the solve --
// with flags[0] set the linearisation, the factorisation and the retraction that are already
// enqueued behind this kernel return at their first instruction, so Twc is not updated from a
// stale or partly written system.  m3r_gn_begin clears both.  EVERY CTA that times out reports (a
// CTA that gave up has not written its rows, whatever CTA 0 saw).
__device__ __forceinline__ void comm_fail(int *ws_flags) {
  if (threadIdx.x == 0) {
    ws_flags[3] = kGnErrExchangeTimeout;
    ws_flags[2] = 1;
    __threadfence();
    ws_flags[0] = 1;
  }
}

// two-phase, phase 1: reduce this rank's rows over all ranks' buffers into red[par][rank]
__global__ void __launch_bounds__(256)
    comm_reduce_scatter_kernel(const CommTable T, const CommSlices S, int rank, int world, int round,
                               int D, int *__restrict__ ws_flags) {
  __shared__ int timed_out;
  const bool stopped = ws_flags[0] != 0;  // read before the wait: comm_fail() of another CTA sets it
  if (!comm_wait_all(T, rank, world, round, 0, &timed_out)) {
    comm_fail(ws_flags);
    return;
  }
  if (stopped) return;  // converged earlier: handshake only
  const int par = round & 1;
  double *dst = T.red[par][rank];
  for (int r = S.begin[rank] + blockIdx.x; r < S.begin[rank + 1]; r += gridDim.x) {
    const int len = comm_row_len(r, D);
    const size_t off = static_cast<size_t>(r) * D;
    for (int c = threadIdx.x; c < len; c += blockDim.x) {
      double s = 0.0;
      for (int p = 0; p < world; p++) s += ld_relaxed_sys_f64(T.buf[par][p] + off + c);
      dst[off + c] = s;
    }
  }
}

// two-phase, phase 2: copy every rank's reduced rows
__global__ void __launch_bounds__(256)
    comm_gather_kernel(const CommTable T, const CommSlices S, int rank, int world, int round,
                       double *__restrict__ out, int D, int *__restrict__ ws_flags) {
  __shared__ int timed_out;
  const bool stopped = ws_flags[0] != 0;
  if (!comm_wait_all(T, rank, world, round, 1, &timed_out)) {
    comm_fail(ws_flags);
    return;
  }
  if (stopped) return;
  const int par = round & 1;
  for (int r = blockIdx.x; r <= D; r += gridDim.x) {
    int p = 0;
    while (r >= S.begin[p + 1]) p++;
    const double *src = T.red[par][p];
    const int len = comm_row_len(r, D);
    const size_t off = static_cast<size_t>(r) * D;
    for (int c = threadIdx.x; c < len; c += blockDim.x) out[off + c] = ld_relaxed_sys_f64(src + off + c);
  }
}

// Wait for every rank's flag of this round, then out = sum over ranks (rank order) of the
// systems: the block-lower triangle of the D x D matrix (the only part the solver reads; a
// block row is 7 scalar rows) and the D right-hand-side entries behind it.  One CTA per
// scalar row, remote rows read as coalesced 8-byte loads over NVLink.
__global__ void __launch_bounds__(256)
    comm_reduce_kernel(const CommTable T, int rank, int world, int round, double *__restrict__ out,
                       int D, int *__restrict__ ws_flags) {
  __shared__ int timed_out;
  const bool stopped = ws_flags[0] != 0;
  if (!comm_wait_all(T, rank, world, round, 0, &timed_out)) {
    comm_fail(ws_flags);
    return;
  }
  if (stopped) return;
  const int par = round & 1;
  for (int r = blockIdx.x; r <= D; r += gridDim.x) {  // row D = the right-hand side
    const int len = (r < D) ? 7 * (r / 7 + 1) : D;
    const size_t off = static_cast<size_t>(r) * D;
    for (int c = threadIdx.x; c < len; c += blockDim.x) {
      double s = 0.0;
      for (int p = 0; p < world; p++) s += ld_relaxed_sys_f64(T.buf[par][p] + off + c);
      out[off + c] = s;
    }
  }
}

}  // namespace m3r

struct m3r_gn_comm {
  int device = 0, rank = 0, world = 1;
  long long cap = 0;          // doubles per system buffer (even)
  unsigned char *area = nullptr;  // [flags | buf0 | buf1 | red0 | red1], cudaMalloc'ed here
  void *peer_area[m3r::kCommMaxWorld] = {nullptr};
  bool opened[m3r::kCommMaxWorld] = {false};  // mapped through cudaIpcOpenMemHandle
  bool connected = false;
  int round = 0;
  m3r::CommTable table{};
};

static std::atomic<int> g_exchange_mode{0};  // 0 by size, 1 one-shot, 2 two-phase

extern "C" void m3r_debug_set_exchange_mode(int mode) { g_exchange_mode.store(mode, std::memory_order_relaxed); }

static size_t comm_area_bytes(long long cap) {
  return static_cast<size_t>(m3r::kCommFlagBytes) + 4 * static_cast<size_t>(cap) * sizeof(double);
}

static void comm_fill_table(m3r_gn_comm *c) {
  for (int p = 0; p < c->world; p++) {
    unsigned char *base = static_cast<unsigned char *>(c->peer_area[p]);
    c->table.flags[p] = reinterpret_cast<int *>(base);
    c->table.buf[0][p] = reinterpret_cast<const double *>(base + m3r::kCommFlagBytes);
    c->table.buf[1][p] = c->table.buf[0][p] + c->cap;
    c->table.red[0][p] = reinterpret_cast<double *>(base + m3r::kCommFlagBytes) + 2 * c->cap;
    c->table.red[1][p] = c->table.red[0][p] + c->cap;
  }
  c->connected = true;
}

extern "C" int m3r_gn_comm_create(m3r_gn_comm **out, int rank, int world_size,
                                  int64_t max_system_doubles) {
  M3R_REQUIRE(out != nullptr, "gn_comm_create: null out pointer");
  M3R_REQUIRE(world_size >= 1 && world_size <= m3r::kCommMaxWorld && rank >= 0 && rank < world_size,
              "gn_comm_create: bad rank %d / world size %d (max %d)", rank, world_size,
              m3r::kCommMaxWorld);
  M3R_REQUIRE(max_system_doubles > 0, "gn_comm_create: empty system");
  m3r_gn_comm *c = new m3r_gn_comm();
  c->rank = rank, c->world = world_size;
  c->cap = (max_system_doubles + 1) & ~1LL;
  cudaError_t e = cudaGetDevice(&c->device);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&c->area), comm_area_bytes(c->cap));
  if (e == cudaSuccess) e = cudaMemset(c->area, 0, comm_area_bytes(c->cap));
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    delete c;
    return m3r::cuda_fail(e, "gn_comm_create");
  }
  c->peer_area[rank] = c->area;
  if (world_size == 1) comm_fill_table(c);
  *out = c;
  return M3R_OK;
}

extern "C" void m3r_gn_comm_destroy(m3r_gn_comm *c) {
  if (!c) return;
  for (int p = 0; p < c->world; p++)
    if (c->opened[p]) cudaIpcCloseMemHandle(c->peer_area[p]);
  cudaFree(c->area);
  delete c;
}

extern "C" int m3r_gn_comm_handle(m3r_gn_comm *c, void *handle_out) {
  M3R_REQUIRE(c && handle_out, "gn_comm_handle: null pointer");
  static_assert(sizeof(cudaIpcMemHandle_t) == M3R_IPC_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t h;
  M3R_CUDA_TRY(cudaIpcGetMemHandle(&h, c->area));
  memcpy(handle_out, &h, sizeof(h));
  return M3R_OK;
}

extern "C" int m3r_gn_comm_connect(m3r_gn_comm *c, const void *handles) {
  M3R_REQUIRE(c && handles, "gn_comm_connect: null pointer");
  for (int p = 0; p < c->world; p++) {
    if (p == c->rank || c->opened[p]) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const unsigned char *>(handles) + static_cast<size_t>(p) * sizeof(h), sizeof(h));
    void *ptr = nullptr;
    M3R_CUDA_TRY(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    c->peer_area[p] = ptr;
    c->opened[p] = true;
  }
  comm_fill_table(c);
  return M3R_OK;
}

extern "C" int m3r_gn_comm_connect_local(m3r_gn_comm **comms, int world_size) {
  M3R_REQUIRE(comms && world_size >= 1 && world_size <= m3r::kCommMaxWorld, "gn_comm_connect_local: bad arguments");
  int cur = 0;
  M3R_CUDA_TRY(cudaGetDevice(&cur));
  for (int a = 0; a < world_size; a++) {
    M3R_REQUIRE(comms[a] && comms[a]->world == world_size && comms[a]->rank == a,
                "gn_comm_connect_local: communicator %d does not match", a);
    M3R_CUDA_TRY(cudaSetDevice(comms[a]->device));
    for (int b = 0; b < world_size; b++) {
      if (comms[b]->device == comms[a]->device) continue;
      cudaError_t e = cudaDeviceEnablePeerAccess(comms[b]->device, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
      else if (e != cudaSuccess) {
        cudaSetDevice(cur);
        return m3r::cuda_fail(e, "cudaDeviceEnablePeerAccess");
      }
    }
  }
  M3R_CUDA_TRY(cudaSetDevice(cur));
  for (int a = 0; a < world_size; a++) {
    for (int b = 0; b < world_size; b++) comms[a]->peer_area[b] = comms[b]->area;
    comm_fill_table(comms[a]);
  }
  return M3R_OK;
}

extern "C" int m3r_gn_comm_next_round(m3r_gn_comm *c, double **system_out) {
  M3R_REQUIRE(c && system_out, "gn_comm_next_round: null pointer");
  M3R_REQUIRE(c->connected, "gn_comm_next_round: communicator is not connected");
  c->round += 1;
  *system_out = const_cast<double *>(c->table.buf[c->round & 1][c->rank]);
  return M3R_OK;
}

extern "C" int m3r_gn_comm_signal(m3r_gn_comm *c, m3r_gn_workspace *ws, void *stream) {
  M3R_REQUIRE(c && ws, "gn_comm_signal: null pointer");
  M3R_REQUIRE(c->connected && c->round > 0, "gn_comm_signal: no round in progress");
  m3r::comm_signal_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(c->table, c->rank, c->world,
                                                                         c->round, 0);
  M3R_LAUNCH_CHECK("comm_signal_kernel");
  return M3R_OK;
}

// true when m3r_gn_comm_reduce would take the reduce-scatter + all-gather form for this system
static bool comm_two_phase(const m3r_gn_comm *c, int num_poses) {
  const int forced = g_exchange_mode.load(std::memory_order_relaxed);  // m3r_debug_set_exchange_mode
  return c->world > 1 &&
         (forced == 2 || (forced == 0 && c->world >= 4 && m3r_gn_system_size(num_poses) >= (1 << 20)));
}

// phase: 0 = the whole reduce (the public entry point); 1 / 2 = its first / second half, for a host
// that drives several ranks itself and orders the halves with events (gn_multi.cuh): in the
// two-phase form 1 = reduce-scatter + signal, 2 = all-gather; in the one-shot form 1 = nothing,
// 2 = the reduce.
static int comm_reduce_phased(m3r_gn_comm *c, m3r_gn_workspace *ws, double *out, int num_poses,
                              void *stream, int phase) {
  M3R_REQUIRE(c && ws && out, "gn_comm_reduce: null pointer");
  M3R_REQUIRE(c->connected && c->round > 0, "gn_comm_reduce: no round in progress");
  M3R_REQUIRE(num_poses >= 2 && m3r_gn_system_size(num_poses) <= c->cap,
              "gn_comm_reduce: system of %d poses exceeds the exchange area", num_poses);
  const int D = 7 * (num_poses - 1);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (comm_two_phase(c, num_poses)) {
    // rows split so that every rank reduces about the same number of elements
    m3r::CommSlices S{};
    long long total = D;  // the right-hand-side row
    for (int r = 0; r < D; r++) total += 7 * (r / 7 + 1);
    long long acc = 0;
    int p = 1;
    S.begin[0] = 0;
    for (int r = 0; r <= D && p < c->world; r++) {
      acc += (r < D) ? 7 * (r / 7 + 1) : D;
      if (acc * c->world >= total * p) S.begin[p++] = r + 1;
    }
    for (; p <= c->world; p++) S.begin[p] = D + 1;
    S.begin[c->world] = D + 1;
    const int mine = S.begin[c->rank + 1] - S.begin[c->rank];
    const int rs_blocks = mine < m3r::kNumSMs * 4 ? (mine > 0 ? mine : 1) : m3r::kNumSMs * 4;
    if (phase != 2) {
      m3r::comm_reduce_scatter_kernel<<<rs_blocks, 256, 0, s>>>(c->table, S, c->rank, c->world, c->round, D,
                                                               ws->flags);
      M3R_LAUNCH_CHECK("comm_reduce_scatter_kernel");
      m3r::comm_signal_kernel<<<1, 32, 0, s>>>(c->table, c->rank, c->world, c->round, 1);
      M3R_LAUNCH_CHECK("comm_signal_kernel");
    }
    if (phase != 1) {
      const int ag_blocks = D + 1 < m3r::kNumSMs * 8 ? D + 1 : m3r::kNumSMs * 8;
      m3r::comm_gather_kernel<<<ag_blocks, 256, 0, s>>>(c->table, S, c->rank, c->world, c->round, out, D,
                                                       ws->flags);
      M3R_LAUNCH_CHECK("comm_gather_kernel");
    }
    return M3R_OK;
  }
  if (phase == 1) return M3R_OK;
  const int blocks = D + 1 < m3r::kNumSMs * 8 ? D + 1 : m3r::kNumSMs * 8;
  m3r::comm_reduce_kernel<<<blocks, 256, 0, s>>>(c->table, c->rank, c->world, c->round, out, D, ws->flags);
  M3R_LAUNCH_CHECK("comm_reduce_kernel");
  return M3R_OK;
}

extern "C" int m3r_gn_comm_reduce(m3r_gn_comm *c, m3r_gn_workspace *ws, double *out, int num_poses,
                                  void *stream) {
  return comm_reduce_phased(c, ws, out, num_poses, stream, 0);
}

extern "C" int m3r_gn_comm_reduce_phase(m3r_gn_comm *c, m3r_gn_workspace *ws, double *out, int num_poses,
                                        int phase, void *stream) {
  M3R_REQUIRE(phase == 1 || phase == 2, "gn_comm_reduce_phase: phase must be 1 or 2 (got %d)", phase);
  return comm_reduce_phased(c, ws, out, num_poses, stream, phase);
}
