// Multi-GPU Gauss-Newton behind ONE host call (include/m3r_backends.h, "Single-process
// multi-GPU").  Included by gn.cu after gn_comm.cuh.
//
// The reference's caller (global_opt.py:142-143, one backend process, main.py:268) hands over
// tensors that live on one device.  This driver lets that unchanged call use G devices:
//   * device 0 of the list (the "primary") is the one the caller's tensors live on; the other
//     devices read them in place over NVLink peer access -- there is no staging copy.  Each
//     device's per-call preparation (m3r_gn_begin) IS the scatter: its pack kernel streams
//     Xs / Cs from the primary once (coalesced) into its local 16-byte records and a local copy
//     of Cs, its compaction kernel streams its shard's rows of idx / valid / Q from the primary
//     and gathers confidences from the LOCAL copy, and from then on every iteration runs on
//     local memory only.  Bytes that cross NVLink per call and per non-primary device:
//     16 N n (pointmaps) + 13 n E / G (its edge rows) + 32 N (poses) + 16 E (ii, jj).
//   * per iteration: linearise the shard into the exchange buffer, signal, reduce over peer
//     memory in rank order (gn_comm.cuh), replicated solve + retraction.  Device 0 works on the
//     caller's Twc itself; the others on replicas that stay bitwise identical to it.
//   * ordering: the secondary streams wait for an event on the caller's stream before they
//     touch the caller's tensors, and the caller's stream waits for their completion events
//     before the call's work counts as done, so the caller may free or overwrite its tensors in
//     stream order as usual.  The host never blocks, and must not: every buffer of every rank is
//     reserved by the m3r_gn_begin calls before the first exchange kernel is enqueued (a
//     device-synchronising cudaFree / cudaMalloc issued while one rank's reduce kernel spins on
//     the flag of a rank whose signal is not enqueued yet would deadlock until the exchange's
//     4 s timeout; gn.cu, m3r_gn_begin).
// A device may appear more than once in the list (loop-back ranks on one GPU: every rank then
// has its own workspace, exchange area and stream); that is how the 1-GPU tests run this path.
#pragma once

struct m3r_gn_multi {
  int G = 0;
  int dev[m3r::kCommMaxWorld] = {0};
  m3r_gn_workspace *ws[m3r::kCommMaxWorld] = {nullptr};
  m3r_gn_comm *comm[m3r::kCommMaxWorld] = {nullptr};
  cudaStream_t st[m3r::kCommMaxWorld] = {nullptr};   // [0]: the caller's stream of the current call
  float *twc[m3r::kCommMaxWorld] = {nullptr};        // pose replicas of ranks 1..G-1
  float *dxr[m3r::kCommMaxWorld] = {nullptr};        // their dx
  float *Kr[m3r::kCommMaxWorld] = {nullptr};         // their copy of the intrinsics
  double *sys[m3r::kCommMaxWorld] = {nullptr};       // reduced system of every rank
  int pose_cap = 0;                                  // poses the buffers above are sized for
  cudaEvent_t ev_start = nullptr;
  cudaEvent_t ev_done[m3r::kCommMaxWorld] = {nullptr};
  cudaEvent_t ev_copied[m3r::kCommMaxWorld] = {nullptr};  // rank r has taken its copy of the caller's Twc / K
  cudaEvent_t ev_sig[m3r::kCommMaxWorld] = {nullptr};     // rank r's signal of the current round is enqueued up to here
  cudaEvent_t ev_sig2[m3r::kCommMaxWorld] = {nullptr};    // ... and its reduce-scatter + second signal (two-phase)
  cudaEvent_t ev_b0[m3r::kCommMaxWorld] = {nullptr}, ev_b1[m3r::kCommMaxWorld] = {nullptr};
  bool timed = false;
};

namespace m3r {

struct DeviceScope {  // restores the caller's current device
  int saved = 0;
  DeviceScope() { cudaGetDevice(&saved); }
  ~DeviceScope() { cudaSetDevice(saved); }
};

static int multi_release_sized(m3r_gn_multi *m) {
  for (int r = 0; r < m->G; r++) {
    cudaSetDevice(m->dev[r]);
    if (m->comm[r]) m3r_gn_comm_destroy(m->comm[r]);
    m->comm[r] = nullptr;
    cudaFree(m->twc[r]), cudaFree(m->dxr[r]), cudaFree(m->sys[r]);
    m->twc[r] = nullptr, m->dxr[r] = nullptr, m->sys[r] = nullptr;
  }
  m->pose_cap = 0;
  return M3R_OK;
}

// exchange areas, pose replicas and system buffers for graphs of up to `num_poses` poses
static int multi_reserve(m3r_gn_multi *m, int num_poses) {
  if (num_poses <= m->pose_cap) return M3R_OK;
  for (int r = 0; r < m->G; r++) {  // nobody may still be reading the old areas
    M3R_CUDA_TRY(cudaSetDevice(m->dev[r]));
    M3R_CUDA_TRY(cudaDeviceSynchronize());
  }
  multi_release_sized(m);
  const int cap = num_poses + num_poses / 4 + 8;  // grow-only with head-room: keyframes arrive one by one
  const int64_t doubles = m3r_gn_system_size(cap);
  for (int r = 0; r < m->G; r++) {
    M3R_CUDA_TRY(cudaSetDevice(m->dev[r]));
    int rc = m3r_gn_comm_create(&m->comm[r], r, m->G, doubles);
    if (rc) return rc;
    M3R_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(&m->sys[r]), static_cast<size_t>(doubles) * sizeof(double)));
    if (r > 0) {
      M3R_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(&m->twc[r]), static_cast<size_t>(cap) * 8 * sizeof(float)));
      M3R_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(&m->dxr[r]), static_cast<size_t>(cap) * 7 * sizeof(float)));
    }
  }
  int rc = m3r_gn_comm_connect_local(m->comm, m->G);
  if (rc) return rc;
  m->pose_cap = cap;
  return M3R_OK;
}

}  // namespace m3r

extern "C" void m3r_gn_multi_destroy(m3r_gn_multi *m);

extern "C" int m3r_gn_multi_create(m3r_gn_multi **out, const int *devices, int num_devices) {
  M3R_REQUIRE(out && devices, "gn_multi_create: null pointer");
  M3R_REQUIRE(num_devices >= 1 && num_devices <= m3r::kCommMaxWorld, "gn_multi_create: %d devices (1..%d)",
              num_devices, m3r::kCommMaxWorld);
  int count = 0;
  M3R_CUDA_TRY(cudaGetDeviceCount(&count));
  for (int r = 0; r < num_devices; r++)
    M3R_REQUIRE(devices[r] >= 0 && devices[r] < count, "gn_multi_create: device %d of %d", devices[r], count);
  m3r::DeviceScope scope;
  m3r_gn_multi *m = new m3r_gn_multi();
  m->G = num_devices;
  for (int r = 0; r < num_devices; r++) m->dev[r] = devices[r];
  int rc = M3R_OK;
  for (int r = 0; r < num_devices && rc == M3R_OK; r++) {
    cudaError_t e = cudaSetDevice(m->dev[r]);
    if (e == cudaSuccess && m->dev[r] != m->dev[0]) {  // this rank reads the caller's tensors in place
      e = cudaDeviceEnablePeerAccess(m->dev[0], 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) e = cudaSuccess, cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_done[r], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_copied[r], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_sig[r], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_sig2[r], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreate(&m->ev_b0[r]);
    if (e == cudaSuccess) e = cudaEventCreate(&m->ev_b1[r]);
    if (e == cudaSuccess && r > 0) e = cudaStreamCreateWithFlags(&m->st[r], cudaStreamNonBlocking);
    if (e == cudaSuccess && r > 0) e = cudaMalloc(reinterpret_cast<void **>(&m->Kr[r]), 16 * sizeof(float));
    if (e != cudaSuccess) rc = m3r::cuda_fail(e, "gn_multi_create");
    if (rc == M3R_OK) rc = m3r_gn_workspace_create(&m->ws[r]);
    if (rc == M3R_OK) m->ws[r]->copy_cs = (r > 0);
  }
  if (rc == M3R_OK) {
    cudaSetDevice(m->dev[0]);
    cudaError_t e = cudaEventCreateWithFlags(&m->ev_start, cudaEventDisableTiming);
    if (e != cudaSuccess) rc = m3r::cuda_fail(e, "gn_multi_create");
  }
  if (rc != M3R_OK) {
    m3r_gn_multi_destroy(m);
    return rc;
  }
  *out = m;
  return M3R_OK;
}

extern "C" void m3r_gn_multi_destroy(m3r_gn_multi *m) {
  if (!m) return;
  m3r::DeviceScope scope;
  for (int r = 0; r < m->G; r++) {
    cudaSetDevice(m->dev[r]);
    cudaDeviceSynchronize();
  }
  m3r::multi_release_sized(m);
  for (int r = 0; r < m->G; r++) {
    cudaSetDevice(m->dev[r]);
    if (m->ws[r]) m3r_gn_workspace_destroy(m->ws[r]);
    if (r > 0 && m->st[r]) cudaStreamDestroy(m->st[r]);  // st[0] is the caller's
    if (m->ev_done[r]) cudaEventDestroy(m->ev_done[r]);
    if (m->ev_copied[r]) cudaEventDestroy(m->ev_copied[r]);
    if (m->ev_sig[r]) cudaEventDestroy(m->ev_sig[r]);
    if (m->ev_sig2[r]) cudaEventDestroy(m->ev_sig2[r]);
    if (m->ev_b0[r]) cudaEventDestroy(m->ev_b0[r]);
    if (m->ev_b1[r]) cudaEventDestroy(m->ev_b1[r]);
    cudaFree(m->Kr[r]);
  }
  if (m->ev_start) cudaEventDestroy(m->ev_start);
  delete m;
}

extern "C" int m3r_gn_multi_solve(m3r_gn_multi *m, const m3r_gn_problem *prob, float *dx, int max_iter,
                                  float delta_thresh, void *stream) {
  M3R_REQUIRE(m && prob && dx, "gn_multi_solve: null pointer");
  M3R_REQUIRE(max_iter >= 0, "gn_multi_solve: negative max_iter");
  m3r::DeviceScope scope;
  const int G = m->G, N = prob->num_poses, n = prob->num_points, E = prob->num_edges;
  int rc = m3r::multi_reserve(m, N);
  if (rc) return rc;
  cudaStream_t s0 = static_cast<cudaStream_t>(stream);
  m->st[0] = s0;
  M3R_CUDA_TRY(cudaSetDevice(m->dev[0]));
  M3R_CUDA_TRY(cudaEventRecord(m->ev_start, s0));  // the caller's tensors are ready from here on

  // ---- per-rank problems: shard of the directed edges, pointers into the caller's tensors
  m3r_gn_problem P[m3r::kCommMaxWorld];
  for (int r = 0; r < G; r++) {
    P[r] = *prob;
    const int base = E / G, rem = E % G;
    const int b0 = r * base + (r < rem ? r : rem);
    const int b1 = b0 + base + (r < rem ? 1 : 0);
    P[r].edge_begin = b0, P[r].edge_end = b1;
    const size_t off = static_cast<size_t>(b0) * n;
    P[r].idx_ii2jj = prob->idx_ii2jj ? prob->idx_ii2jj + off : nullptr;
    P[r].valid_match = prob->valid_match ? prob->valid_match + off : nullptr;
    P[r].Q = prob->Q ? prob->Q + off : nullptr;
    if (r > 0) {
      P[r].Twc = m->twc[r];
      if (prob->K) P[r].K = m->Kr[r];
    }
  }
  // ---- preparation = scatter (header comment): remote reads happen here and only here
  for (int r = 0; r < G; r++) {
    M3R_CUDA_TRY(cudaSetDevice(m->dev[r]));
    if (r > 0) {
      M3R_CUDA_TRY(cudaStreamWaitEvent(m->st[r], m->ev_start, 0));
      M3R_CUDA_TRY(cudaMemcpyPeerAsync(m->twc[r], m->dev[r], prob->Twc, m->dev[0],
                                       static_cast<size_t>(N) * 8 * sizeof(float), m->st[r]));
      if (prob->K)
        M3R_CUDA_TRY(cudaMemcpyPeerAsync(m->Kr[r], m->dev[r], prob->K, m->dev[0], 9 * sizeof(float), m->st[r]));
      M3R_CUDA_TRY(cudaEventRecord(m->ev_copied[r], m->st[r]));
    }
    M3R_CUDA_TRY(cudaEventRecord(m->ev_b0[r], m->st[r]));
    if ((rc = m3r_gn_begin(m->ws[r], &P[r], m->st[r]))) return rc;
    M3R_CUDA_TRY(cudaEventRecord(m->ev_b1[r], m->st[r]));
  }
  m->timed = true;
  // rank 0 retracts the caller's Twc in place: not before every replica has been copied from it
  M3R_CUDA_TRY(cudaSetDevice(m->dev[0]));
  for (int r = 1; r < G; r++) M3R_CUDA_TRY(cudaStreamWaitEvent(s0, m->ev_copied[r], 0));
  // ---- iterations.  The device-side handshake of the exchange (flags, gn_comm.cuh) runs as in
  // the one-process-per-GPU form, but here ONE host thread drives every rank, so it also orders the
  // streams with events: a rank's reduce is enqueued behind every rank's signal.  The reduce
  // kernels then find their flags set when they start and never spin -- which matters, because a
  // spinning kernel is only safe when the kernels it waits for are guaranteed to run beside it,
  // and CUDA gives no such guarantee for two streams of one device (loop-back ranks), nor to a
  // host thread that could block between the two enqueues.
  for (int it = 0; it < max_iter; it++) {
    for (int r = 0; r < G; r++) {
      M3R_CUDA_TRY(cudaSetDevice(m->dev[r]));
      double *buf = nullptr;
      if ((rc = m3r_gn_comm_next_round(m->comm[r], &buf))) return rc;
      if ((rc = m3r_gn_linearize(m->ws[r], &P[r], buf, m->st[r]))) return rc;
      if ((rc = m3r_gn_comm_signal(m->comm[r], m->ws[r], m->st[r]))) return rc;
      M3R_CUDA_TRY(cudaEventRecord(m->ev_sig[r], m->st[r]));
    }
    const bool two_phase = comm_two_phase(m->comm[0], N);
    if (two_phase) {  // reduce-scatter + second signal of every rank, behind everybody's first signal
      for (int r = 0; r < G; r++) {
        M3R_CUDA_TRY(cudaSetDevice(m->dev[r]));
        for (int q = 0; q < G; q++)
          if (q != r) M3R_CUDA_TRY(cudaStreamWaitEvent(m->st[r], m->ev_sig[q], 0));
        if ((rc = comm_reduce_phased(m->comm[r], m->ws[r], m->sys[r], N, m->st[r], 1))) return rc;
        M3R_CUDA_TRY(cudaEventRecord(m->ev_sig2[r], m->st[r]));
      }
    }
    for (int r = 0; r < G; r++) {
      M3R_CUDA_TRY(cudaSetDevice(m->dev[r]));
      for (int q = 0; q < G; q++)
        if (q != r) M3R_CUDA_TRY(cudaStreamWaitEvent(m->st[r], two_phase ? m->ev_sig2[q] : m->ev_sig[q], 0));
      if ((rc = comm_reduce_phased(m->comm[r], m->ws[r], m->sys[r], N, m->st[r], 2))) return rc;
      if ((rc = m3r_gn_solve_retract(m->ws[r], &P[r], m->sys[r], r == 0 ? dx : m->dxr[r], delta_thresh,
                                     m->st[r])))
        return rc;
    }
  }
  // ---- the call's work on the caller's stream ends when every rank has finished
  for (int r = 1; r < G; r++) {
    M3R_CUDA_TRY(cudaSetDevice(m->dev[r]));
    M3R_CUDA_TRY(cudaEventRecord(m->ev_done[r], m->st[r]));
  }
  M3R_CUDA_TRY(cudaSetDevice(m->dev[0]));
  for (int r = 1; r < G; r++) M3R_CUDA_TRY(cudaStreamWaitEvent(s0, m->ev_done[r], 0));
  return M3R_OK;
}

// After a solve: iterations executed (synchronises every rank), the time every rank's
// preparation took (begin_ms[num_devices]: remote reads + packing + compaction = the scatter
// cost of the single-process mode), whether the pose replicas are bitwise equal to the
// caller's Twc.  For tests and benchmarks.
extern "C" int m3r_gn_multi_report(m3r_gn_multi *m, const float *Twc, int num_poses, float *begin_ms,
                                   int *iters, int *replicas_equal) {
  M3R_REQUIRE(m && m->timed, "gn_multi_report: no solve has run");
  m3r::DeviceScope scope;
  int rc = M3R_OK;
  for (int r = 0; r < m->G; r++) {
    M3R_CUDA_TRY(cudaSetDevice(m->dev[r]));
    int it = 0;
    if ((rc = m3r_gn_iterations_done(m->ws[r], m->st[r], &it))) return rc;
    if (r == 0 && iters) *iters = it;
    if (begin_ms) M3R_CUDA_TRY(cudaEventElapsedTime(&begin_ms[r], m->ev_b0[r], m->ev_b1[r]));
  }
  if (replicas_equal && Twc) {
    *replicas_equal = 1;
    const size_t bytes = static_cast<size_t>(num_poses) * 8 * sizeof(float);
    std::vector<unsigned char> a(bytes), b(bytes);
    M3R_CUDA_TRY(cudaSetDevice(m->dev[0]));
    M3R_CUDA_TRY(cudaMemcpy(a.data(), Twc, bytes, cudaMemcpyDeviceToHost));
    for (int r = 1; r < m->G; r++) {
      M3R_CUDA_TRY(cudaSetDevice(m->dev[r]));
      M3R_CUDA_TRY(cudaMemcpy(b.data(), m->twc[r], bytes, cudaMemcpyDeviceToHost));
      if (memcmp(a.data(), b.data(), bytes) != 0) *replicas_equal = 0;
    }
  }
  return M3R_OK;
}
