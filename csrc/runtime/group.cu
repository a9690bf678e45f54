// group.cu -- single-process multi-GPU: one Solver per device, all devices peer-mapped, one persistent
// kernel per device launched back to back.  This is what the CLI and `solve(..., gpus=G)` use when the
// program is not running under torchrun; the multi-process path (parallel/fabric.py) wires the same
// Solver objects through CUDA IPC instead.
//
// Replaces the reference's process model (MPI_Init ... one rank per GPU selected by the launcher's
// CUDA_VISIBLE_DEVICES) for the one-node case: device enumeration, peer access and
// the Px x Py assignment happen here.
#include "group.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <exception>
#include <thread>

#include "../api/error.h"

namespace pb {

// Runs fn(i) for i in [0, n) on n host threads and rethrows the first exception.
template <class F>
static void parallel_for_devices(int n, F&& fn) {
  std::vector<std::thread> threads;
  std::vector<std::exception_ptr> errors(n);
  for (int i = 0; i < n; ++i)
    threads.emplace_back([&, i] {
      try {
        fn(i);
      } catch (...) {
        errors[i] = std::current_exception();
      }
    });
  for (auto& t : threads) t.join();
  for (auto& e : errors)
    if (e) std::rethrow_exception(e);
}

void bootstrap_devices(const std::vector<int>& devices) {
  const int n = (int)devices.size();
  // 1. primary contexts, all at once
  parallel_for_devices(n, [&](int i) {
    PB_CUDA_CHECK(cudaSetDevice(devices[i]));
    PB_CUDA_CHECK(cudaFree(nullptr));
  });
  // 2. peer access from every device to every other one
  parallel_for_devices(n, [&](int i) {
    const int a = devices[i];
    PB_CUDA_CHECK(cudaSetDevice(a));
    for (int b : devices) {
      if (a == b) continue;
      int can = 0;
      PB_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, a, b));
      if (!can)
        throw std::runtime_error("devices " + std::to_string(a) + " and " + std::to_string(b) + " have no peer access");
      const cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled)
        cudaGetLastError();
      else
        PB_CUDA_CHECK(e);
    }
  });
}

void p2p_matrix(const std::vector<int>& devices, int* perf, int* access, int* atomics) {
  const int n = (int)devices.size();
  for (int a = 0; a < n; ++a)
    for (int b = 0; b < n; ++b) {
      int pr = 0, ac = 1, at = 1;
      if (a != b) {
        PB_CUDA_CHECK(cudaDeviceGetP2PAttribute(&pr, cudaDevP2PAttrPerformanceRank, devices[a], devices[b]));
        PB_CUDA_CHECK(cudaDeviceGetP2PAttribute(&ac, cudaDevP2PAttrAccessSupported, devices[a], devices[b]));
        PB_CUDA_CHECK(cudaDeviceGetP2PAttribute(&at, cudaDevP2PAttrNativeAtomicSupported, devices[a], devices[b]));
      }
      if (perf) perf[a * n + b] = pr;
      if (access) access[a * n + b] = ac;
      if (atomics) atomics[a * n + b] = at;
    }
}

// PB_TRACE_INIT=1: where a cold start spends its time, on stderr
static double trace_now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
static void trace_init(const char* what, double& t_prev) {
  static const bool on = std::getenv("PB_TRACE_INIT") != nullptr;
  const double t = trace_now();
  if (on) std::fprintf(stderr, "[pb init] %-28s %.3f s\n", what, t - t_prev);
  t_prev = t;
}

DeviceGroup::DeviceGroup(int M, int N, int Px, int Py, const std::vector<int>& devices) : M_(M), N_(N) {
  const int n = Px * Py;
  if ((int)devices.size() != n) throw std::runtime_error("DeviceGroup: need one device per rank");
  grid_ = make_grid(M, N);
  double t_prev = trace_now();
  bootstrap_devices(devices);
  trace_init("bootstrap (again)", t_prev);
  // one Solver per device (arena allocation and zeroing), all devices at once
  solvers_.resize(n);
  streams_.resize(n);
  parallel_for_devices(n, [&](int r) {
    solvers_[r].reset(new Solver(M, N, Px, Py, r, 1, devices[r]));
    PB_CUDA_CHECK(cudaSetDevice(devices[r]));
    PB_CUDA_CHECK(cudaStreamCreateWithFlags(&streams_[r], cudaStreamNonBlocking));
  });
  trace_init("solvers (arenas)", t_prev);
  for (int r = 0; r < n; ++r)
    for (int q = 0; q < n; ++q)
      if (q != r) solvers_[r]->set_peer_base(q, solvers_[q]->arena(0));
  parallel_for_devices(n, [&](int r) { solvers_[r]->finalize(); });
  trace_init("finalize (maps, contexts)", t_prev);
}

DeviceGroup::~DeviceGroup() {
  for (size_t r = 0; r < solvers_.size(); ++r) {
    cudaSetDevice(solvers_[r]->device());
    cudaDeviceSynchronize();
    cudaStreamDestroy(streams_[r]);
  }
}

void DeviceGroup::solve(const SolveOptions& opt, pb_stats* out) {
  if (opt.method != kMethodPersistent && opt.method != kMethodPersistentLdg && opt.method != kMethodResident &&
      opt.method != kMethodPersistentTma && solvers_.size() > 1)
    throw std::runtime_error("DeviceGroup: only the persistent method runs on several devices");
  double t_prev = trace_now();
  for (size_t r = 0; r < solvers_.size(); ++r) solvers_[r]->enqueue(opt, streams_[r]);
  trace_init("enqueue (first: module load)", t_prev);
  std::vector<pb_stats> st(solvers_.size());
  for (size_t r = 0; r < solvers_.size(); ++r) solvers_[r]->collect(streams_[r], &st[r]);
  trace_init("collect (device time)", t_prev);
  if (!out) return;
  *out = st[0];
  for (size_t r = 1; r < st.size(); ++r) {
    const pb_stats& s = st[r];
    if (s.status == PB_STATUS_WATCHDOG) out->status = s.status;
    out->sum_e2_in_D += s.sum_e2_in_D;
    out->sum_e2_all += s.sum_e2_all;
    out->nodes_in_D += s.nodes_in_D;
    out->max_abs = std::max(out->max_abs, s.max_abs);
    out->max_w = std::max(out->max_w, s.max_w);
    // timers: max over ranks (S4:962-966)
    out->seconds = std::max(out->seconds, s.seconds);
    out->t_gpu = std::max(out->t_gpu, s.t_gpu);
    out->t_copy = std::max(out->t_copy, s.t_copy);
    out->t_comm = std::max(out->t_comm, s.t_comm);
    out->t_precond = std::max(out->t_precond, s.t_precond);
    out->t_dot = std::max(out->t_dot, s.t_dot);
    out->t_assemble = std::max(out->t_assemble, s.t_assemble);
    out->t_iterations = std::max(out->t_iterations, s.t_iterations);
    out->t_phase_ab = std::max(out->t_phase_ab, s.t_phase_ab), out->t_phase_c = std::max(out->t_phase_c, s.t_phase_c);
    out->t_red1 = std::max(out->t_red1, s.t_red1), out->t_red2 = std::max(out->t_red2, s.t_red2);
  }
  out->l2_in_D = std::sqrt(grid_.h1h2 * out->sum_e2_in_D);
  out->l2_all = std::sqrt(grid_.h1h2 * out->sum_e2_all);
}

void DeviceGroup::copy_solution_to_host(double* full) {
  std::memset(full, 0, sizeof(double) * (size_t)(M_ + 1) * (size_t)(N_ + 1));
  for (size_t r = 0; r < solvers_.size(); ++r) solvers_[r]->copy_solution_to_host(full, streams_[r]);
}

}  // namespace pb
