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

#include "../api/error.h"

namespace pb {

DeviceGroup::DeviceGroup(int M, int N, int Px, int Py, const std::vector<int>& devices) : M_(M), N_(N) {
  const int n = Px * Py;
  if ((int)devices.size() != n) throw std::runtime_error("DeviceGroup: need one device per rank");
  grid_ = make_grid(M, N);
  // peer access between every pair of distinct devices
  for (int a : devices)
    for (int b : devices) {
      if (a == b) continue;
      int can = 0;
      PB_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, a, b));
      if (!can) throw std::runtime_error("DeviceGroup: devices " + std::to_string(a) + " and " + std::to_string(b) +
                                         " have no peer access");
      PB_CUDA_CHECK(cudaSetDevice(a));
      cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled)
        cudaGetLastError();
      else
        PB_CUDA_CHECK(e);
    }
  for (int r = 0; r < n; ++r) {
    solvers_.emplace_back(new Solver(M, N, Px, Py, r, 1, devices[r]));
    cudaStream_t s;
    PB_CUDA_CHECK(cudaSetDevice(devices[r]));
    PB_CUDA_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    streams_.push_back(s);
  }
  for (int r = 0; r < n; ++r)
    for (int q = 0; q < n; ++q)
      if (q != r) solvers_[r]->set_peer_base(q, solvers_[q]->arena(0));
  for (auto& s : solvers_) s->finalize();
}

DeviceGroup::~DeviceGroup() {
  for (size_t r = 0; r < solvers_.size(); ++r) {
    cudaSetDevice(solvers_[r]->device());
    cudaDeviceSynchronize();
    cudaStreamDestroy(streams_[r]);
  }
}

void DeviceGroup::solve(const SolveOptions& opt, pb_stats* out) {
  if (opt.method != kMethodPersistent && opt.method != kMethodPersistentLdg && solvers_.size() > 1)
    throw std::runtime_error("DeviceGroup: only the persistent method runs on several devices");
  for (size_t r = 0; r < solvers_.size(); ++r) solvers_[r]->enqueue(opt, streams_[r]);
  std::vector<pb_stats> st(solvers_.size());
  for (size_t r = 0; r < solvers_.size(); ++r) solvers_[r]->collect(streams_[r], &st[r]);
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
