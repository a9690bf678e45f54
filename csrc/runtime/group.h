// group.h -- single-process multi-GPU group of Solvers (see group.cu).
#pragma once
#include <memory>
#include <vector>

#include "solver.h"

namespace pb {

// Contexts and peer access for `devices`, one host thread per device (context creation and the n - 1
// cudaDeviceEnablePeerAccess calls of a device are what a cold start spends its time on; serially they cost
// 2.3 s + 5.3 s at 8 GPUs).  Idempotent.
void bootstrap_devices(const std::vector<int>& devices);
// cudaDevP2PAttr{PerformanceRank, AccessSupported, NativeAtomicSupported} for every ordered pair, row-major n x n
void p2p_matrix(const std::vector<int>& devices, int* perf, int* access, int* atomics);

class DeviceGroup {
 public:
  DeviceGroup(int M, int N, int Px, int Py, const std::vector<int>& devices);
  ~DeviceGroup();
  int size() const { return (int)solvers_.size(); }
  Solver& solver(int r) { return *solvers_[r]; }
  void solve(const SolveOptions& opt, pb_stats* out);
  void copy_solution_to_host(double* full);  // (M+1) x (N+1), boundary zero

 private:
  int M_, N_;
  Grid grid_;
  std::vector<std::unique_ptr<Solver>> solvers_;
  std::vector<cudaStream_t> streams_;
};

}  // namespace pb
