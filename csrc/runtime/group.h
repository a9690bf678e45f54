// group.h -- single-process multi-GPU group of Solvers (see group.cu).
#pragma once
#include <memory>
#include <vector>

#include "solver.h"

namespace pb {

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
