// binding.cpp -- thin PyTorch extension over the C ABI: solve(M, N, ...) -> (report dict, w tensor).
//
// The north-star asks for "a thin PyTorch extension binding solve(M,N,...)"; the reference itself has no
// library interface (argv only, S4:995-999).  The returned tensor is a dense (M+1) x (N+1) float64 CUDA
// tensor (boundary of the rectangle = 0) filled by the native gather kernel; the solver's own padded
// arrays stay inside the native arena.
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/extension.h>

#include <map>
#include <memory>
#include <stdexcept>
#include <tuple>

#include "poisson_b200.h"

namespace {

struct Deleter {
  void operator()(pb_solver* s) const { pb_solver_destroy(s); }
};
std::map<std::tuple<int, int, int, int, int>, std::unique_ptr<pb_solver, Deleter>> g_cache;

pb_solver* get_solver(int M, int N, int Px, int Py, int device) {
  auto key = std::make_tuple(M, N, Px, Py, device);
  auto it = g_cache.find(key);
  if (it != g_cache.end()) return it->second.get();
  pb_solver* s = pb_solver_create(M, N, Px, Py, 0, Px * Py, device);
  if (!s) throw std::runtime_error(std::string("pb_solver_create: ") + pb_last_error());
  if (pb_solver_finalize(s) != 0) throw std::runtime_error(std::string("pb_solver_finalize: ") + pb_last_error());
  g_cache[key].reset(s);
  return s;
}

py::tuple solve(int M, int N, int Px, int Py, const std::string& norm, double delta, long long max_iter,
                const std::string& method, bool exact_div, bool want_solution) {
  const int device = c10::cuda::current_device();
  c10::cuda::CUDAGuard guard(device);
  auto stream = c10::cuda::getCurrentCUDAStream(device);
  pb_solver* s = get_solver(M, N, Px, Py, device);
  pb_solve_options opt{};
  opt.method = method == "persistent" ? PB_METHOD_PERSISTENT
                                      : (method == "step_fused" ? PB_METHOD_STEP_FUSED : PB_METHOD_STEP_CLASSIC);
  opt.norm = norm == "unweighted" ? PB_NORM_UNWEIGHTED : PB_NORM_WEIGHTED;
  opt.delta = delta;
  opt.max_iter = max_iter;
  opt.exact_div = exact_div ? 2 : 0;  // 2 = Markstein sequence, bit-identical to IEEE division
  opt.error_report = 1;
  pb_stats st{};
  if (pb_solver_solve(s, &opt, stream.stream(), &st) != 0)
    throw std::runtime_error(std::string("pb_solver_solve: ") + pb_last_error());
  py::dict rep;
  rep["iters"] = st.iters;
  rep["status"] = st.status;
  rep["diff_norm"] = st.diff_norm;
  rep["l2_in_D"] = st.l2_in_D;
  rep["l2_all"] = st.l2_all;
  rep["max_abs"] = st.max_abs;
  rep["max_w"] = st.max_w;
  rep["nodes_in_D"] = st.nodes_in_D;
  rep["seconds"] = st.seconds;
  rep["t_assemble"] = st.t_assemble;
  rep["t_iterations"] = st.t_iterations;
  rep["t_gpu"] = st.t_gpu;
  rep["t_comm"] = st.t_comm;
  rep["t_dot"] = st.t_dot;
  rep["t_precond"] = st.t_precond;
  rep["t_copy"] = st.t_copy;
  torch::Tensor w;
  if (want_solution) {
    w = torch::zeros({M + 1, N + 1}, torch::dtype(torch::kFloat64).device(torch::kCUDA, device));
    for (int l = 0; l < Px * Py; ++l) {
      pb_block b{};
      pb_solver_block(s, l, &b);
      double* dst = w.data_ptr<double>() + (size_t)b.i0 * (N + 1) + b.j0;
      if (pb_solver_gather_block(s, l, dst, N + 1, stream.stream()) != 0)
        throw std::runtime_error(std::string("pb_solver_gather_block: ") + pb_last_error());
    }
  }
  return py::make_tuple(rep, w);
}

void clear() { g_cache.clear(); }

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "poisson_b200 PyTorch binding";
  m.def("solve", &solve, py::arg("M"), py::arg("N"), py::arg("Px") = 1, py::arg("Py") = 1,
        py::arg("norm") = "weighted", py::arg("delta") = 1e-6, py::arg("max_iter") = -1,
        py::arg("method") = "persistent", py::arg("exact_div") = true, py::arg("want_solution") = true);
  m.def("clear", &clear);
}
