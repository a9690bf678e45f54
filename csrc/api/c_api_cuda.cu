// c_api_cuda.cu -- C ABI of the GPU solver (see include/poisson_b200.h).
#include <exception>
#include <memory>
#include <vector>

#include "../../include/poisson_b200.h"
#include "../cuda/launch.h"
#include "../runtime/group.h"
#include "../runtime/solver.h"
#include "error.h"

// A pb_solver either owns its Solver or borrows one that lives inside a pb_group.
struct pb_solver {
  std::unique_ptr<pb::Solver> owned;
  pb::Solver& impl;
  pb_solver(int M, int N, int Px, int Py, int first, int nlocal, int dev)
      : owned(new pb::Solver(M, N, Px, Py, first, nlocal, dev)), impl(*owned) {}
  explicit pb_solver(pb::Solver& borrowed) : impl(borrowed) {}
};

struct pb_group {
  pb::DeviceGroup impl;
  std::vector<std::unique_ptr<pb_solver>> views;
  pb_group(int M, int N, int Px, int Py, const std::vector<int>& devs) : impl(M, N, Px, Py, devs) {
    for (int r = 0; r < impl.size(); ++r) views.emplace_back(new pb_solver(impl.solver(r)));
  }
};

namespace {
template <class F>
int guarded(F&& f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    pb::set_last_error(e.what());
    return -1;
  }
}
inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
}  // namespace

extern "C" {

int pb_cuda_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

pb_solver* pb_solver_create(int M, int N, int Px, int Py, int first_rank, int nlocal, int device) {
  try {
    return new pb_solver(M, N, Px, Py, first_rank, nlocal, device);
  } catch (const std::exception& e) {
    pb::set_last_error(e.what());
    return nullptr;
  }
}

void pb_solver_destroy(pb_solver* s) { delete s; }

int pb_solver_ipc_handle(pb_solver* s, int l, void* out64) {
  return guarded([&] { s->impl.ipc_handle(l, out64); });
}
int pb_solver_open_peer(pb_solver* s, int rank, const void* h) {
  return guarded([&] { s->impl.open_peer(rank, h); });
}
int pb_solver_set_peer_base(pb_solver* s, int rank, void* base) {
  return guarded([&] { s->impl.set_peer_base(rank, base); });
}
void* pb_solver_arena(pb_solver* s, int l) { return s->impl.arena(l); }
int pb_solver_finalize(pb_solver* s) {
  return guarded([&] { s->impl.finalize(); });
}

int pb_solver_solve(pb_solver* s, const pb_solve_options* o, void* stream, pb_stats* stats) {
  return guarded([&] {
    pb::SolveOptions opt;
    opt.method = o->method;
    opt.norm = o->norm;
    opt.delta = o->delta;
    opt.max_iter = o->max_iter;
    opt.exact_div = o->exact_div;
    opt.ctas_per_sm = o->ctas_per_sm;
    opt.error_report = o->error_report;
    opt.tile_rows = o->tile_rows;
    opt.l2_hints = o->l2_hints;
    opt.compressed = o->full_coefficients ? 0 : 1;
    opt.ap_mode = o->ap_mode;
    opt.no_balance = o->no_balance;
    opt.ab_mode = o->ab_mode;
    opt.watchdog_s = o->watchdog_s;
    s->impl.solve(opt, as_stream(stream), stats);
  });
}

int pb_solver_block(pb_solver* s, int l, pb_block* out) {
  return guarded([&] {
    const pb::LocalBlock& b = s->impl.block(l);
    out->rank = b.rank, out->px = b.px, out->py = b.py, out->i0 = b.i0, out->j0 = b.j0;
    out->nx = b.nx, out->ny = b.ny, out->pitch = b.pitch;
    for (int d = 0; d < 4; ++d) out->nbr[d] = b.nbr[d];
  });
}

void* pb_solver_array(pb_solver* s, int l, int which) {
  try {
    return s->impl.array(l, which);
  } catch (const std::exception& e) {
    pb::set_last_error(e.what());
    return nullptr;
  }
}
void* pb_solver_control_red(pb_solver* s, int l) { return s->impl.control(l)->red; }
void* pb_solver_control_err(pb_solver* s, int l) { return s->impl.control(l)->err; }

int pb_solver_copy_solution(pb_solver* s, double* host_full, void* stream) {
  return guarded([&] { s->impl.copy_solution_to_host(host_full, as_stream(stream)); });
}
int pb_solver_gather_block(pb_solver* s, int l, void* dev_out, long long ld, void* stream) {
  return guarded([&] { s->impl.gather_block(l, static_cast<double*>(dev_out), ld, as_stream(stream)); });
}
int pb_solver_info(pb_solver* s, int* ctas, long long* launches, long long* h2d, long long* d2h) {
  if (ctas) *ctas = s->impl.last_ctas_per_rank();
  if (launches) *launches = s->impl.last_launches();
  if (h2d) *h2d = (long long)s->impl.h2d_bytes_per_solve();
  if (d2h) *d2h = (long long)s->impl.d2h_bytes_per_solve();
  return 0;
}

int pb_solver_cta_times(pb_solver* s, int local_idx, double* host_out, int max_ctas) {
  int n = 0;
  const int rc = guarded([&] { n = s->impl.cta_times(local_idx, host_out, max_ctas); });
  return rc != 0 ? rc : n;
}
int pb_solver_cta_report(pb_solver* s, int local_idx, double* host_out, int max_ctas) {
  int n = 0;
  const int rc = guarded([&] { n = s->impl.cta_report(local_idx, host_out, max_ctas); });
  return rc != 0 ? rc : n;
}
const char* pb_solver_last_kernel(pb_solver* s) { return s->impl.last_kernel(); }
int pb_solver_tile_counts(pb_solver* s, long long* out3) {
  return guarded([&] { s->impl.tile_counts(out3); });
}
int pb_solver_ap_recompute(pb_solver* s, int ap_mode) { return s->impl.ap_recompute(ap_mode) ? 1 : 0; }

// ---- single-process multi-GPU group ----
int pb_runtime_bootstrap(const int* devices, int n) {
  return guarded([&] { pb::bootstrap_devices(std::vector<int>(devices, devices + n)); });
}
int pb_p2p_matrix(const int* devices, int n, int* perf, int* access, int* atomics) {
  return guarded([&] { pb::p2p_matrix(std::vector<int>(devices, devices + n), perf, access, atomics); });
}

pb_group* pb_group_create(int M, int N, int Px, int Py, const int* devices, int ndevices) {
  try {
    return new pb_group(M, N, Px, Py, std::vector<int>(devices, devices + ndevices));
  } catch (const std::exception& e) {
    pb::set_last_error(e.what());
    return nullptr;
  }
}
void pb_group_destroy(pb_group* g) { delete g; }
int pb_group_solve(pb_group* g, const pb_solve_options* o, pb_stats* stats) {
  return guarded([&] {
    pb::SolveOptions opt;
    opt.method = o->method;
    opt.norm = o->norm;
    opt.delta = o->delta;
    opt.max_iter = o->max_iter;
    opt.exact_div = o->exact_div;
    opt.ctas_per_sm = o->ctas_per_sm;
    opt.error_report = o->error_report;
    opt.tile_rows = o->tile_rows;
    opt.l2_hints = o->l2_hints;
    opt.compressed = o->full_coefficients ? 0 : 1;
    opt.ap_mode = o->ap_mode;
    opt.no_balance = o->no_balance;
    opt.ab_mode = o->ab_mode;
    opt.watchdog_s = o->watchdog_s;
    g->impl.solve(opt, stats);
  });
}
int pb_group_copy_solution(pb_group* g, double* host_full) {
  return guarded([&] { g->impl.copy_solution_to_host(host_full); });
}
pb_solver* pb_group_solver(pb_group* g, int rank) {
  return rank >= 0 && rank < (int)g->views.size() ? g->views[rank].get() : nullptr;
}

// ---- kernel-level step API ----
#define PB_STEP(expr) return guarded([&] { PB_CUDA_CHECK(cudaSetDevice(s->impl.device())); PB_CUDA_CHECK(expr); })

int pb_step_set_options(pb_solver* s, int norm, double delta, long long max_iter, void* stream) {
  return guarded([&] {
    PB_CUDA_CHECK(cudaSetDevice(s->impl.device()));
    const pb::Grid g = s->impl.host_ctx(0).g;
    pb::DeviceOptions o;
    o.delta = delta;
    o.max_iter = max_iter >= 0 ? max_iter : (long long)(g.M - 1) * (long long)(g.N - 1);
    o.norm = norm;
    o.pad = 0;
    o.watchdog_ns = 0;
    for (int l = 0; l < s->impl.nlocal(); ++l)
      PB_CUDA_CHECK(cudaMemcpyAsync(&s->impl.control(l)->delta, &o, sizeof(o), cudaMemcpyHostToDevice, as_stream(stream)));
    PB_CUDA_CHECK(cudaStreamSynchronize(as_stream(stream)));
  });
}
int pb_step_assemble(pb_solver* s, int l, void* st) {
  PB_STEP(pb::launch_assemble(s->impl.device_ctx(l), s->impl.host_ctx(l), as_stream(st)));
}
int pb_step_begin(pb_solver* s, int l, void* st) { PB_STEP(pb::launch_step_begin(s->impl.device_ctx(l), as_stream(st))); }
int pb_step_dot_zr(pb_solver* s, int l, void* st) {
  PB_STEP(pb::launch_step_dot_zr(s->impl.device_ctx(l), s->impl.host_ctx(l), as_stream(st)));
}
int pb_step_stencil(pb_solver* s, int l, int fuse_p, int div, void* st) {
  PB_STEP(pb::launch_step_stencil(s->impl.device_ctx(l), s->impl.host_ctx(l), fuse_p != 0, div, as_stream(st)));
}
int pb_step_update(pb_solver* s, int l, int fuse_p, void* st) {
  PB_STEP(pb::launch_step_update(s->impl.device_ctx(l), s->impl.host_ctx(l), fuse_p != 0, as_stream(st)));
}
int pb_step_update_p(pb_solver* s, int l, void* st) {
  PB_STEP(pb::launch_step_update_p(s->impl.device_ctx(l), s->impl.host_ctx(l), as_stream(st)));
}
int pb_step_reduce_local(pb_solver* s, int l, int nv, void* st) {
  PB_STEP(pb::launch_step_reduce_local(s->impl.device_ctx(l), s->impl.host_ctx(l), nv, as_stream(st)));
}
int pb_step_scalars(pb_solver* s, int l, int phase, void* st) {
  PB_STEP(pb::launch_step_scalars(s->impl.device_ctx(l), phase, as_stream(st)));
}
int pb_step_pack_edges(pb_solver* s, int l, void* buf, int stride, void* st) {
  PB_STEP(pb::launch_pack_edges(s->impl.device_ctx(l), s->impl.host_ctx(l), static_cast<double*>(buf), stride, as_stream(st)));
}
int pb_step_unpack_edges(pb_solver* s, int l, const void* buf, int stride, void* st) {
  PB_STEP(pb::launch_unpack_edges(s->impl.device_ctx(l), s->impl.host_ctx(l), static_cast<const double*>(buf), stride,
                                  as_stream(st)));
}
int pb_step_error_report(pb_solver* s, int l, void* st) {
  PB_STEP(pb::launch_error_report(s->impl.device_ctx(l), s->impl.host_ctx(l), as_stream(st)));
}
int pb_solver_xgpu_bench(pb_solver* s, int iters, int mode, int edge, int ctas, void* st, double* out2) {
  return guarded([&] { s->impl.xgpu_bench(iters, mode, edge, ctas, as_stream(st), out2); });
}
int pb_debug_div_check(double h, unsigned long long n, unsigned long long seed, void* counts, void* st) {
  return guarded([&] { PB_CUDA_CHECK(pb::launch_div_check(h, n, seed, static_cast<unsigned long long*>(counts), as_stream(st))); });
}
int pb_debug_tma_probe(pb_solver* s, int l, int idx, int mode, int x, int y, unsigned int nbytes, void* out, void* status,
                       void* st) {
  PB_STEP(pb::launch_tma_probe(s->impl.host_ctx(l).tmaps, s->impl.host_maps(l), idx, mode, x, y, nbytes, static_cast<double*>(out),
                               static_cast<int*>(status), as_stream(st)));
}
int pb_step_read_control(pb_solver* s, int l, int* iter, int* status, double* diff_norm, void* st) {
  return guarded([&] {
    PB_CUDA_CHECK(cudaSetDevice(s->impl.device()));
    pb::Control c;
    PB_CUDA_CHECK(cudaMemcpyAsync(&c, s->impl.control(l), sizeof(c), cudaMemcpyDeviceToHost, as_stream(st)));
    PB_CUDA_CHECK(cudaStreamSynchronize(as_stream(st)));
    if (iter) *iter = c.iter;
    if (status) *status = c.status;
    if (diff_norm) *diff_norm = c.diff_norm;
  });
}

}  // extern "C"
