// dist_solver.cpp -- see dist_solver.h.
//
// Iteration (same recurrence as csrc/cpu/oracle.cpp and the GPU kernels; order of S2:400-457):
//   p = r/D + beta p  ->  halo(p)  ->  Ap = A p, (Ap, p) -> all-reduce -> alpha  ->  w += alpha p, r -= alpha Ap,
//   sum (dw)^2 and (r/D, r) -> ONE all-reduce of two values (the reference issues two, S2:435 and S2:439) -> stop test,
//   beta.
#include "dist_solver.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <stdexcept>
#include <vector>

namespace pb {

namespace {
double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
}

int cpu_dist_solve(int M, int N, ProcGrid pg, ShmComm& comm, const CpuOptions& opt, double* w_block, SolveStats* stats,
                   DistTimes* times) {
  if (comm.size() != pg.Px * pg.Py) throw std::runtime_error("cpu_dist_solve: communicator size != Px * Py");
  const Grid g = make_grid(M, N);
  const LocalBlock blk = make_local_block(M, N, pg, comm.rank());
  const int T = opt.threads > 0 ? opt.threads : 1;
  const size_t n = blk.elems();
  std::vector<double> a(n), b(n), D(n), r(n), w(n, 0.0), p(n, 0.0), Ap(n, 0.0);
  // edge staging: x-edges (rows) are contiguous in the padded layout, y-edges (columns) are gathered
  std::vector<double> s_ym(blk.nx), s_yp(blk.nx), r_xm(blk.ny), r_xp(blk.ny), r_ym(blk.nx), r_yp(blk.nx);
  const long long max_iter = opt.max_iter >= 0 ? opt.max_iter : (long long)(M - 1) * (long long)(N - 1);
  DistTimes tm;

  comm.barrier();
  const double t0 = now_s();
  cpu_assemble_local(g, blk, a.data(), b.data(), D.data(), r.data(), T);
  tm.assemble = now_s() - t0;
  double zr_old;
  {
    double v[1] = {cpu_dot_zr_local(blk, r.data(), D.data(), T)};
    comm.allreduce_sum(v, 1);
    zr_old = v[0] * g.h1h2;
  }
  auto halo = [&]() {
    for (int i = 0; i < blk.nx; ++i) {
      s_ym[i] = p[blk.index(i + 1, 1)];
      s_yp[i] = p[blk.index(i + 1, blk.ny)];
    }
    const double* send[4] = {&p[blk.index(1, 1)], &p[blk.index(blk.nx, 1)], s_ym.data(), s_yp.data()};
    double* recv[4] = {r_xm.data(), r_xp.data(), r_ym.data(), r_yp.data()};
    const int count[4] = {blk.ny, blk.ny, blk.nx, blk.nx};
    comm.exchange(blk.nbr, send, recv, count);
    if (blk.nbr[kXm] >= 0) std::copy(r_xm.begin(), r_xm.end(), &p[blk.index(0, 1)]);
    if (blk.nbr[kXp] >= 0) std::copy(r_xp.begin(), r_xp.end(), &p[blk.index(blk.nx + 1, 1)]);
    if (blk.nbr[kYm] >= 0)
      for (int i = 0; i < blk.nx; ++i) p[blk.index(i + 1, 0)] = r_ym[i];
    if (blk.nbr[kYp] >= 0)
      for (int i = 0; i < blk.nx; ++i) p[blk.index(i + 1, blk.ny + 1)] = r_yp[i];
  };
  double beta = 0.0, diff_norm = 0.0;
  int status = 0;
  long long k = 0;
  for (k = 1; k <= max_iter; ++k) {
    cpu_update_p_local(blk, p.data(), r.data(), D.data(), beta, T);
    double tc = now_s();
    halo();
    tm.comm += now_s() - tc;
    double v[2] = {cpu_apply_A_local(g, blk, p.data(), a.data(), b.data(), Ap.data(), T), 0.0};
    tc = now_s();
    comm.allreduce_sum(v, 1);
    tm.comm += now_s() - tc;
    const double denom = v[0] * g.h1h2;
    if (opt.signed_breakdown ? denom < kBreakdownTol : std::fabs(denom) < kBreakdownTol) {
      status = 2;
      break;
    }
    const double alpha = zr_old / denom;
    cpu_update_wr_local(blk, w.data(), r.data(), p.data(), Ap.data(), D.data(), alpha, v, T);
    tc = now_s();
    comm.allreduce_sum(v, 2);
    tm.comm += now_s() - tc;
    diff_norm = opt.norm == kNormWeighted ? std::sqrt(v[0] * g.h1h2) : std::sqrt(v[0]);
    const double zr_new = v[1] * g.h1h2;
    if (diff_norm < opt.delta) {
      status = 1;
      break;
    }
    beta = zr_new / zr_old;
    zr_old = zr_new;
  }
  if (k > max_iter) k = max_iter;
  comm.barrier();
  const double seconds = now_s() - t0;

  double e[5];
  cpu_error_local(g, blk, w.data(), e);
  double sums[3] = {e[0], e[1], e[4]}, maxs[3] = {e[2], e[3], tm.comm};
  comm.allreduce_sum(sums, 3);
  comm.allreduce_max(maxs, 3);
  if (stats) {
    stats->iters = (int)k;
    stats->converged = status;
    stats->diff_norm = diff_norm;
    stats->l2_in_D = std::sqrt(g.h1h2 * sums[0]);
    stats->l2_all = std::sqrt(g.h1h2 * sums[1]);
    stats->max_abs = maxs[0];
    stats->max_w = maxs[1];
    stats->nodes_in_D = (long long)sums[2];
    stats->seconds = seconds;
  }
  if (times) {
    *times = tm;
    times->comm = maxs[2];
  }
  if (w_block) std::copy(w.begin(), w.end(), w_block);
  return 0;
}

}  // namespace pb
