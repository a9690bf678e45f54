// c_api_cpu.cpp -- C ABI for the host-only parts (decomposition, problem definition, CPU solver).
#include <cstring>
#include <exception>
#include <string>

#include "../../include/poisson_b200.h"
#include "../core/decomp.h"
#include "../core/problem.h"
#include <algorithm>

#include "../cpu/dist_solver.h"
#include "../cpu/oracle.h"
#include "error.h"

namespace pb {
thread_local std::string g_last_error;
void set_last_error(const std::string& msg) { g_last_error = msg; }
}  // namespace pb

namespace {

pb::LocalBlock to_block(const pb_block* b) {
  pb::LocalBlock o;
  o.rank = b->rank;
  o.px = b->px;
  o.py = b->py;
  o.i0 = b->i0;
  o.j0 = b->j0;
  o.nx = b->nx;
  o.ny = b->ny;
  o.pitch = b->pitch;
  for (int d = 0; d < 4; ++d) o.nbr[d] = b->nbr[d];
  return o;
}

void fill_stats(const pb::SolveStats& s, pb_stats* out) {
  std::memset(out, 0, sizeof(*out));
  out->iters = s.iters;
  out->status = s.converged;
  out->diff_norm = s.diff_norm;
  out->l2_in_D = s.l2_in_D;
  out->l2_all = s.l2_all;
  out->max_abs = s.max_abs;
  out->max_w = s.max_w;
  out->nodes_in_D = s.nodes_in_D;
  out->seconds = s.seconds;
}

}  // namespace

extern "C" {

int pb_version(void) { return 100; }
const char* pb_last_error(void) { return pb::g_last_error.c_str(); }

void pb_choose_proc_grid(int nranks, int* Px, int* Py) {
  const pb::ProcGrid g = pb::choose_proc_grid(nranks);
  *Px = g.Px;
  *Py = g.Py;
}

int pb_choose_grid_topology(int n, const int* perf, int M, int* Px, int* Py, int* order) {
  if (n < 1 || n > 64 || !perf || !Px || !Py || !order) {
    pb::set_last_error("pb_choose_grid_topology: 1..64 devices and non-null arguments are required");
    return -1;
  }
  pb::ProcGrid g;
  const int kind = pb::choose_grid_from_topology(n, perf, M, 32, &g, order);
  *Px = g.Px;
  *Py = g.Py;
  return kind;
}

int pb_local_block(int M, int N, int Px, int Py, int rank, pb_block* out) {
  if (M < 2 || N < 2 || Px < 1 || Py < 1 || rank < 0 || rank >= Px * Py) {
    pb::set_last_error("pb_local_block: invalid arguments");
    return -1;
  }
  if (Px > M - 1 || Py > N - 1) {
    pb::set_last_error("pb_local_block: more blocks than interior nodes along an axis");
    return -1;
  }
  const pb::LocalBlock b = pb::make_local_block(M, N, pb::ProcGrid{Px, Py}, rank);
  out->rank = b.rank;
  out->px = b.px;
  out->py = b.py;
  out->i0 = b.i0;
  out->j0 = b.j0;
  out->nx = b.nx;
  out->ny = b.ny;
  out->pitch = b.pitch;
  for (int d = 0; d < 4; ++d) out->nbr[d] = b.nbr[d];
  return 0;
}

int pb_col_offset(void) { return pb::kColOff; }

double pb_coef_a(int M, int N, int i, int j) { return pb::coef_a(pb::make_grid(M, N), i, j); }
double pb_coef_b(int M, int N, int i, int j) { return pb::coef_b(pb::make_grid(M, N), i, j); }
double pb_rhs(int M, int N, int i, int j) { return pb::rhs_value(pb::make_grid(M, N), i, j); }
double pb_diag(int M, int N, int i, int j) { return pb::diag_value(pb::make_grid(M, N), i, j); }
double pb_chord(double fixed, double lo, double hi, int vertical) {
  return vertical ? pb::chord_vertical(fixed, lo, hi) : pb::chord_horizontal(fixed, lo, hi);
}
double pb_exact(double x, double y) { return pb::exact_solution(x, y); }

int pb_cpu_solve(int M, int N, int norm, double delta, long long max_iter, int threads, int signed_breakdown,
                 double* w_out, pb_stats* stats) {
  try {
    if (M < 2 || N < 2) {
      pb::set_last_error("pb_cpu_solve: M and N must be >= 2");
      return -1;
    }
    pb::CpuOptions opt;
    opt.norm = norm;
    opt.delta = delta;
    opt.max_iter = max_iter;
    opt.threads = threads;
    opt.signed_breakdown = signed_breakdown;
    pb::SolveStats st;
    pb::cpu_solve(M, N, opt, w_out, &st);
    if (stats) fill_stats(st, stats);
    return 0;
  } catch (const std::exception& e) {
    pb::set_last_error(e.what());
    return -2;
  }
}

int pb_cpu_dist_solve(int M, int N, int Px, int Py, int rank, const char* job, int norm, double delta, long long max_iter,
                      int threads, double* w_block, pb_stats* stats) {
  try {
    if (M < 2 || N < 2 || Px < 1 || Py < 1 || rank < 0 || rank >= Px * Py || !job) {
      pb::set_last_error("pb_cpu_dist_solve: invalid arguments");
      return -1;
    }
    const pb::ProcGrid pg{Px, Py};
    size_t line = 0;
    for (int r = 0; r < Px * Py; ++r) {
      const pb::LocalBlock b = pb::make_local_block(M, N, pg, r);
      line = std::max(line, (size_t)std::max(b.nx, b.ny));
    }
    pb::ShmComm comm(job, rank, Px * Py, line);
    pb::CpuOptions opt;
    opt.norm = norm;
    opt.delta = delta;
    opt.max_iter = max_iter;
    opt.threads = threads;
    pb::SolveStats st;
    pb::DistTimes tm;
    pb::cpu_dist_solve(M, N, pg, comm, opt, w_block, &st, &tm);
    if (stats) {
      fill_stats(st, stats);
      stats->t_assemble = tm.assemble;
      stats->t_comm = tm.comm;
      stats->t_iterations = st.seconds - tm.assemble;
    }
    return 0;
  } catch (const std::exception& e) {
    pb::set_last_error(e.what());
    return -2;
  }
}

int pb_cpu_error_report(int M, int N, const double* w, pb_stats* stats) {
  pb::SolveStats st;
  pb::cpu_error_report(M, N, w, &st);
  fill_stats(st, stats);
  return 0;
}

int pb_cpu_assemble_local(int M, int N, const pb_block* blk, double* a, double* b, double* D, double* rhs,
                          int threads) {
  pb::cpu_assemble_local(pb::make_grid(M, N), to_block(blk), a, b, D, rhs, threads);
  return 0;
}

double pb_cpu_apply_A_local(int M, int N, const pb_block* blk, const double* p, const double* a, const double* b,
                            double* Ap, int threads) {
  return pb::cpu_apply_A_local(pb::make_grid(M, N), to_block(blk), p, a, b, Ap, threads);
}

int pb_cpu_update_wr_local(const pb_block* blk, double* w, double* r, const double* p, const double* Ap,
                           const double* D, double alpha, double* out2, int threads) {
  pb::cpu_update_wr_local(to_block(blk), w, r, p, Ap, D, alpha, out2, threads);
  return 0;
}

int pb_cpu_update_p_local(const pb_block* blk, double* p, const double* r, const double* D, double beta,
                          int threads) {
  pb::cpu_update_p_local(to_block(blk), p, r, D, beta, threads);
  return 0;
}

double pb_cpu_dot_zr_local(const pb_block* blk, const double* r, const double* D, int threads) {
  return pb::cpu_dot_zr_local(to_block(blk), r, D, threads);
}

int pb_cpu_error_local(int M, int N, const pb_block* blk, const double* w, double* out5) {
  pb::cpu_error_local(pb::make_grid(M, N), to_block(blk), w, out5);
  return 0;
}

}  // extern "C"
