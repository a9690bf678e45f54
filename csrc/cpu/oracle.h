// oracle.h -- CPU implementation of the solver: the correctness oracle and the timed CPU baseline.
//
// Covers the reference's CPU stages: sequential (S0:106-172), OpenMP (S1:133-199) and, through the
// *_local functions driven by the Python gloo backend, the distributed stages (S2:356-460, S3:364-470).
#pragma once
#include <cstddef>

#include "../core/decomp.h"
#include "../core/problem.h"

namespace pb {

struct CpuOptions {
  int norm = kNormWeighted;       // stop norm: weighted (stages 1-4) or unweighted (stage 0)
  double delta = 1e-6;            // stop tolerance
  long long max_iter = -1;        // <0: (M-1)(N-1) as in S4:1006
  int threads = 1;                // OpenMP threads (1 = serial)
  int signed_breakdown = 0;       // 1: stop when denom < 1e-15 (stages 0/1), 0: |denom| < 1e-15
};

struct SolveStats {
  int iters = 0;
  int converged = 0;  // 1 converged, 0 hit max_iter, 2 breakdown
  double diff_norm = 0.0;
  double l2_in_D = 0.0, l2_all = 0.0, max_abs = 0.0, max_w = 0.0;
  long long nodes_in_D = 0;
  double seconds = 0.0;  // solver wall time (assembly + iterations)
};

// Whole-domain solve. `w_out`, if not null, receives the (M+1) x (N+1) nodal field (row = i, x
// direction; zero on the boundary of Omega).
int cpu_solve(int M, int N, const CpuOptions& opt, double* w_out, SolveStats* stats);

// Error of a nodal field (same layout as w_out) against the analytic solution.
void cpu_error_report(int M, int N, const double* w, SolveStats* stats);

// ---- local-block operations (padded LocalBlock layout, ring included) ----
void cpu_assemble_local(const Grid& g, const LocalBlock& blk, double* a, double* b, double* D, double* rhs,
                        int threads);
// Ap = A p on the owned nodes; returns sum(Ap * p) (unscaled).
double cpu_apply_A_local(const Grid& g, const LocalBlock& blk, const double* p, const double* a, const double* b,
                         double* Ap, int threads);
// w += alpha p, r -= alpha Ap; out[0] = sum (w_new - w_old)^2, out[1] = sum (r/D) r (both unscaled).
void cpu_update_wr_local(const LocalBlock& blk, double* w, double* r, const double* p, const double* Ap,
                         const double* D, double alpha, double out[2], int threads);
// p = r/D + beta p on the owned nodes.
void cpu_update_p_local(const LocalBlock& blk, double* p, const double* r, const double* D, double beta,
                        int threads);
// sum (r/D) r over owned nodes (unscaled).
double cpu_dot_zr_local(const LocalBlock& blk, const double* r, const double* D, int threads);
// error sums over owned nodes: out = {sum e^2 in D, sum e^2 all, max |e|, max w, nodes in D}
void cpu_error_local(const Grid& g, const LocalBlock& blk, const double* w, double out[5]);

}  // namespace pb
