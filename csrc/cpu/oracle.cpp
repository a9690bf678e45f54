// oracle.cpp -- CPU solver (serial + OpenMP) and local-block operations.  See oracle.h.
//
// The whole-domain solver follows the reference iteration exactly:
//   init  w=0, r=B, z=r/D, p=z, zr=(z,r)                               S0:113-121, S2:384-396
//   loop  Ap=Ap; denom=(Ap,p); break on tiny denom; alpha=zr/denom     S0:126-131, S2:407-416
//         w+=alpha p; r-=alpha Ap; diff^2 from the two stored w values S2:418-427, S4:649-657
//         z=r/D; zr_new=(z,r); stop on ||dw|| < delta                  S0:145-160, S2:429-448
//         beta=zr_new/zr; p=z+beta p                                   S0:162-168
// Unlike the reference, storage is flat (one allocation per field, none inside the loop) and the
// diagonal D is assembled once.
#include "oracle.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <vector>

#if defined(_OPENMP)
#include <omp.h>
#endif

namespace pb {

namespace {
inline int clamp_threads(int t) {
#if defined(_OPENMP)
  return t < 1 ? 1 : t;
#else
  (void)t;
  return 1;
#endif
}
}  // namespace

void cpu_error_report(int M, int N, const double* w, SolveStats* st) {
  const Grid g = make_grid(M, N);
  const int S = N + 1;
  double s_in = 0.0, s_all = 0.0, mx = 0.0, mw = -1e300;
  long long cnt = 0;
  for (int i = 1; i < M; ++i) {
    const double x = node_x(g, i);
    for (int j = 1; j < N; ++j) {
      const double y = node_y(g, j);
      const bool inside = in_domain(x, y);
      const double e = w[(size_t)i * S + j] - exact_solution(x, y);
      s_all += e * e;
      if (inside) {
        s_in += e * e;
        ++cnt;
      }
      mx = std::max(mx, std::fabs(e));
      mw = std::max(mw, w[(size_t)i * S + j]);
    }
  }
  st->l2_in_D = std::sqrt(g.h1h2 * s_in);
  st->l2_all = std::sqrt(g.h1h2 * s_all);
  st->max_abs = mx;
  st->max_w = mw;
  st->nodes_in_D = cnt;
}

int cpu_solve(int M, int N, const CpuOptions& opt, double* w_out, SolveStats* stats) {
  if (M < 2 || N < 2) return -1;
  const auto t0 = std::chrono::steady_clock::now();
  const Grid g = make_grid(M, N);
  const int S = N + 1;
  const size_t n = (size_t)(M + 1) * (size_t)S;
  const int nthreads = clamp_threads(opt.threads);
  const long long max_iter = opt.max_iter >= 0 ? opt.max_iter : (long long)(M - 1) * (long long)(N - 1);
  (void)nthreads;

  std::vector<double> a(n, 0.0), b(n, 0.0), D(n, 1.0), w(n, 0.0), r(n, 0.0), p(n, 0.0), Ap(n, 0.0);

  // assembly: a,b on 1..M x 1..N (S0:47-58), rhs and diagonal on the interior
#pragma omp parallel for num_threads(nthreads) schedule(static)
  for (int i = 1; i <= M; ++i)
    for (int j = 1; j <= N; ++j) {
      a[(size_t)i * S + j] = coef_a(g, i, j);
      b[(size_t)i * S + j] = coef_b(g, i, j);
    }
#pragma omp parallel for num_threads(nthreads) schedule(static)
  for (int i = 1; i < M; ++i)
    for (int j = 1; j < N; ++j) {
      const size_t c = (size_t)i * S + j;
      D[c] = diag_from_faces(g, a[c], a[c + S], b[c], b[c + 1]);
      r[c] = rhs_value(g, i, j);
    }

  double zr_old = 0.0;
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(+ : zr_old)
  for (int i = 1; i < M; ++i)
    for (int j = 1; j < N; ++j) {
      const size_t c = (size_t)i * S + j;
      const double z = D[c] != 0.0 ? r[c] / D[c] : 0.0;
      p[c] = z;
      zr_old += z * r[c];
    }
  zr_old *= g.h1h2;

  const double h1 = g.h1, h2 = g.h2;
  const double c1 = -1.0 / h1, c2 = -1.0 / h2;
  int iters = 0, status = 0;
  double diff_norm = 0.0;

  for (long long k = 1; k <= max_iter; ++k) {
    iters = (int)k;
    double denom = 0.0;
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(+ : denom)
    for (int i = 1; i < M; ++i)
      for (int j = 1; j < N; ++j) {
        const size_t c = (size_t)i * S + j;
        const double v = stencil_literal(p[c], p[c - S], p[c + S], p[c - 1], p[c + 1], a[c], a[c + S], b[c],
                                         b[c + 1], h1, h2, c1, c2);
        Ap[c] = v;
        denom += v * p[c];
      }
    denom *= g.h1h2;
    if (opt.signed_breakdown ? (denom < kBreakdownTol) : (std::fabs(denom) < kBreakdownTol)) {
      status = 2;
      break;
    }
    const double alpha = zr_old / denom;

    double diff2 = 0.0, zr_new = 0.0;
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(+ : diff2, zr_new)
    for (int i = 1; i < M; ++i)
      for (int j = 1; j < N; ++j) {
        const size_t c = (size_t)i * S + j;
        const double w_old = w[c];
        const double w_new = w_old + alpha * p[c];
        const double r_new = r[c] - alpha * Ap[c];
        w[c] = w_new;
        r[c] = r_new;
        const double d = w_new - w_old;
        diff2 += d * d;
        const double z = D[c] != 0.0 ? r_new / D[c] : 0.0;
        zr_new += z * r_new;
      }
    zr_new *= g.h1h2;
    diff_norm = opt.norm == kNormWeighted ? std::sqrt(diff2 * g.h1h2) : std::sqrt(diff2);
    if (diff_norm < opt.delta) {
      status = 1;
      break;
    }
    const double beta = zr_new / zr_old;
    zr_old = zr_new;
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int i = 1; i < M; ++i)
      for (int j = 1; j < N; ++j) {
        const size_t c = (size_t)i * S + j;
        const double z = D[c] != 0.0 ? r[c] / D[c] : 0.0;
        p[c] = z + beta * p[c];
      }
  }
  const auto t1 = std::chrono::steady_clock::now();

  if (stats) {
    stats->iters = iters;
    stats->converged = status;
    stats->diff_norm = diff_norm;
    stats->seconds = std::chrono::duration<double>(t1 - t0).count();
    cpu_error_report(M, N, w.data(), stats);
  }
  if (w_out) std::copy(w.begin(), w.end(), w_out);
  return iters;
}

// ------------------------------------------------------------------------------------------------
// local-block operations
// ------------------------------------------------------------------------------------------------

void cpu_assemble_local(const Grid& g, const LocalBlock& blk, double* a, double* b, double* D, double* rhs,
                        int threads) {
  const int nthreads = clamp_threads(threads);
  (void)nthreads;
  // a, b on the block plus its ring, from global indices (S4:154-178); D, rhs on block plus ring, with
  // the Dirichlet ring (global boundary) set to rhs = 0 and D = the diagonal of a cell far outside the
  // ellipse (any non-zero value would do: it only ever divides zero), so that r/D is 0 there.
  const double inv_eps = 1.0 / g.eps;
  const double d_ring = diag_from_faces(g, inv_eps, inv_eps, inv_eps, inv_eps);
#pragma omp parallel for num_threads(nthreads) schedule(static)
  for (int li = 0; li <= blk.nx + 1; ++li)
    for (int lj = 0; lj <= blk.ny + 1; ++lj) {
      const int gi = blk.i0 + li - 1, gj = blk.j0 + lj - 1;
      const size_t c = blk.index(li, lj);
      a[c] = coef_a(g, gi, gj);
      b[c] = coef_b(g, gi, gj);
      const bool interior = gi >= 1 && gi <= g.M - 1 && gj >= 1 && gj <= g.N - 1;
      D[c] = interior ? diag_value(g, gi, gj) : d_ring;
      rhs[c] = interior ? rhs_value(g, gi, gj) : 0.0;
    }
}

double cpu_apply_A_local(const Grid& g, const LocalBlock& blk, const double* p, const double* a, const double* b,
                         double* Ap, int threads) {
  const int nthreads = clamp_threads(threads);
  (void)nthreads;
  const double h1 = g.h1, h2 = g.h2, c1 = -1.0 / h1, c2 = -1.0 / h2;
  const size_t S = (size_t)blk.pitch;
  double acc = 0.0;
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(+ : acc)
  for (int li = 1; li <= blk.nx; ++li)
    for (int lj = 1; lj <= blk.ny; ++lj) {
      const size_t c = blk.index(li, lj);
      const double v = stencil_literal(p[c], p[c - S], p[c + S], p[c - 1], p[c + 1], a[c], a[c + S], b[c],
                                       b[c + 1], h1, h2, c1, c2);
      Ap[c] = v;
      acc += v * p[c];
    }
  return acc;
}

void cpu_update_wr_local(const LocalBlock& blk, double* w, double* r, const double* p, const double* Ap,
                         const double* D, double alpha, double out[2], int threads) {
  const int nthreads = clamp_threads(threads);
  (void)nthreads;
  double diff2 = 0.0, zr = 0.0;
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(+ : diff2, zr)
  for (int li = 1; li <= blk.nx; ++li)
    for (int lj = 1; lj <= blk.ny; ++lj) {
      const size_t c = blk.index(li, lj);
      const double w_old = w[c];
      const double w_new = w_old + alpha * p[c];
      const double r_new = r[c] - alpha * Ap[c];
      w[c] = w_new;
      r[c] = r_new;
      const double d = w_new - w_old;
      diff2 += d * d;
      const double z = D[c] != 0.0 ? r_new / D[c] : 0.0;
      zr += z * r_new;
    }
  out[0] = diff2;
  out[1] = zr;
}

void cpu_update_p_local(const LocalBlock& blk, double* p, const double* r, const double* D, double beta,
                        int threads) {
  const int nthreads = clamp_threads(threads);
  (void)nthreads;
#pragma omp parallel for num_threads(nthreads) schedule(static)
  for (int li = 1; li <= blk.nx; ++li)
    for (int lj = 1; lj <= blk.ny; ++lj) {
      const size_t c = blk.index(li, lj);
      const double z = D[c] != 0.0 ? r[c] / D[c] : 0.0;
      p[c] = z + beta * p[c];
    }
}

double cpu_dot_zr_local(const LocalBlock& blk, const double* r, const double* D, int threads) {
  const int nthreads = clamp_threads(threads);
  (void)nthreads;
  double zr = 0.0;
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(+ : zr)
  for (int li = 1; li <= blk.nx; ++li)
    for (int lj = 1; lj <= blk.ny; ++lj) {
      const size_t c = blk.index(li, lj);
      const double z = D[c] != 0.0 ? r[c] / D[c] : 0.0;
      zr += z * r[c];
    }
  return zr;
}

void cpu_error_local(const Grid& g, const LocalBlock& blk, const double* w, double out[5]) {
  double s_in = 0.0, s_all = 0.0, mx = 0.0, mw = -1e300, cnt = 0.0;
  for (int li = 1; li <= blk.nx; ++li) {
    const double x = node_x(g, blk.i0 + li - 1);
    for (int lj = 1; lj <= blk.ny; ++lj) {
      const double y = node_y(g, blk.j0 + lj - 1);
      const double v = w[blk.index(li, lj)];
      const double e = v - exact_solution(x, y);
      s_all += e * e;
      if (in_domain(x, y)) {
        s_in += e * e;
        cnt += 1.0;
      }
      mx = std::max(mx, std::fabs(e));
      mw = std::max(mw, v);
    }
  }
  out[0] = s_in;
  out[1] = s_all;
  out[2] = mx;
  out[3] = mw;
  out[4] = cnt;
}

}  // namespace pb
