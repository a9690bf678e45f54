// problem.h -- the mathematical contract of the fictitious-domain Poisson problem, written once for
// host and device.
//
// Parity notes (reference = /root/reference, tags as in SURVEY.md):
//   constants            S4:11-13   (A1,B1,A2,B2,F_VAL)
//   ellipse membership   S4:32-34   (is_in_D)
//   chord clipping       S4:46-76   (calculate_segment_length_in_D)
//   face coefficients    S4:162-177 (a_ij, b_ij with the 1e-9 tolerances)
//   right-hand side      S4:181-191
//   diagonal             S4:558-561
//
// Every expression below is evaluated with individually rounded IEEE operations (no FMA contraction),
// on the CPU because we build with -ffp-contract=off and on the GPU because the arithmetic goes through
// __dmul_rn/__dadd_rn/__dsub_rn.  fp64 division and sqrt are correctly rounded on both sides, therefore
// the device assembly kernel is bit-identical to the CPU oracle.
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define PB_HD __host__ __device__ __forceinline__
#else
#define PB_HD inline
#endif

namespace pb {

// Rectangle Omega = [A1,B1] x [A2,B2], right-hand side constant.
constexpr double kA1 = -1.0, kB1 = 1.0;
constexpr double kA2 = -0.6, kB2 = 0.6;
constexpr double kF = 1.0;
constexpr double kFaceTol = 1e-9;       // tolerance of the "face fully inside / fully outside" tests
constexpr double kBreakdownTol = 1e-15;  // |(Ap,p)| below this stops the iteration

// Individually rounded arithmetic (see header comment).
PB_HD double mul_rn(double x, double y) {
#if defined(__CUDA_ARCH__)
  return __dmul_rn(x, y);
#else
  return x * y;
#endif
}
PB_HD double add_rn(double x, double y) {
#if defined(__CUDA_ARCH__)
  return __dadd_rn(x, y);
#else
  return x + y;
#endif
}
PB_HD double sub_rn(double x, double y) {
#if defined(__CUDA_ARCH__)
  return __dsub_rn(x, y);
#else
  return x - y;
#endif
}
PB_HD double max_d(double x, double y) { return x < y ? y : x; }
PB_HD double min_d(double x, double y) { return y < x ? y : x; }

// Grid description shared by every backend.
struct Grid {
  int M, N;       // number of cells in x and y; nodes are 0..M, 0..N
  double h1, h2;  // mesh widths
  double eps;     // fictitious-domain parameter max(h1,h2)^2
  double h1h2;    // quadrature weight of the discrete inner product
};

PB_HD Grid make_grid(int M, int N) {
  Grid g;
  g.M = M;
  g.N = N;
  g.h1 = (kB1 - kA1) / M;
  g.h2 = (kB2 - kA2) / N;
  double hm = max_d(g.h1, g.h2);
  g.eps = mul_rn(hm, hm);
  g.h1h2 = mul_rn(g.h1, g.h2);
  return g;
}

PB_HD double node_x(const Grid& g, int i) { return add_rn(kA1, mul_rn((double)i, g.h1)); }
PB_HD double node_y(const Grid& g, int j) { return add_rn(kA2, mul_rn((double)j, g.h2)); }

// (x,y) strictly inside the ellipse x^2 + 4 y^2 < 1.
PB_HD bool in_domain(double x, double y) {
  return add_rn(mul_rn(x, x), mul_rn(mul_rn(4.0, y), y)) < 1.0;
}

// Length of {x = x0, y in [lo,hi]} inside the ellipse.
PB_HD double chord_vertical(double x0, double lo, double hi) {
  if (fabs(x0) >= 1.0) return 0.0;
  double val = sub_rn(1.0, mul_rn(x0, x0)) / 4.0;
  double root = sqrt(max_d(0.0, val));
  return max_d(0.0, sub_rn(min_d(hi, root), max_d(lo, -root)));
}

// Length of {y = y0, x in [lo,hi]} inside the ellipse.
PB_HD double chord_horizontal(double y0, double lo, double hi) {
  if (fabs(mul_rn(2.0, y0)) >= 1.0) return 0.0;
  double val = sub_rn(1.0, mul_rn(mul_rn(4.0, y0), y0));
  double root = sqrt(max_d(0.0, val));
  return max_d(0.0, sub_rn(min_d(hi, root), max_d(lo, -root)));
}

// Blend 1 (inside) and 1/eps (outside) by the fraction of the face of length h that lies in D.
PB_HD double face_coefficient(double len, double h, double eps) {
  if (fabs(sub_rn(len, h)) < kFaceTol) return 1.0;
  if (len < kFaceTol) return 1.0 / eps;
  double frac = len / h;
  return add_rn(frac, sub_rn(1.0, frac) / eps);
}

// a_ij: face x = x_i - h1/2, y in [y_j - h2/2, y_j + h2/2].
PB_HD double coef_a(const Grid& g, int i, int j) {
  double x = node_x(g, i), y = node_y(g, j);
  double hh1 = mul_rn(0.5, g.h1), hh2 = mul_rn(0.5, g.h2);
  double len = chord_vertical(sub_rn(x, hh1), sub_rn(y, hh2), add_rn(y, hh2));
  return face_coefficient(len, g.h2, g.eps);
}

// b_ij: face y = y_j - h2/2, x in [x_i - h1/2, x_i + h1/2].
PB_HD double coef_b(const Grid& g, int i, int j) {
  double x = node_x(g, i), y = node_y(g, j);
  double hh1 = mul_rn(0.5, g.h1), hh2 = mul_rn(0.5, g.h2);
  double len = chord_horizontal(sub_rn(y, hh2), sub_rn(x, hh1), add_rn(x, hh1));
  return face_coefficient(len, g.h1, g.eps);
}

// Right-hand side at node (i,j): F inside D, 0 outside (no cut-cell weighting).
PB_HD double rhs_value(const Grid& g, int i, int j) {
  return in_domain(node_x(g, i), node_y(g, j)) ? kF : 0.0;
}

// Diagonal of the operator from its four face coefficients.
PB_HD double diag_from_faces(const Grid& g, double a_ij, double a_ip, double b_ij, double b_jp) {
  return add_rn(add_rn(a_ip, a_ij) / mul_rn(g.h1, g.h1), add_rn(b_jp, b_ij) / mul_rn(g.h2, g.h2));
}

PB_HD double diag_value(const Grid& g, int i, int j) {
  return diag_from_faces(g, coef_a(g, i, j), coef_a(g, i + 1, j), coef_b(g, i, j), coef_b(g, i, j + 1));
}

// Analytic solution used only for the error report: u = (1 - x^2 - 4 y^2)/10 in D, 0 elsewhere.
PB_HD double exact_solution(double x, double y) {
  double s = x * x + 4.0 * y * y;
  return s < 1.0 ? (1.0 - s) / 10.0 : 0.0;
}

// Literal five-point operator, same association as the reference (S4:526-535):
//   -1/h1 * ( a_ip (p_ip - p_c)/h1 - a_c (p_c - p_im)/h1 ) + (same in y).
// c1 = -1.0/h1 and c2 = -1.0/h2 are passed in (loop invariants, correctly rounded divisions).
PB_HD double stencil_literal(double pc, double pim, double pip, double pjm, double pjp, double a_c,
                             double a_ip, double b_c, double b_jp, double h1, double h2, double c1,
                             double c2) {
  double ax = c1 * (a_ip * (pip - pc) / h1 - a_c * (pc - pim) / h1);
  double ay = c2 * (b_jp * (pjp - pc) / h2 - b_c * (pc - pjm) / h2);
  return ax + ay;
}

enum NormKind : int { kNormWeighted = 0, kNormUnweighted = 1 };

}  // namespace pb
