// phases.cuh -- the two bandwidth-bound phases of one PCG iteration, written per tile so that the
// step-wise kernels and the persistent cooperative kernel share them.
//
//   phase AB  (replaces update_p_kernel S4:663-676, the halo exchange S4:331-500, apply_A_kernel
//              S4:507-536 and dot_kernel(Ap,p) S4:574-598):
//       p_new = r/D + beta * p_old on the tile AND its one-cell ring (the ring of the block itself is
//       recomputed from the neighbour's r values, which the neighbour stored into our ring during its
//       phase C), Ap = A p_new on the tile, partial (Ap, p_new).
//   phase C   (replaces update_w_r_kernel S4:626-660, apply_Dinv_kernel S4:541-562, dot_kernel(z,r)):
//       w += alpha p, r -= alpha Ap, partial sum (w_new - w_old)^2, partial (r/D, r); boundary nodes
//       additionally store r_new straight into the neighbour GPU's ring (peer store over NVLink).
//
// z = r/D is never stored and always a true division.  The divisions by h1/h2 inside the stencil follow
// the DivMode template parameter (see below).
#pragma once
#include "device_utils.cuh"

namespace pb {

constexpr int kThreads = 512;                 // CTA size of the tile kernels
constexpr int kSmemPitch = kTileY + 2;        // doubles per row of the p_new tile in shared memory
constexpr int kSmemTile = (kTileX + 2) * kSmemPitch;

struct TileGeom {
  int li0, lj0;  // local index of the tile's first owned node
  int cx, cy;    // owned rows / columns in this tile (partial tiles at the block edge)
};

__device__ __forceinline__ TileGeom tile_geom(const RankCtx& c, int tile) {
  TileGeom t;
  const int tx = tile / c.tiles_y, ty = tile - tx * c.tiles_y;
  t.li0 = 1 + tx * kTileX;
  t.lj0 = 1 + ty * kTileY;
  t.cx = min(kTileX, c.nx - t.li0 + 1);
  t.cy = min(kTileY, c.ny - t.lj0 + 1);
  return t;
}

__device__ __forceinline__ double safe_div(double r, double d) { return d != 0.0 ? r / d : 0.0; }

// How the six divisions by h1 / h2 inside the stencil are evaluated.
//   kDivTrue      x / h, the compiler's IEEE division (about a dozen fp64 instructions each)
//   kDivMarkstein q = x*y with y = RN(1/h), then one residual correction q' = fma(fma(-q, h, x), y, q)
//                 (Markstein): the correctly rounded quotient, i.e. the same bits as kDivTrue, in three
//                 instructions and without the range fix-up path of the general division (operands here are
//                 far from overflow / underflow).  bench/div_check.py: 0 mismatches in 1.4e9 samples.
//   kDivRecip     x * y, one rounding differs from x / h in the last bit now and then
enum DivMode : int { kDivRecip = 0, kDivTrue = 1, kDivMarkstein = 2 };

struct StencilConsts {
  double h1, h2, c1, c2, ih1, ih2;  // c = -1/h, ih = RN(1/h)
};
__device__ __forceinline__ StencilConsts make_stencil_consts(const Grid& g) {
  StencilConsts k;
  k.h1 = g.h1, k.h2 = g.h2;
  k.c1 = -1.0 / g.h1, k.c2 = -1.0 / g.h2;
  k.ih1 = 1.0 / g.h1, k.ih2 = 1.0 / g.h2;
  return k;
}

template <int DIV>
__device__ __forceinline__ double div_const(double x, double h, double y) {
  if (DIV == kDivTrue) return x / h;
  const double q = __dmul_rn(x, y);
  if (DIV == kDivRecip) return q;
  return fma(fma(-q, h, x), y, q);
}

// z = r / D with a per-node divisor.  kDivTrue: IEEE division (on this data its out-of-line slow path runs
// often: residuals are zero or tiny next to D ~ 1e7, and the profile showed 13 % of all issued instructions
// there).  Otherwise: y = RN(1/D) from __drcp_rn -- its range check looks at D only, which is always an
// ordinary number -- followed by the same Markstein correction as div_const.  Same bits as the division
// unless |r/D| is near the subnormal range (never the case before convergence: r >= 1e-20, D <= 1e9).
template <int DIV>
__device__ __forceinline__ double div_var(double r, double D) {
  if (DIV == kDivTrue) return D != 0.0 ? r / D : 0.0;
  const double y = __drcp_rn(D);
  const double q = __dmul_rn(r, y);
  const double z = fma(fma(-q, D, r), y, q);
  return D != 0.0 ? z : 0.0;
}

// z = r / D when RN(1/D) is already known (D is one of the two diagonal constants of the coefficient classes).
template <int DIV>
__device__ __forceinline__ double div_known(double r, double D, double y) {
  if (DIV == kDivTrue) return r / D;
  const double q = __dmul_rn(r, y);
  return fma(fma(-q, D, r), y, q);
}

// (A p) at one node, association of the reference (S4:526-535).
template <int DIV>
__device__ __forceinline__ double stencil_value(const StencilConsts& k, double pc, double pim, double pip, double pjm,
                                                double pjp, double a_c, double a_ip, double b_c, double b_jp) {
  // every rounding is pinned (no compiler-chosen contraction), so all DivModes and all kernels that use
  // this function produce the same bits for the same inputs; the one fused step is Ax + Ay = fma(c1, dx, Ay),
  // which is what nvcc makes of the reference's expression
  const double dx = __dsub_rn(div_const<DIV>(__dmul_rn(a_ip, __dsub_rn(pip, pc)), k.h1, k.ih1),
                              div_const<DIV>(__dmul_rn(a_c, __dsub_rn(pc, pim)), k.h1, k.ih1));
  const double dy = __dsub_rn(div_const<DIV>(__dmul_rn(b_jp, __dsub_rn(pjp, pc)), k.h2, k.ih2),
                              div_const<DIV>(__dmul_rn(b_c, __dsub_rn(pc, pjm)), k.h2, k.ih2));
  return fma(k.c1, dx, __dmul_rn(k.c2, dy));
}

// Phase AB on one tile; returns this thread's contribution to sum(Ap * p_new).
// FUSE_P = false: `pin` already holds the direction including its ring (classic path); nothing is
// written to `pout`.
template <bool FUSE_P, int DIV>
__device__ __forceinline__ double phase_ab_tile(const RankCtx& c, int tile, double beta, const double* __restrict__ pin,
                                                double* __restrict__ pout, double* sp) {
  const TileGeom t = tile_geom(c, tile);
  const int S = c.pitch;
  // 1. extended tile -> shared memory
  const int ext_rows = t.cx + 2, ext_cols = t.cy + 2;
  for (int e = threadIdx.x; e < ext_rows * kSmemPitch; e += blockDim.x) {
    const int row = e / kSmemPitch, col = e - row * kSmemPitch;
    if (col >= ext_cols) continue;
    const bool row_in = row >= 1 && row <= t.cx, col_in = col >= 1 && col <= t.cy;
    if (!(row_in || col_in)) continue;  // corners of the extended tile are not needed
    const int li = t.li0 - 1 + row, lj = t.lj0 - 1 + col;
    const size_t g = (size_t)li * S + (lj + kColOff);
    double v;
    if (FUSE_P) {
      v = fma(beta, __ldcg(&pin[g]), div_var<DIV>(__ldcg(&c.r[g]), __ldg(&c.D[g])));
      const bool ring = (li == 0) || (li == c.nx + 1) || (lj == 0) || (lj == c.ny + 1);
      if ((row_in && col_in) || ring) pout[g] = v;
    } else {
      v = __ldcg(&pin[g]);
    }
    sp[row * kSmemPitch + col] = v;
  }
  __syncthreads();
  // 2. stencil on the owned nodes
  const StencilConsts K = make_stencil_consts(c.g);
  double acc = 0.0;
  const int col = threadIdx.x % kTileY + 1;
  for (int row = threadIdx.x / kTileY + 1; row <= t.cx; row += blockDim.x / kTileY) {
    if (col > t.cy) continue;
    const int li = t.li0 - 1 + row, lj = t.lj0 - 1 + col;
    const size_t g = (size_t)li * S + (lj + kColOff);
    const double* s = sp + row * kSmemPitch + col;
    const double pc = s[0], pim = s[-kSmemPitch], pip = s[kSmemPitch], pjm = s[-1], pjp = s[1];
    const double a_c = __ldg(&c.a[g]), a_ip = __ldg(&c.a[g + S]);
    const double b_c = __ldg(&c.b[g]), b_jp = __ldg(&c.b[g + 1]);
    const double v = stencil_value<DIV>(K, pc, pim, pip, pjm, pjp, a_c, a_ip, b_c, b_jp);
    c.Ap[g] = v;
    acc = fma(v, pc, acc);
  }
  __syncthreads();  // the tile buffer is reused by the next tile
  return acc;
}

// Phase C on one tile; adds this thread's contributions to diff2 and zr.
__device__ __forceinline__ void phase_c_tile(const RankCtx& c, int tile, double alpha, const double* __restrict__ p,
                                             double& diff2, double& zr) {
  const TileGeom t = tile_geom(c, tile);
  const int S = c.pitch;
  const int col = threadIdx.x % kTileY + 1;
  if (col > t.cy) return;
  for (int row = threadIdx.x / kTileY + 1; row <= t.cx; row += blockDim.x / kTileY) {
    const int li = t.li0 - 1 + row, lj = t.lj0 - 1 + col;
    const size_t g = (size_t)li * S + (lj + kColOff);
    const double w_old = c.w[g];
    const double w_new = fma(alpha, __ldcg(&p[g]), w_old);
    const double r_new = fma(-alpha, c.Ap[g], c.r[g]);
    c.w[g] = w_new;
    c.r[g] = r_new;
    const double d = __dsub_rn(w_new, w_old);
    diff2 = fma(d, d, diff2);
    zr = fma(safe_div(r_new, __ldg(&c.D[g])), r_new, zr);
    // halo push: our boundary values become the neighbour's ring
    if (li == 1 && c.nbr_r[kXm])
      c.nbr_r[kXm][(size_t)(c.nbr_nx[kXm] + 1) * c.nbr_pitch[kXm] + (lj + kColOff)] = r_new;
    if (li == c.nx && c.nbr_r[kXp]) c.nbr_r[kXp][(size_t)0 * c.nbr_pitch[kXp] + (lj + kColOff)] = r_new;
    if (lj == 1 && c.nbr_r[kYm])
      c.nbr_r[kYm][(size_t)li * c.nbr_pitch[kYm] + (c.nbr_ny[kYm] + 1 + kColOff)] = r_new;
    if (lj == c.ny && c.nbr_r[kYp]) c.nbr_r[kYp][(size_t)li * c.nbr_pitch[kYp] + (0 + kColOff)] = r_new;
  }
}

// Initial (z, r) with z = r/D over one tile.
__device__ __forceinline__ double dot_zr_tile(const RankCtx& c, int tile) {
  const TileGeom t = tile_geom(c, tile);
  const int S = c.pitch;
  const int col = threadIdx.x % kTileY + 1;
  double zr = 0.0;
  if (col > t.cy) return zr;
  for (int row = threadIdx.x / kTileY + 1; row <= t.cx; row += blockDim.x / kTileY) {
    const size_t g = (size_t)(t.li0 - 1 + row) * S + (t.lj0 - 1 + col + kColOff);
    const double r = c.r[g];
    zr = fma(safe_div(r, c.D[g]), r, zr);
  }
  return zr;
}

}  // namespace pb
