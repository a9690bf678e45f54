// resident.cu -- KR: the persistent cooperative PCG loop for blocks that fit into the shared memory of the SMs.
//
// A block of up to ~1.2 M nodes per GPU (800 x 1200 on one GPU, the blocks of any grid up to about 1000^2 per GPU)
// is cut into one sub-block per CTA; the CTA loads r once and keeps p (with its one-cell ring), r and w of its
// sub-block in shared memory for the whole solve and Ap in registers between the two phases of an iteration.  An
// iteration then moves no field data through HBM or L2 at all, except
//   * the r values on the rim of a sub-block, which the owner also stores into the global r array in phase C and the
//     neighbouring CTAs read from there (one L2 round trip, ordered by the release / acquire chain of reduction 2)
//     when they recompute p on their ring in the next phase AB -- the same "the ring of p is recomputed, never
//     exchanged" scheme as in the TMA kernel, so an iteration still has exactly two grid-wide synchronisation points;
//   * the r values on the rim of the whole block, which travel to the neighbour GPU as sequence-tagged LL records
//     (same lines, same protocol as persistent_tma.cu) and are polled by the ring threads directly;
//   * a, b, D of the few sub-blocks the ellipse's boundary crosses (read-only, from L2).
// What remains per iteration is arithmetic out of shared memory (about 50 fp64 operations per node) and the two
// in-kernel reductions (star_sum): the regime SURVEY.md section 7.4 describes for 800 x 1200.
//
// Same recurrence, same individually rounded expressions (stencil_value, div_known / div_var) and the same control
// block as the other kernels; the summation order differs, so results agree with them to rounding, not bitwise.
//
// Replaces, for small blocks, the same reference pieces as persistent_tma.cu: the host loop S4:846-942 with its
// kernels S4:507-676, the halo exchange S4:331-500 and the three MPI_Allreduce per iteration.
#include <type_traits>

#include "launch.h"
#include "phases.cuh"

namespace pb {

constexpr int kResMaxPerThread = 16;  // nodes per thread: sub-blocks of up to 8192 nodes

// p with ring, r, w, D on the ring, and one class byte per cell of the sub-block and its ring
__host__ __device__ inline size_t resident_smem_doubles(int sx, int sy) {
  return (size_t)(sx + 2) * (sy + 2) + 2 * (size_t)sx * sy + 2 * (size_t)(sx + sy) + ((size_t)(sx + 2) * (sy + 2) + 7) / 8;
}

template <int DIV>
__global__ void __launch_bounds__(kThreads, 1)
pcg_resident_kernel(const RankCtx* __restrict__ ctxs, int ctas_per_rank, int sx, int sy_log2, int bx, int by, int pair_mode) {
  __shared__ RankCtx c;
  __shared__ double s_red[80];
  __shared__ unsigned long long s_wd_limit;
  extern __shared__ __align__(16) double s_res[];

  const int vrank = blockIdx.x / ctas_per_rank;
  const int cta = blockIdx.x - vrank * ctas_per_rank;
  const int ncta = ctas_per_rank;
  {
    const int* src = reinterpret_cast<const int*>(ctxs + vrank);
    int* dst = reinterpret_cast<int*>(&c);
    for (int i = threadIdx.x; i < (int)(sizeof(RankCtx) / sizeof(int)); i += blockDim.x) dst[i] = src[i];
  }
  if (threadIdx.x == 0) {
    const unsigned long long wd = ctxs[vrank].ctl->watchdog_ns;
    s_wd_limit = wd ? wd : kDefaultWatchdogNs;
  }
  __syncthreads();
  star_pair_init(pair_mode & 1);
  Control* ctl = c.ctl;
  const int S = c.pitch;
  const int sy = 1 << sy_log2, PP = sy + 2;
  // sub-block of this CTA: rows li0 .. li0 + cx - 1, columns lj0 .. lj0 + cy - 1 (cx = 0: nothing to do but to take
  // part in the reductions -- blocks of a decomposition differ by one row)
  const int ix = cta / by, iy = cta - ix * by;
  const int li0 = 1 + ix * sx, lj0 = 1 + iy * sy;
  int cx = min(sx, c.nx - li0 + 1), cy = min(sy, c.ny - lj0 + 1);
  if (cx <= 0 || cy <= 0 || ix >= bx) cx = cy = 0;
  double* P = s_res;                      // (sx + 2) x PP, ring included; cell (row, col) of the sub-block at (row + 1) * PP + col + 1
  double* R = P + (size_t)(sx + 2) * PP;  // sx x sy
  double* W = R + (size_t)sx * sy;        // sx x sy
  double* Dr = W + (size_t)sx * sy;       // D on the ring: top cy | bottom cy | left cx | right cx (offsets sy, 2 sy, 2 sy + sx)
  // class bytes of the sub-block and its ring (same (row + 1) * PP + col + 1 indexing as P): a mixed sub-block takes
  // a, b, D of a cell from its class (1, 1/eps, the two diagonals) and reads the arrays only for the few cells whose
  // class says "anything else" -- the cut faces -- instead of 6 values per node and iteration from L2
  unsigned char* C8 = reinterpret_cast<unsigned char*>(Dr + 2 * (size_t)(sx + sy));

  const double delta = ctl->delta;
  const long long max_iter = ctl->max_iter;
  const int norm = ctl->norm;
  const double h1h2 = c.g.h1h2;
  const StencilConsts K = make_stencil_consts(c.g);
  const double inv_eps = 1.0 / c.g.eps;
  const double diag_in = diag_from_faces(c.g, 1.0, 1.0, 1.0, 1.0);
  const double diag_out = diag_from_faces(c.g, inv_eps, inv_eps, inv_eps, inv_eps);
  unsigned int seq = ctl->red_seq, arrive_target = 0;
  const unsigned int hseq = ctl->halo_seq;
  const bool multi_rank = c.nranks > 1 && !(pair_mode & 2);  // several ranks: cross_rank_sum measured faster than the star (persistent_tma.cu)
#define PB_RES_RED(nv)                                                                                          \
  (arrive_target += ncta, multi_rank ? cross_rank_sum<nv>(c, v, seq, arrive_target, cta, ncta, s_red, false, false) \
                                     : star_sum<nv>(c, v, seq, cta, ncta, s_wd_limit, s_red, nullptr, pair_mode & 1))

  // ---- coefficient class of the sub-block and its ring (corners excluded): 0 inside, 1 outside, 2 mixed ----
  int mode;
  {
    bool all_in = true, all_out = true;
    for (int e = threadIdx.x; e < (cx + 2) * (cy + 2) && cx > 0; e += kThreads) {
      const int row = e / (cy + 2), col = e - row * (cy + 2);
      if ((row == 0 || row == cx + 1) && (col == 0 || col == cy + 1)) continue;
      const unsigned int b = c.cls[(size_t)(li0 - 1 + row) * S + (lj0 - 1 + col + kColOff)];
      all_in = all_in && (b == 0x00u);
      all_out = all_out && (b == 0x15u);
    }
    const int in = __syncthreads_and(all_in ? 1 : 0), out = __syncthreads_and(all_out ? 1 : 0);
    mode = in ? 0 : (out ? 1 : 2);
  }
  const double cf = mode == 1 ? inv_eps : 1.0;
  const double dg = mode == 1 ? diag_out : diag_in, rc = __drcp_rn(dg);

  // ---- load: r -> R, w = 0, p = 0 (ring included), D on the ring, class bytes ----
  for (int e = threadIdx.x; e < (sx + 2) * PP; e += kThreads) {
    P[e] = 0.0;
    const int row = e / PP, col = e - row * PP;
    C8[e] = (cx > 0 && row <= cx + 1 && col <= cy + 1) ? c.cls[(size_t)(li0 - 1 + row) * S + (lj0 - 1 + col + kColOff)] : 0;
  }
  for (int e = threadIdx.x; e < sx * sy; e += kThreads) {
    const int row = e >> sy_log2, col = e & (sy - 1);
    R[e] = (row < cx && col < cy) ? c.r[(size_t)(li0 + row) * S + (lj0 + col + kColOff)] : 0.0;
    W[e] = 0.0;
  }
  // Ring cell `t` of this CTA: 0 .. cy-1 top, cy .. 2cy-1 bottom, then cx left, cx right.  One thread each.
  const int nring = 2 * cy + 2 * cx;
  int ring_li = 0, ring_lj = 0, ring_e = 0, ring_dir = -1, ring_t = 0;  // block-local node, cell in P, side, index along the block edge
  if ((int)threadIdx.x < nring) {
    int t = threadIdx.x;
    if (t < cy) {
      ring_li = li0 - 1, ring_lj = lj0 + t, ring_e = t + 1, ring_dir = kXm;
    } else if (t < 2 * cy) {
      t -= cy;
      ring_li = li0 + cx, ring_lj = lj0 + t, ring_e = (cx + 1) * PP + t + 1, ring_dir = kXp;
    } else if (t < 2 * cy + cx) {
      t -= 2 * cy;
      ring_li = li0 + t, ring_lj = lj0 - 1, ring_e = (t + 1) * PP, ring_dir = kYm;
    } else {
      t -= 2 * cy + cx;
      ring_li = li0 + t, ring_lj = lj0 + cy, ring_e = (t + 1) * PP + cy + 1, ring_dir = kYp;
    }
    ring_t = (ring_dir < 2 ? ring_lj : ring_li) - 1;
    Dr[threadIdx.x] = c.D[(size_t)ring_li * S + (ring_lj + kColOff)];
  }
  // a ring cell that lies on the rim of the whole block next to another rank is fed by that rank's halo records
  const bool ring_remote = (int)threadIdx.x < nring &&
                           ((ring_dir == kXm && ring_li == 0) || (ring_dir == kXp && ring_li == c.nx + 1) ||
                            (ring_dir == kYm && ring_lj == 0) || (ring_dir == kYp && ring_lj == c.ny + 1)) &&
                           c.nbr_r[ring_dir] != nullptr;
  __syncthreads();

  const bool timing = threadIdx.x == 0;
  unsigned long long t_prev = timing ? globaltimer_ns() : 0ull, t_start = t_prev;
  unsigned long long t_ab = 0, t_c = 0, t_r1 = 0, t_r2 = 0, t_init = 0, t_halo = 0;
#define PB_STAMP(acc)                                 \
  if (timing) {                                       \
    const unsigned long long now_ = globaltimer_ns(); \
    acc += now_ - t_prev;                             \
    t_prev = now_;                                    \
  }
  // value of a coefficient from its 2-bit class: 0 -> v_in, 1 -> v_out, else the array
  auto by_class = [&](unsigned int k, double v_in, double v_out, const double* arr, size_t g) -> double {
    return k == 0u ? v_in : (k == 1u ? v_out : __ldg(arr + g));
  };
  // D of the owned node at P-cell `pe`: the sub-block's constant, or by the cell's class for a mixed sub-block
  auto node_D = [&](int pe, size_t g) -> double {
    return mode == 2 ? by_class((C8[pe] >> 4) & 3u, diag_in, diag_out, c.D, g) : dg;
  };
  auto div_D = [&](double r, int pe, size_t g) -> double {
    return mode == 2 ? div_var<DIV>(r, by_class((C8[pe] >> 4) & 3u, diag_in, diag_out, c.D, g)) : div_known<DIV>(r, dg, rc);
  };

  // ---- init: zr_old = (D^-1 r, r) ----
  double zr_old;
  {
    double v[1] = {0.0};
    for (int q = 0; q < kResMaxPerThread; ++q) {
      const int e = threadIdx.x + q * kThreads;
      const int row = e >> sy_log2, col = e & (sy - 1);
      if (row < cx && col < cy) {
        const double r = R[e];
        v[0] = fma(safe_div(r, node_D((row + 1) * PP + col + 1, (size_t)(li0 + row) * S + (lj0 + col + kColOff))), r, v[0]);
      }
    }
    ++seq;
    const RedResult rr = PB_RES_RED(1);
    if (!rr.ok) {
      if (threadIdx.x == 0 && cta == 0) {
        ctl->status = PB_STATUS_WATCHDOG;
        ctl->red_seq = seq;
      }
      star_pair_exit(pair_mode & 1);
      return;
    }
    zr_old = rr.v[0] * h1h2;
  }
  PB_STAMP(t_init);

  double beta = 0.0, alpha = 0.0, denom = 0.0, diff_norm = 0.0;
  int status = PB_STATUS_MAXITER;
  bool halo_timeout = false;
  long long k = 0;
  for (k = 1; k <= max_iter; ++k) {
    // ======================= phase AB =======================
    // 1. p on the ring from the neighbours' r (iteration k - 1), p on the sub-block from R
    if ((int)threadIdx.x < nring) {
      double r;
      if (ring_remote && k > 1) {
        const unsigned long long th0 = timing ? globaltimer_ns() : 0ull;
        const unsigned int want = hseq + (unsigned int)(k - 1);
        const unsigned long long* src = c.halo_ll + (((size_t)ring_dir * 2 + (want & 1u)) * (size_t)c.ll_cap + (size_t)ring_t) * 2;
        unsigned long long w0, w1, tstart = 0;
        unsigned int spins = 0;
        while (true) {
          w0 = ld_relaxed_sys_u64(src), w1 = ld_relaxed_sys_u64(src + 1);
          if ((unsigned int)(w0 >> 32) == want && (unsigned int)(w1 >> 32) == want) break;
          if ((++spins & 0xffu) == 0u) {
            const unsigned long long now = globaltimer_ns();
            if (tstart == 0) tstart = now;
            if (now - tstart > s_wd_limit) {
              halo_timeout = true;
              break;
            }
          }
        }
        r = ll_decode(w0, w1);
        if (timing) t_halo += globaltimer_ns() - th0;
      } else {
        // inside the block: the neighbouring CTA stored it in its phase C; on the boundary of Omega: 0 for ever;
        // first iteration: the right-hand side the assembly kernel wrote on the ring
        r = __ldcg(&c.r[(size_t)ring_li * S + (ring_lj + kColOff)]);
      }
      P[ring_e] = fma(beta, P[ring_e], div_var<DIV>(r, Dr[threadIdx.x]));
    }
#pragma unroll 4
    for (int q = 0; q < kResMaxPerThread; ++q) {
      const int e = threadIdx.x + q * kThreads;
      const int row = e >> sy_log2, col = e & (sy - 1);
      if (row < cx && col < cy) {
        const int pe = (row + 1) * PP + col + 1;
        P[pe] = fma(beta, P[pe], div_D(R[e], pe, (size_t)(li0 + row) * S + (lj0 + col + kColOff)));
      }
    }
    __syncthreads();
    // 2. Ap = A p (kept in registers for phase C), partial (Ap, p)
    double ap[kResMaxPerThread];
    double acc_ab = 0.0;
#pragma unroll
    for (int q = 0; q < kResMaxPerThread; ++q) {
      const int e = threadIdx.x + q * kThreads;
      const int row = e >> sy_log2, col = e & (sy - 1);
      ap[q] = 0.0;
      if (row < cx && col < cy) {
        const double* s = P + (row + 1) * PP + col + 1;
        double a_c = cf, a_ip = cf, b_c = cf, b_jp = cf;
        if (mode == 2) {
          const size_t g = (size_t)(li0 + row) * S + (lj0 + col + kColOff);
          const int pe = (row + 1) * PP + col + 1;
          const unsigned int k0 = C8[pe], kx = C8[pe + PP], ky = C8[pe + 1];
          a_c = by_class(k0 & 3u, 1.0, inv_eps, c.a, g), a_ip = by_class(kx & 3u, 1.0, inv_eps, c.a, g + S);
          b_c = by_class((k0 >> 2) & 3u, 1.0, inv_eps, c.b, g), b_jp = by_class((ky >> 2) & 3u, 1.0, inv_eps, c.b, g + 1);
        }
        ap[q] = stencil_value<DIV>(K, s[0], s[-PP], s[PP], s[-1], s[1], a_c, a_ip, b_c, b_jp);
        acc_ab = fma(ap[q], s[0], acc_ab);
      }
    }
    PB_STAMP(t_ab);
    {
      double v[1] = {acc_ab};
      ++seq;
      const RedResult rr = PB_RES_RED(1);
      PB_STAMP(t_r1);
      if (!rr.ok) {
        status = PB_STATUS_WATCHDOG;
        break;
      }
      denom = rr.v[0] * h1h2;
    }
    if (fabs(denom) < kBreakdownTol) {
      status = PB_STATUS_BREAKDOWN;
      break;
    }
    alpha = zr_old / denom;
    // ======================= phase C =======================
    double acc_d = 0.0, acc_zr = 0.0;
    const unsigned int hs = hseq + (unsigned int)k;
#pragma unroll
    for (int q = 0; q < kResMaxPerThread; ++q) {
      const int e = threadIdx.x + q * kThreads;
      const int row = e >> sy_log2, col = e & (sy - 1);
      if (row < cx && col < cy) {
        const int li = li0 + row, lj = lj0 + col;
        const size_t g = (size_t)li * S + (lj + kColOff);
        const double pc = P[(row + 1) * PP + col + 1];
        const double w_old = W[e];
        const double w_new = fma(alpha, pc, w_old);
        const double r_new = fma(-alpha, ap[q], R[e]);
        W[e] = w_new;
        R[e] = r_new;
        const double dw = __dsub_rn(w_new, w_old);
        acc_d = fma(dw, dw, acc_d);
        acc_zr = fma(div_D(r_new, (row + 1) * PP + col + 1, g), r_new, acc_zr);
        // rim of the sub-block: the neighbouring CTAs read it from the global array in their next phase AB
        if (row == 0 || row == cx - 1 || col == 0 || col == cy - 1) {
          c.r[g] = r_new;
          // rim of the block: sequence-tagged records into the neighbour GPU's halo lines (peer stores over NVLink)
          if (li == 1 || li == c.nx || lj == 1 || lj == c.ny) {
            unsigned long long w0, w1;
            ll_encode(r_new, hs, w0, w1);
            auto push = [&](int d, int t) {
              unsigned long long* dst = c.nbr_ll[d] + ((size_t)(hs & 1u) * (size_t)c.nbr_ll_cap[d] + (size_t)t) * 2;
              st_relaxed_sys_u64(dst, w0);
              st_relaxed_sys_u64(dst + 1, w1);
            };
            if (li == 1 && c.nbr_r[kXm]) push(kXm, lj - 1);
            if (li == c.nx && c.nbr_r[kXp]) push(kXp, lj - 1);
            if (lj == 1 && c.nbr_r[kYm]) push(kYm, li - 1);
            if (lj == c.ny && c.nbr_r[kYp]) push(kYp, li - 1);
          }
        }
      }
    }
    PB_STAMP(t_c);
    double zr_new;
    {
      double v[2] = {acc_d, acc_zr};
      ++seq;
      const RedResult rr = PB_RES_RED(2);
      PB_STAMP(t_r2);
      if (!rr.ok) {
        status = PB_STATUS_WATCHDOG;
        break;
      }
      diff_norm = norm == PB_NORM_WEIGHTED ? sqrt(rr.v[0] * h1h2) : sqrt(rr.v[0]);
      zr_new = rr.v[1] * h1h2;
    }
    if (diff_norm < delta) {
      status = PB_STATUS_CONVERGED;
      break;
    }
    beta = zr_new / zr_old;
    zr_old = zr_new;
  }
  if (k > max_iter) k = max_iter;

  // ---- write back: w, r and the last direction (into both p buffers' owner, p[1]: see Control::cur) ----
  for (int e = threadIdx.x; e < sx * sy; e += kThreads) {
    const int row = e >> sy_log2, col = e & (sy - 1);
    if (row < cx && col < cy) {
      const size_t g = (size_t)(li0 + row) * S + (lj0 + col + kColOff);
      c.w[g] = W[e];
      c.r[g] = R[e];
      c.p[1][g] = P[(row + 1) * PP + col + 1];
    }
  }
  if (halo_timeout) ctl->red[2] = 1.0;  // sticky marker, zeroed by the host before the launch
  if (threadIdx.x == 0) {
    double* mine = reinterpret_cast<double*>(reinterpret_cast<char*>(ctl) + kCtaTimesOffset) + kCtaTimeWords * cta;
    mine[0] = (double)t_ab, mine[1] = (double)t_c, mine[2] = (double)t_r1, mine[3] = (double)t_r2;
    mine[4] = (double)mode, mine[5] = (double)(cx * cy), mine[6] = 0.0;  // (diagnostics: class and size of the sub-block)
    mine[7] = (double)t_halo;
  }
  __syncthreads();
  if (cta == 0 && threadIdx.x == 0) {
    if (ctl->red[2] != 0.0) status = PB_STATUS_WATCHDOG;
    ctl->iter = (int)k;
    ctl->status = status;
    ctl->diff_norm = diff_norm;
    ctl->alpha = alpha;
    ctl->beta = beta;
    ctl->denom = denom;
    ctl->zr_old = zr_old;
    ctl->cur = 1;
    ctl->red_seq = seq;
    ctl->halo_seq = hseq + (unsigned int)k + 1u;
    ctl->t_ab = t_ab;
    ctl->t_c = t_c;
    ctl->t_red1 = t_r1;
    ctl->t_red2 = t_r2;
    ctl->t_init = t_init;
    ctl->t_total = globaltimer_ns() - t_start;
    ctl->t_halo = t_halo;
  }
  star_pair_exit(pair_mode & 1);
#undef PB_RES_RED
#undef PB_STAMP
}

// ---------------------------------------------------------------------------------------------------
// Host side: the cut of a block into one sub-block per CTA.  Columns come in powers of two (32, 64, 128: the index
// arithmetic of 16 nodes per thread and phase is shifts), rows are free; among the cuts that give every CTA at most
// 8192 nodes and fit into shared memory the one with the fewest nodes per CTA wins.
bool resident_partition(int nx, int ny, int max_ctas, size_t max_smem_bytes, ResidentCut* out) {
  ResidentCut best{};
  long long best_nodes = -1;
  for (int lg = 5; lg <= 7; ++lg) {
    const int sy = 1 << lg;
    const int by = (ny + sy - 1) / sy;
    if (by > max_ctas) continue;
    const int bx = std::min(max_ctas / by, nx);
    if (bx < 1) continue;
    const int sx = (nx + bx - 1) / bx;
    const long long nodes = (long long)sx * sy;
    if (nodes > (long long)kResMaxPerThread * kThreads) continue;
    if (2 * sy + 2 * sx > kThreads) continue;
    if (resident_smem_doubles(sx, sy) * sizeof(double) > max_smem_bytes) continue;
    const int bx_used = (nx + sx - 1) / sx;  // fewer row blocks can cover the block once sx is rounded up
    if (best_nodes < 0 || nodes < best_nodes) {
      best_nodes = nodes;
      best = ResidentCut{sx, lg, bx_used, by};
    }
  }
  if (best_nodes < 0) return false;
  *out = best;
  return true;
}

size_t resident_smem_bytes(const ResidentCut& cut) { return resident_smem_doubles(cut.sx, 1 << cut.sy_log2) * sizeof(double); }

cudaError_t launch_resident(const RankCtx* d_ctxs, int nvirt, const ResidentCut& cut, int div, cudaStream_t stream) {
  const void* fn = div == kDivTrue ? (const void*)pcg_resident_kernel<kDivTrue>
                                   : (div == kDivMarkstein ? (const void*)pcg_resident_kernel<kDivMarkstein>
                                                           : (const void*)pcg_resident_kernel<kDivRecip>);
  const size_t smem = resident_smem_bytes(cut);
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int ctas = cut.bx * cut.by, sx = cut.sx, lg = cut.sy_log2, bx = cut.bx, by = cut.by, pair = 0;
  void* args[] = {(void*)&d_ctxs, (void*)&ctas, (void*)&sx, (void*)&lg, (void*)&bx, (void*)&by, (void*)&pair};
  return launch_cooperative_pairs(fn, nvirt, ctas, kThreads, args, &pair, smem, stream);
}

}  // namespace pb
