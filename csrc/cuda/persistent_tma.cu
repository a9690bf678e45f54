// persistent_tma.cu -- KP-TMA: the persistent cooperative PCG loop with TMA-fed, multi-stage tile pipelines.
//
// Same algorithm and cross-rank protocol as persistent.cu; what changes is how a CTA moves data.  Each
// phase is a software pipeline over the CTA's tiles: one elected thread issues `cp.async.bulk.tensor.2d`
// loads for tile j+S-1 (all operand boxes of the tile land in one shared-memory stage and complete on one
// mbarrier) while all warps compute tile j out of shared memory.  A stage holds
//     phase AB: r, D, p_old boxes of (TX+2) x (TY+4) (tile + ring; one more column on each side because a
//               box must start on a 16-byte boundary), a box (TX+1) x TY, b box TX x (TY+2)
//     phase C : w, p, r, Ap, D boxes of TX x TY
// so one SM keeps S-1 complete tiles (45-88 KB each) in flight without spending registers on it, which is
// what a bandwidth-bound fp64 stencil needs on a part with ~40 KB/SM of latency-bandwidth product.
// The TMA unit also does the address generation for the misaligned ring boxes and zero-fills the part of
// a box that lies outside the array (partial tiles at the block edge).
//
// Replaces the same reference pieces as persistent.cu (S4:846-942, kernels S4:507-676).
#include "launch.h"
#include "phases.cuh"
#include "tma_utils.cuh"

namespace pb {

enum TmaMap : int {
  kMapRExt = 0, kMapDExt, kMapP0Ext, kMapP1Ext, kMapA, kMapB, kMapW, kMapP0, kMapP1, kMapR, kMapAp, kMapD,
  kMapClsAB,  // class bytes, (TX+2) rows x 80 bytes, box starts at a multiple of 16 columns
  kMapClsC    // class bytes, TX rows x 80 bytes
};
constexpr int kClsBoxBytes = 80;  // 64 tile columns + ring + the up to 15 bytes in front of the aligned start

// Shared-memory stage layout.  CC = "compressed coefficients": a, b and D are not streamed as fp64 boxes;
// one class byte per cell (RankCtx::cls) selects 1, 1/eps or the respective diagonal constant, and only the
// O(M+N) cells cut by the ellipse read the fp64 arrays.  82 instead of 112 B/node/iteration, 5 instead of 8
// arrays in the streamed working set, and three pipeline stages instead of two.
template <int TX, bool CC>
struct TmaCfg {
  static constexpr int TY = kTileY;
  static constexpr int ER = TX + 2, EC = TY + 4;  // ring boxes start one column early: TMA wants 16-byte starts
  static constexpr int BC = TY + 2;               // b box width
  static constexpr int kExtBytes = ((ER * EC * 8 + 127) / 128) * 128;
  static constexpr int kABytes = (TX + 1) * TY * 8;
  static constexpr int kBBytes = TX * BC * 8;
  static constexpr int kTileBytes = TX * TY * 8;
  static constexpr int kClsABBytes = ((ER * kClsBoxBytes + 127) / 128) * 128;
  static constexpr int kClsCBytes = ((TX * kClsBoxBytes + 127) / 128) * 128;
  // phase AB offsets
  static constexpr int kOffR = 0;
  static constexpr int kOffD = CC ? 0 : kExtBytes;  // unused with CC
  static constexpr int kOffP = CC ? kExtBytes : 2 * kExtBytes;
  static constexpr int kOffA = 3 * kExtBytes;       // unused with CC
  static constexpr int kOffB = 3 * kExtBytes + kABytes;
  static constexpr int kOffClsAB = 2 * kExtBytes;   // CC only
  static constexpr int kStageAB = CC ? 2 * kExtBytes + kClsABBytes : 3 * kExtBytes + kABytes + kBBytes;
  // phase C offsets: w, p, r, Ap, then D (fp64 box) or the class bytes
  static constexpr int kOffCD = 4 * kTileBytes;
  static constexpr int kStageC = CC ? 4 * kTileBytes + kClsCBytes : 5 * kTileBytes;
  static constexpr int kStageBytes = kStageAB > kStageC ? kStageAB : kStageC;
  static constexpr int kStages = CC ? (TX >= 32 ? 3 : 6) : (TX >= 32 ? 2 : 4);
  // transaction bytes = full boxes, out-of-bounds parts included
  static constexpr uint32_t kTxAB = CC ? 2u * ER * EC * 8u + (uint32_t)(ER * kClsBoxBytes)
                                       : 3u * ER * EC * 8u + (uint32_t)kABytes + (uint32_t)kBBytes;
  static constexpr uint32_t kTxC = CC ? 4u * (uint32_t)kTileBytes + (uint32_t)(TX * kClsBoxBytes) : 5u * (uint32_t)kTileBytes;
  static constexpr size_t kSmemBytes = (size_t)kStages * kStageBytes + 128;
  static_assert(kSmemBytes <= 227 * 1024, "stages do not fit into shared memory");
};

template <int TX, int DIV, bool CC>
__global__ void __launch_bounds__(kThreads, 1)
pcg_persistent_tma_kernel(const RankCtx* __restrict__ ctxs, int ctas_per_rank, int map_variant, int l2_hints) {
  using Cfg = TmaCfg<TX, CC>;
  constexpr int TY = Cfg::TY, EC = Cfg::EC, BC = Cfg::BC, S = Cfg::kStages;
  __shared__ RankCtx c;
  __shared__ double s_red[80];
  __shared__ __align__(8) uint64_t s_bar[S];
  extern __shared__ __align__(128) unsigned char s_dyn[];

  const int vrank = blockIdx.x / ctas_per_rank;
  const int cta = blockIdx.x - vrank * ctas_per_rank;
  const int ncta = ctas_per_rank;
  {
    const int* src = reinterpret_cast<const int*>(ctxs + vrank);
    int* dst = reinterpret_cast<int*>(&c);
    for (int i = threadIdx.x; i < (int)(sizeof(RankCtx) / sizeof(int)); i += blockDim.x) dst[i] = src[i];
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) mbar_init(&s_bar[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  // s_dyn is declared __align__(128); deriving every stage pointer from it by plain pointer arithmetic keeps
  // the accesses in the shared address space (LDS/STS instead of generic LD/ST)
  unsigned char* const stage_base = s_dyn;
  const CUtensorMap* maps = reinterpret_cast<const CUtensorMap*>(c.tmaps) + map_variant * kMapsPerVariant;

  Control* ctl = c.ctl;
  const int tiles_y = (c.ny + TY - 1) / TY, tiles_x = (c.nx + TX - 1) / TX;
  const int ntiles = tiles_x * tiles_y;
  const int nmine = cta < ntiles ? (ntiles - cta + ncta - 1) / ncta : 0;
  // Tile order: row-major, except that the last tile row is visited right after the first one.  The tiles
  // that store halo rows into the x-neighbours (phase C) and the tiles that need the neighbours' rows
  // (phase AB) then come first on both sides, so the system-scope release of the stores and the NVLink
  // latency hide behind the remaining tiles instead of sitting in front of a reduction.  (Enumerating the
  // whole perimeter first was measured 7 % slower per phase: it breaks the streaming order of the sweep.)
  const bool has_nbr = c.nbr_r[0] || c.nbr_r[1] || c.nbr_r[2] || c.nbr_r[3];
  const int npre = min(S - 1, nmine);  // tiles of a phase that are issued ahead of the phase
  const bool timing = (cta == 0 && threadIdx.x == 0);
  const int S_pitch = c.pitch;

  unsigned int seq = ctl->red_seq, arrive_target = 0;
  const double delta = ctl->delta;
  const long long max_iter = ctl->max_iter;
  const int norm = ctl->norm;
  const double h1h2 = c.g.h1h2;
  const StencilConsts K = make_stencil_consts(c.g);
  // values behind the coefficient classes (same device functions as the assembly kernel -> same bits)
  const double inv_eps = 1.0 / c.g.eps;
  const double d_in = diag_from_faces(c.g, 1.0, 1.0, 1.0, 1.0), d_out = diag_from_faces(c.g, inv_eps, inv_eps, inv_eps, inv_eps);
  const double y_in = __drcp_rn(d_in), y_out = __drcp_rn(d_out);
  auto face_value = [&](unsigned int k, const double* cell) { return k == 0u ? 1.0 : (k == 1u ? inv_eps : __ldg(cell)); };

  // L2 policy per operand: what both phases of every iteration re-read (r, D, p) is kept, what is touched once
  // per iteration (a, b, w, Ap after its single use) is marked evict-first, so that on a working set around
  // the L2 size (4096^2 over 8 GPUs: 134 MB per GPU vs 126 MB) the re-read part stays resident.
  const uint64_t pol_keep = l2_hints ? kL2EvictLast : kL2EvictNormal;
  const uint64_t pol_stream = l2_hints ? kL2EvictFirst : kL2EvictNormal;

  unsigned int n_issued = 0, n_consumed = 0;  // uniform across the CTA

  auto tile_origin = [&](int j, int& li0, int& lj0) {
    const int t = cta + j * ncta;
    int tx = t / tiles_y;
    const int ty = t - tx * tiles_y;
    if (has_nbr && tiles_x > 2) tx = tx == 0 ? 0 : (tx == 1 ? tiles_x - 1 : tx - 1);
    li0 = 1 + tx * TX;
    lj0 = 1 + ty * TY;
  };
  // index of my last tile that touches an edge shared with a neighbour (-1: none)
  int last_halo_j = -1;
  if (has_nbr) {
    for (int j = 0; j < nmine; ++j) {
      int li0, lj0;
      tile_origin(j, li0, lj0);
      const int cx = min(TX, c.nx - li0 + 1), cy = min(TY, c.ny - lj0 + 1);
      if ((li0 == 1 && c.nbr_r[kXm]) || (li0 + cx - 1 == c.nx && c.nbr_r[kXp]) || (lj0 == 1 && c.nbr_r[kYm]) ||
          (lj0 + cy - 1 == c.ny && c.nbr_r[kYp]))
        last_halo_j = j;
    }
  }
  // Halo completion counters (see ctx.h): value at launch, and by how much one phase C of the neighbour
  // raises ours (= the number of its tiles along the shared edge).
  unsigned long long halo_base[4];
  int edge_tiles[4];
#pragma unroll
  for (int d = 0; d < 4; ++d) {
    halo_base[d] = ld_relaxed_sys_u64(c.halo_cnt + d);
    edge_tiles[d] = c.nbr_r[d] ? (d < 2 ? (c.nbr_ny[d] + TY - 1) / TY : (c.nbr_nx[d] + TX - 1) / TX) : 0;
  }
  // thread 0: wait until the ring cells of the tile at (li0, lj0) hold the neighbours' r of round `rounds`
  auto wait_halo = [&](int li0, int lj0, long long rounds) -> bool {
    const int cx = min(TX, c.nx - li0 + 1), cy = min(TY, c.ny - lj0 + 1);
    const bool touch[4] = {li0 == 1, li0 + cx - 1 == c.nx, lj0 == 1, lj0 + cy - 1 == c.ny};
    bool waited = false;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      if (!touch[d] || !c.nbr_r[d]) continue;
      waited = true;
      const unsigned long long want = halo_base[d] + (unsigned long long)rounds * (unsigned long long)edge_tiles[d];
      unsigned long long t0 = 0;
      unsigned int spins = 0;
      while (ld_acquire_sys_u64(c.halo_cnt + d) < want) {
        if ((++spins & 0xffu) == 0u) {
          const unsigned long long now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          if (now - t0 > (ctl->watchdog_ns ? ctl->watchdog_ns : kDefaultWatchdogNs)) {
            ctl->red[2] = 1.0;  // sticky marker (zeroed by the host before the launch): halo wait timed out
            break;
          }
        }
      }
    }
    return waited;
  };
  // Phase-AB loads of tile j into stage `st`.  part 1 = everything that does not depend on the other CTAs'
  // phase C (D, p, a, b; arms the barrier with the full byte count), part 2 = the r box.
  auto issue_ab_part = [&](int j, unsigned int st, int pbuf, bool part1, bool part2, long long rounds) {
    if (threadIdx.x == 0) {
      int li0, lj0;
      tile_origin(j, li0, lj0);
      unsigned char* sb = stage_base + (size_t)st * Cfg::kStageBytes;
      uint64_t* bar = &s_bar[st];
      const int x = kColOff + lj0 - 2, y = li0 - 1;  // even column: 16-byte aligned box start
      if (part1) {
        fence_proxy_async_smem();
        mbar_arrive_expect_tx(bar, Cfg::kTxAB);
        tma_load_2d_hint(sb + Cfg::kOffP, maps + (pbuf ? kMapP1Ext : kMapP0Ext), x, y, bar, pol_keep);
        if constexpr (CC) {
          tma_load_2d_hint(sb + Cfg::kOffClsAB, maps + kMapClsAB, lj0 - 1, y, bar, pol_keep);  // lj0 - 1 = 64 * ty
        } else {
          tma_load_2d_hint(sb + Cfg::kOffD, maps + kMapDExt, x, y, bar, pol_keep);
          tma_load_2d_hint(sb + Cfg::kOffA, maps + kMapA, x + 2, y + 1, bar, pol_stream);
          tma_load_2d_hint(sb + Cfg::kOffB, maps + kMapB, x + 2, y + 1, bar, pol_stream);
        }
      }
      if (part2) {
        // only a tile on a shared edge reads cells another GPU wrote since the phase-start fence
        if (has_nbr && wait_halo(li0, lj0, rounds)) fence_proxy_async();  // acquired generic-proxy stores -> TMA read
        tma_load_2d_hint(sb, maps + kMapRExt, x, y, bar, pol_keep);
      }
    }
  };
  auto issue_ab = [&](int j, int pbuf, long long rounds) {
    issue_ab_part(j, n_issued % S, pbuf, true, true, rounds);
    ++n_issued;
  };
  auto issue_c = [&](int j, int pbuf) {
    if (threadIdx.x == 0) {
      int li0, lj0;
      tile_origin(j, li0, lj0);
      const unsigned int st = n_issued % S;
      unsigned char* sb = stage_base + (size_t)st * Cfg::kStageBytes;
      uint64_t* bar = &s_bar[st];
      fence_proxy_async_smem();
      mbar_arrive_expect_tx(bar, Cfg::kTxC);
      const int x = kColOff + lj0, y = li0;
      tma_load_2d_hint(sb, maps + kMapW, x, y, bar, pol_stream);
      tma_load_2d_hint(sb + Cfg::kTileBytes, maps + (pbuf ? kMapP1 : kMapP0), x, y, bar, pol_keep);
      tma_load_2d_hint(sb + 2 * Cfg::kTileBytes, maps + kMapR, x, y, bar, pol_keep);
      tma_load_2d_hint(sb + 3 * Cfg::kTileBytes, maps + kMapAp, x, y, bar, pol_stream);
      if constexpr (CC)
        tma_load_2d_hint(sb + Cfg::kOffCD, maps + kMapClsC, lj0 - 1, y, bar, pol_keep);
      else
        tma_load_2d_hint(sb + Cfg::kOffCD, maps + kMapD, x, y, bar, pol_keep);
    }
    ++n_issued;
  };
  bool tma_ok = true;
  auto wait_stage = [&]() -> unsigned char* {
    const unsigned int st = n_consumed % S;
    if (!mbar_wait(&s_bar[st], (n_consumed / S) & 1u)) {
      if (threadIdx.x == 0) {  // internal error: leave a trace and abandon; the other CTAs run into the watchdog
        ctl->status = PB_STATUS_TMA_TIMEOUT;
        ctl->red[0] = (double)n_consumed;
        ctl->red[1] = (double)cta;
      }
      tma_ok = false;
    }
    return stage_base + (size_t)st * Cfg::kStageBytes;
  };
  // nothing may be in flight into shared memory when the CTA exits
  auto drain = [&]() {
    while (n_consumed < n_issued && tma_ok) {
      wait_stage();
      ++n_consumed;
    }
  };

  unsigned long long t_prev = timing ? globaltimer_ns() : 0ull, t_start = t_prev;
  unsigned long long t_ab = 0, t_c = 0, t_r1 = 0, t_r2 = 0, t_init = 0;
#define PB_STAMP(acc)                                 \
  if (timing) {                                       \
    const unsigned long long now_ = globaltimer_ns(); \
    acc += now_ - t_prev;                             \
    t_prev = now_;                                    \
  }

  int cur = 0;
  // Prefetch for the first phase AB.  Everything it reads was written by the assembly kernel (an earlier
  // launch), so both parts can go out before the initial reduction.
  for (int j = 0; j < npre; ++j) issue_ab(j, cur, 0);

  // ---- init: zr_old = (D^-1 r, r); plain loads, executed once ----
  double zr_old;
  {
    double v[1] = {0.0};
    const int nt32 = c.tiles_x * c.tiles_y;
    for (int t = cta; t < nt32; t += ncta) v[0] += dot_zr_tile(c, t);
    ++seq;
    arrive_target += ncta;
    const RedResult rr = cross_rank_sum<1>(c, v, seq, arrive_target, cta, ncta, s_red, false, false);
    if (!rr.ok) {
      drain();
      if (threadIdx.x == 0 && cta == 0) {
        ctl->status = PB_STATUS_WATCHDOG;
        ctl->red_seq = seq;
      }
      return;
    }
    zr_old = rr.v[0] * h1h2;
  }
  PB_STAMP(t_init);

  double beta = 0.0, alpha = 0.0, denom = 0.0, diff_norm = 0.0;
  int status = PB_STATUS_MAXITER;
  long long k = 0;
  for (k = 1; k <= max_iter; ++k) {
    double* pout = c.p[cur ^ 1];
    // ======================= phase AB (tiles 0..npre-1 are already in flight) =======================
    double acc_ab = 0.0;
    for (int j = 0; j < nmine && tma_ok; ++j) {
      if (j + S - 1 < nmine) issue_ab(j + S - 1, cur, k - 1);
      unsigned char* sb = wait_stage();
      if (!tma_ok) break;
      const double* s_r = reinterpret_cast<const double*>(sb);
      const double* s_D = reinterpret_cast<const double*>(sb + Cfg::kOffD);
      double* s_p = reinterpret_cast<double*>(sb + Cfg::kOffP);
      const double* s_a = reinterpret_cast<const double*>(sb + Cfg::kOffA);
      const double* s_b = reinterpret_cast<const double*>(sb + Cfg::kOffB);
      const unsigned char* s_cls = sb + Cfg::kOffClsAB;  // CC: byte of ext cell (row, col) at row * 80 + 3 + col
      int li0, lj0;
      tile_origin(j, li0, lj0);
      const int cx = min(TX, c.nx - li0 + 1), cy = min(TY, c.ny - lj0 + 1);
      // 1. p_new = r/D + beta p_old on the tile and its ring, in place in shared memory.
      //    Full tiles take fixed-trip loops that the compiler unrolls, so the dependent fp64 chains of several
      //    cells are in flight per thread (16 warps per SM do not hide them otherwise).
      const bool full = (cx == TX) && (cy == TY);
      auto ab_cell = [&](int e) {
        const int row = e / EC, col = e - row * EC - 1;  // col 0 = left ring column of the tile
        if (col < 0 || col >= cy + 2) return;
        const bool row_in = row >= 1 && row <= cx, col_in = col >= 1 && col <= cy;
        if (!(row_in || col_in)) return;
        const int li = li0 - 1 + row, lj = lj0 - 1 + col;
        double z;
        if constexpr (CC) {
          const unsigned int kd = (s_cls[row * kClsBoxBytes + 3 + col] >> 4) & 3u;
          if (kd == 2u)
            z = div_var<DIV>(s_r[e], __ldg(&c.D[(size_t)li * S_pitch + (lj + kColOff)]));
          else
            z = div_known<DIV>(s_r[e], kd ? d_out : d_in, kd ? y_out : y_in);
        } else {
          z = div_var<DIV>(s_r[e], s_D[e]);
        }
        const double v = fma(beta, s_p[e], z);
        s_p[e] = v;
        const bool ring = (li == 0) || (li == c.nx + 1) || (lj == 0) || (lj == c.ny + 1);
        if ((row_in && col_in) || ring) pout[(size_t)li * S_pitch + (lj + kColOff)] = v;
      };
      if (full) {
        constexpr int kCells = (TX + 2) * EC;
#pragma unroll
        for (int k = 0; k < (kCells + kThreads - 1) / kThreads; ++k) {
          const int e = threadIdx.x + k * kThreads;
          if (e < kCells) ab_cell(e);
        }
      } else {
        for (int e = threadIdx.x; e < (cx + 2) * EC; e += kThreads) ab_cell(e);
      }
      __syncthreads();
      // 2. Ap = A p_new on the owned nodes of the tile
      {
        const int col = threadIdx.x % TY + 1;
        auto ab_node = [&](int row, double& pc) {
          const double* s = s_p + row * EC + col + 1;
          pc = s[0];
          const size_t g = (size_t)(li0 - 1 + row) * S_pitch + (lj0 - 1 + col + kColOff);
          double a_c, a_ip, b_c, b_jp;
          if constexpr (CC) {
            const unsigned char* q = s_cls + row * kClsBoxBytes + 3 + col;
            const unsigned int k_c = q[0], k_ip = q[kClsBoxBytes], k_jp = q[1];
            a_c = face_value(k_c & 3u, c.a + g), a_ip = face_value(k_ip & 3u, c.a + g + S_pitch);
            b_c = face_value((k_c >> 2) & 3u, c.b + g), b_jp = face_value((k_jp >> 2) & 3u, c.b + g + 1);
          } else {
            a_c = s_a[(row - 1) * TY + (col - 1)], a_ip = s_a[row * TY + (col - 1)];
            b_c = s_b[(row - 1) * BC + (col - 1)], b_jp = s_b[(row - 1) * BC + col];
          }
          const double v = stencil_value<DIV>(K, pc, s[-EC], s[EC], s[-1], s[1], a_c, a_ip, b_c, b_jp);
          c.Ap[g] = v;
          return v;
        };
        if (full) {
          constexpr int kPer = TX / (kThreads / TY);  // nodes per thread
          double v[kPer], pc[kPer];
#pragma unroll
          for (int k = 0; k < kPer; ++k) v[k] = ab_node(threadIdx.x / TY + 1 + k * (kThreads / TY), pc[k]);
#pragma unroll
          for (int k = 0; k < kPer; ++k) acc_ab = fma(v[k], pc[k], acc_ab);  // same order as the generic loop
        } else if (col <= cy) {
          for (int row = threadIdx.x / TY + 1; row <= cx; row += kThreads / TY) {
            double pc;
            const double v = ab_node(row, pc);
            acc_ab = fma(v, pc, acc_ab);
          }
        }
      }
      __syncthreads();  // stage free
      ++n_consumed;
    }
    if (!tma_ok) break;
    // Prefetch for phase C: it reads only what this CTA itself wrote (p_new, Ap) or owns (w, r, D), so its
    // first tiles can be in flight while the reduction below waits for the slowest CTA / GPU.
    if (threadIdx.x == 0) fence_proxy_async();  // our generic-proxy global stores -> TMA reads
    for (int j = 0; j < npre; ++j) issue_c(j, cur ^ 1);
    PB_STAMP(t_ab);
    {
      double v[1] = {acc_ab};
      ++seq;
      arrive_target += ncta;
      const RedResult rr = cross_rank_sum<1>(c, v, seq, arrive_target, cta, ncta, s_red, false, false);
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
    int pushed[4] = {0, 0, 0, 0};  // tiles of this CTA whose edge values went into each neighbour's ring
    for (int j = 0; j < nmine && tma_ok; ++j) {
      if (j + S - 1 < nmine) issue_c(j + S - 1, cur ^ 1);
      unsigned char* sb = wait_stage();
      if (!tma_ok) break;
      const double* s_w = reinterpret_cast<const double*>(sb);
      const double* s_p = reinterpret_cast<const double*>(sb + Cfg::kTileBytes);
      const double* s_r = reinterpret_cast<const double*>(sb + 2 * Cfg::kTileBytes);
      const double* s_Ap = reinterpret_cast<const double*>(sb + 3 * Cfg::kTileBytes);
      const double* s_D = reinterpret_cast<const double*>(sb + Cfg::kOffCD);
      const unsigned char* s_cls = sb + Cfg::kOffCD;  // CC: byte of node (row, col) at row * 80 + 4 + col
      int li0, lj0;
      tile_origin(j, li0, lj0);
      const int cx = min(TX, c.nx - li0 + 1), cy = min(TY, c.ny - lj0 + 1);
      const int col = threadIdx.x % TY;
      const bool full = (cx == TX) && (cy == TY);
      // loads and arithmetic of one node; stores and halo stores included
      auto c_node = [&](int row, double& d_out, double& z_out, double& r_out) {
        const int e = row * TY + col;
        const int li = li0 + row, lj = lj0 + col;
        const size_t g = (size_t)li * S_pitch + (lj + kColOff);
        const double w_old = s_w[e];
        const double w_new = fma(alpha, s_p[e], w_old);
        const double r_new = fma(-alpha, s_Ap[e], s_r[e]);
        if (l2_hints)
          __stcs(&c.w[g], w_new);  // streaming store: w is not read again before the next iteration's phase C
        else
          c.w[g] = w_new;
        c.r[g] = r_new;
        d_out = __dsub_rn(w_new, w_old);
        if constexpr (CC) {
          const unsigned int kd = (s_cls[row * kClsBoxBytes + 4 + col] >> 4) & 3u;
          z_out = kd == 2u ? div_var<DIV>(r_new, __ldg(&c.D[g])) : div_known<DIV>(r_new, kd ? d_out : d_in, kd ? y_out : y_in);
        } else {
          z_out = div_var<DIV>(r_new, s_D[e]);
        }
        r_out = r_new;
        if (li == 1 && c.nbr_r[kXm])
          c.nbr_r[kXm][(size_t)(c.nbr_nx[kXm] + 1) * c.nbr_pitch[kXm] + (lj + kColOff)] = r_new;
        if (li == c.nx && c.nbr_r[kXp]) c.nbr_r[kXp][(size_t)(lj + kColOff)] = r_new;
        if (lj == 1 && c.nbr_r[kYm])
          c.nbr_r[kYm][(size_t)li * c.nbr_pitch[kYm] + (c.nbr_ny[kYm] + 1 + kColOff)] = r_new;
        if (lj == c.ny && c.nbr_r[kYp]) c.nbr_r[kYp][(size_t)li * c.nbr_pitch[kYp] + kColOff] = r_new;
      };
      if (full) {
        constexpr int kPer = TX / (kThreads / TY);
        double d[kPer], z[kPer], rn[kPer];
#pragma unroll
        for (int k = 0; k < kPer; ++k) c_node(threadIdx.x / TY + k * (kThreads / TY), d[k], z[k], rn[k]);
#pragma unroll
        for (int k = 0; k < kPer; ++k) {
          acc_d = fma(d[k], d[k], acc_d);
          acc_zr = fma(z[k], rn[k], acc_zr);
        }
      } else if (col < cy) {
        for (int row = threadIdx.x / TY; row < cx; row += kThreads / TY) {
          double d, z, rn;
          c_node(row, d, z, rn);
          acc_d = fma(d, d, acc_d);
          acc_zr = fma(z, rn, acc_zr);
        }
      }
      if (has_nbr && j <= last_halo_j) {
        pushed[kXm] += (li0 == 1), pushed[kXp] += (li0 + cx - 1 == c.nx);
        pushed[kYm] += (lj0 == 1), pushed[kYp] += (lj0 + cy - 1 == c.ny);
      }
      __syncthreads();  // stage free; all halo stores of this tile are issued
      // Release the halo stores of this CTA's edge tiles system-wide and count them in at the neighbour --
      // one tile after the last edge tile (or at the end of the phase), when the stores have long been
      // acknowledged and the fence costs little; the whole CTA would otherwise wait for it at the next barrier.
      if (has_nbr && last_halo_j >= 0 && (j == last_halo_j + 1 || (j == nmine - 1 && last_halo_j == nmine - 1)) &&
          threadIdx.x == 32) {
        fence_acq_rel_sys();
#pragma unroll
        for (int d = 0; d < 4; ++d)
          if (pushed[d] && c.nbr_r[d]) red_add_relaxed_sys_u64(c.nbr_halo_cnt[d], (unsigned long long)pushed[d]);
      }
      ++n_consumed;
    }
    if (!tma_ok) break;
    // Prefetch for the next phase AB, part 1: D, a, b are constant and p_new (buffer cur^1, ring included)
    // was completed before the previous reduction.  The r box waits for the reduction below, because its
    // ring cells are written by other CTAs / GPUs during this very phase.
    const unsigned int pre_base = n_issued;
    if (threadIdx.x == 0) fence_proxy_async();
    for (int j = 0; j < npre; ++j) {
      issue_ab_part(j, n_issued % S, cur ^ 1, true, false, k);
      ++n_issued;
    }
    PB_STAMP(t_c);
    double zr_new;
    bool stop = false;
    {
      double v[2] = {acc_d, acc_zr};
      ++seq;
      arrive_target += ncta;
      const RedResult rr = cross_rank_sum<2>(c, v, seq, arrive_target, cta, ncta, s_red, false, false);
      PB_STAMP(t_r2);
      if (!rr.ok) {
        status = PB_STATUS_WATCHDOG;
        stop = true;
      }
      diff_norm = norm == PB_NORM_WEIGHTED ? sqrt(rr.v[0] * h1h2) : sqrt(rr.v[0]);
      zr_new = rr.v[1] * h1h2;
    }
    // part 2 of the prefetch: everybody's r (ours, our neighbours' ring values) is final and acquired
    if (threadIdx.x == 0) fence_proxy_async();
    for (int j = 0; j < npre; ++j) issue_ab_part(j, (pre_base + j) % S, cur ^ 1, false, true, k);
    cur ^= 1;
    if (stop) break;
    if (diff_norm < delta) {
      status = PB_STATUS_CONVERGED;
      break;
    }
    beta = zr_new / zr_old;
    zr_old = zr_new;
  }
  if (k > max_iter) k = max_iter;
  drain();

  if (cta == 0 && threadIdx.x == 0 && tma_ok) {
    if (ctl->red[2] != 0.0) status = PB_STATUS_WATCHDOG;
    ctl->iter = (int)k;
    ctl->status = status;
    ctl->diff_norm = diff_norm;
    ctl->alpha = alpha;
    ctl->beta = beta;
    ctl->denom = denom;
    ctl->zr_old = zr_old;
    ctl->cur = cur;
    ctl->red_seq = seq;
    ctl->t_ab = t_ab;
    ctl->t_c = t_c;
    ctl->t_red1 = t_r1;
    ctl->t_red2 = t_r2;
    ctl->t_init = t_init;
    ctl->t_total = globaltimer_ns() - t_start;
  }
#undef PB_STAMP
}

// ---------------------------------------------------------------------------------------------------
template <int TX, int DIV, bool CC>
static cudaError_t launch_variant(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int variant, int l2_hints,
                                  cudaStream_t stream) {
  const void* fn = (const void*)pcg_persistent_tma_kernel<TX, DIV, CC>;
  const size_t smem = TmaCfg<TX, CC>::kSmemBytes;
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  void* args[] = {(void*)&d_ctxs, (void*)&ctas_per_rank, (void*)&variant, (void*)&l2_hints};
  return cudaLaunchCooperativeKernel(fn, dim3(nvirt * ctas_per_rank), dim3(kThreads), args, smem, stream);
}

template <int TX, bool CC>
static cudaError_t launch_div(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int variant, int div, int l2_hints,
                              cudaStream_t stream) {
  if (div == kDivTrue) return launch_variant<TX, kDivTrue, CC>(d_ctxs, nvirt, ctas_per_rank, variant, l2_hints, stream);
  if (div == kDivMarkstein) return launch_variant<TX, kDivMarkstein, CC>(d_ctxs, nvirt, ctas_per_rank, variant, l2_hints, stream);
  return launch_variant<TX, kDivRecip, CC>(d_ctxs, nvirt, ctas_per_rank, variant, l2_hints, stream);
}

cudaError_t launch_persistent_tma(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int tile_rows, int div,
                                  int l2_hints, bool compressed, cudaStream_t stream) {
  if (tile_rows == 32)
    return compressed ? launch_div<32, true>(d_ctxs, nvirt, ctas_per_rank, 0, div, l2_hints, stream)
                      : launch_div<32, false>(d_ctxs, nvirt, ctas_per_rank, 0, div, l2_hints, stream);
  return compressed ? launch_div<16, true>(d_ctxs, nvirt, ctas_per_rank, 1, div, l2_hints, stream)
                    : launch_div<16, false>(d_ctxs, nvirt, ctas_per_rank, 1, div, l2_hints, stream);
}

// Debug / unit-test probe: one box load through a tensor map at (x, y) into shared memory, copied out to
// `out` (nbytes), out_status = 1 on completion, 2 on timeout.  mode 0: map in global memory; mode 1: same,
// after a tensormap-proxy acquire fence; mode 2: map passed as a __grid_constant__ kernel parameter.
template <int MODE>
__global__ void tma_probe_kernel(const CUtensorMap* gmap, const __grid_constant__ CUtensorMap pmap, int x, int y,
                                 unsigned int nbytes, double* out, int* out_status) {
  extern __shared__ __align__(128) unsigned char s_dyn[];
  __shared__ __align__(8) uint64_t bar;
  unsigned char* buf = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(s_dyn) + 127) & ~uintptr_t(127));
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const CUtensorMap* map = MODE == 2 ? &pmap : gmap;
    if (MODE == 1)
      asm volatile("fence.proxy.tensormap::generic.acquire.gpu [%0], 128;" ::"l"(reinterpret_cast<uint64_t>(gmap)) : "memory");
    mbar_arrive_expect_tx(&bar, nbytes);
    tma_load_2d(buf, map, x, y, &bar);
  }
  const bool ok = mbar_wait(&bar, 0);
  if (ok)
    for (unsigned int i = threadIdx.x; i < nbytes / 8; i += blockDim.x) out[i] = reinterpret_cast<double*>(buf)[i];
  if (threadIdx.x == 0) *out_status = ok ? 1 : 2;
}

cudaError_t launch_tma_probe(const void* dmaps, const void* hmaps, int idx, int mode, int x, int y, unsigned int nbytes,
                             double* out, int* out_status, cudaStream_t stream) {
  const CUtensorMap* g = static_cast<const CUtensorMap*>(dmaps) + idx;
  const CUtensorMap p = static_cast<const CUtensorMap*>(hmaps)[idx];
  const size_t smem = nbytes + 256;
  cudaError_t e = cudaSuccess;
  if (mode == 0) {
    e = cudaFuncSetAttribute(tma_probe_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) tma_probe_kernel<0><<<1, 128, smem, stream>>>(g, p, x, y, nbytes, out, out_status);
  } else if (mode == 1) {
    e = cudaFuncSetAttribute(tma_probe_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) tma_probe_kernel<1><<<1, 128, smem, stream>>>(g, p, x, y, nbytes, out, out_status);
  } else {
    e = cudaFuncSetAttribute(tma_probe_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) tma_probe_kernel<2><<<1, 128, smem, stream>>>(g, p, x, y, nbytes, out, out_status);
  }
  return e != cudaSuccess ? e : cudaGetLastError();
}

// Host side: the twelve tensor maps of one rank for one tile height.  The driver entry point is fetched
// through the runtime so that the library has no link-time dependency on libcuda.
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

static bool encode_u8(CUtensorMap* m, unsigned char* base, int rows, int pitch, int box_cols, int box_rows) {
  if (!encode_fn()) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static bool encode(CUtensorMap* m, double* base, int rows, int pitch, int box_cols, int box_rows) {
  if (!encode_fn()) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch * sizeof(double)};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, base, dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool build_tensor_maps(const RankCtx& h, int tile_rows, void* out12) {
  CUtensorMap* m = static_cast<CUtensorMap*>(out12);
  const int rows = h.nx + 2, P = h.pitch, TX = tile_rows, TY = kTileY;
  bool ok = true;
  ok &= encode(m + kMapRExt, h.r, rows, P, TY + 4, TX + 2);
  ok &= encode(m + kMapDExt, h.D, rows, P, TY + 4, TX + 2);
  ok &= encode(m + kMapP0Ext, h.p[0], rows, P, TY + 4, TX + 2);
  ok &= encode(m + kMapP1Ext, h.p[1], rows, P, TY + 4, TX + 2);
  ok &= encode(m + kMapA, h.a, rows, P, TY, TX + 1);
  ok &= encode(m + kMapB, h.b, rows, P, TY + 2, TX);
  ok &= encode(m + kMapW, h.w, rows, P, TY, TX);
  ok &= encode(m + kMapP0, h.p[0], rows, P, TY, TX);
  ok &= encode(m + kMapP1, h.p[1], rows, P, TY, TX);
  ok &= encode(m + kMapR, h.r, rows, P, TY, TX);
  ok &= encode(m + kMapAp, h.Ap, rows, P, TY, TX);
  ok &= encode(m + kMapD, h.D, rows, P, TY, TX);
  ok &= encode_u8(m + kMapClsAB, h.cls, rows, P, kClsBoxBytes, TX + 2);
  ok &= encode_u8(m + kMapClsC, h.cls, rows, P, kClsBoxBytes, TX);
  return ok;
}

size_t tensor_map_bytes() { return sizeof(CUtensorMap); }

}  // namespace pb
