// persistent_tma.cu -- KP-TMA: the persistent cooperative PCG loop with TMA-fed, multi-stage tile pipelines.
//
// Same algorithm and cross-rank protocol as persistent.cu; what changes is how a CTA moves data.  Each
// phase is a software pipeline over the CTA's tiles: one elected thread issues `cp.async.bulk.tensor.2d`
// loads for tile j+S-1 (all operand boxes of the tile land in one shared-memory stage and complete on one
// mbarrier) while all warps compute tile j out of shared memory.  A stage holds
//     phase AB: r, p_old boxes of (TX+2) x (TY+4) (tile + ring; one more column on each side because a box
//               must start on a 16-byte boundary) and, for tiles with non-constant coefficients, the D box
//               of the same shape, the a box (TX+1) x TY and the b box TX x (TY+2)
//     phase C : w, p, r, Ap boxes of TX x TY (+ the D box for tiles with non-constant coefficients); with
//               `recompute_ap`, full constant-coefficient tiles load w, r and the (TX+2) x (TY+4) box of p_new
//               instead and evaluate the stencil a second time, so Ap is neither stored nor reloaded for them
// so one SM keeps S-1 complete tiles (up to 88 KB each) in flight without spending registers on it, which is
// what a bandwidth-bound fp64 stencil needs on a part with ~40 KB/SM of latency-bandwidth product.
// The TMA unit also does the address generation for the misaligned ring boxes and zero-fills the part of
// a box that lies outside the array (partial tiles at the block edge).
//
// Coefficient classes (CC = true, the default): a and b take the value 1 on faces inside the ellipse, 1/eps
// outside, and something else only on the O(M+N) faces the ellipse cuts; D follows.  The assembly kernel
// stores one class byte per cell; at launch every CTA classifies its tiles once.  A tile whose cells (ring
// included) are all "inside" or all "outside" needs no coefficient boxes at all -- a, b, D are CTA-uniform
// constants -- which removes 24 of the 112 bytes per node and iteration for all but the few tiles crossed by
// the ellipse or touching the boundary of Omega; those keep the full fp64 boxes.  Values are selected, never
// recomputed, so the results are bit-identical to streaming the coefficient arrays (tests compare them).
//
// Work items and who computes what: the block is cut into super-tiles of 64 x 64 nodes (tall mode) whose two 32-row
// halves travel as ONE item when both are full and share a constant coefficient class; a greedy pass at launch gives
// the items to the CTAs by estimated cost (same result on every CTA).  Phase AB of a full constant-coefficient item
// runs by warp-private rows (run_rows: p_new in registers, neighbours by shuffle, no barrier between the p update
// and the stencil); mixed and partial tiles take the tile path (run_tile: p update in place in shared memory,
// barrier, stencil).  Both evaluate the same expressions on the same operands.
//
// Replaces the same reference pieces as persistent.cu (S4:846-942, kernels S4:507-676).
#include <type_traits>

#include "launch.h"
#include "phases.cuh"
#include "tma_utils.cuh"

namespace pb {

enum TmaMap : int { kMapRExt = 0, kMapDExt, kMapP0Ext, kMapP1Ext, kMapA, kMapB, kMapW, kMapP0, kMapP1, kMapR, kMapAp, kMapD };
constexpr int kMaxItems = 768;  // work items per CTA (the host refuses blocks that would need more)

// One unit of work of a CTA in both phases: a tile of TX rows, or (tall) two vertically adjacent full tiles with the
// same constant coefficients that travel and are processed as one 2 TX x TY tile.  A tall item halves the per-tile
// fixed cost of a phase (issue path, barrier waits, mbarrier wait: about 1000 of the 3300 cycles a 32 x 64 tile of
// phase AB takes) for the constant-coefficient tiles that make up all but a few per cent of a large block.
struct TileItem {
  unsigned short li0, lj0;  // first owned node
  unsigned char flags;      // bits 0-1 coefficient class (0 inside, 1 outside, 2 mixed), bit 2 Ap recomputed, bit 3 tall
  unsigned char pad;
};
constexpr int kItemRecomp = 4, kItemTall = 8;
constexpr int kTallVariant = 2;  // tensor-map set with boxes of 2 * 32 rows
constexpr int kCostTall = 59, kCostHalf = 38;  // balancer: a tall item, a full 32-row item with recomputed Ap (0.1 us)

template <int TX>
struct TmaCfg {
  static constexpr int TY = kTileY;
  static constexpr int ER = TX + 2, EC = TY + 4;  // ring boxes start one column early: TMA wants 16-byte starts
  static constexpr int BC = TY + 2;               // b box width
  static constexpr int kExtBytes = ((ER * EC * 8 + 127) / 128) * 128;
  static constexpr int kABytes = (TX + 1) * TY * 8;
  static constexpr int kBBytes = TX * BC * 8;
  static constexpr int kTileBytes = TX * TY * 8;
  // phase AB: r | p | D | a | b      phase C: w | p | r | Ap | D
  static constexpr int kOffR = 0, kOffP = kExtBytes, kOffD = 2 * kExtBytes, kOffA = 3 * kExtBytes;
  static constexpr int kOffB = 3 * kExtBytes + kABytes;
  static constexpr int kOffCD = 4 * kTileBytes;
  static constexpr int kOffCPe = 3 * kTileBytes;  // phase C with recomputed Ap: w | - | r | p_new (tile + ring)
  static constexpr int kStageAB = 3 * kExtBytes + kABytes + kBBytes;
  static constexpr int kStageC = 5 * kTileBytes;
  static_assert(kOffCPe + kExtBytes <= kStageC, "recompute layout must fit into a phase-C stage");
  // tall items (TX = 32 only; always constant coefficients and recomputed Ap): phase AB r | p, phase C w | r | p_new,
  // boxes of 2 TX (+ 2) rows
  static constexpr bool kTall = TX == 32;
  static constexpr int ER2 = 2 * TX + 2;
  static constexpr int kExtBytes2 = ((ER2 * EC * 8 + 127) / 128) * 128;
  static constexpr int kTileBytes2 = 2 * kTileBytes;
  static constexpr int kStageTall = kTall ? 2 * kTileBytes2 + kExtBytes2 : 0;  // phase C is the larger one
  static_assert(2 * kExtBytes2 <= 2 * kTileBytes2 + kExtBytes2, "tall phase-AB stage must fit");
  static constexpr uint32_t kTxTallAB = 2u * ER2 * EC * 8u;
  static constexpr uint32_t kTxTallC = 2u * (uint32_t)kTileBytes2 + 1u * ER2 * EC * 8u;
  static constexpr int kStage32 = kStageAB > kStageC ? kStageAB : kStageC;
  static constexpr int kStageBytes = kStage32 > kStageTall ? kStage32 : kStageTall;
  static constexpr int kStages = TX >= 32 ? 2 : 4;
  // transaction bytes = full boxes, out-of-bounds parts included
  static constexpr uint32_t kTxABCore = 2u * ER * EC * 8u;                                       // r, p
  static constexpr uint32_t kTxABCoef = 1u * ER * EC * 8u + (uint32_t)kABytes + (uint32_t)kBBytes;  // D, a, b
  static constexpr uint32_t kTxCCore = 4u * (uint32_t)kTileBytes;                                // w, p, r, Ap
  static constexpr uint32_t kTxCCoef = (uint32_t)kTileBytes;                                     // D
  static constexpr uint32_t kTxCRecomp = 2u * (uint32_t)kTileBytes + 1u * ER * EC * 8u;          // w, r, p_new + ring
  static constexpr size_t kSmemBytes = (size_t)kStages * kStageBytes + 128;
  static_assert(kSmemBytes <= 227 * 1024, "stages do not fit into shared memory");
};

template <int TX, int DIV, bool CC>
__global__ void __launch_bounds__(kThreads, 1)
pcg_persistent_tma_kernel(const RankCtx* __restrict__ ctxs, int ctas_per_rank, int map_variant, int l2_hints,
                          int recompute_ap, int tall_mode, int balance_mode, int pair_mode) {
  using Cfg = TmaCfg<TX>;
  constexpr int TY = Cfg::TY, EC = Cfg::EC, BC = Cfg::BC, S = Cfg::kStages;
  __shared__ RankCtx c;
  __shared__ double s_red[80];
  __shared__ __align__(8) uint64_t s_bar[S];
  __shared__ TileItem s_items[kMaxItems];    // work items of phase AB
  __shared__ TileItem s_items_c[kMaxItems];  // work items of phase C: the same nodes; a tall item whose Ap is stored, not
                                             // recomputed, appears as its two 32-row halves here (a tall phase-C stage
                                             // has no room for p and Ap boxes)
  __shared__ int s_filled[S];  // per stage: ring cells were unpacked for the tile in flight (thread 0 only)
  __shared__ volatile int s_abort;  // some thread's TMA wait timed out; read after a barrier, so CTA-uniform
  __shared__ unsigned long long s_wd_limit;  // device-side wait limit in ns (kept out of the registers)
  __shared__ unsigned long long s_t_halo;    // thread 0: ns spent unpacking (= waiting for) the neighbours' halo records
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
    for (int s = 0; s < S; ++s) {
      mbar_init(&s_bar[s], 1);
      s_filled[s] = 0;
    }
    s_abort = 0;
    s_t_halo = 0;
    const unsigned long long wd = ctxs[vrank].ctl->watchdog_ns;
    s_wd_limit = wd ? wd : kDefaultWatchdogNs;
    mbar_fence_init();
  }
  __syncthreads();
  star_pair_init(pair_mode & 1);  // clusters of two CTAs: the partner's mbarrier exists from here on
  // s_dyn is declared __align__(128); deriving every stage pointer from it by plain pointer arithmetic keeps
  // the accesses in the shared address space (LDS/STS instead of generic LD/ST)
  unsigned char* const stage_base = s_dyn;
  const CUtensorMap* maps = reinterpret_cast<const CUtensorMap*>(c.tmaps) + map_variant * kMapsPerVariant;

  Control* ctl = c.ctl;
  const CUtensorMap* maps_tall = reinterpret_cast<const CUtensorMap*>(c.tmaps) + kTallVariant * kMapsPerVariant;
  // The block is cut into "super-tiles" of srows x TY nodes, srows = 2 TX when tall items are allowed, else TX;
  // super-tile t belongs to CTA t mod ncta in both phases.
  const bool tall_ok = Cfg::kTall && (tall_mode & 1) != 0;
  const bool rows_ab = TX == 32 && (tall_mode & 2) != 0;  // phase AB of full constant-coefficient items by warp-private rows
  const int srows = tall_ok ? 2 * TX : TX;
  const int tiles_y = (c.ny + TY - 1) / TY, tiles_x = (c.nx + srows - 1) / srows;
  const int ntiles = tiles_x * tiles_y;
  const int nsuper = cta < ntiles ? (ntiles - cta + ncta - 1) / ncta : 0;
  const bool has_nbr = c.nbr_r[0] || c.nbr_r[1] || c.nbr_r[2] || c.nbr_r[3];
  const bool timing = threadIdx.x == 0;  // every CTA keeps its own phase clocks; CTA 0's go into the report
  const int S_pitch = c.pitch;

  unsigned int seq = ctl->red_seq;
  // One rank: the record star (3.3-3.6 us against 4.2-4.4 us).  Several ranks: the counter + elected aggregator of
  // cross_rank_sum, which measured faster there (2 GPUs: 5.3 against 5.6 us, 8 GPUs: 6.4 against 9.6 us): the CTAs
  // that wait for the other ranks' records poll their slot lines with nranks lanes each, and CTA 0's gather loop of the
  // star has to share the L2 with that traffic, while an elected last arriver reads the partials just once.
  unsigned int arrive_target = 0;
  const bool multi_rank = c.nranks > 1 && !(pair_mode & 2);  // (bit 1 of pair_mode: the star for several ranks too, A/B switch)
#define PB_RED_PRE arrive_target += ncta;
#define PB_RED(nv)                                                                                      \
  (multi_rank ? cross_rank_sum<nv>(c, v, seq, arrive_target, cta, ncta, s_red, false, false)            \
              : star_sum<nv>(c, v, seq, cta, ncta, s_wd_limit, s_red, nullptr, pair_mode & 1))
  const double delta = ctl->delta;
  const long long max_iter = ctl->max_iter;
  const int norm = ctl->norm;
  const double h1h2 = c.g.h1h2;
  const StencilConsts K = make_stencil_consts(c.g);
  // values behind the coefficient classes (same device functions as the assembly kernel -> same bits)
  const double inv_eps = 1.0 / c.g.eps;
  const double diag_in = diag_from_faces(c.g, 1.0, 1.0, 1.0, 1.0);
  const double diag_out = diag_from_faces(c.g, inv_eps, inv_eps, inv_eps, inv_eps);
  const double rcp_in = __drcp_rn(diag_in), rcp_out = __drcp_rn(diag_out);

  // L2 policy per operand: what both phases of every iteration re-read (r, D, p) is kept, what is touched once
  // per iteration (a, b, w, Ap after its single use) is marked evict-first (optional).
  const uint64_t pol_keep = l2_hints ? kL2EvictLast : kL2EvictNormal;
  const uint64_t pol_stream = l2_hints ? kL2EvictFirst : kL2EvictNormal;

  unsigned int n_issued = 0, n_consumed = 0;  // uniform across the CTA

  // ---- work items of this CTA (one pass over the class bytes per launch) ----
  // Coefficient classes: 0 = every cell of the tile and its ring is "inside" (a = b = 1, D = diag_in), 1 = all
  // "outside" (1/eps, diag_out), 2 = mixed (full coefficient boxes).  Full tiles with constant coefficients can get Ap
  // in phase C by evaluating the stencil again on the p_new box (the same device function on the same operands: the
  // same bits), which saves the store and the reload of Ap.
  // Which CTA works on which super-tile.  Cyclic (t mod ncta) when every tile costs the same; but tall items are
  // cheaper per node than 32-row items and mixed tiles dearer, and the CTAs that get the ellipse's boundary end up
  // 10 % behind the others, who then wait for them in both reductions (measured at 4096^2: 27 us of a 215 us
  // iteration).  So: every CTA classifies its cyclic share and publishes one code byte per super-tile, one grid-wide
  // barrier, then warp 0 of every CTA walks the super-tiles in sweep order and gives each to the CTA with the least
  // cost so far (ties: lowest CTA).  Same integers everywhere, so every CTA arrives at the same assignment; with equal
  // costs it is the cyclic one; the sweep order -- and with it the "edge rows first" order of the halo -- is kept.
  int nmine = 0;  // items; uniform across the CTA
  {
    __shared__ int s_nmine;
    auto classify = [&](int li0, int lj0, int cx, int cy) -> int {
      if constexpr (!CC) return 2;
      bool all_in = true, all_out = true;
      for (int e = threadIdx.x; e < (cx + 2) * (cy + 2); e += kThreads) {
        const int row = e / (cy + 2), col = e - row * (cy + 2);
        if ((row == 0 || row == cx + 1) && (col == 0 || col == cy + 1)) continue;
        const unsigned int b = c.cls[(size_t)(li0 - 1 + row) * S_pitch + (lj0 - 1 + col + kColOff)];
        all_in = all_in && (b == 0x00u);
        all_out = all_out && (b == 0x15u);
      }
      const int in = __syncthreads_and(all_in ? 1 : 0), out = __syncthreads_and(all_out ? 1 : 0);
      return in ? 0 : (out ? 1 : 2);
    };
    // first node of the super-tile in row tx, column ty of the sweep (sweep index t = tx * tiles_y + ty)
    auto origin_at = [&](int tx, int ty, int& li0, int& lj0) {
      // Row order: first, last, second, ...: the tiles that store halo rows for the x-neighbours (phase C) and the
      // tiles that need the neighbours' rows (phase AB) then come first on both sides, so the NVLink latency of the
      // halo records hides behind the remaining tiles instead of sitting in front of a reduction.  (Enumerating the
      // whole perimeter first was measured 7 % slower per phase: it breaks the streaming order of the sweep.)
      if (has_nbr && tiles_x > 2) tx = tx == 0 ? 0 : (tx == 1 ? tiles_x - 1 : tx - 1);
      li0 = 1 + tx * srows, lj0 = 1 + ty * TY;
    };
    auto origin = [&](int t, int& li0, int& lj0) {
      const int tx = t / tiles_y;
      origin_at(tx, t - tx * tiles_y, li0, lj0);
    };
    // code of a super-tile: bits 0-1 class of the first half, 2-3 of the second, 4 second half exists, 5 tall,
    // 6 / 7 first / second half is a full tile
    auto super_code = [&](int li0, int lj0) -> int {
      const int cy = min(TY, c.ny - lj0 + 1);
      const int cx0 = min(TX, c.nx - li0 + 1);
      const int m0 = classify(li0, lj0, cx0, cy);
      const bool full0 = cx0 == TX && cy == TY;
      const bool second = srows > TX && li0 + TX <= c.nx;
      int m1 = 2, cx1 = 0;
      if (second) {
        cx1 = min(TX, c.nx - (li0 + TX) + 1);
        m1 = classify(li0 + TX, lj0, cx1, cy);
      }
      const bool full1 = second && cx1 == TX && cy == TY;
      const bool tall = second && full0 && full1 && m0 != 2 && m0 == m1;
      return m0 | (m1 << 2) | (second ? 16 : 0) | (tall ? 32 : 0) | (full0 ? 64 : 0) | (full1 ? 128 : 0);
    };
    // items of a super-tile, appended by `writer` (one thread); returns how many
    // `part`: 0 the whole super-tile, 1 / 2 only its first / second half (a tall super-tile cut in two by the balancer)
    auto emit = [&](int li0, int lj0, int code, int at, bool writer, int part) -> int {
      const int m0 = code & 3, m1 = (code >> 2) & 3;
      auto put = [&](int k, int li, int flags) {
        if (writer && k < kMaxItems) {
          s_items[k].li0 = (unsigned short)li, s_items[k].lj0 = (unsigned short)lj0;
          s_items[k].flags = (unsigned char)flags, s_items[k].pad = 0;
        }
      };
      if (part != 0) {  // halves of a tall super-tile: full tiles, constant coefficients
        put(at, part == 1 ? li0 : li0 + TX, m0 | (recompute_ap ? kItemRecomp : 0));
        return 1;
      }
      if (code & 32) {
        put(at, li0, m0 | (recompute_ap ? kItemRecomp : 0) | kItemTall);
        return 1;
      }
      put(at, li0, m0 | ((recompute_ap && m0 != 2 && (code & 64)) ? kItemRecomp : 0));
      if (!(code & 16)) return 1;
      put(at + 1, li0 + TX, m1 | ((recompute_ap && m1 != 2 && (code & 128)) ? kItemRecomp : 0));
      return 2;
    };
    // relative cost of a super-tile, phases AB + C, in 0.1 us (least-squares fit of the per-CTA busy times at 4096^2,
    // 2048 x 4096 and 1024 x 4096 with phase AB by rows, bench/experiments/cta_balance.py: tall item 5.9 us, full 32-row
    // item with constant coefficients and recomputed Ap 3.8 us, with stored Ap or partial 4.5 us, mixed 5.3 us)
    auto half_cost = [&](int m, bool full_tile) -> int { return m == 2 ? 53 : ((recompute_ap && full_tile) ? kCostHalf : 45); };
    auto cost = [&](int code) -> int {
      if (code & 32) return kCostTall;
      return half_cost(code & 3, (code & 64) != 0) + ((code & 16) ? half_cost((code >> 2) & 3, (code & 128) != 0) : 0);
    };
    unsigned char* tcode = reinterpret_cast<unsigned char*>(ctl) + kTileCodeOffset;
    const bool balance = CC && balance_mode != 0 && ncta > 1 && ntiles <= kTileCodeBytes;
    for (int j = 0; j < nsuper; ++j) {
      const int t = cta + j * ncta;
      int li0, lj0;
      origin(t, li0, lj0);
      const int code = super_code(li0, lj0);
      if (balance) {
        if (threadIdx.x == 0) tcode[t] = (unsigned char)code;
      } else {
        nmine += emit(li0, lj0, code, nmine, threadIdx.x == 0, 0);
      }
    }
    if (balance) {
      {  // grid-wide barrier (a reduction of nothing): everybody's codes are written and released
        double v[1] = {0.0};
        ++seq;
        PB_RED_PRE
        (void)PB_RED(1);
      }
      if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        // Cost so far per CTA.  Up to 160 CTAs (one GPU: 148) the costs of CTAs lane, lane + 32, ... live in this lane's
        // registers and the minimum is one redux instruction; the walk over 4096 super-tiles is a serial chain, and
        // with the table in shared memory (the fallback) it took 2.5 ms at the start of every 4096^2 solve.
        const bool in_regs = ncta <= 160;
        int ld0 = 0, ld1 = 0, ld2 = 0, ld3 = 0, ld4 = 0;
        int* load = reinterpret_cast<int*>(s_dyn);  // the stage memory is free until the first load is issued
        if (!in_regs) {
          for (int i = lane; i < ncta; i += 32) load[i] = 0;
          __syncwarp();
        }
        // the CTA with the least cost so far (key = cost * 1024 + CTA: the minimum is unique) and that cost
        auto least_loaded = [&](int& cost_so_far) -> int {
          int best = 0x7fffffff;
          if (in_regs) {
            if (lane < ncta) best = ld0 * 1024 + lane;
            if (lane + 32 < ncta) best = min(best, ld1 * 1024 + lane + 32);
            if (lane + 64 < ncta) best = min(best, ld2 * 1024 + lane + 64);
            if (lane + 96 < ncta) best = min(best, ld3 * 1024 + lane + 96);
            if (lane + 128 < ncta) best = min(best, ld4 * 1024 + lane + 128);
            best = __reduce_min_sync(0xffffffffu, best);
          } else {
            for (int i = lane; i < ncta; i += 32) best = min(best, load[i] * 1024 + i);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
          }
          cost_so_far = best >> 10;
          return best & 1023;
        };
        auto charge = [&](int who, int add) {
          if (in_regs) {
            ld0 += who == lane ? add : 0;
            ld1 += who == lane + 32 ? add : 0;
            ld2 += who == lane + 64 ? add : 0;
            ld3 += who == lane + 96 ? add : 0;
            ld4 += who == lane + 128 ? add : 0;
          } else {
            if (lane == 0) load[who] += add;
            __syncwarp();
          }
        };
        // Even share of the total cost.  A tall item that would take its CTA more than a third of an item above that
        // share is cut into its two 32-row halves, which go to the least loaded CTAs separately: with few super-tiles
        // per CTA (the blocks of a multi-GPU run: 512 on 148 CTAs at 4096^2 on 8 GPUs) that is the difference between
        // "4 tall items against 3" and "3 tall items + one half everywhere".
        int total = 0, chunk = 0;  // chunk: codes of super-tiles t0 .. t0 + 31, one per lane
        for (int t0 = 0; t0 < ntiles; t0 += 32) total += t0 + lane < ntiles ? cost((int)__ldcg(tcode + t0 + lane)) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
        const int share = (total + ncta - 1) / ncta;
        int mine = 0;
        for (int t = 0, tx = 0, ty = 0; t < ntiles; ++t) {
          if ((t & 31) == 0) chunk = t + lane < ntiles ? (int)__ldcg(tcode + t + lane) : 0;
          const int code = __shfl_sync(0xffffffffu, chunk, t & 31);
          int li0, lj0, sofar;
          origin_at(tx, ty, li0, lj0);  // = origin(t), without the division
          if (++ty == tiles_y) ty = 0, ++tx;
          int who = least_loaded(sofar);
          if ((code & 32) && sofar + kCostTall > share + kCostTall / 3) {
            charge(who, recompute_ap ? kCostHalf : 45);
            if (who == cta) mine += emit(li0, lj0, code, mine, lane == 0, 1);
            who = least_loaded(sofar);
            charge(who, recompute_ap ? kCostHalf : 45);
            if (who == cta) mine += emit(li0, lj0, code, mine, lane == 0, 2);
          } else {
            charge(who, cost(code));
            if (who == cta) mine += emit(li0, lj0, code, mine, lane == 0, 0);
          }
        }
        if (lane == 0) s_nmine = mine;
      }
      __syncthreads();
      nmine = s_nmine;
    }
    nmine = min(nmine, kMaxItems);  // the host sizes the launch so that this never cuts anything off
    __syncthreads();
  }
  // phase-C list (every thread counts, thread 0 writes)
  int nmine_c = 0;
  for (int j = 0; j < nmine; ++j) {
    const TileItem it = s_items[j];
    const bool split = (it.flags & kItemTall) && !(it.flags & kItemRecomp);
    if (threadIdx.x == 0 && nmine_c + (split ? 2 : 1) <= kMaxItems) {
      s_items_c[nmine_c] = it;
      if (split) {
        s_items_c[nmine_c].flags = it.flags & 3;
        s_items_c[nmine_c + 1] = s_items_c[nmine_c];
        s_items_c[nmine_c + 1].li0 = (unsigned short)(it.li0 + TX);
      }
    }
    nmine_c += split ? 2 : 1;
  }
  nmine_c = min(nmine_c, kMaxItems);
  __syncthreads();
  const int npre = min(S - 1, nmine);      // items of phase AB that are issued ahead of the phase
  const int npre_c = min(S - 1, nmine_c);  // the same for phase C
  auto item_rows = [&](int flags) -> int { return (flags & kItemTall) ? 2 * TX : TX; };

  // Halo rows travel as LL records (payload32 | sequence32 word pairs, see ctx.h): the producer needs no fence and
  // no counter, the consumer recognises complete values by their sequence number.  Round k of this launch
  // carries sequence hseq + k; the number persists across launches like red_seq.
  const unsigned int hseq = ctl->halo_seq;
  // warp 0: copy the neighbours' r values of round `rounds` for the ring cells of the tile at (li0, lj0) from the
  // halo lines into our r ring; returns whether the tile touches a shared edge at all
  auto unpack_halo = [&](int li0, int lj0, int rows, long long rounds) -> bool {
    const int cx = min(rows, c.nx - li0 + 1), cy = min(TY, c.ny - lj0 + 1);
    const bool touch[4] = {li0 == 1, li0 + cx - 1 == c.nx, lj0 == 1, lj0 + cy - 1 == c.ny};
    const unsigned int want = hseq + (unsigned int)rounds;
    const unsigned long long limit = s_wd_limit;
    const int lane = threadIdx.x & 31;
    bool any = false;
#pragma unroll
    for (int d = 0; d < 4; ++d) any = any || (touch[d] && c.nbr_r[d]);
    if (!any) return false;
    const unsigned long long th0 = threadIdx.x == 0 ? globaltimer_ns() : 0ull;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      if (!touch[d] || !c.nbr_r[d]) continue;
      const unsigned long long* line = c.halo_ll + ((size_t)d * 2 + (want & 1u)) * (size_t)c.ll_cap * 2;
      const int t0 = d < 2 ? lj0 - 1 : li0 - 1, n = d < 2 ? cy : cx;
      for (int i = lane; i < n; i += 32) {
        const unsigned long long* src = line + (size_t)(t0 + i) * 2;
        unsigned long long w0, w1, tstart = 0;
        unsigned int spins = 0;
        while (true) {
          w0 = ld_relaxed_sys_u64(src), w1 = ld_relaxed_sys_u64(src + 1);
          if ((unsigned int)(w0 >> 32) == want && (unsigned int)(w1 >> 32) == want) break;
          if ((++spins & 0xffu) == 0u) {
            const unsigned long long now = globaltimer_ns();
            if (tstart == 0) tstart = now;
            if (now - tstart > limit) {
              ctl->red[2] = 1.0;  // sticky marker (zeroed by the host before the launch): halo wait timed out
              break;
            }
          }
        }
        const double v = __longlong_as_double((long long)((w0 & 0xffffffffull) | (w1 << 32)));
        const int li = d == kXm ? 0 : (d == kXp ? c.nx + 1 : t0 + i + 1);
        const int lj = d == kYm ? 0 : (d == kYp ? c.ny + 1 : t0 + i + 1);
        c.r[(size_t)li * S_pitch + (lj + kColOff)] = v;
      }
    }
    // every lane orders the ring cells it wrote (generic proxy) before the TMA read of the r box that lane 0 issues
    fence_proxy_async_global();
    __syncwarp();
    if (threadIdx.x == 0) s_t_halo += globaltimer_ns() - th0;
    return true;
  };
  // Phase-AB loads of tile j into stage `st`.  part 1 = everything that does not depend on the other CTAs'
  // phase C (p and, for mixed tiles, D, a, b; arms the barrier with the full byte count), part 2 = the r box.
  auto issue_ab_part = [&](int j, unsigned int st, int pbuf, bool part1, bool part2, long long rounds) {
    if (threadIdx.x < 32) {
      const TileItem it = s_items[j];
      const int li0 = it.li0, lj0 = it.lj0, mode = it.flags & 3;
      const bool tall = (it.flags & kItemTall) != 0;
      unsigned char* sb = stage_base + (size_t)st * Cfg::kStageBytes;
      uint64_t* bar = &s_bar[st];
      const int x = kColOff + lj0 - 2, y = li0 - 1;  // even column: 16-byte aligned box start
      const CUtensorMap* mp = tall ? maps_tall : maps;
      // A tile on a shared edge gets its ring cells filled from the halo lines by the whole warp.  This depends
      // only on the neighbour GPU, so it belongs to part 1 and its latency hides in the reduction that follows.
      const bool filled = part1 && has_nbr && rounds > 0 && unpack_halo(li0, lj0, item_rows(it.flags), rounds);
      if (part1 && threadIdx.x == 0) {
        if (filled) s_filled[st] = 1;
        const bool coef = mode == 2;
        fence_proxy_async_smem();
        mbar_arrive_expect_tx(bar, tall ? Cfg::kTxTallAB : Cfg::kTxABCore + (coef ? Cfg::kTxABCoef : 0u));
        tma_load_2d_hint(sb + (tall ? Cfg::kExtBytes2 : Cfg::kOffP), mp + (pbuf ? kMapP1Ext : kMapP0Ext), x, y, bar, pol_keep);
        if (coef) {
          tma_load_2d_hint(sb + Cfg::kOffD, maps + kMapDExt, x, y, bar, pol_keep);
          tma_load_2d_hint(sb + Cfg::kOffA, maps + kMapA, x + 2, y + 1, bar, pol_stream);
          tma_load_2d_hint(sb + Cfg::kOffB, maps + kMapB, x + 2, y + 1, bar, pol_stream);
        }
      }
      if (part2) {
        if (threadIdx.x == 0) {
          // the elected lane orders the ring cells written above (generic proxy) before the TMA read of the box
          if (s_filled[st]) {
            fence_proxy_async();
            s_filled[st] = 0;
          }
          tma_load_2d_hint(sb + Cfg::kOffR, mp + kMapRExt, x, y, bar, pol_keep);
        }
      }
    }
  };
  auto issue_ab = [&](int j, int pbuf, long long rounds) {
    issue_ab_part(j, n_issued % S, pbuf, true, true, rounds);
    ++n_issued;
  };
  // Phase-C loads of tile j into stage `st`.  part 1 = what this CTA itself wrote or owns (arms the barrier with
  // the full byte count); part 2 = the p_new box with its ring for a tile that recomputes Ap: the ring cells
  // were written by other CTAs during phase AB, so that box has to wait for the reduction after it.
  auto issue_c_part = [&](int j, unsigned int st, int pbuf, bool part1, bool part2) {
    if (threadIdx.x == 0) {
      const TileItem it = s_items_c[j];
      const int li0 = it.li0, lj0 = it.lj0;
      unsigned char* sb = stage_base + (size_t)st * Cfg::kStageBytes;
      uint64_t* bar = &s_bar[st];
      const bool coef = (it.flags & 3) == 2;
      const bool recomp = (it.flags & kItemRecomp) != 0, tall = (it.flags & kItemTall) != 0;
      const int x = kColOff + lj0, y = li0;
      if (tall) {  // w | r | p_new with ring, boxes of 2 TX rows
        if (part1) {
          fence_proxy_async_smem();
          mbar_arrive_expect_tx(bar, Cfg::kTxTallC);
          tma_load_2d_hint(sb, maps_tall + kMapW, x, y, bar, pol_stream);
          tma_load_2d_hint(sb + Cfg::kTileBytes2, maps_tall + kMapR, x, y, bar, pol_keep);
        }
        if (part2)
          tma_load_2d_hint(sb + 2 * Cfg::kTileBytes2, maps_tall + (pbuf ? kMapP1Ext : kMapP0Ext), x - 2, y - 1, bar, pol_keep);
        return;
      }
      if (part1) {
        fence_proxy_async_smem();
        mbar_arrive_expect_tx(bar, recomp ? Cfg::kTxCRecomp : Cfg::kTxCCore + (coef ? Cfg::kTxCCoef : 0u));
        tma_load_2d_hint(sb, maps + kMapW, x, y, bar, pol_stream);
        tma_load_2d_hint(sb + 2 * Cfg::kTileBytes, maps + kMapR, x, y, bar, pol_keep);
        if (!recomp) {
          tma_load_2d_hint(sb + Cfg::kTileBytes, maps + (pbuf ? kMapP1 : kMapP0), x, y, bar, pol_keep);
          tma_load_2d_hint(sb + 3 * Cfg::kTileBytes, maps + kMapAp, x, y, bar, pol_stream);
          if (coef) tma_load_2d_hint(sb + Cfg::kOffCD, maps + kMapD, x, y, bar, pol_keep);
        }
      }
      if (part2 && recomp)
        tma_load_2d_hint(sb + Cfg::kOffCPe, maps + (pbuf ? kMapP1Ext : kMapP0Ext), x - 2, y - 1, bar, pol_keep);
    }
  };
  auto issue_c = [&](int j, int pbuf) {
    issue_c_part(j, n_issued % S, pbuf, true, true);
    ++n_issued;
  };
  // A stage wait that times out (an internal error: a transaction count that never completes) must not split the
  // CTA around a barrier: the thread only raises s_abort, stops waiting for further stages (`waiting`, per thread)
  // and carries on with whatever the stages hold; every thread reads the flag once after the last barrier of the
  // phase, so the whole CTA leaves the iteration together.
  bool tma_ok = true, waiting = true;
  auto wait_stage = [&]() -> unsigned char* {
    const unsigned int st = n_consumed % S;
    if (waiting && !mbar_wait(&s_bar[st], (n_consumed / S) & 1u)) {
      waiting = false;
      s_abort = 1;
      if (threadIdx.x == 0) {  // leave a trace; the other CTAs run into the watchdog
        ctl->status = PB_STATUS_TMA_TIMEOUT;
        ctl->red[0] = (double)n_consumed;
        ctl->red[1] = (double)cta;
      }
    }
    return stage_base + (size_t)st * Cfg::kStageBytes;
  };
  // nothing may be in flight into shared memory when the CTA exits (no barrier inside: every thread for itself)
  auto drain = [&]() {
    while (n_consumed < n_issued && !s_abort) {
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
    PB_RED_PRE
    const RedResult rr = PB_RED(1);
    if (!rr.ok) {
      drain();
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
  long long k = 0;
  for (k = 1; k <= max_iter; ++k) {
    double* pout = c.p[cur ^ 1];
    // ======================= phase AB (tiles 0..npre-1 are already in flight) =======================
    double acc_ab = 0.0;
    for (int j = 0; j < nmine; ++j) {
      if (j + S - 1 < nmine) issue_ab(j + S - 1, cur, k - 1);
      unsigned char* sb = wait_stage();
      const TileItem item = s_items[j];  // CTA-uniform
      const int li0 = item.li0, lj0 = item.lj0, tmode = item.flags & 3;
      const bool tall = (item.flags & kItemTall) != 0;
      const int rows = tall ? 2 * TX : TX;  // rows of the box in shared memory (a tall item is always full)
      const double* s_r = reinterpret_cast<const double*>(sb + Cfg::kOffR);
      double* s_p = reinterpret_cast<double*>(sb + (tall ? Cfg::kExtBytes2 : Cfg::kOffP));
      const double* s_D = reinterpret_cast<const double*>(sb + Cfg::kOffD);
      const double* s_a = reinterpret_cast<const double*>(sb + Cfg::kOffA);
      const double* s_b = reinterpret_cast<const double*>(sb + Cfg::kOffB);
      const int cx = min(rows, c.nx - li0 + 1), cy = min(TY, c.ny - lj0 + 1);
      const bool full = (cx == rows) && (cy == TY);
      const double cf = tmode == 1 ? inv_eps : 1.0;  // uniform tiles: every face coefficient
      const double dg = tmode == 1 ? diag_out : diag_in, rc = tmode == 1 ? rcp_out : rcp_in;
      const bool keep_ap = (item.flags & kItemRecomp) == 0;  // phase C reloads Ap (else it evaluates the stencil again)
      const int nhalf = rows / TX;  // full tiles are processed in slices of TX rows, 4 (TX = 32) nodes per thread each

      // The two halves of the phase on one tile.  `uniform` is a compile-time tag, so the constant-coefficient
      // path carries no shared-memory coefficient loads and the mixed path no selects.
      auto run_tile = [&](auto uniform) {
        constexpr bool U = decltype(uniform)::value;
        // 1. p_new = r/D + beta p_old on the tile and its ring, in place in shared memory.  The owned cells
        //    map to threads without any division or validity test (4 cells per thread and slice, unrolled for full
        //    tiles so that their dependent fp64 chains overlap); the ring (top, bottom, left, right: at most
        //    2 TY + 2 rows cells) is one extra cell for the first threads.  A cell of shared-memory row `er`,
        //    tile column `ec` (0 = left ring) sits at er * EC + ec + 1 (the box starts one column early).
        auto p_cell = [&](int er, int ec, bool store) {
          const int e = er * EC + ec + 1;
          const double z = U ? div_known<DIV>(s_r[e], dg, rc) : div_var<DIV>(s_r[e], s_D[e]);
          const double v = fma(beta, s_p[e], z);
          s_p[e] = v;
          if (store) pout[(size_t)(li0 - 1 + er) * S_pitch + (lj0 - 1 + ec + kColOff)] = v;
        };
        {
          const int c0 = threadIdx.x % TY, r00 = threadIdx.x / TY;
          if (full) {
            // same as p_cell(r00 + 8 q + 1, c0 + 1, true), with the index arithmetic hoisted out of the unrolled body
            constexpr int kRows = kThreads / TY;
            for (int h = 0; h < nhalf; ++h) {
              const double* sr = s_r + (h * TX + r00 + 1) * EC + c0 + 2;
              double* sp = s_p + (h * TX + r00 + 1) * EC + c0 + 2;
              const double* sd = s_D + (h * TX + r00 + 1) * EC + c0 + 2;
              double* po = pout + (size_t)(li0 + h * TX + r00) * S_pitch + (lj0 + c0 + kColOff);
              const size_t gstep = (size_t)kRows * S_pitch;
#pragma unroll
              for (int q = 0; q < TX / kRows; ++q) {
                const double z = U ? div_known<DIV>(sr[q * kRows * EC], dg, rc) : div_var<DIV>(sr[q * kRows * EC], sd[q * kRows * EC]);
                const double v = fma(beta, sp[q * kRows * EC], z);
                sp[q * kRows * EC] = v;
                po[q * gstep] = v;
              }
            }
          } else if (c0 < cy) {
            for (int r0 = r00; r0 < cx; r0 += kThreads / TY) p_cell(r0 + 1, c0 + 1, true);
          }
          // ring cells are stored to global memory only where they are ring cells of the whole block
          if (threadIdx.x < 2 * TY) {
            const int side = threadIdx.x / TY, cc = threadIdx.x - side * TY;
            if (cc < cy) p_cell(side ? cx + 1 : 0, cc + 1, side ? (li0 + cx - 1 == c.nx) : (li0 == 1));
          } else if (threadIdx.x < 2 * TY + 2 * rows) {
            const int t = threadIdx.x - 2 * TY, side = t / rows, rr = t - side * rows;
            if (rr < cx) p_cell(rr + 1, side ? cy + 1 : 0, side ? (lj0 + cy - 1 == c.ny) : (lj0 == 1));
          }
        }
        __syncthreads();
        // 2. Ap = A p_new on the owned nodes of the tile
        const int col = threadIdx.x % TY + 1;
        auto ab_node = [&](int row, double& pc) {
          const double* s = s_p + row * EC + col + 1;
          pc = s[0];
          double a_c, a_ip, b_c, b_jp;
          if constexpr (U) {
            a_c = a_ip = b_c = b_jp = cf;
          } else {
            a_c = s_a[(row - 1) * TY + (col - 1)], a_ip = s_a[row * TY + (col - 1)];
            b_c = s_b[(row - 1) * BC + (col - 1)], b_jp = s_b[(row - 1) * BC + col];
          }
          const double v = stencil_value<DIV>(K, pc, s[-EC], s[EC], s[-1], s[1], a_c, a_ip, b_c, b_jp);
          c.Ap[(size_t)(li0 - 1 + row) * S_pitch + (lj0 - 1 + col + kColOff)] = v;
          return v;
        };
        if (full) {
          // same as ab_node(r00 + 8 q + 1), index arithmetic hoisted
          constexpr int kRows = kThreads / TY, kPer = TX / kRows;  // nodes per thread and slice
          const int r00 = threadIdx.x / TY;
          for (int h = 0; h < nhalf; ++h) {
            const double* sp = s_p + (h * TX + r00 + 1) * EC + col + 1;
            const double* sa = s_a + (h * TX + r00) * TY + (col - 1);
            const double* sb2 = s_b + (h * TX + r00) * BC + (col - 1);
            double* ap = c.Ap + (size_t)(li0 + h * TX + r00) * S_pitch + (lj0 - 1 + col + kColOff);
            const size_t gstep = (size_t)kRows * S_pitch;
            double v[kPer], pc[kPer];
#pragma unroll
            for (int q = 0; q < kPer; ++q) {
              const double* s = sp + q * kRows * EC;
              pc[q] = s[0];
              double a_c, a_ip, b_c, b_jp;
              if constexpr (U) {
                a_c = a_ip = b_c = b_jp = cf;
              } else {
                a_c = sa[q * kRows * TY], a_ip = sa[q * kRows * TY + TY];
                b_c = sb2[q * kRows * BC], b_jp = sb2[q * kRows * BC + 1];
              }
              v[q] = stencil_value<DIV>(K, pc[q], s[-EC], s[EC], s[-1], s[1], a_c, a_ip, b_c, b_jp);
              if (keep_ap) ap[q * gstep] = v[q];
            }
#pragma unroll
            for (int q = 0; q < kPer; ++q) acc_ab = fma(v[q], pc[q], acc_ab);  // same order as the generic loop
          }
        } else if (col <= cy) {
          for (int row = threadIdx.x / TY + 1; row <= cx; row += kThreads / TY) {
            double pc;
            const double v = ab_node(row, pc);
            acc_ab = fma(v, pc, acc_ab);
          }
        }
      };
      // The same on a full item with constant coefficients, without the barrier in the middle: warp w owns NR = rows / 16
      // consecutive rows of the item and all of its 64 columns (two per lane).  Every lane evaluates p_new for its two
      // columns on the warp's rows and on the row above and below them (in registers; the two extra rows are computed
      // twice in the CTA, which costs two multiply-adds per node and saves the round trip of p_new through shared
      // memory with the barrier behind it), gets the left / right neighbours from the adjacent lanes by shuffle (the
      // outermost lanes evaluate the ring column), and goes straight on to the stencil.  Same expressions on the same
      // operands as p_cell / ab_node: the same bits for p_new and Ap; only the order of the per-thread partial sums of
      // (Ap, p) differs from the tile path.
      auto run_rows = [&](auto nr_tag) {
        constexpr int NR = decltype(nr_tag)::value;
        // 16-byte accesses: a lane's two nodes start at an even column of the box in shared memory and at an even
        // column of the padded row in global memory (lj0 is odd, kColOff is odd, the pitch is a multiple of 16 doubles)
        static_assert(EC % 2 == 0 && TY % 2 == 0 && kColOff % 2 == 1, "rows path: pairs of nodes must be 16-byte aligned");
        static_assert(NR * (kThreads / 32) == TX || NR * (kThreads / 32) == 2 * TX, "rows path: the warps' rows must tile the item");
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        const int row0 = 1 + warp * NR;  // first own row of the box in shared memory (row 0 is the ring)
        const int ec = 2 * lane + 2;     // box column of the lane's first node (the box starts one column early)
        double pn[NR + 2][2];
#pragma unroll
        for (int i = 0; i < NR + 2; ++i) {
          const int e = (row0 - 1 + i) * EC + ec;
          const double2 rr = *reinterpret_cast<const double2*>(s_r + e);
          const double2 pp = *reinterpret_cast<const double2*>(s_p + e);
          pn[i][0] = fma(beta, pp.x, div_known<DIV>(rr.x, dg, rc));
          pn[i][1] = fma(beta, pp.y, div_known<DIV>(rr.y, dg, rc));
        }
        double edge[NR];  // lanes 0 / 31: p_new in the ring column left / right of the item, own rows
#pragma unroll
        for (int i = 0; i < NR; ++i) edge[i] = 0.0;
        const bool outer = lane == 0 || lane == 31;
        if (outer) {
          const int ecol = lane == 0 ? 1 : EC - 2;
#pragma unroll
          for (int i = 0; i < NR; ++i) {
            const int e = (row0 + i) * EC + ecol;
            edge[i] = fma(beta, s_p[e], div_known<DIV>(s_r[e], dg, rc));
          }
        }
        double* po = pout + (size_t)(li0 - 1 + row0) * S_pitch + (lj0 + 2 * lane + kColOff);
        double* ap = c.Ap + (size_t)(li0 - 1 + row0) * S_pitch + (lj0 + 2 * lane + kColOff);
#pragma unroll
        for (int i = 0; i < NR; ++i) {
          double left = __shfl_up_sync(0xffffffffu, pn[i + 1][1], 1);
          double right = __shfl_down_sync(0xffffffffu, pn[i + 1][0], 1);
          if (lane == 0) left = edge[i];
          if (lane == 31) right = edge[i];
          const double v0 = stencil_value<DIV>(K, pn[i + 1][0], pn[i][0], pn[i + 2][0], left, pn[i + 1][1], cf, cf, cf, cf);
          const double v1 = stencil_value<DIV>(K, pn[i + 1][1], pn[i][1], pn[i + 2][1], pn[i + 1][0], right, cf, cf, cf, cf);
          *reinterpret_cast<double2*>(po + (size_t)i * S_pitch) = make_double2(pn[i + 1][0], pn[i + 1][1]);
          if (keep_ap) *reinterpret_cast<double2*>(ap + (size_t)i * S_pitch) = make_double2(v0, v1);
          acc_ab = fma(v0, pn[i + 1][0], acc_ab);
          acc_ab = fma(v1, pn[i + 1][1], acc_ab);
        }
        // ring cells of the whole block keep their p_new in global memory (the next phase AB reads them as p_old)
        if (warp == 0 && li0 == 1)
          *reinterpret_cast<double2*>(po - S_pitch) = make_double2(pn[0][0], pn[0][1]);
        if (warp == kThreads / 32 - 1 && li0 + rows - 1 == c.nx)
          *reinterpret_cast<double2*>(po + (size_t)NR * S_pitch) = make_double2(pn[NR + 1][0], pn[NR + 1][1]);
        if (lane == 0 && lj0 == 1) {
#pragma unroll
          for (int i = 0; i < NR; ++i) po[(size_t)i * S_pitch - 1] = edge[i];
        }
        if (lane == 31 && lj0 + TY - 1 == c.ny) {
#pragma unroll
          for (int i = 0; i < NR; ++i) po[(size_t)i * S_pitch + 2] = edge[i];
        }
      };
      if (rows_ab && full && tmode != 2) {
        if (tall)
          run_rows(std::integral_constant<int, 2 * TX / (kThreads / 32)>{});
        else
          run_rows(std::integral_constant<int, TX / (kThreads / 32)>{});
      } else if (tmode != 2)
        run_tile(std::true_type{});
      else
        run_tile(std::false_type{});
      // last tile: every thread orders its global stores of the phase (p_new, Ap) before the TMA reads that
      // thread 0 issues for phase C right after the barrier
      if (j == nmine - 1) fence_proxy_async_global();
      __syncthreads();  // stage free
      ++n_consumed;
    }
    if (s_abort) {
      tma_ok = false;
      break;
    }
    // Prefetch for phase C: it reads only what this CTA itself wrote (p_new, Ap) or owns (w, r, D), so its
    // first tiles can be in flight while the reduction below waits for the slowest CTA / GPU.  The exception is
    // the p_new box with ring of a tile that recomputes Ap (part 2, issued right after the reduction).
    const unsigned int c_pre_base = n_issued;
    if (threadIdx.x == 0) fence_proxy_async();  // our generic-proxy global stores -> TMA reads
    for (int j = 0; j < npre_c; ++j) {
      issue_c_part(j, n_issued % S, cur ^ 1, true, false);
      ++n_issued;
    }
    PB_STAMP(t_ab);
    {
      double v[1] = {acc_ab};
      ++seq;
      PB_RED_PRE
      const RedResult rr = PB_RED(1);
      PB_STAMP(t_r1);
      // part 2 of the prefetch: every CTA's p_new is complete and acquired (issued on every path, the barriers
      // of these stages are already armed for the bytes)
      if (recompute_ap) {
        if (threadIdx.x == 0) fence_proxy_async();
        for (int j = 0; j < npre_c; ++j) issue_c_part(j, (c_pre_base + j) % S, cur ^ 1, false, true);
      }
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
    for (int j = 0; j < nmine_c; ++j) {
      if (j + S - 1 < nmine_c) issue_c(j + S - 1, cur ^ 1);
      unsigned char* sb = wait_stage();
      const TileItem item = s_items_c[j];  // CTA-uniform
      const int li0 = item.li0, lj0 = item.lj0, tmode = item.flags & 3;
      const bool tall = (item.flags & kItemTall) != 0;
      const int rows = tall ? 2 * TX : TX;
      // 32-row items: w | p | r | Ap | D, with recomputed Ap w | - | r | p_new + ring; tall items: w | r | p_new + ring
      const double* s_w = reinterpret_cast<const double*>(sb);
      const double* s_p = reinterpret_cast<const double*>(sb + Cfg::kTileBytes);
      const double* s_r = reinterpret_cast<const double*>(sb + (tall ? Cfg::kTileBytes2 : 2 * Cfg::kTileBytes));
      const double* s_Ap = reinterpret_cast<const double*>(sb + 3 * Cfg::kTileBytes);
      const double* s_D = reinterpret_cast<const double*>(sb + Cfg::kOffCD);
      const double* s_pe = reinterpret_cast<const double*>(sb + (tall ? 2 * Cfg::kTileBytes2 : Cfg::kOffCPe));  // recompute only
      const int cx = min(rows, c.nx - li0 + 1), cy = min(TY, c.ny - lj0 + 1);
      const int col = threadIdx.x % TY;
      const bool full = (cx == rows) && (cy == TY);
      const double dg = tmode == 1 ? diag_out : diag_in, rc = tmode == 1 ? rcp_out : rcp_in;
      const double cf = tmode == 1 ? inv_eps : 1.0;
      const bool recomp = (item.flags & kItemRecomp) != 0;  // CTA-uniform; implies full and constant coefficients
      const bool edge_tile = has_nbr && (li0 == 1 || li0 + cx - 1 == c.nx || lj0 == 1 || lj0 + cy - 1 == c.ny);
      const int nhalf = rows / TX;
      auto run_tile = [&](auto uniform) {
        constexpr bool U = decltype(uniform)::value;
        // p_new and Ap of the node in tile row `row` (0-based): reloaded, or the stencil of phase AB again
        auto p_and_ap = [&](int row, double& pc, double& ap) {
          if (U && recomp) {
            const double* s = s_pe + (row + 1) * EC + col + 2;
            pc = s[0];
            ap = stencil_value<DIV>(K, pc, s[-EC], s[EC], s[-1], s[1], cf, cf, cf, cf);
          } else {
            pc = s_p[row * TY + col];
            ap = s_Ap[row * TY + col];
          }
        };
        // loads and arithmetic of one node; stores and halo stores included
        auto c_node = [&](int row, double& dw, double& z, double& rn) {
          const int e = row * TY + col;
          const int li = li0 + row, lj = lj0 + col;
          const size_t g = (size_t)li * S_pitch + (lj + kColOff);
          double pc, ap;
          p_and_ap(row, pc, ap);
          const double w_old = s_w[e];
          const double w_new = fma(alpha, pc, w_old);
          const double r_new = fma(-alpha, ap, s_r[e]);
          if (l2_hints)
            __stcs(&c.w[g], w_new);  // streaming store: w is not read again before the next iteration's phase C
          else
            c.w[g] = w_new;
          c.r[g] = r_new;
          dw = __dsub_rn(w_new, w_old);
          z = U ? div_known<DIV>(r_new, dg, rc) : div_var<DIV>(r_new, s_D[e]);
          rn = r_new;
          if (edge_tile) {  // halo: our edge values go into the neighbour's halo line (peer stores over NVLink)
            const unsigned int hs = hseq + (unsigned int)k;
            const unsigned long long bits = (unsigned long long)__double_as_longlong(r_new);
            const unsigned long long w0 = (bits & 0xffffffffull) | ((unsigned long long)hs << 32);
            const unsigned long long w1 = (bits >> 32) | ((unsigned long long)hs << 32);
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
        };
        if (full && !edge_tile) {
          // same as c_node(r00 + 8 q) without the halo stores, index arithmetic hoisted
          constexpr int kRows = kThreads / TY, kPer = TX / kRows;
          const int r00 = threadIdx.x / TY;
          for (int h = 0; h < nhalf; ++h) {
            const int e0 = (h * TX + r00) * TY + col;
            double* pw = c.w + (size_t)(li0 + h * TX + r00) * S_pitch + (lj0 + col + kColOff);
            double* pr = c.r + (size_t)(li0 + h * TX + r00) * S_pitch + (lj0 + col + kColOff);
            const size_t gstep = (size_t)kRows * S_pitch;
            double dw[kPer], z[kPer], rn[kPer];
#pragma unroll
            for (int q = 0; q < kPer; ++q) {
              const int e = e0 + q * kRows * TY;
              double pc, ap;
              p_and_ap(h * TX + r00 + q * kRows, pc, ap);
              const double w_old = s_w[e];
              const double w_new = fma(alpha, pc, w_old);
              rn[q] = fma(-alpha, ap, s_r[e]);
              if (l2_hints)
                __stcs(pw + q * gstep, w_new);
              else
                pw[q * gstep] = w_new;
              pr[q * gstep] = rn[q];
              dw[q] = __dsub_rn(w_new, w_old);
              z[q] = U ? div_known<DIV>(rn[q], dg, rc) : div_var<DIV>(rn[q], s_D[e]);
            }
#pragma unroll
            for (int q = 0; q < kPer; ++q) {
              acc_d = fma(dw[q], dw[q], acc_d);
              acc_zr = fma(z[q], rn[q], acc_zr);
            }
          }
        } else if (full) {
          constexpr int kPer = TX / (kThreads / TY);
          for (int h = 0; h < nhalf; ++h) {
            double dw[kPer], z[kPer], rn[kPer];
#pragma unroll
            for (int q = 0; q < kPer; ++q) c_node(h * TX + threadIdx.x / TY + q * (kThreads / TY), dw[q], z[q], rn[q]);
#pragma unroll
            for (int q = 0; q < kPer; ++q) {
              acc_d = fma(dw[q], dw[q], acc_d);
              acc_zr = fma(z[q], rn[q], acc_zr);
            }
          }
        } else if (col < cy) {
          for (int row = threadIdx.x / TY; row < cx; row += kThreads / TY) {
            double dw, z, rn;
            c_node(row, dw, z, rn);
            acc_d = fma(dw, dw, acc_d);
            acc_zr = fma(z, rn, acc_zr);
          }
        }
      };
      if (tmode != 2)
        run_tile(std::true_type{});
      else
        run_tile(std::false_type{});
      if (j == nmine_c - 1) fence_proxy_async_global();  // as in phase AB: w, r before the TMA reads of the next phase
      __syncthreads();  // stage free
      ++n_consumed;
    }
    if (s_abort) {
      tma_ok = false;
      break;
    }
    // Prefetch for the next phase AB, part 1: coefficients are constant and p_new (buffer cur^1, ring included)
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
      PB_RED_PRE
      const RedResult rr = PB_RED(2);
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

  if (threadIdx.x == 0 && tma_ok) {
    // where this CTA spent the solve (ns): phase AB, phase C, waiting in reduction 1, waiting in reduction 2
    double* mine = reinterpret_cast<double*>(reinterpret_cast<char*>(ctl) + kCtaTimesOffset) + kCtaTimeWords * cta;
    mine[0] = (double)t_ab, mine[1] = (double)t_c, mine[2] = (double)t_r1, mine[3] = (double)t_r2;
    // tiles of this CTA by what they move per node and iteration (the bench's byte model counts these)
    // (a tall item counts as the two tiles it is made of)
    int n_recomp = 0, n_mixed = 0, n_tiles = 0;
    for (int j = 0; j < nmine; ++j) {
      const int f = s_items[j].flags, w = (f & kItemTall) ? 2 : 1;
      n_tiles += w;
      n_recomp += (f & kItemRecomp) ? w : 0;
      n_mixed += (f & 3) == 2 ? 1 : 0;
    }
    mine[4] = (double)n_recomp, mine[5] = (double)(n_tiles - n_mixed - n_recomp), mine[6] = (double)n_mixed;
    mine[7] = (double)s_t_halo;
  }
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
    ctl->halo_seq = hseq + (unsigned int)k + 1u;
    ctl->t_ab = t_ab;
    ctl->t_c = t_c;
    ctl->t_red1 = t_r1;
    ctl->t_red2 = t_r2;
    ctl->t_init = t_init;
    ctl->t_total = globaltimer_ns() - t_start;
    ctl->t_halo = s_t_halo;
  }
  star_pair_exit(pair_mode & 1);
#undef PB_STAMP
#undef PB_RED
#undef PB_RED_PRE
}

// ---------------------------------------------------------------------------------------------------
template <int TX, int DIV, bool CC>
static cudaError_t launch_variant(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int variant, int l2_hints,
                                  int recompute_ap, int tall, int balance, cudaStream_t stream) {
  const void* fn = (const void*)pcg_persistent_tma_kernel<TX, DIV, CC>;
  const size_t smem = TmaCfg<TX>::kSmemBytes;
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int pair = 0;
  void* args[] = {(void*)&d_ctxs, (void*)&ctas_per_rank, (void*)&variant, (void*)&l2_hints, (void*)&recompute_ap, (void*)&tall,
                  (void*)&balance, (void*)&pair};
  return launch_cooperative_pairs(fn, nvirt, ctas_per_rank, kThreads, args, &pair, smem, stream);
}

template <int TX, bool CC>
static cudaError_t launch_div(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int variant, int div, int l2_hints,
                              int recompute_ap, int tall, int balance, cudaStream_t stream) {
  if (div == kDivTrue)
    return launch_variant<TX, kDivTrue, CC>(d_ctxs, nvirt, ctas_per_rank, variant, l2_hints, recompute_ap, tall, balance, stream);
  if (div == kDivMarkstein)
    return launch_variant<TX, kDivMarkstein, CC>(d_ctxs, nvirt, ctas_per_rank, variant, l2_hints, recompute_ap, tall, balance, stream);
  return launch_variant<TX, kDivRecip, CC>(d_ctxs, nvirt, ctas_per_rank, variant, l2_hints, recompute_ap, tall, balance, stream);
}

cudaError_t launch_persistent_tma(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int tile_rows, int div,
                                  int l2_hints, bool compressed, bool recompute_ap, bool balance, bool rows_ab,
                                  cudaStream_t stream) {
  const int ra = (compressed && recompute_ap) ? 1 : 0, bal = balance ? 1 : 0;  // without coefficient classes no tile is known to be constant
  // tile_rows: 16, 32, or 64 = 32-row tiles with pairs of full constant-coefficient tiles merged into tall items (in
  // both phases when Ap is recomputed, in phase AB only when it is stored)
  if (tile_rows == 16)
    return compressed ? launch_div<16, true>(d_ctxs, nvirt, ctas_per_rank, 1, div, l2_hints, ra, 0, bal, stream)
                      : launch_div<16, false>(d_ctxs, nvirt, ctas_per_rank, 1, div, l2_hints, ra, 0, 0, stream);
  const int tall = (tile_rows == 64 ? 1 : 0) | (rows_ab ? 2 : 0);  // bit 0: tall items, bit 1: phase AB by warp-private rows
  return compressed ? launch_div<32, true>(d_ctxs, nvirt, ctas_per_rank, 0, div, l2_hints, ra, tall, bal, stream)
                    : launch_div<32, false>(d_ctxs, nvirt, ctas_per_rank, 0, div, l2_hints, ra, 0, 0, stream);
}

int persistent_tma_max_items() { return kMaxItems; }

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
  return ok;
}

size_t tensor_map_bytes() { return sizeof(CUtensorMap); }

}  // namespace pb
