// persistent.cu -- KP: the whole PCG iteration loop as one cooperative, persistent kernel.
//
// Replaces the host loop of the reference (S4:846-942: 6 kernel launches, 6 cudaDeviceSynchronize,
// 3 x 256 KiB D2H, up to 8 halo memcpys, 4 MPI_Sendrecv and 3 MPI_Allreduce per iteration) by
//
//   loop:  phase AB  ->  cross-rank sum {(Ap,p)}            -> alpha
//          phase C   ->  cross-rank sum {sum dw^2, (z,r)}   -> stop test, beta
//
// with alpha, beta and the stop decision computed redundantly (and bit-identically) by every CTA of
// every rank.  Halo data (the neighbour's new r on our ring) is stored by the neighbour during its
// phase C and becomes visible through the release/acquire chain of the second reduction, so an
// iteration has exactly two cross-GPU synchronisation points and no separate halo exchange.
//
// One launch serves `nvirt` rank contexts: blockIdx.x / ctas_per_rank selects the context.  Real
// multi-GPU runs use nvirt = 1 per process; tests place several virtual ranks on one GPU, where the
// cooperative launch guarantees that all CTAs that wait for each other are co-resident.
#include <cooperative_groups.h>

#include <cstdlib>

#include "launch.h"
#include "phases.cuh"

namespace pb {

template <int DIV>
__global__ void __launch_bounds__(kThreads, 2)
pcg_persistent_kernel(const RankCtx* __restrict__ ctxs, int ctas_per_rank) {
  __shared__ RankCtx c;
  __shared__ double s_red[80];
  extern __shared__ double s_tile[];

  const int vrank = blockIdx.x / ctas_per_rank;
  const int cta = blockIdx.x - vrank * ctas_per_rank;
  const int ncta = ctas_per_rank;
  {
    const int* src = reinterpret_cast<const int*>(ctxs + vrank);
    int* dst = reinterpret_cast<int*>(&c);
    for (int i = threadIdx.x; i < (int)(sizeof(RankCtx) / sizeof(int)); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();

  Control* ctl = c.ctl;
  const int ntiles = c.tiles_x * c.tiles_y;
  const bool has_nbr = c.nbr_r[0] || c.nbr_r[1] || c.nbr_r[2] || c.nbr_r[3];
  const bool timing = (cta == 0 && threadIdx.x == 0);

  // Launch-time state.  red_seq persists across launches so that slot contents left over from an
  // earlier solve can never be mistaken for a value of this one; the arrival counter is zeroed by the
  // host before the launch.
  unsigned int seq = ctl->red_seq;
  unsigned int arrive_target = 0;
  const double delta = ctl->delta;
  const long long max_iter = ctl->max_iter;
  const int norm = ctl->norm;
  const double h1h2 = c.g.h1h2;

  unsigned long long t_prev = timing ? globaltimer_ns() : 0ull, t_start = t_prev;
  unsigned long long t_ab = 0, t_c = 0, t_r1 = 0, t_r2 = 0, t_init = 0;
#define PB_STAMP(acc)                           \
  if (timing) {                                 \
    const unsigned long long now_ = globaltimer_ns(); \
    acc += now_ - t_prev;                       \
    t_prev = now_;                              \
  }

  // ---- init: zr_old = (D^-1 r, r) ----
  double zr_old;
  {
    double v[1] = {0.0};
    for (int t = cta; t < ntiles; t += ncta) v[0] += dot_zr_tile(c, t);
    ++seq;
    arrive_target += ncta;
    const RedResult rr = cross_rank_sum<1>(c, v, seq, arrive_target, cta, ncta, s_red, false, c.sys_scope != 0);
    if (!rr.ok) {
      if (threadIdx.x == 0 && cta == 0) {
        ctl->status = PB_STATUS_WATCHDOG;
        ctl->red_seq = seq;
      }
      return;
    }
    zr_old = rr.v[0] * h1h2;
  }
  PB_STAMP(t_init);

  int cur = 0;
  double beta = 0.0, alpha = 0.0, denom = 0.0, diff_norm = 0.0;
  int status = PB_STATUS_MAXITER;
  long long k = 0;
  for (k = 1; k <= max_iter; ++k) {
    const double* pin = c.p[cur];
    double* pout = c.p[cur ^ 1];
    // ---- phase AB ----
    {
      double v[1] = {0.0};
      for (int t = cta; t < ntiles; t += ncta) v[0] += phase_ab_tile<true, DIV>(c, t, beta, pin, pout, s_tile);
      PB_STAMP(t_ab);
      ++seq;
      arrive_target += ncta;
      const RedResult rr = cross_rank_sum<1>(c, v, seq, arrive_target, cta, ncta, s_red, false, c.sys_scope != 0);
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
    // ---- phase C ----
    double zr_new;
    {
      double v[2] = {0.0, 0.0};
      for (int t = cta; t < ntiles; t += ncta) phase_c_tile(c, t, alpha, pout, v[0], v[1]);
      PB_STAMP(t_c);
      ++seq;
      arrive_target += ncta;
      const RedResult rr = cross_rank_sum<2>(c, v, seq, arrive_target, cta, ncta, s_red, has_nbr, c.sys_scope != 0);
      PB_STAMP(t_r2);
      if (!rr.ok) {
        status = PB_STATUS_WATCHDOG;
        break;
      }
      diff_norm = norm == PB_NORM_WEIGHTED ? sqrt(rr.v[0] * h1h2) : sqrt(rr.v[0]);
      zr_new = rr.v[1] * h1h2;
    }
    cur ^= 1;
    if (diff_norm < delta) {
      status = PB_STATUS_CONVERGED;
      break;
    }
    beta = zr_new / zr_old;
    zr_old = zr_new;
  }
  if (k > max_iter) k = max_iter;

  if (cta == 0 && threadIdx.x == 0) {
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
// Micro-benchmark of the communication primitives alone (BASELINE.json config 5): `iters` rounds of
//   mode 0: cross-rank reduction of two doubles, no halo traffic
//   mode 1: every rank stores `edge` doubles into each neighbour's r ring (spread over all CTAs), then
//           the reduction whose system-scope release/acquire chain publishes them (protocol of persistent.cu)
//   mode 2: the first CTAs store the edges, release them with one system fence each and bump the
//           neighbour's halo counter; the reduction stays at GPU scope; CTA 0 finally checks its own
//           counters (the TMA kernel's first protocol, kept for comparison)
//   mode 3: the reduction alone through star_sum (sequence-tagged records, no atomics; persistent_tma.cu)
//   mode 4: the first CTAs store the edges as sequence-tagged LL records into the neighbours' halo lines (no fence,
//           no counter), warp 0 of the same CTAs polls the incoming records, then star_sum: the protocol of
//           persistent_tma.cu
//   mode 5: mode 3 with clock64 stamps between the steps of star_sum; thread 0 of every CTA leaves the mean cycles
//           per step in the per-CTA report area (pb_solver_cta_report)
// CTA 0 of each rank measures the whole loop with %globaltimer -> Control::t_total (ns).
__global__ void __launch_bounds__(kThreads, 1)
xgpu_bench_kernel(const RankCtx* __restrict__ ctxs, int ctas_per_rank, int iters, int mode, int edge, int pair_mode) {
  __shared__ RankCtx c;
  __shared__ double s_red[80];
  const int vrank = blockIdx.x / ctas_per_rank;
  const int cta = blockIdx.x - vrank * ctas_per_rank, ncta = ctas_per_rank;
  {
    const int* src = reinterpret_cast<const int*>(ctxs + vrank);
    int* dst = reinterpret_cast<int*>(&c);
    for (int i = threadIdx.x; i < (int)(sizeof(RankCtx) / sizeof(int)); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  star_pair_init(pair_mode & 1);
  unsigned int seq = c.ctl->red_seq, arrive_target = 0;
  const unsigned int hseq = c.ctl->halo_seq;
  const unsigned long long wd_limit = c.ctl->watchdog_ns ? c.ctl->watchdog_ns : kDefaultWatchdogNs;
  const bool has_nbr = c.nbr_r[0] || c.nbr_r[1] || c.nbr_r[2] || c.nbr_r[3];
  double check = 0.0;
  bool ok = true;
  long long tr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int npush = min(ncta, 16);  // mode 2: CTAs that act as "perimeter" CTAs
  unsigned long long base[4];
  for (int d = 0; d < 4; ++d) base[d] = ld_relaxed_sys_u64(c.halo_cnt + d);
  {  // nobody bumps a counter before everybody has read its base
    double v[2] = {0.0, 0.0};
    ++seq;
    arrive_target += ncta;
    ok = cross_rank_sum<2>(c, v, seq, arrive_target, cta, ncta, s_red, false, false).ok;
  }
  const unsigned long long t0 = globaltimer_ns();
  for (int it = 0; it < iters && ok; ++it) {
    if (mode == 2 && cta < npush) {
      const int gtid = cta * kThreads + threadIdx.x, gsz = npush * kThreads;
      const double val = (double)(it + 1);
      for (int i = gtid; i < edge; i += gsz) {
        if (c.nbr_r[kXm] && i < c.ny) c.nbr_r[kXm][(size_t)(c.nbr_nx[kXm] + 1) * c.nbr_pitch[kXm] + (i + 1 + kColOff)] = val;
        if (c.nbr_r[kXp] && i < c.ny) c.nbr_r[kXp][(size_t)(i + 1 + kColOff)] = val;
        if (c.nbr_r[kYm] && i < c.nx) c.nbr_r[kYm][(size_t)(i + 1) * c.nbr_pitch[kYm] + (c.nbr_ny[kYm] + 1 + kColOff)] = val;
        if (c.nbr_r[kYp] && i < c.nx) c.nbr_r[kYp][(size_t)(i + 1) * c.nbr_pitch[kYp] + kColOff] = val;
      }
      __syncthreads();
      if (threadIdx.x == 32 && has_nbr) {
        fence_acq_rel_sys();
        for (int d = 0; d < 4; ++d)
          if (c.nbr_r[d]) red_add_relaxed_sys_u64(c.nbr_halo_cnt[d], 1ull);
      }
    }
    if (mode == 1) {
      const int gtid = cta * kThreads + threadIdx.x, gsz = ncta * kThreads;
      const double val = (double)(it + 1);
      for (int i = gtid; i < edge; i += gsz) {
        if (c.nbr_r[kXm] && i < c.ny) c.nbr_r[kXm][(size_t)(c.nbr_nx[kXm] + 1) * c.nbr_pitch[kXm] + (i + 1 + kColOff)] = val;
        if (c.nbr_r[kXp] && i < c.ny) c.nbr_r[kXp][(size_t)(i + 1 + kColOff)] = val;
        if (c.nbr_r[kYm] && i < c.nx) c.nbr_r[kYm][(size_t)(i + 1) * c.nbr_pitch[kYm] + (c.nbr_ny[kYm] + 1 + kColOff)] = val;
        if (c.nbr_r[kYp] && i < c.nx) c.nbr_r[kYp][(size_t)(i + 1) * c.nbr_pitch[kYp] + kColOff] = val;
      }
    }
    if (mode == 4 && cta < npush && has_nbr) {
      const unsigned int hs = hseq + (unsigned int)it + 1u;
      const int gtid = cta * kThreads + threadIdx.x, gsz = npush * kThreads;
      unsigned long long w0, w1;
      ll_encode((double)(it + 1), hs, w0, w1);
      for (int i = gtid; i < edge; i += gsz) {
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          if (!c.nbr_r[d] || i >= (d < 2 ? c.ny : c.nx)) continue;
          unsigned long long* dst = c.nbr_ll[d] + ((size_t)(hs & 1u) * (size_t)c.nbr_ll_cap[d] + (size_t)i) * 2;
          st_relaxed_sys_u64(dst, w0);
          st_relaxed_sys_u64(dst + 1, w1);
        }
      }
      if (threadIdx.x < 32) {  // warp 0: this CTA's share of the incoming records
        for (int d = 0; d < 4; ++d) {
          if (!c.nbr_r[d]) continue;
          const unsigned long long* line = c.halo_ll + ((size_t)d * 2 + (hs & 1u)) * (size_t)c.ll_cap * 2;
          const int n = min(edge, d < 2 ? c.ny : c.nx);
          for (int i = cta * 32 + (int)threadIdx.x; i < n; i += npush * 32) {
            double got[1];
            // a record that never arrives shows up in the checksum, not in the (CTA-uniform) loop condition
            if (!ll_poll<1, true>(line + (size_t)i * 2, 2, hs, wd_limit, got) || got[0] != (double)(it + 1)) check = -1e30;
          }
        }
      }
    }
    double v[2] = {threadIdx.x == 0 ? 1.0 : 0.0, 0.0};
    ++seq;
    arrive_target += ncta;
    const RedResult rr = mode == 5 ? star_sum<2, true>(c, v, seq, cta, ncta, wd_limit, s_red, tr, pair_mode & 1)
                         : mode >= 3 ? star_sum<2>(c, v, seq, cta, ncta, wd_limit, s_red, nullptr, pair_mode & 1)
                                   : cross_rank_sum<2>(c, v, seq, arrive_target, cta, ncta, s_red, mode == 1 && has_nbr,
                                                       mode == 1 && c.sys_scope != 0);
    ok = rr.ok;
    check += rr.v[0];
    if (mode == 2 && cta == 0 && threadIdx.x == 0) {
      for (int d = 0; d < 4; ++d) {
        if (!c.nbr_r[d]) continue;
        // every rank has the same number of pushing CTAs, so each neighbour adds npush per round
        const unsigned long long want = base[d] + (unsigned long long)(it + 1) * (unsigned long long)npush;
        unsigned long long spins = 0;
        while (ld_acquire_sys_u64(c.halo_cnt + d) < want && ++spins < (1ull << 26)) {
        }
      }
    }
  }
  if (mode == 5 && threadIdx.x == 0) {
    double* mine = reinterpret_cast<double*>(reinterpret_cast<char*>(c.ctl) + kCtaTimesOffset) + kCtaTimeWords * cta;
    for (int i = 0; i < 8; ++i) mine[i] = (double)tr[i] / (double)max(1, iters);
  }
  star_pair_exit(pair_mode & 1);
  if (cta == 0 && threadIdx.x == 0) {
    c.ctl->t_total = globaltimer_ns() - t0;
    c.ctl->red_seq = seq;
    c.ctl->halo_seq = hseq + (unsigned int)iters + 1u;
    c.ctl->red[0] = check;  // = iters * nranks * ncta when every reduction was complete
    c.ctl->status = ok ? PB_STATUS_CONVERGED : PB_STATUS_WATCHDOG;
  }
}

cudaError_t launch_xgpu_bench(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int iters, int mode, int edge,
                              cudaStream_t stream) {
  int pair = 0;
  void* args[] = {(void*)&d_ctxs, (void*)&ctas_per_rank, (void*)&iters, (void*)&mode, (void*)&edge, (void*)&pair};
  if (mode < 3)  // the counter + aggregator reduction knows nothing about clusters
    return cudaLaunchCooperativeKernel((const void*)xgpu_bench_kernel, dim3(nvirt * ctas_per_rank), dim3(kThreads), args, 0, stream);
  return launch_cooperative_pairs((const void*)xgpu_bench_kernel, nvirt, ctas_per_rank, kThreads, args, &pair, 0, stream);
}

// Cooperative launch of nvirt * ctas_per_rank CTAs; with PB_CLUSTERS=1 in the environment as thread-block clusters of
// two when every rank has an even number of CTAs (*pair_arg, which must be one of the kernel's arguments, is set to 1
// then), falling back to the plain launch when the clustered one is refused.  Off by default: measured on a B200
// the pair level makes the reduction slower inside the real kernels (3.52 against 3.31 us per reduction with 148 CTAs
// of 512 threads; 800 x 1200: 23.5 against 22.3 us per iteration; 4096 x 512: 36.9 against 35.7), although the
// stand-alone model of it (bench/experiments/hop_latency.cu, one warp per CTA) gains 0.6-0.8 us per round.
cudaError_t launch_cooperative_pairs(const void* fn, int nvirt, int ctas_per_rank, int threads, void** args, int* pair_arg,
                                     size_t smem, cudaStream_t stream) {
  static const bool enabled = std::getenv("PB_CLUSTERS") != nullptr;
  static const int star_everywhere = std::getenv("PB_MULTIRANK_STAR") != nullptr ? 2 : 0;  // A/B switch, bit 1 of the flag
  *pair_arg = star_everywhere;
  if (enabled && ctas_per_rank >= 2 && (ctas_per_rank & 1) == 0) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nvirt * ctas_per_rank);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = 2, attr[1].val.clusterDim.y = 1, attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    *pair_arg = 1 | star_everywhere;
    const cudaError_t e = cudaLaunchKernelExC(&cfg, fn, args);
    if (e == cudaSuccess) return e;
    cudaGetLastError();  // refused (e.g. not all clusters co-resident): fall back to single CTAs
    *pair_arg = star_everywhere;
  }
  return cudaLaunchCooperativeKernel(fn, dim3(nvirt * ctas_per_rank), dim3(threads), args, smem, stream);
}

size_t persistent_smem_bytes() { return sizeof(double) * kSmemTile; }

static const void* persistent_fn(int div) {
  return div == kDivTrue ? (const void*)pcg_persistent_kernel<kDivTrue>
                         : (div == kDivMarkstein ? (const void*)pcg_persistent_kernel<kDivMarkstein>
                                                 : (const void*)pcg_persistent_kernel<kDivRecip>);
}

int persistent_max_ctas_per_sm(int div) {
  int n = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, persistent_fn(div), kThreads, persistent_smem_bytes());
  return n;
}

cudaError_t launch_persistent(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int div, cudaStream_t stream) {
  void* args[] = {(void*)&d_ctxs, (void*)&ctas_per_rank};
  const dim3 grid(nvirt * ctas_per_rank), block(kThreads);
  const void* fn = persistent_fn(div);
  return cudaLaunchCooperativeKernel(fn, grid, block, args, persistent_smem_bytes(), stream);
}

}  // namespace pb
