// solver_run.cu -- Solver::solve and friends (the part of solver.h that enqueues work).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../api/error.h"
#include "../cuda/launch.h"
#include "solver.h"

namespace pb {

// How phase C gets Ap on constant-coefficient tiles.  Evaluating the stencil again instead of storing and
// reloading Ap trades 16 B/node/iteration for one more stencil evaluation: measured -9 % to -12 % per
// iteration once the arrays an iteration touches (w, r, p, p', Ap) no longer fit into L2 (3.1 M ... 67 M nodes
// per GPU), +3.5 % when they do (2.1 M nodes and below: the phase is latency-bound there, not byte-bound).
bool Solver::ap_recompute(int ap_mode) const {
  if (ap_mode != 0) return ap_mode == 2;
  size_t touched = 0;
  for (const LocalBlock& b : blocks_) touched += (size_t)(b.nx + 2) * (size_t)b.pitch * sizeof(double) * 5;
  return touched > (size_t)l2_bytes_ / 5 * 4;
}

bool Solver::resident_cut(ResidentCut* cut) const {
  // one cut for all ranks of the solve (every rank must take the same decision, and one launch serves all local
  // ranks): sized for the largest block of the whole grid, which is block 0
  const LocalBlock b0 = make_local_block(grid_.M, grid_.N, pg_, 0);
  const int nx = b0.nx, ny = b0.ny;
  return resident_partition(nx, ny, std::min(sm_count_ / nlocal(), kMaxCtas), (size_t)(227 - 10) * 1024, cut);
}

void Solver::solve_persistent(const SolveOptions& opt, cudaStream_t stream) {
  if (!fabric_complete_) throw std::runtime_error("persistent solve needs every rank's arena (open_peer) first");
  const int exact = opt.exact_div;
  // Small blocks run in the resident kernel: no field data moves through L2 / HBM during the iteration.  It is
  // taken automatically up to 128 Ki nodes per rank, where it measured faster than the TMA kernel (100 x 1200: 11.7
  // against 12.1 us per iteration, 160 x 160: 10.5 against 11.9); above that the sub-blocks the ellipse's boundary
  // crosses (a third of the CTAs at 800 x 1200) take twice as long as the others and the TMA kernel wins (22.2
  // against 30.9 us), so there it runs only on request (method = resident).  Options that only exist for the TMA
  // kernel (tile height, Ap mode, streamed coefficients, L2 hints) pin that kernel, as does method = persistent_tma.
  ResidentCut cut{};
  const LocalBlock b0 = make_local_block(grid_.M, grid_.N, pg_, 0);
  const bool small = (long long)b0.nx * b0.ny <= 128 * 1024;
  const bool want_resident = opt.method == kMethodResident ||
                             (opt.method == kMethodPersistent && small && opt.tile_rows == 0 && opt.ap_mode == 0 &&
                              opt.compressed != 0 && opt.l2_hints == 0 && std::getenv("PB_NO_RESIDENT") == nullptr);
  if (want_resident && resident_cut(&cut)) {
    for (int l = 0; l < nlocal(); ++l) {
      PB_CUDA_CHECK(cudaMemsetAsync(hctx_[l].arrive, 0, 128, stream));
      PB_CUDA_CHECK(cudaMemsetAsync(control(l)->red, 0, sizeof(double) * 4, stream));
    }
    last_ctas_per_rank_ = cut.bx * cut.by;
    last_tile_rows_ = 0;
    last_kernel_ = "resident";
    PB_CUDA_CHECK(launch_resident(dctx_, nlocal(), cut, exact, stream));
    last_launches_ += 1;
    return;
  }
  if (opt.method == kMethodResident)
    throw std::runtime_error("method 'resident': the block does not fit into the shared memory of the SMs (at most 8192 "
                             "nodes per CTA); use method 'persistent'");
  const bool tma = opt.method != kMethodPersistentLdg;
  last_kernel_ = tma ? "tma" : "ldg";
  int per_sm = tma ? 1 : persistent_max_ctas_per_sm(exact);
  if (per_sm < 1) throw std::runtime_error("persistent kernel does not fit on an SM");
  if (!tma && opt.ctas_per_sm > 0) per_sm = std::min(per_sm, opt.ctas_per_sm);
  int max_tiles = 1;
  for (const RankCtx& c : hctx_) max_tiles = std::max(max_tiles, c.tiles_x * c.tiles_y);
  int ctas = (per_sm * sm_count_) / nlocal();
  ctas = std::min(ctas, kMaxCtas);
  ctas = std::min(ctas, max_tiles);
  if (tma && ctas > 2) ctas &= ~1;  // an even number of CTAs per rank: the kernel then runs as clusters of two
  if (ctas < 1) throw std::runtime_error("too many virtual ranks for one cooperative launch");
  last_ctas_per_rank_ = ctas;
  for (int l = 0; l < nlocal(); ++l) {
    PB_CUDA_CHECK(cudaMemsetAsync(hctx_[l].arrive, 0, 128, stream));
    PB_CUDA_CHECK(cudaMemsetAsync(control(l)->red, 0, sizeof(double) * 4, stream));
  }
  if (tma) {
    // Tile height: 16 and 32 as asked; 64 = 32-row tiles with pairs of full constant-coefficient tiles merged into
    // tall items, which halves the fixed cost per tile of a phase: in both phases when Ap is recomputed, in phase AB
    // only when it is stored (a tall phase-C stage has no room for p and Ap boxes; phase C then works on the two
    // halves).  Auto (0) takes 64 when Ap is recomputed and every CTA gets at least four 64-row super-tiles -- the
    // balancer cuts tall items in two where a whole one would overshoot a CTA's share -- else 32.  Measured: the
    // 4-GPU block of 4096^2 (1024 x 4096, 6.9 super-tiles per CTA) 62.0 against 64.9 us per iteration; with stored Ap
    // (L2-resident blocks) tall items in phase AB alone gain nothing (512 x 4096: 34.8 against 34.9 us on one GPU,
    // 41.6 against 40.0 us in the 8-GPU run), so they are only taken on request there.
    const bool recompute = ap_recompute(opt.ap_mode) && opt.compressed != 0;
    long long super = 0;
    for (const RankCtx& c : hctx_) super = std::max<long long>(super, (long long)((c.nx + 63) / 64) * ((c.ny + kTileY - 1) / kTileY));
    int tile_rows = opt.tile_rows == 16 ? 16 : (opt.tile_rows == 64 ? 64 : 32);
    if (opt.tile_rows == 0 && recompute && super >= 4LL * ctas) tile_rows = 64;
    const long long per_cta = (max_tiles * (tile_rows == 16 ? 2 : 1) + ctas - 1) / ctas;
    if (per_cta > persistent_tma_max_items())
      throw std::runtime_error("block too large for one GPU: " + std::to_string(per_cta) + " tiles per CTA, the persistent "
                               "kernel holds " + std::to_string(persistent_tma_max_items()) + "; use more GPUs");
    last_tile_rows_ = tile_rows;
    static const bool env_no_balance = std::getenv("PB_NO_BALANCE") != nullptr;  // A/B switch for measurements
    const char* env_rows = std::getenv("PB_AB_ROWS");  // A/B switch for measurements (read per launch): "0" = tile path
    const bool rows_ab = opt.ab_mode == 0 && !(env_rows && env_rows[0] == '0');
    PB_CUDA_CHECK(launch_persistent_tma(dctx_, nlocal(), ctas, tile_rows, exact, opt.l2_hints, opt.compressed != 0,
                                        ap_recompute(opt.ap_mode), !(env_no_balance || opt.no_balance), rows_ab, stream));
  } else
    PB_CUDA_CHECK(launch_persistent(dctx_, nlocal(), ctas, exact, stream));
  last_launches_ += 1;
}

// Separate launches per phase; only for a solver whose ranks are all local and alone (nranks == 1):
// distributed step-wise runs are driven from Python with a library collective between the launches.
void Solver::solve_steps(const SolveOptions& opt, cudaStream_t stream) {
  if (nranks() != 1) throw std::runtime_error("step-wise solve from C++ supports a single rank only");
  const bool fused = opt.method == kMethodStepFused;
  const int exact = opt.exact_div;
  const RankCtx& h = hctx_[0];
  const RankCtx* d = dctx_;
  const long long max_iter = pinned_opts_->max_iter;
  PB_CUDA_CHECK(launch_step_begin(d, stream));
  PB_CUDA_CHECK(launch_step_dot_zr(d, h, stream));
  PB_CUDA_CHECK(launch_step_reduce_local(d, h, 1, stream));
  PB_CUDA_CHECK(launch_step_scalars(d, kStepInit, stream));
  last_launches_ += 4;
  const int chunk = 64;
  int* pinned_status = reinterpret_cast<int*>(pinned_ctl_);  // reused as a 4-byte mailbox here
  for (long long k = 0; k < max_iter; k += chunk) {
    const long long n = std::min<long long>(chunk, max_iter - k);
    for (long long i = 0; i < n; ++i) {
      if (!fused) {  // p = r/D + beta p (beta = 0, p = 0 in the first iteration)
        PB_CUDA_CHECK(launch_step_update_p(d, h, stream));
        last_launches_ += 1;
      }
      PB_CUDA_CHECK(launch_step_stencil(d, h, fused, exact, stream));
      PB_CUDA_CHECK(launch_step_reduce_local(d, h, 1, stream));
      PB_CUDA_CHECK(launch_step_scalars(d, kStepAlpha, stream));
      PB_CUDA_CHECK(launch_step_update(d, h, fused, stream));
      PB_CUDA_CHECK(launch_step_reduce_local(d, h, 2, stream));
      PB_CUDA_CHECK(launch_step_scalars(d, kStepBeta, stream));
      last_launches_ += fused ? 7 : 6;
    }
    PB_CUDA_CHECK(cudaMemcpyAsync(pinned_status, &control(0)->status, sizeof(int), cudaMemcpyDeviceToHost, stream));
    PB_CUDA_CHECK(cudaStreamSynchronize(stream));
    if (*pinned_status != PB_STATUS_RUNNING) break;
  }
}

void Solver::enqueue(const SolveOptions& opt, cudaStream_t stream) {
  if (!finalized_) throw std::runtime_error("Solver::solve before finalize()");
  PB_CUDA_CHECK(cudaSetDevice(device_));
  last_launches_ = 0;
  last_method_ = opt.method;
  // 1. options: pinned host -> device (the only input of a solve besides M, N fixed at creation)
  pinned_opts_->delta = opt.delta;
  pinned_opts_->max_iter = opt.max_iter >= 0 ? opt.max_iter : (long long)(grid_.M - 1) * (long long)(grid_.N - 1);
  pinned_opts_->norm = opt.norm;
  pinned_opts_->pad = 0;
  pinned_opts_->watchdog_ns = opt.watchdog_s > 0.0 ? (unsigned long long)(opt.watchdog_s * 1e9) : 0ull;
  for (int l = 0; l < nlocal(); ++l)
    PB_CUDA_CHECK(cudaMemcpyAsync(&control(l)->delta, pinned_opts_, sizeof(DeviceOptions), cudaMemcpyHostToDevice,
                                  stream));
  // 2. assembly, iteration, error report -- all enqueued, no host synchronisation in between
  PB_CUDA_CHECK(cudaEventRecord(ev_[0], stream));
  for (int l = 0; l < nlocal(); ++l) PB_CUDA_CHECK(launch_assemble(dctx_ + l, hctx_[l], stream));
  last_launches_ += nlocal();
  PB_CUDA_CHECK(cudaEventRecord(ev_[1], stream));
  if (opt.method == kMethodPersistent || opt.method == kMethodPersistentLdg || opt.method == kMethodResident ||
      opt.method == kMethodPersistentTma)
    solve_persistent(opt, stream);
  else
    solve_steps(opt, stream);
  PB_CUDA_CHECK(cudaEventRecord(ev_[2], stream));
  if (opt.error_report) {
    for (int l = 0; l < nlocal(); ++l) PB_CUDA_CHECK(launch_error_report(dctx_ + l, hctx_[l], stream));
    last_launches_ += nlocal();
  }
  PB_CUDA_CHECK(cudaEventRecord(ev_[3], stream));
  // 3. results: device -> pinned host
  for (int l = 0; l < nlocal(); ++l)
    PB_CUDA_CHECK(cudaMemcpyAsync(pinned_ctl_ + l, control(l), sizeof(Control), cudaMemcpyDeviceToHost, stream));
}

void Solver::collect(cudaStream_t stream, pb_stats* out) {
  PB_CUDA_CHECK(cudaSetDevice(device_));
  PB_CUDA_CHECK(cudaStreamSynchronize(stream));
  if (!out) return;
  std::memset(out, 0, sizeof(*out));
  float ms_asm = 0, ms_iter = 0, ms_err = 0;
  PB_CUDA_CHECK(cudaEventElapsedTime(&ms_asm, ev_[0], ev_[1]));
  PB_CUDA_CHECK(cudaEventElapsedTime(&ms_iter, ev_[1], ev_[2]));
  PB_CUDA_CHECK(cudaEventElapsedTime(&ms_err, ev_[2], ev_[3]));
  const Control& c0 = pinned_ctl_[0];
  out->iters = c0.iter;
  out->status = c0.status;
  out->diff_norm = c0.diff_norm;
  double s_in = 0, s_all = 0, mx = 0, mw = -1e300, cnt = 0;
  unsigned long long t_ab = 0, t_c = 0, t_r1 = 0, t_r2 = 0, t_init = 0, t_halo = 0;
  for (int l = 0; l < nlocal(); ++l) {
    const Control& c = pinned_ctl_[l];
    if (c.status == PB_STATUS_WATCHDOG) out->status = PB_STATUS_WATCHDOG;
    s_in += c.err[0], s_all += c.err[1], cnt += c.err[4];
    mx = std::max(mx, c.err[2]), mw = std::max(mw, c.err[3]);
    t_ab = std::max(t_ab, c.t_ab), t_c = std::max(t_c, c.t_c), t_r1 = std::max(t_r1, c.t_red1);
    t_r2 = std::max(t_r2, c.t_red2), t_init = std::max(t_init, c.t_init), t_halo = std::max(t_halo, c.t_halo);
  }
  out->sum_e2_in_D = s_in;
  out->sum_e2_all = s_all;
  out->l2_in_D = std::sqrt(grid_.h1h2 * s_in);
  out->l2_all = std::sqrt(grid_.h1h2 * s_all);
  out->max_abs = mx;
  out->max_w = mw;
  out->nodes_in_D = (long long)cnt;
  out->seconds = 1e-3 * (double)(ms_asm + ms_iter + ms_err);
  out->t_assemble = 1e-3 * ms_asm;
  out->t_iterations = 1e-3 * ms_iter;
  if (last_method_ == kMethodPersistent || last_method_ == kMethodPersistentLdg || last_method_ == kMethodResident ||
      last_method_ == kMethodPersistentTma) {
    // The reference's five buckets (S4:695-700), fed from CTA 0's device clocks; every nanosecond of the loop is in
    // exactly one of them.  gpu = the two phases minus the halo wait inside them; comm = waiting for and unpacking
    // the neighbours' halo records (TMA kernel; the plain-load kernel publishes its halo inside reduction 2, so its
    // share cannot be separated and is 0 here); dot = the in-kernel reductions, cross-GPU part included; precond = 0:
    // r/D is a few instructions inside both phases and has no time of its own; copy = 0: nothing is copied.
    out->t_gpu = 1e-9 * (double)(t_ab + t_c - std::min(t_halo, t_ab + t_c));
    out->t_dot = 1e-9 * (double)(t_r1 + t_r2 + t_init);
    out->t_comm = 1e-9 * (double)t_halo;
    out->t_precond = 0.0;
    out->t_phase_ab = 1e-9 * (double)t_ab, out->t_phase_c = 1e-9 * (double)t_c;
    out->t_red1 = 1e-9 * (double)t_r1, out->t_red2 = 1e-9 * (double)t_r2;
  } else {
    out->t_gpu = 1e-3 * ms_iter;
  }
  out->t_copy = 0.0;
}

void Solver::solve(const SolveOptions& opt, cudaStream_t stream, pb_stats* out) {
  enqueue(opt, stream);
  collect(stream, out);
}

void Solver::xgpu_bench(int iters, int mode, int edge, int ctas, cudaStream_t stream, double out[2]) {
  if (!fabric_complete_) throw std::runtime_error("xgpu_bench needs every rank's arena");
  PB_CUDA_CHECK(cudaSetDevice(device_));
  if (ctas <= 0) ctas = sm_count_ / nlocal();
  ctas = std::min(ctas, kMaxCtas);
  if (ctas > 2) ctas &= ~1;
  last_ctas_per_rank_ = ctas;
  for (int l = 0; l < nlocal(); ++l) PB_CUDA_CHECK(cudaMemsetAsync(hctx_[l].arrive, 0, 128, stream));
  PB_CUDA_CHECK(launch_xgpu_bench(dctx_, nlocal(), ctas, iters, mode, edge, stream));
  PB_CUDA_CHECK(cudaMemcpyAsync(pinned_ctl_, control(0), sizeof(Control), cudaMemcpyDeviceToHost, stream));
  PB_CUDA_CHECK(cudaStreamSynchronize(stream));
  if (pinned_ctl_[0].status == PB_STATUS_WATCHDOG) throw std::runtime_error("xgpu_bench: watchdog fired");
  out[0] = (double)pinned_ctl_[0].t_total / (double)std::max(1, iters);
  out[1] = pinned_ctl_[0].red[0];
}

int Solver::cta_times(int l, double* host_out, int max_ctas) {
  PB_CUDA_CHECK(cudaSetDevice(device_));
  const int n = std::min(last_ctas_per_rank_, max_ctas);
  if (n <= 0) return 0;
  PB_CUDA_CHECK(cudaMemcpy2D(host_out, sizeof(double) * 4, arenas_[l] + kCtaTimesOffset, sizeof(double) * kCtaTimeWords,
                             sizeof(double) * 4, (size_t)n, cudaMemcpyDeviceToHost));
  for (int i = 0; i < 4 * n; ++i) host_out[i] *= 1e-9;
  return n;
}

// (after a resident-kernel solve words 4 and 5 hold the sub-block's coefficient class and node count instead)
int Solver::cta_report(int l, double* host_out, int max_ctas) {
  PB_CUDA_CHECK(cudaSetDevice(device_));
  const int n = std::min(last_ctas_per_rank_, max_ctas);
  if (n <= 0) return 0;
  PB_CUDA_CHECK(cudaMemcpy(host_out, arenas_[l] + kCtaTimesOffset, sizeof(double) * kCtaTimeWords * (size_t)n,
                           cudaMemcpyDeviceToHost));
  return n;
}

void Solver::tile_counts(long long out[3]) {
  PB_CUDA_CHECK(cudaSetDevice(device_));
  out[0] = out[1] = out[2] = 0;
  const int n = last_ctas_per_rank_;
  if (n <= 0 || std::string(last_kernel_) != "tma") return;
  std::vector<double> buf((size_t)n * kCtaTimeWords);
  for (int l = 0; l < nlocal(); ++l) {
    PB_CUDA_CHECK(cudaMemcpy(buf.data(), arenas_[l] + kCtaTimesOffset, sizeof(double) * buf.size(), cudaMemcpyDeviceToHost));
    for (int i = 0; i < n; ++i)
      for (int k = 0; k < 3; ++k) out[k] += (long long)buf[(size_t)i * kCtaTimeWords + 4 + k];
  }
}

void Solver::gather_block(int l, double* dev_out, long long ld, cudaStream_t stream) {
  const LocalBlock& b = blocks_[l];
  PB_CUDA_CHECK(launch_gather_interior(array(l, 0), b.nx, b.ny, b.pitch, dev_out, ld, stream));
}

void Solver::copy_solution_to_host(double* full, cudaStream_t stream) {
  PB_CUDA_CHECK(cudaSetDevice(device_));
  const int S = grid_.N + 1;
  for (int l = 0; l < nlocal(); ++l) {
    const LocalBlock& b = blocks_[l];
    // strided D2H: rows li = 1..nx, columns lj = 1..ny of w -> full[(i0 + li - 1) * S + j0 + lj - 1]
    PB_CUDA_CHECK(cudaMemcpy2DAsync(full + (size_t)b.i0 * S + b.j0, sizeof(double) * S,
                                    array(l, 0) + b.index(1, 1), sizeof(double) * b.pitch, sizeof(double) * b.ny,
                                    b.nx, cudaMemcpyDeviceToHost, stream));
  }
  PB_CUDA_CHECK(cudaStreamSynchronize(stream));
}

}  // namespace pb
