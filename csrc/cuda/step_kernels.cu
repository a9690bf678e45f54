// step_kernels.cu -- the iteration as separate launches (one per phase), with all scalars on the
// device.  Used by the library-collective baseline (NCCL / gloo halo exchange and all-reduce between
// the launches), by the single-GPU step path and by the kernel unit tests.
//
// Kernel map to the reference: stencil = apply_A_kernel S4:507-536 (+ dot_kernel(Ap,p) S4:865),
// update = update_w_r_kernel S4:626-660 + apply_Dinv_kernel S4:541-562 + dot_kernel(z,r) S4:919,
// update_p = update_p_kernel S4:663-676, reduce_local = the host-side partial sums S4:771-796,
// scalars = the host arithmetic on alpha / beta / stop (S4:875-876, S4:896-905, S4:930-931),
// pack/unpack = the edge staging of exchange_halos_2d_gpu S4:358-402, S4:431-499.
#include "launch.h"
#include "phases.cuh"

namespace pb {

namespace {
__device__ __forceinline__ bool running(const RankCtx& c) { return ld_relaxed_gpu_s32(&c.ctl->status) == PB_STATUS_RUNNING; }

// per-CTA partials -> partials[blockIdx]
template <int NV>
__device__ __forceinline__ void store_partials(const RankCtx& c, double (&v)[NV], double* scratch) {
  block_sum<NV>(v, scratch);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) c.partials[(size_t)blockIdx.x * kPartialStride + k] = v[k];
  }
}
}  // namespace

__global__ void step_begin_kernel(const RankCtx* ctx) {
  Control* ctl = ctx->ctl;
  ctl->status = PB_STATUS_RUNNING;
  ctl->iter = 0;
  ctl->beta = 0.0;
  ctl->alpha = 0.0;
  ctl->cur = 0;
  ctl->diff_norm = 0.0;
}

__global__ void __launch_bounds__(kThreads) step_dot_zr_kernel(const RankCtx* __restrict__ ctx) {
  __shared__ double s_red[64];
  const RankCtx& c = *ctx;
  double v[1] = {0.0};
  for (int t = blockIdx.x; t < c.tiles_x * c.tiles_y; t += gridDim.x) v[0] += dot_zr_tile(c, t);
  store_partials<1>(c, v, s_red);
}

template <bool FUSE_P, int DIV>
__global__ void __launch_bounds__(kThreads) step_stencil_kernel(const RankCtx* __restrict__ ctx) {
  extern __shared__ double s_tile[];
  __shared__ double s_red[64];
  const RankCtx& c = *ctx;
  if (!running(c)) return;
  const int cur = c.ctl->cur;
  // fused: read p[cur], write p[cur^1]; classic: p[cur] is complete (ring filled by the exchange)
  const double beta = c.ctl->beta;
  double v[1] = {0.0};
  for (int t = blockIdx.x; t < c.tiles_x * c.tiles_y; t += gridDim.x)
    v[0] += phase_ab_tile<FUSE_P, DIV>(c, t, beta, c.p[cur], c.p[cur ^ 1], s_tile);
  store_partials<1>(c, v, s_red);
}

template <bool FUSE_P>
__global__ void __launch_bounds__(kThreads) step_update_kernel(const RankCtx* __restrict__ ctx) {
  __shared__ double s_red[64];
  const RankCtx& c = *ctx;
  if (!running(c)) return;
  const int cur = c.ctl->cur;
  double v[2] = {0.0, 0.0};
  const double alpha = c.ctl->alpha;
  for (int t = blockIdx.x; t < c.tiles_x * c.tiles_y; t += gridDim.x)
    phase_c_tile(c, t, alpha, FUSE_P ? c.p[cur ^ 1] : c.p[cur], v[0], v[1]);
  store_partials<2>(c, v, s_red);
}

__global__ void __launch_bounds__(kThreads) step_update_p_kernel(const RankCtx* __restrict__ ctx) {
  const RankCtx& c = *ctx;
  if (!running(c)) return;
  double* p = c.p[c.ctl->cur];
  const double beta = c.ctl->beta;
  const int col = threadIdx.x % kTileY + 1;
  for (int tile = blockIdx.x; tile < c.tiles_x * c.tiles_y; tile += gridDim.x) {
    const TileGeom t = tile_geom(c, tile);
    if (col > t.cy) continue;
    for (int row = threadIdx.x / kTileY + 1; row <= t.cx; row += blockDim.x / kTileY) {
      const size_t g = (size_t)(t.li0 - 1 + row) * c.pitch + (t.lj0 - 1 + col + kColOff);
      p[g] = fma(beta, p[g], safe_div(c.r[g], c.D[g]));
    }
  }
}

// One CTA: partials[0..n) -> ctl->red[0..NV)
__global__ void __launch_bounds__(kThreads) step_reduce_local_kernel(const RankCtx* __restrict__ ctx, int n, int nv) {
  __shared__ double s_red[64];
  const RankCtx& c = *ctx;
  if (!running(c)) return;
  double acc[2] = {0.0, 0.0};
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    acc[0] += c.partials[(size_t)i * kPartialStride];
    if (nv > 1) acc[1] += c.partials[(size_t)i * kPartialStride + 1];
  }
  block_sum<2>(acc, s_red);
  if (threadIdx.x == 0) {
    c.ctl->red[0] = acc[0];
    c.ctl->red[1] = nv > 1 ? acc[1] : 0.0;
  }
}

__global__ void step_scalars_kernel(const RankCtx* ctx, int phase) {
  const RankCtx& c = *ctx;
  Control* ctl = c.ctl;
  if (ctl->status != PB_STATUS_RUNNING) return;
  const double h1h2 = c.g.h1h2;
  if (phase == kStepInit) {
    ctl->zr_old = ctl->red[0] * h1h2;
  } else if (phase == kStepAlpha) {
    ctl->iter += 1;
    const double denom = ctl->red[0] * h1h2;
    ctl->denom = denom;
    if (fabs(denom) < kBreakdownTol) {
      ctl->status = PB_STATUS_BREAKDOWN;
      return;
    }
    ctl->alpha = ctl->zr_old / denom;
  } else {
    const double diff2 = ctl->red[0];
    const double zr_new = ctl->red[1] * h1h2;
    ctl->diff_norm = ctl->norm == PB_NORM_WEIGHTED ? sqrt(diff2 * h1h2) : sqrt(diff2);
    ctl->zr_new = zr_new;
    if (ctl->diff_norm < ctl->delta) {
      ctl->status = PB_STATUS_CONVERGED;
      return;
    }
    if ((long long)ctl->iter >= ctl->max_iter) {
      ctl->status = PB_STATUS_MAXITER;
      return;
    }
    ctl->beta = zr_new / ctl->zr_old;
    ctl->zr_old = zr_new;
  }
}

// fused-path bookkeeping: after phase C the new direction lives in the other buffer
__global__ void step_flip_kernel(const RankCtx* ctx) {
  Control* ctl = ctx->ctl;
  if (ctl->status == PB_STATUS_RUNNING) ctl->cur ^= 1;
}

// Edge staging for library halo exchanges.  Buffer layout: [x- | x+ | y- | y+], each of max(nx,ny).
__global__ void pack_edges_kernel(const RankCtx* __restrict__ ctx, double* __restrict__ buf, int stride) {
  const RankCtx& c = *ctx;
  const double* p = c.p[c.ctl->cur];
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < c.ny) {
    buf[0 * stride + t] = p[(size_t)1 * c.pitch + (t + 1 + kColOff)];
    buf[1 * stride + t] = p[(size_t)c.nx * c.pitch + (t + 1 + kColOff)];
  }
  if (t < c.nx) {
    buf[2 * stride + t] = p[(size_t)(t + 1) * c.pitch + (1 + kColOff)];
    buf[3 * stride + t] = p[(size_t)(t + 1) * c.pitch + (c.ny + kColOff)];
  }
}

__global__ void unpack_edges_kernel(const RankCtx* __restrict__ ctx, const double* __restrict__ buf, int stride) {
  const RankCtx& c = *ctx;
  double* p = c.p[c.ctl->cur];
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < c.ny) {
    if (c.nbr[kXm] >= 0) p[(size_t)0 * c.pitch + (t + 1 + kColOff)] = buf[0 * stride + t];
    if (c.nbr[kXp] >= 0) p[(size_t)(c.nx + 1) * c.pitch + (t + 1 + kColOff)] = buf[1 * stride + t];
  }
  if (t < c.nx) {
    if (c.nbr[kYm] >= 0) p[(size_t)(t + 1) * c.pitch + (0 + kColOff)] = buf[2 * stride + t];
    if (c.nbr[kYp] >= 0) p[(size_t)(t + 1) * c.pitch + (c.ny + 1 + kColOff)] = buf[3 * stride + t];
  }
}

// Debug / unit test: how often do the cheap division sequences differ from x / h in the last bit?
// counts[0] = samples, [1] = one-correction mismatches, [2] = two-correction mismatches, [3] = x*y mismatches.
__global__ void div_check_kernel(double h, unsigned long long n, unsigned long long seed, unsigned long long* counts) {
  const double y = 1.0 / h;
  unsigned long long bad1 = 0, bad2 = 0, bad0 = 0, cnt = 0;
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    unsigned long long s = (i + seed) * 6364136223846793005ull + 1442695040888963407ull;
    s ^= s >> 29;
    s *= 0xbf58476d1ce4e5b9ull;
    s ^= s >> 32;
    // mantissa random, exponent in [-40, 40], random sign
    const double m = 1.0 + (double)(s >> 12) * (1.0 / 4503599627370496.0);
    const int e = (int)((s >> 3) % 81) - 40;
    const double x = ((s & 1) ? -1.0 : 1.0) * scalbn(m, e);
    const double t = x / h;
    const double q0 = x * y;
    const double q1 = fma(fma(-q0, h, x), y, q0);
    const double q2 = fma(fma(-q1, h, x), y, q1);
    bad0 += (q0 != t), bad1 += (q1 != t), bad2 += (q2 != t), ++cnt;
  }
  atomicAdd(&counts[0], cnt), atomicAdd(&counts[1], bad1), atomicAdd(&counts[2], bad2), atomicAdd(&counts[3], bad0);
}

cudaError_t launch_div_check(double h, unsigned long long n, unsigned long long seed, unsigned long long* counts,
                             cudaStream_t s) {
  div_check_kernel<<<592, 256, 0, s>>>(h, n, seed, counts);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
static int step_grid(const RankCtx& h) {
  const int n = h.tiles_x * h.tiles_y;
  return n < 592 ? n : 592;
}

cudaError_t launch_step_begin(const RankCtx* d_ctx, cudaStream_t s) {
  step_begin_kernel<<<1, 1, 0, s>>>(d_ctx);
  return cudaGetLastError();
}
cudaError_t launch_step_dot_zr(const RankCtx* d_ctx, const RankCtx& h, cudaStream_t s) {
  step_dot_zr_kernel<<<step_grid(h), kThreads, 0, s>>>(d_ctx);
  return cudaGetLastError();
}
template <bool FUSE_P>
static void launch_stencil_div(const RankCtx* d_ctx, int n, size_t sm, int div, cudaStream_t s) {
  if (div == kDivTrue)
    step_stencil_kernel<FUSE_P, kDivTrue><<<n, kThreads, sm, s>>>(d_ctx);
  else if (div == kDivMarkstein)
    step_stencil_kernel<FUSE_P, kDivMarkstein><<<n, kThreads, sm, s>>>(d_ctx);
  else
    step_stencil_kernel<FUSE_P, kDivRecip><<<n, kThreads, sm, s>>>(d_ctx);
}
cudaError_t launch_step_stencil(const RankCtx* d_ctx, const RankCtx& h, bool fuse_p, int div, cudaStream_t s) {
  const int n = step_grid(h);
  const size_t sm = sizeof(double) * kSmemTile;
  if (fuse_p)
    launch_stencil_div<true>(d_ctx, n, sm, div, s);
  else
    launch_stencil_div<false>(d_ctx, n, sm, div, s);
  return cudaGetLastError();
}
cudaError_t launch_step_update(const RankCtx* d_ctx, const RankCtx& h, bool fuse_p, cudaStream_t s) {
  const int n = step_grid(h);
  if (fuse_p) {
    step_update_kernel<true><<<n, kThreads, 0, s>>>(d_ctx);
    step_flip_kernel<<<1, 1, 0, s>>>(d_ctx);
  } else {
    step_update_kernel<false><<<n, kThreads, 0, s>>>(d_ctx);
  }
  return cudaGetLastError();
}
cudaError_t launch_step_update_p(const RankCtx* d_ctx, const RankCtx& h, cudaStream_t s) {
  step_update_p_kernel<<<step_grid(h), kThreads, 0, s>>>(d_ctx);
  return cudaGetLastError();
}
cudaError_t launch_step_reduce_local(const RankCtx* d_ctx, const RankCtx& h, int nvalues, cudaStream_t s) {
  step_reduce_local_kernel<<<1, kThreads, 0, s>>>(d_ctx, step_grid(h), nvalues);
  return cudaGetLastError();
}
cudaError_t launch_step_scalars(const RankCtx* d_ctx, int phase, cudaStream_t s) {
  step_scalars_kernel<<<1, 1, 0, s>>>(d_ctx, phase);
  return cudaGetLastError();
}
cudaError_t launch_pack_edges(const RankCtx* d_ctx, const RankCtx& h, double* buf, int stride, cudaStream_t s) {
  const int n = h.nx > h.ny ? h.nx : h.ny;
  pack_edges_kernel<<<(n + 255) / 256, 256, 0, s>>>(d_ctx, buf, stride);
  return cudaGetLastError();
}
cudaError_t launch_unpack_edges(const RankCtx* d_ctx, const RankCtx& h, const double* buf, int stride,
                                cudaStream_t s) {
  const int n = h.nx > h.ny ? h.nx : h.ny;
  unpack_edges_kernel<<<(n + 255) / 256, 256, 0, s>>>(d_ctx, buf, stride);
  return cudaGetLastError();
}

}  // namespace pb
