// launch.h -- host-callable launchers of the device kernels (one declaration per kernel family).
#pragma once
#include <cuda_runtime.h>

#include "ctx.h"

namespace pb {

// K0 (assemble.cu): a, b, D, r = B, w = 0, p = 0, Ap = 0 on the block and its ring.
cudaError_t launch_assemble(const RankCtx* d_ctx, const RankCtx& h_ctx, cudaStream_t stream);
// error report partials -> Control::err (assemble.cu)
cudaError_t launch_error_report(const RankCtx* d_ctx, const RankCtx& h_ctx, cudaStream_t stream);
// gather the owned nodes of w into a dense (nx x ny) buffer with row stride `ld` (assemble.cu)
cudaError_t launch_gather_interior(const double* w, int nx, int ny, int pitch, double* out, long long ld,
                                   cudaStream_t stream);

// cooperative launch, as clusters of two CTAs where possible (persistent.cu); *pair_arg (one of the kernel's
// arguments) tells the kernel which it got
cudaError_t launch_cooperative_pairs(const void* fn, int nvirt, int ctas_per_rank, int threads, void** args, int* pair_arg,
                                     size_t smem, cudaStream_t stream);
// communication micro-benchmark (persistent.cu): reductions alone (mode 0) or halo stores + reduction (mode 1)
cudaError_t launch_xgpu_bench(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int iters, int mode, int edge,
                              cudaStream_t stream);
// `div` everywhere: 0 reciprocal multiply, 1 true division, 2 Markstein (bit-identical to 1), see phases.cuh
// KP (persistent.cu)
size_t persistent_smem_bytes();
int persistent_max_ctas_per_sm(int div);
cudaError_t launch_persistent(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int div, cudaStream_t stream);

// KP-TMA (persistent_tma.cu): tile_rows is 32 (2 stages), 16 (4 stages) or 64 (32-row tiles, pairs of full
// constant-coefficient tiles merged into tall items)
int persistent_tma_max_items();  // work items one CTA can hold
cudaError_t launch_persistent_tma(const RankCtx* d_ctxs, int nvirt, int ctas_per_rank, int tile_rows, int div,
                                  int l2_hints, bool compressed, bool recompute_ap, bool balance, bool rows_ab,
                                  cudaStream_t stream);
cudaError_t launch_tma_probe(const void* dmaps, const void* hmaps, int idx, int mode, int x, int y, unsigned int nbytes,
                             double* out, int* out_status, cudaStream_t stream);
bool build_tensor_maps(const RankCtx& h_ctx, int tile_rows, void* out12);
size_t tensor_map_bytes();

// KR (resident.cu): blocks that fit into the SMs' shared memory, one sub-block per CTA
struct ResidentCut {
  int sx, sy_log2, bx, by;  // sub-block rows, log2 of its columns, sub-blocks along x and y (bx * by CTAs per rank)
};
bool resident_partition(int nx, int ny, int max_ctas, size_t max_smem_bytes, ResidentCut* out);
size_t resident_smem_bytes(const ResidentCut& cut);
cudaError_t launch_resident(const RankCtx* d_ctxs, int nvirt, const ResidentCut& cut, int div, cudaStream_t stream);

// step-wise kernels (step_kernels.cu); every kernel returns immediately when Control::status is not
// PB_STATUS_RUNNING, so the host may enqueue iterations ahead of the stop test.
enum StepPhase : int { kStepInit = 0, kStepAlpha = 1, kStepBeta = 2 };
cudaError_t launch_step_begin(const RankCtx* d_ctx, cudaStream_t stream);  // status = RUNNING, iter = 0
cudaError_t launch_step_dot_zr(const RankCtx* d_ctx, const RankCtx& h_ctx, cudaStream_t stream);
cudaError_t launch_step_stencil(const RankCtx* d_ctx, const RankCtx& h_ctx, bool fuse_p, int div,
                                cudaStream_t stream);
cudaError_t launch_step_update(const RankCtx* d_ctx, const RankCtx& h_ctx, bool fuse_p, cudaStream_t stream);
cudaError_t launch_step_update_p(const RankCtx* d_ctx, const RankCtx& h_ctx, cudaStream_t stream);
// local partial sums -> Control::red[0..1]
cudaError_t launch_step_reduce_local(const RankCtx* d_ctx, const RankCtx& h_ctx, int nvalues, cudaStream_t stream);
// Control::red (global sums) -> alpha / beta / stop decision
cudaError_t launch_step_scalars(const RankCtx* d_ctx, int phase, cudaStream_t stream);
// pack / unpack the four edges of p for a library (NCCL / gloo) halo exchange
cudaError_t launch_pack_edges(const RankCtx* d_ctx, const RankCtx& h_ctx, double* sendbuf, int stride,
                              cudaStream_t stream);
cudaError_t launch_unpack_edges(const RankCtx* d_ctx, const RankCtx& h_ctx, const double* recvbuf, int stride,
                                cudaStream_t stream);

// stand-alone Ap = A p with fused (Ap, p) (apply_a_op.cu); scratch: apply_a_tiles(nx, ny) doubles
int apply_a_tiles(int nx, int ny);
cudaError_t launch_apply_a_dot(int M, int N, int nx, int ny, int pitch, const double* p, const double* a,
                               const double* b, double* Ap, double* dot_out, double* scratch, cudaStream_t stream);
// unit test of the division sequences (step_kernels.cu)
cudaError_t launch_div_check(double h, unsigned long long n, unsigned long long seed, unsigned long long* counts,
                             cudaStream_t stream);

}  // namespace pb
