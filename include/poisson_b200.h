/* poisson_b200.h -- C ABI of libpoisson_b200.
 *
 * Everything the Python package, the CLI and third-party callers use goes through these functions.
 * All functions return 0 on success and a negative value on failure unless stated otherwise;
 * pb_last_error() gives the message of the last failure on the calling thread.
 *
 * Reference parity tags (SURVEY.md legend): S0..S4 are the stage sources of
 * mxy-kit/poisson-ellipse-openmp-mpi-cuda; each declaration names what it replaces.
 */
#ifndef POISSON_B200_H
#define POISSON_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PB_NORM_WEIGHTED 0   /* sqrt(h1 h2 sum dw^2)  -- stages 1-4 (S4:656-659, S4:896) */
#define PB_NORM_UNWEIGHTED 1 /* sqrt(sum dw^2)        -- stage 0   (S0:149-154)         */

#define PB_STATUS_MAXITER 0
#define PB_STATUS_CONVERGED 1
#define PB_STATUS_BREAKDOWN 2 /* |(Ap,p)| < 1e-15 (S4:875) */
#define PB_STATUS_RUNNING 3
#define PB_STATUS_WATCHDOG 4 /* a device-side wait timed out; the solve was abandoned */
#define PB_STATUS_TMA_TIMEOUT 5 /* a TMA pipeline stage never completed (internal error) */

/* Result of a solve.  Timers are seconds; for multi-rank runs the caller takes the max over ranks. */
typedef struct pb_stats {
  int iters;
  int status;
  double diff_norm;  /* last ||w(k+1)-w(k)|| in the selected norm */
  double l2_in_D;    /* sqrt(h1 h2 sum_{nodes in D} (w-u)^2) */
  double l2_all;     /* same over all interior nodes (u = 0 outside D) */
  double max_abs;    /* max |w-u| */
  double max_w;      /* max w */
  long long nodes_in_D;
  double seconds;    /* solver time: assembly + iterations (+ error report), device-timed on GPU */
  /* the reference's five buckets (S4:695-700), fed from device-side clocks on the GPU paths */
  double t_gpu;      /* stencil + update phases */
  double t_copy;     /* host<->device copies inside the solve (0 on the fused paths) */
  double t_comm;     /* waiting for / unpacking the neighbours' halo records (CPU stages: halo exchanges + all-reduces) */
  double t_precond;  /* 0 on the fused paths: z = r/D is a few instructions inside both phases, it has no time of its own */
  double t_dot;      /* the in-kernel reductions, cross-GPU part included */
  double t_assemble; /* coefficient / right-hand-side assembly */
  double t_iterations; /* the PCG loop alone (CUDA events) */
  double sum_e2_in_D; /* raw sums of squared errors over the local ranks, for cross-process reduction */
  double sum_e2_all;
  /* persistent kernels: device-clock split of the iteration (CTA 0 of the slowest rank) */
  double t_phase_ab, t_phase_c, t_red1, t_red2;
} pb_stats;

/* Local block of the 2-D decomposition in the padded layout (see csrc/core/decomp.h). */
typedef struct pb_block {
  int rank, px, py;
  int i0, j0; /* global indices of local (1,1) */
  int nx, ny; /* owned nodes */
  int pitch;  /* doubles per row */
  int nbr[4]; /* x-, x+, y-, y+ neighbour ranks or -1 */
} pb_block;

int pb_version(void);
const char* pb_last_error(void);

/* ---- decomposition (replaces choose_process_grid S4:82-86 and decompose_2d S4:97-133) ---- */
void pb_choose_proc_grid(int nranks, int* Px, int* Py);
/* Process grid and device order from a peer-to-peer performance matrix (perf[a*n+b], lower = better; see
 * csrc/core/decomp.h): uniform fabric -> n x 1 slabs, equal islands -> islands x members with the x-cuts between the
 * islands, else the reference's grid.  order[rank] = index of the device that runs `rank`.  Returns 0 / 1 / 2 for
 * these three cases, < 0 on bad arguments. */
int pb_choose_grid_topology(int n, const int* perf, int M, int* Px, int* Py, int* order);
int pb_local_block(int M, int N, int Px, int Py, int rank, pb_block* out);
int pb_col_offset(void); /* column of ring cell lj=0 inside a padded row */

/* ---- scalar problem definition (replaces S4:32-76, S4:162-191) ---- */
double pb_coef_a(int M, int N, int i, int j);
double pb_coef_b(int M, int N, int i, int j);
double pb_rhs(int M, int N, int i, int j);
double pb_diag(int M, int N, int i, int j);
double pb_chord(double fixed, double lo, double hi, int vertical);
double pb_exact(double x, double y);

/* ---- CPU solver: oracle and CPU baseline (replaces stages 0/1: S0:106-172, S1:133-199) ---- */
int pb_cpu_solve(int M, int N, int norm, double delta, long long max_iter, int threads,
                 int signed_breakdown, double* w_out /* (M+1)*(N+1) or NULL */, pb_stats* stats);
int pb_cpu_error_report(int M, int N, const double* w, pb_stats* stats);
/* Distributed CPU solver, host loop in C++ (replaces solve_mpi of stages 2 and 3: S2:356-460, S3:364-470).
 * Collective over the Px*Py processes of one node that pass the same `job` string (the transport is a POSIX
 * shared-memory segment named after it, csrc/runtime/shm_comm.h; the image has no MPI): this process is `rank` and
 * owns block `rank` of the grid.  threads > 1 = the MPI + OpenMP stage.  w_block, if not NULL, receives this rank's
 * block in the padded layout of pb_local_block ((nx+2) x pitch doubles).  stats are global (same on every rank):
 * seconds = barrier-to-barrier wall time of assembly + iterations, t_comm = halo exchanges + all-reduces (max over
 * ranks). */
int pb_cpu_dist_solve(int M, int N, int Px, int Py, int rank, const char* job, int norm, double delta, long long max_iter,
                      int threads, double* w_block /* or NULL */, pb_stats* stats);

/* ---- CPU local-block operations for the distributed CPU backend (replaces the per-rank loops of
 *      stages 2/3: S2:124-232, S2:418-456) ---- */
int pb_cpu_assemble_local(int M, int N, const pb_block* blk, double* a, double* b, double* D, double* rhs,
                          int threads);
double pb_cpu_apply_A_local(int M, int N, const pb_block* blk, const double* p, const double* a,
                            const double* b, double* Ap, int threads);
int pb_cpu_update_wr_local(const pb_block* blk, double* w, double* r, const double* p, const double* Ap,
                           const double* D, double alpha, double* out2, int threads);
int pb_cpu_update_p_local(const pb_block* blk, double* p, const double* r, const double* D, double beta,
                          int threads);
double pb_cpu_dot_zr_local(const pb_block* blk, const double* r, const double* D, int threads);
int pb_cpu_error_local(int M, int N, const pb_block* blk, const double* w, double* out5);

/* ---- GPU solver (replaces gradient_solver_mpi S4:687-982 and its kernels S4:507-676) ----
 * A pb_solver owns the device memory ("arena") of `nlocal` consecutive ranks of a Px x Py process grid on
 * one device.  Multi-GPU: one process per GPU creates a solver with nlocal = 1, exchanges the 64-byte IPC
 * handles with its peers (pb_solver_ipc_handle / pb_solver_open_peer), then calls pb_solver_finalize.
 * Single-GPU "virtual ranks": nlocal = Px*Py, no exchange needed. */
typedef struct pb_solver pb_solver;

#define PB_METHOD_PERSISTENT 0   /* one cooperative launch per solve, fused halo + reductions: the resident kernel for blocks
                                    of up to 128 Ki nodes per rank (unless tile_rows / ap_mode / l2_hints /
                                    full_coefficients are set), else the TMA kernel */
#define PB_METHOD_STEP_FUSED 1   /* separate launches, fused p-update + stencil (single rank) */
#define PB_METHOD_STEP_CLASSIC 2 /* separate launches, reference-like 3-kernel structure (single rank) */
#define PB_METHOD_PERSISTENT_LDG 3 /* persistent kernel with plain global loads instead of TMA pipelines */
#define PB_METHOD_RESIDENT 4       /* persistent kernel that keeps p, r, w of a small block in shared memory for the whole
                                      solve (<= 8192 nodes per CTA, i.e. ~1.2 M nodes per GPU); an error when it does not fit */
#define PB_METHOD_PERSISTENT_TMA 5 /* the TMA kernel even where PB_METHOD_PERSISTENT would take the resident one */

typedef struct pb_solve_options {
  int method;
  int norm;
  double delta;
  long long max_iter; /* < 0: (M-1)(N-1) as in S4:1006 */
  int exact_div;      /* divisions by h1/h2 in the stencil: 1 true IEEE division, 2 Markstein sequence (same bits as
                         1, three instructions), 0 multiplication by the reciprocal (last-bit differences) */
  int ctas_per_sm;    /* 0 = auto */
  int error_report;   /* 1: compute the error against the analytic solution after the solve */
  int tile_rows;      /* TMA persistent kernel: 32 (2 pipeline stages), 16 (4 stages), 64 (32-row tiles with pairs of full
                         constant-coefficient tiles merged into 64-row items; needs the recomputed Ap, else 32 is
                         used); 0 = auto */
  double watchdog_s;  /* 0 = default (20 s): device-side waits give up after this long */
  int l2_hints;       /* 1: TMA loads carry L2 eviction priorities (keep r, D, p; stream a, b, w, Ap) */
  int full_coefficients; /* 1: stream a, b, D as fp64 boxes instead of the 1-byte coefficient classes (A/B testing) */
  int ap_mode;        /* TMA persistent kernel, constant-coefficient tiles: 1 = phase AB stores Ap and phase C reloads
                         it; 2 = phase C recomputes Ap from the p box (same bits, 16 B/node less traffic);
                         0 = the library's default */
  int no_balance;     /* TMA persistent kernel: 1 = tiles go to CTAs cyclically; 0 = cost-balanced assignment (the
                         default; the sums are then taken in a different order, so results differ in the last bits from
                         a cyclic run) */
  int ab_mode;        /* TMA persistent kernel, phase AB on full tiles with constant coefficients.  0 (default): every
                         warp owns whole rows of the tile, evaluates p_new for them and for the row above and below in
                         registers and goes straight on to the stencil (no barrier between the two halves of the
                         phase, no round trip of p_new through shared memory).  1: the tile path (p update in place in
                         shared memory, barrier, stencil), which is what mixed and partial tiles always take.  Same
                         bits for p_new and Ap either way; the order of the per-thread partial sums of (Ap, p)
                         differs, so the two agree to rounding. */
} pb_solve_options;

int pb_cuda_device_count(void);
/* Runtime bootstrap for `n` devices, the counterpart of the reference's MPI_Init + launcher GPU binding (S4:986-990,
 * SURVEY.md section 0.7): creates the primary contexts and enables peer access between every pair, one host thread
 * per device.  Idempotent; pb_group_create calls it itself. */
int pb_runtime_bootstrap(const int* devices, int n);
/* Peer-to-peer topology of `n` devices as CUDA reports it: perf[a*n+b] = cudaDevP2PAttrPerformanceRank of a -> b
 * (lower is better), access / atomics = cudaDevP2PAttrAccessSupported / NativeAtomicSupported; diagonals are 0 / 1 / 1. */
int pb_p2p_matrix(const int* devices, int n, int* perf, int* access, int* atomics);
pb_solver* pb_solver_create(int M, int N, int Px, int Py, int first_rank, int nlocal, int device);
void pb_solver_destroy(pb_solver* s);
int pb_solver_ipc_handle(pb_solver* s, int local_idx, void* out64);
int pb_solver_open_peer(pb_solver* s, int rank, const void* handle64);
int pb_solver_set_peer_base(pb_solver* s, int rank, void* arena_base);
void* pb_solver_arena(pb_solver* s, int local_idx);
int pb_solver_finalize(pb_solver* s);
int pb_solver_solve(pb_solver* s, const pb_solve_options* opt, void* cuda_stream, pb_stats* stats);
int pb_solver_block(pb_solver* s, int local_idx, pb_block* out);
/* device pointer of a local array: 0 w, 1 r, 2 p0, 3 p1, 4 Ap, 5 a, 6 b, 7 D; (nx+2) x pitch doubles */
void* pb_solver_array(pb_solver* s, int local_idx, int which);
void* pb_solver_control_red(pb_solver* s, int local_idx); /* device double[4]: staging of step reductions */
void* pb_solver_control_err(pb_solver* s, int local_idx); /* device double[5]: error report partials */
int pb_solver_copy_solution(pb_solver* s, double* host_full /* (M+1)*(N+1) */, void* cuda_stream);
int pb_solver_gather_block(pb_solver* s, int local_idx, void* dev_out, long long ld, void* cuda_stream);
int pb_solver_info(pb_solver* s, int* ctas_per_rank, long long* launches, long long* h2d_bytes, long long* d2h_bytes);
/* Per-CTA device clocks of the last persistent (TMA) solve: out[4 * cta + {0,1,2,3}] = seconds in phase AB, phase C,
   waiting in reduction 1, waiting in reduction 2.  Returns the number of CTAs written (<= max_ctas) or < 0. */
int pb_solver_cta_times(pb_solver* s, int local_idx, double* host_out, int max_ctas);
/* The raw per-CTA report area: 8 doubles per CTA (after a solve: ns in AB, C, reduction 1, reduction 2, the three tile
   counts below, ns waiting for halo records; after pb_solver_xgpu_bench mode 5: mean clock cycles per step of the
   reduction).  Returns the number of CTAs written or < 0. */
int pb_solver_cta_report(pb_solver* s, int local_idx, double* host_out, int max_ctas);
/* Which kernel the last persistent solve ran: "resident", "tma" or "ldg" ("" before the first solve). */
const char* pb_solver_last_kernel(pb_solver* s);
/* Tiles of the last persistent (TMA) solve by what they move per node and iteration, summed over the local ranks:
   out3 = {constant coefficients with Ap recomputed (64 B), constant coefficients with Ap stored (80 B), mixed tiles
   that stream a, b, D (112 B)}.  Counted by the kernel itself; the benchmark's roofline uses these. */
int pb_solver_tile_counts(pb_solver* s, long long* out3);
/* 1 if a solve with this ap_mode recomputes Ap in phase C on this solver's blocks (0 resolves by working-set size) */
int pb_solver_ap_recompute(pb_solver* s, int ap_mode);

/* ---- single-process multi-GPU group: one rank per listed device, peer access inside the process.
 *      Used by the CLI and by solve(..., gpus=G) when not running under a multi-process launcher. ---- */
typedef struct pb_group pb_group;
pb_group* pb_group_create(int M, int N, int Px, int Py, const int* devices, int ndevices);
void pb_group_destroy(pb_group* g);
int pb_group_solve(pb_group* g, const pb_solve_options* opt, pb_stats* stats);
int pb_group_copy_solution(pb_group* g, double* host_full /* (M+1)*(N+1) */);
pb_solver* pb_group_solver(pb_group* g, int rank); /* borrowed */

/* kernel-level step API (library-collective baseline: the caller runs NCCL / gloo between the launches) */
int pb_step_set_options(pb_solver* s, int norm, double delta, long long max_iter, void* cuda_stream);
int pb_step_assemble(pb_solver* s, int local_idx, void* cuda_stream);
int pb_step_begin(pb_solver* s, int local_idx, void* cuda_stream);
int pb_step_dot_zr(pb_solver* s, int local_idx, void* cuda_stream);
int pb_step_stencil(pb_solver* s, int local_idx, int fuse_p, int exact_div, void* cuda_stream);
int pb_step_update(pb_solver* s, int local_idx, int fuse_p, void* cuda_stream);
int pb_step_update_p(pb_solver* s, int local_idx, void* cuda_stream);
int pb_step_reduce_local(pb_solver* s, int local_idx, int nvalues, void* cuda_stream);
int pb_step_scalars(pb_solver* s, int local_idx, int phase /* 0 init, 1 alpha, 2 beta */, void* cuda_stream);
int pb_step_pack_edges(pb_solver* s, int local_idx, void* dev_buf, int stride, void* cuda_stream);
int pb_step_unpack_edges(pb_solver* s, int local_idx, const void* dev_buf, int stride, void* cuda_stream);
int pb_step_error_report(pb_solver* s, int local_idx, void* cuda_stream);
/* ---- stand-alone operator: Ap = A p and dot = h1 h2 sum(Ap * p) on one block in the padded layout
 *      (replaces apply_A_kernel S4:507-536 + dot_gpu(Ap, p) S4:771-786).  All pointers are device pointers;
 *      p, a, b, Ap are (nx+2) x pitch doubles, dot_out is one double.  Stream-ordered; the per-tile scratch
 *      (one double per 32 x 64 tile) is taken from the stream's memory pool for the duration of the call. ---- */
int pb_apply_A_dot(int M, int N, int nx, int ny, int pitch, const void* p, const void* a, const void* b, void* Ap,
                   void* dot_out, void* cuda_stream);

/* pb_apply_A_dot for callers that must not allocate (no counterpart in the reference, which allocates all its
 * device memory up front with cudaMalloc and offers no library entry points at all).  Same arguments and
 * results as pb_apply_A_dot; the per-tile scratch lives in a 512 KiB array built into the library instead of
 * coming from the stream's memory pool, and the call makes no cudaMalloc / cudaFree / pool request.  Meant
 * for programs that own all device memory through their own pool, that sit on a path where the driver's
 * allocator lock is not acceptable, or that may not allocate after start-up.  Limit: ceil(nx/32) *
 * ceil(ny/64) <= 65536 tiles, larger blocks return -1 with an explanatory pb_last_error().  The first call
 * on each device looks up the array's device address (do it at start-up, before any stream capture) and keeps
 * it for the life of the process: after a cudaDeviceReset on that device the kept address is stale, so a
 * program that resets a device must not call this entry point on it again.  Calls on one device share the
 * array, so calls that overlap on the GPU race -- keep them stream-ordered. */
int pb_apply_A_dot_noalloc(int M, int N, int nx, int ny, int pitch, const void* p, const void* a, const void* b,
                           void* Ap, void* dot_out, void* cuda_stream);

/* communication micro-benchmark (collective over all ranks): `iters` rounds of the in-kernel cross-rank
 * reduction (mode 0: counter + elected aggregator, the plain-load kernel's; mode 3: the record star of the TMA
 * kernel; mode 5: mode 3 with per-step clocks) or of halo stores of `edge` doubles per neighbour + reduction
 * (modes 1, 2: fenced stores; mode 4: halo rows as sequence-tagged records + record star, the TMA kernel's protocol);
 * out2 = {ns per round measured on this rank, checksum}.  ctas = 0: one CTA per SM. */
int pb_solver_xgpu_bench(pb_solver* s, int iters, int mode, int edge, int ctas, void* cuda_stream, double* out2);
/* unit test: compares x*y, one- and two-correction Markstein quotients with x / h on n pseudo-random x;
 * dev_counts = u64[4] {samples, one-correction mismatches, two-correction mismatches, x*y mismatches} */
int pb_debug_div_check(double h, unsigned long long n, unsigned long long seed, void* dev_counts, void* cuda_stream);
/* unit-test probe: load one TMA box (map index 0..23, see csrc/cuda/persistent_tma.cu) at element coordinates
 * (x, y) into dev_out; *dev_status = 1 done, 2 timed out */
int pb_debug_tma_probe(pb_solver* s, int local_idx, int map_index, int mode, int x, int y, unsigned int nbytes, void* dev_out,
                       void* dev_status, void* cuda_stream);
int pb_step_read_control(pb_solver* s, int local_idx, int* iter, int* status, double* diff_norm, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* POISSON_B200_H */
