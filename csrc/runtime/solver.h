// solver.h -- host-side driver of the GPU solver: owns the arenas of the local ranks, knows the peers,
// builds the device contexts and runs a solve.
//
// Replaces gradient_solver_mpi's host part (S4:687-982): allocation (S4:737-768), the iteration loop
// (now one cooperative launch), the timer collection (S4:955-979).
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/poisson_b200.h"
#include "../cuda/ctx.h"
#include "../cuda/launch.h"

namespace pb {

enum Method : int {
  kMethodPersistent = 0,     // TMA kernel; blocks that fit on chip go to the resident kernel unless options pin the TMA one
  kMethodStepFused = 1,
  kMethodStepClassic = 2,
  kMethodPersistentLdg = 3,
  kMethodResident = 4,       // shared-memory-resident kernel (refused when the block does not fit)
  kMethodPersistentTma = 5   // TMA kernel, never the resident one
};

struct SolveOptions {
  int method = kMethodPersistent;
  int norm = PB_NORM_WEIGHTED;
  double delta = 1e-6;
  long long max_iter = -1;  // <0: (M-1)(N-1)
  int exact_div = 2;        // stencil divisions: 0 reciprocal multiply, 1 true division, 2 Markstein (= 1 bitwise)
  int ctas_per_sm = 0;      // 0 = auto
  int error_report = 1;
  int compressed = 1;       // TMA kernel: coefficient classes (1 byte/cell) instead of streaming a, b, D
  int no_balance = 0;       // TMA kernel: 1 = cyclic tile -> CTA map instead of the cost-balanced one
  int ab_mode = 0;          // TMA kernel, phase AB of full constant-coefficient items: 0 = warp-private rows, 1 = tile path
  int ap_mode = 0;          // TMA kernel, uniform tiles: 1 store + reload Ap, 2 recompute it in phase C, 0 = default
  int l2_hints = 0;         // TMA loads carry L2 eviction priorities (helps ~2 % when the working set ~ L2 size)
  int tile_rows = 0;        // TMA persistent kernel: 32 (2 stages), 16 (4 stages), 64 (32 + tall items); 0 = auto
  double watchdog_s = 0.0;  // 0 = default (20 s)
};

// Options as they travel host -> device every solve (pinned staging buffer -> Control).
struct DeviceOptions {
  double delta;
  long long max_iter;
  int norm;
  int pad;
  unsigned long long watchdog_ns;
};

class Solver {
 public:
  Solver(int M, int N, int Px, int Py, int first_rank, int nlocal, int device);
  ~Solver();
  Solver(const Solver&) = delete;
  Solver& operator=(const Solver&) = delete;

  int nlocal() const { return (int)blocks_.size(); }
  int nranks() const { return pg_.Px * pg_.Py; }
  int first_rank() const { return first_rank_; }
  int device() const { return device_; }
  const LocalBlock& block(int l) const { return blocks_[l]; }
  const RankCtx& host_ctx(int l) const { return hctx_[l]; }
  const RankCtx* device_ctx(int l) const { return dctx_ + l; }
  char* arena(int l) const { return arenas_[l]; }
  size_t arena_size(int l) const { return arena_bytes(blocks_[l]); }
  double* array(int l, int which) const;  // 0 w, 1 r, 2 p0, 3 p1, 4 Ap, 5 a, 6 b, 7 D
  const void* host_maps(int l) const;
  Control* control(int l) const { return reinterpret_cast<Control*>(arenas_[l] + kCtlOffset); }

  void ipc_handle(int l, void* out64) const;
  void open_peer(int rank, const void* handle64);
  void set_peer_base(int rank, void* base);
  void finalize();  // all peers known -> build and upload the contexts

  void solve(const SolveOptions& opt, cudaStream_t stream, pb_stats* out);
  // the two halves of solve(): enqueue everything without host synchronisation / wait and read the report
  void enqueue(const SolveOptions& opt, cudaStream_t stream);
  void collect(cudaStream_t stream, pb_stats* out);
  // communication micro-benchmark: returns {ns per round (this rank), checksum}; collective over all ranks
  void xgpu_bench(int iters, int mode, int edge, int ctas, cudaStream_t stream, double out[2]);
  // dense (M+1) x (N+1) host field from the local ranks' blocks (other ranks' nodes are left untouched)
  void copy_solution_to_host(double* full, cudaStream_t stream);
  // dense nx x ny device copy of one local block's owned nodes
  void gather_block(int l, double* dev_out, long long ld, cudaStream_t stream);

  int last_ctas_per_rank() const { return last_ctas_per_rank_; }
  int last_tile_rows() const { return last_tile_rows_; }  // tile height the last persistent TMA solve ran with
  const char* last_kernel() const { return last_kernel_; }
  long long last_launches() const { return last_launches_; }
  int cta_report(int l, double* host_out, int max_ctas);  // the raw kCtaTimeWords doubles per CTA
  // tiles of the last persistent TMA solve by traffic class, summed over the local ranks: constant coefficients
  // with Ap recomputed, constant coefficients with Ap stored, mixed (coefficient boxes streamed)
  void tile_counts(long long out[3]);
  int cta_times(int l, double* host_out, int max_ctas);  // seconds per CTA: AB, C, reduction 1, reduction 2
  bool ap_recompute(int ap_mode) const;  // resolves pb_solve_options.ap_mode (0 = by working-set size)
  size_t h2d_bytes_per_solve() const { return sizeof(DeviceOptions) * blocks_.size(); }
  size_t d2h_bytes_per_solve() const { return sizeof(Control) * blocks_.size(); }

 private:
  void solve_persistent(const SolveOptions& opt, cudaStream_t stream);
  bool resident_cut(ResidentCut* cut) const;  // the cut of the local blocks into one sub-block per CTA, if they fit
  void solve_steps(const SolveOptions& opt, cudaStream_t stream);

  Grid grid_;
  ProcGrid pg_;
  int first_rank_, device_;
  bool finalized_ = false;
  bool fabric_complete_ = false;
  std::vector<LocalBlock> blocks_;
  std::vector<char*> arenas_;
  std::vector<void*> peer_base_;     // per global rank: arena base as seen from this device
  std::vector<void*> opened_ipc_;    // bases obtained from cudaIpcOpenMemHandle (to close)
  std::vector<RankCtx> hctx_;
  RankCtx* dctx_ = nullptr;
  std::vector<char> hmaps_store_;
  char* hmaps_ = nullptr;  // host copy of the tensor maps (64-byte aligned inside hmaps_store_)
  char* dmaps_ = nullptr;  // tensor maps, [nlocal][2 variants][12]
  DeviceOptions* pinned_opts_ = nullptr;
  Control* pinned_ctl_ = nullptr;
  cudaEvent_t ev_[4];
  int sm_count_ = 148;
  int l2_bytes_ = 126 << 20;
  int last_ctas_per_rank_ = 0;
  int last_tile_rows_ = 0;
  const char* last_kernel_ = "";  // "tma", "ldg" or "resident": what the last persistent solve ran
  int last_method_ = 0;
  long long last_launches_ = 0;
};

}  // namespace pb
