// ctx.h -- device-visible description of one rank's part of the solve.
//
// One `RankCtx` describes the local block of a rank, its arrays, its four neighbours and the
// cross-rank reduction workspace.  A real multi-GPU run has one context per process/GPU with peer
// pointers obtained through CUDA IPC; a single-GPU "virtual rank" run (tests, decomposition checks)
// places several contexts on one device and one cooperative launch serves all of them.
//
// Replaces, in one structure, what the reference passes around as loose arguments: the local ranges
// (S4:706-709), the eight device arrays (S4:737-748), the neighbour ranks (S4:341-344) and the
// partial-sum buffer (S4:766-768).
#pragma once
#include <cstddef>
#include <cstdint>

#include "../core/decomp.h"
#include "../core/problem.h"

namespace pb {

constexpr int kMaxRanks = 16;
constexpr int kMaxCtas = 1024;      // per rank
constexpr int kRedValues = 2;       // doubles carried by one cross-rank reduction
constexpr int kSlotWords = 4;       // 2 doubles -> 4 (payload32 | seq32) words

// Arena layout (identical on every rank, so a peer needs only the arena base address).
constexpr size_t kCtlOffset = 0;                   // Control
constexpr size_t kSlotsOffset = 4096;              // [2 parities][kMaxRanks][kSlotWords] u64
constexpr size_t kArriveOffset = 8192;             // arrival counter (u32) + padding
constexpr size_t kHaloCntOffset = 8192 + 1024;     // u64[4]: halo tiles received per direction (monotonic)
constexpr int kPartialStride = 8;                  // doubles per CTA in the partials area
constexpr size_t kPartialsOffset = 16384;          // [kMaxCtas][kPartialStride] doubles
constexpr int kCtaTimeWords = 8;                   // doubles per CTA in the per-CTA report area
constexpr size_t kCtaTimesOffset = 81920;          // [kMaxCtas][kCtaTimeWords] doubles: ns in AB, C, reduction 1, reduction 2,
                                                   // tiles by kind (constant coefficients with Ap recomputed / stored, mixed), ns halo wait
constexpr size_t kLl1Offset = 147456;              // [kRedValues][kMaxCtas][2] u64: per-CTA records of star_sum (CTA -> CTA 0)
constexpr size_t kResOffset = 180224;              // [kRedValues][2] u64: result record of star_sum (CTA 0 -> every CTA)
constexpr size_t kTileCodeOffset = 180352;         // one code byte per super-tile (TMA kernel, cost-balanced tile assignment)
constexpr int kTileCodeBytes = 16256;              // up to the end of the header
constexpr size_t kHeaderBytes = 196608;            // arrays start here; r is the first array

// Scalars of the iteration, status and device-side timers.  Lives at the start of the arena.
struct Control {
  // iteration state
  double zr_old, alpha, beta, denom, diff2, zr_new, diff_norm;
  double red[4];             // staging for step-wise reductions (local sums, then global sums)
  int iter;                  // index of the current / final iteration
  int status;                // PB_STATUS_*
  unsigned int red_seq;      // sequence number of the last cross-rank reduction (persists across solves)
  unsigned int halo_seq;     // sequence number of the last halo round (persists across solves, like red_seq)
  int cur;                   // which p buffer holds the current direction
  int pad0;
  // options
  double delta;
  long long max_iter;
  int norm;
  int pad1;
  unsigned long long watchdog_ns;  // 0 = default
  // device timers (ns, accumulated by CTA 0 of the persistent kernel)
  unsigned long long t_ab, t_c, t_red1, t_red2, t_init, t_total;
  unsigned long long t_halo;  // CTA 0: ns spent waiting for / unpacking the neighbours' halo records (inside t_c)
  // error report partials {sum e^2 in D, sum e^2 all, max |e|, max w, nodes in D}
  double err[5];
};

struct RankCtx {
  Grid g;
  int i0, j0, nx, ny, pitch;
  int rank, nranks;
  int nbr[4];  // neighbour ranks (x-, x+, y-, y+), -1 at the boundary of Omega
  // local arrays, all (nx+2) x pitch
  double *w, *r, *p[2], *Ap, *a, *b, *D;
  // Coefficient classes, one byte per cell, same (row, column) indexing as the fp64 arrays with row pitch
  // `pitch` bytes: bits 0-1 class of a, bits 2-3 class of b, bits 4-5 class of D; class 0 = the value of a cell
  // whose faces lie inside the ellipse (a = b = 1), 1 = outside (1/eps), 2 = anything else (read the array).
  // All but O(M+N) cells are class 0 or 1; the TMA kernel classifies whole tiles from these bytes once per
  // launch and loads coefficient boxes only for tiles that are not uniformly class 0 or class 1.
  unsigned char* cls;
  // neighbour r arrays as seen from this GPU (peer mapped), with their row pitch and extents
  double* nbr_r[4];
  int nbr_pitch[4];
  int nbr_nx[4], nbr_ny[4];
  // reduction workspace (local) and every rank's slot area as seen from this GPU
  double* partials;
  unsigned int* arrive;
  unsigned long long* slots;
  unsigned long long* peer_slots[kMaxRanks];
  Control* ctl;
  // halo completion counters: ours (incremented by the neighbours) and, per direction, the neighbour's
  // counter that our stores into its ring must bump (its opposite direction).  Used by the plain-load
  // persistent kernel (persistent.cu) only; the TMA kernel sends halo rows as LL records (halo_ll below)
  unsigned long long* halo_cnt;
  unsigned long long* nbr_halo_cnt[4];
  // Fence-free halo lines (TMA kernel): per direction and parity a line of (payload32 | seq32) word pairs, one pair
  // per edge node, written by the neighbour and unpacked into our r ring by the warp that issues the tile's loads.
  // halo_ll = [4 directions][2 parities][ll_cap][2] u64 of this rank; nbr_ll[d] = the neighbour's line block for
  // ITS direction d^1 (what our edge d feeds), nbr_ll_cap[d] its capacity.
  unsigned long long* halo_ll;
  int ll_cap;
  unsigned long long* nbr_ll[4];
  int nbr_ll_cap[4];
  const void* tmaps;  // CUtensorMap[kMapVariants][kMapsPerVariant] in device memory (TMA persistent kernel)
  int sys_scope;  // 1: some peer lives on another device, fences must be system-wide
  // tiling of the block for the stencil phase
  int tiles_x, tiles_y;
};

constexpr int kTileX = 32;  // rows per stencil tile
constexpr int kTileY = 64;  // columns per stencil tile

// 8 fp64 arrays + one byte per cell of coefficient classes (see RankCtx::cls)
inline size_t cls_bytes(const LocalBlock& b) { return ((b.elems() + 255) / 256) * 256; }
inline int ll_capacity(const LocalBlock& b) { return (b.nx > b.ny ? b.nx : b.ny) + 2; }
inline size_t halo_ll_offset(const LocalBlock& b) { return kHeaderBytes + 8 * b.elems() * sizeof(double) + cls_bytes(b); }
inline size_t halo_ll_bytes(const LocalBlock& b) { return (size_t)4 * 2 * ll_capacity(b) * 2 * sizeof(unsigned long long); }
inline size_t arena_bytes(const LocalBlock& b) { return halo_ll_offset(b) + halo_ll_bytes(b); }
constexpr int kMapsPerVariant = 16;  // tensor-map slots per tile height (12 used, see persistent_tma.cu)
constexpr int kMapVariants = 3;      // box heights 32, 16 and 64 rows

}  // namespace pb
