// decomp.h -- process grid and 2-D block decomposition of the interior nodes, plus the padded local
// array layout used by every device kernel.
//
// Parity: choose_process_grid S4:82-86 (Px <= Py, Px the largest divisor of P not above sqrt(P));
// decompose_2d S4:97-133 (rank = py*Px + px, the first `rem` blocks get one extra node);
// local storage S4:712-724 (one ring of halo cells) -- here the pitch is padded for 128-byte rows and
// TMA (base 16-byte aligned, row stride a multiple of 16 bytes).
#pragma once
#include <cstddef>
#include <cstdint>

namespace pb {

struct ProcGrid {
  int Px, Py;
};

// Same answer as the reference for every P (1->1x1, 2->1x2, 4->2x2, 8->2x4, 20->4x5, 32->4x8).
ProcGrid choose_proc_grid(int nranks);

// Process grid and device order derived from the box's peer-to-peer topology (replaces choose_process_grid,
// S4:82-86, for GPUs: the reference knows nothing about the machine it runs on).
//   perf[a * n + b] = performance rank of the link a -> b as CUDA reports it (cudaDevP2PAttrPerformanceRank: lower is
//   better; the diagonal is ignored), min_slab_rows = the thinnest slab the tile kernels accept (32).
// * every pair equally ranked (NVSwitch): n x 1 slabs along x -- every shared edge is then a contiguous row, halo bytes
//   are negligible next to HBM bytes and every link is as good as any other; devices in their given order.  Slabs
//   thinner than min_slab_rows fall back to the reference's near-square grid.
// * the best-ranked links form m equal islands of k devices (two NVLink-connected quads joined by a slower link):
//   m x k grid with the cuts of the x direction between the islands, so that every y-neighbour pair sits inside an
//   island and only the m - 1 x-cuts cross the slow links: rank (px, py) = py * m + px runs on member py of island px.
// * anything else: the reference's grid, devices in their given order.
// order[rank] = index (into the caller's device list) of the device that runs `rank`.  Returns 0 uniform, 1 islands,
// 2 fallback.
int choose_grid_from_topology(int n, const int* perf, int M, int min_slab_rows, ProcGrid* grid, int* order);

// Closed interval of global interior indices owned by block `k` of `parts` along an axis with
// `cells` cells (interior nodes 1..cells-1).
struct Range {
  int start, end;  // inclusive; count = end - start + 1
  int count() const { return end - start + 1; }
};
Range split_axis(int cells, int parts, int k);

// Column offset of the ring cell lj = 0 inside a padded row: interior lj = 1 lands on a 32-byte
// boundary, rows start on 128-byte boundaries.
constexpr int kColOff = 3;
constexpr int kPitchAlign = 16;  // doubles

struct LocalBlock {
  int rank, px, py;  // position in the process grid
  int i0, j0;        // global index of local (li=1, lj=1)
  int nx, ny;        // owned interior nodes
  int pitch;         // doubles per padded row
  // neighbour ranks, -1 on the boundary of Omega: x- ("left"), x+ ("right"), y- ("down"), y+ ("up")
  int nbr[4];
  size_t elems() const { return (size_t)(nx + 2) * (size_t)pitch; }
  size_t index(int li, int lj) const { return (size_t)li * (size_t)pitch + (size_t)(lj + kColOff); }
};

int padded_pitch(int ny);
LocalBlock make_local_block(int M, int N, ProcGrid pg, int rank);

enum Dir : int { kXm = 0, kXp = 1, kYm = 2, kYp = 3 };

}  // namespace pb
