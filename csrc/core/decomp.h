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
