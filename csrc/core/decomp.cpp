// decomp.cpp -- see decomp.h.
#include "decomp.h"

#include <cmath>

namespace pb {

ProcGrid choose_proc_grid(int nranks) {
  ProcGrid g{1, nranks < 1 ? 1 : nranks};
  if (nranks < 1) return g;
  int px = (int)std::sqrt((double)nranks);
  while ((px + 1) * (px + 1) <= nranks) ++px;  // guard against sqrt rounding down
  while (px * px > nranks) --px;
  while (px > 1 && nranks % px != 0) --px;
  g.Px = px;
  g.Py = nranks / px;
  return g;
}

Range split_axis(int cells, int parts, int k) {
  const int total = cells - 1;  // interior nodes
  const int base = total / parts;
  const int rem = total % parts;
  // blocks 0..rem-1 hold base+1 nodes
  const int before = k * base + (k < rem ? k : rem);
  const int mine = base + (k < rem ? 1 : 0);
  Range r;
  r.start = 1 + before;
  r.end = r.start + mine - 1;
  return r;
}

int padded_pitch(int ny) {
  const int need = ny + 2 + kColOff;
  return (need + kPitchAlign - 1) / kPitchAlign * kPitchAlign;
}

LocalBlock make_local_block(int M, int N, ProcGrid pg, int rank) {
  LocalBlock b;
  b.rank = rank;
  b.px = rank % pg.Px;
  b.py = rank / pg.Px;
  const Range ri = split_axis(M, pg.Px, b.px);
  const Range rj = split_axis(N, pg.Py, b.py);
  b.i0 = ri.start;
  b.j0 = rj.start;
  b.nx = ri.count();
  b.ny = rj.count();
  b.pitch = padded_pitch(b.ny);
  b.nbr[kXm] = b.px > 0 ? rank - 1 : -1;
  b.nbr[kXp] = b.px < pg.Px - 1 ? rank + 1 : -1;
  b.nbr[kYm] = b.py > 0 ? rank - pg.Px : -1;
  b.nbr[kYp] = b.py < pg.Py - 1 ? rank + pg.Px : -1;
  return b;
}

}  // namespace pb
