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

int choose_grid_from_topology(int n, const int* perf, int M, int min_slab_rows, ProcGrid* grid, int* order) {
  for (int r = 0; r < n; ++r) order[r] = r;
  *grid = choose_proc_grid(n);
  if (n <= 1) return 0;
  int best = perf[1], worst = perf[1];
  for (int a = 0; a < n; ++a)
    for (int b = 0; b < n; ++b) {
      if (a == b) continue;
      best = perf[a * n + b] < best ? perf[a * n + b] : best;
      worst = perf[a * n + b] > worst ? perf[a * n + b] : worst;
    }
  if (best == worst) {  // uniform fabric
    if ((M - 1) / n >= min_slab_rows) *grid = ProcGrid{n, 1};
    return 0;
  }
  // islands: connected components of the best-ranked links (treated as undirected)
  int comp[64];
  if (n > 64) return 2;
  for (int a = 0; a < n; ++a) comp[a] = -1;
  int ncomp = 0;
  for (int a = 0; a < n; ++a) {
    if (comp[a] >= 0) continue;
    int stack[64], top = 0;
    stack[top++] = a;
    comp[a] = ncomp;
    while (top > 0) {
      const int u = stack[--top];
      for (int v = 0; v < n; ++v)
        if (v != u && comp[v] < 0 && (perf[u * n + v] == best || perf[v * n + u] == best)) {
          comp[v] = ncomp;
          stack[top++] = v;
        }
    }
    ++ncomp;
  }
  if (ncomp < 2 || n % ncomp != 0) return 2;
  const int k = n / ncomp;
  int count[64] = {0};
  for (int a = 0; a < n; ++a) ++count[comp[a]];
  for (int c = 0; c < ncomp; ++c)
    if (count[c] != k) return 2;
  // rank (px, py) = py * ncomp + px -> member py of island px (members in their given order)
  int seen[64] = {0};
  for (int a = 0; a < n; ++a) {
    const int px = comp[a], py = seen[px]++;
    order[py * ncomp + px] = a;
  }
  *grid = ProcGrid{ncomp, k};
  return 1;
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
