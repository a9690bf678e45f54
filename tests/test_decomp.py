"""Process grid and block decomposition: parity with the reference's algorithm (restated literally from the
description in SURVEY.md section 2.3, C7/C8: S4:82-86 and S4:97-133) and structural properties."""
import math

import pytest

from poisson_ellipse_openmp_mpi_cuda_b200.parallel import decomp


def ref_choose(size):
    px = int(math.sqrt(size))
    while px > 1 and size % px != 0:
        px -= 1
    return px, size // px


def ref_decompose(M, N, Px, Py, rank):
    px, py = rank % Px, rank // Px

    def axis(total, parts, k):
        base, rem = total // parts, total % parts
        off = 1
        for q in range(k):
            off += base + (1 if q < rem else 0)
        n = base + (1 if k < rem else 0)
        return off, off + n - 1

    i0, i1 = axis(M - 1, Px, px)
    j0, j1 = axis(N - 1, Py, py)
    return i0, i1, j0, j1


@pytest.mark.parametrize("P", list(range(1, 33)))
def test_process_grid_equals_reference(P):
    assert decomp.choose_process_grid(P) == ref_choose(P)


def test_known_grids():
    expect = {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4), 16: (4, 4), 20: (4, 5), 32: (4, 8)}
    for P, g in expect.items():
        assert decomp.choose_process_grid(P) == g


@pytest.mark.parametrize("grid", [(40, 40), (800, 1200), (4096, 4096), (16384, 32768), (37, 91)])
@pytest.mark.parametrize("P", [1, 2, 3, 4, 6, 8, 12, 16])
def test_blocks_equal_reference_and_tile_the_interior(grid, P):
    M, N = grid
    Px, Py = decomp.choose_process_grid(P)
    covered = 0
    sizes_x, sizes_y = set(), set()
    for rank in range(P):
        b = decomp.local_block(M, N, Px, Py, rank)
        i0, i1, j0, j1 = ref_decompose(M, N, Px, Py, rank)
        assert (b.i0, b.i0 + b.nx - 1, b.j0, b.j0 + b.ny - 1) == (i0, i1, j0, j1)
        assert (b.px, b.py) == (rank % Px, rank // Px)
        covered += b.nx * b.ny
        sizes_x.add(b.nx)
        sizes_y.add(b.ny)
        # neighbours: rank +- 1 in x, rank +- Px in y (S4:341-344)
        assert b.nbr[0] == (rank - 1 if b.px > 0 else -1)
        assert b.nbr[1] == (rank + 1 if b.px < Px - 1 else -1)
        assert b.nbr[2] == (rank - Px if b.py > 0 else -1)
        assert b.nbr[3] == (rank + Px if b.py < Py - 1 else -1)
        # layout: 128-byte rows, room for the ring
        assert b.pitch % 16 == 0 and b.pitch >= b.ny + 2 + decomp.col_offset()
    assert covered == (M - 1) * (N - 1)
    assert max(sizes_x) - min(sizes_x) <= 1 and max(sizes_y) - min(sizes_y) <= 1


def test_edge_sizes_match_survey_table():
    """halo message sizes computed with the reference's decompose_2d."""
    b = decomp.local_block(800, 1200, 2, 4, 0)
    assert (b.ny + 2) * 8 in (2416, 2408, 2400, 2424) and (b.nx + 2) * 8 in (3216, 3208, 3200, 3224)
    b = decomp.local_block(16384, 32768, 2, 4, 3)
    assert abs((b.nx + 2) * 8 - 65536) <= 16 and abs((b.ny + 2) * 8 - 65536) <= 16


def test_grid_parsing_and_slab_grid():
    assert decomp.parse_grid(None, 8) == (2, 4)
    assert decomp.parse_grid("reference", 8) == (2, 4)
    assert decomp.parse_grid("4x2", 8) == (4, 2)
    assert decomp.parse_grid((1, 8), 8) == (1, 8)
    assert decomp.parse_grid("auto", 8, 4096) == (8, 1)
    assert decomp.parse_grid("auto", 8, 100) == (2, 4)  # slabs thinner than a tile: fall back
    with pytest.raises(ValueError):
        decomp.parse_grid("3x3", 8)


def test_invalid_decompositions_are_rejected(native):
    import ctypes as C

    b = native.Block()
    assert native.lib().pb_local_block(4, 4, 8, 1, 0, C.byref(b)) != 0  # more blocks than interior nodes
    assert "more blocks" in native.last_error()
