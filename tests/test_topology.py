"""Process grid from a peer-to-peer performance matrix (csrc/core/decomp.cpp, shared by the CLI and Python): mocked
matrices, no GPU needed.  Replaces choose_process_grid (S4:82-86), which looks at the rank count only."""
from poisson_ellipse_openmp_mpi_cuda_b200.parallel.topology import grid_from_topology


def _uniform(n, rank=0):
    return [[0 if a == b else rank for b in range(n)] for a in range(n)]


def test_uniform_fabric_gives_slabs():
    for n in (2, 4, 8):
        px, py, order, kind = grid_from_topology(_uniform(n), 4096)
        assert (px, py) == (n, 1) and order == list(range(n)) and kind == "uniform"


def test_uniform_fabric_thin_slabs_fall_back_to_reference_grid():
    # 8 slabs of a 200-row grid would be 24 rows thin (< one 32-row tile): the reference's 2 x 4 instead
    px, py, order, kind = grid_from_topology(_uniform(8), 200)
    assert (px, py) == (2, 4) and kind == "uniform"


def test_two_islands_put_the_cut_between_them():
    # devices 0-3 and 4-7 are NVLink quads (rank 0), the quads see each other through a slower link (rank 1)
    perf = [[0 if a == b else (0 if a // 4 == b // 4 else 1) for b in range(8)] for a in range(8)]
    px, py, order, kind = grid_from_topology(perf, 4096)
    assert (px, py) == (2, 4) and kind == "islands"
    # rank = py * Px + px: px selects the island, so x-neighbours (rank, rank + 1) are the only cross-island pairs
    for rank, dev in enumerate(order):
        assert dev // 4 == rank % 2
    for rank in range(0, 8 - 2):  # y-neighbours rank, rank + Px share an island
        assert order[rank] // 4 == order[rank + 2] // 4
    assert sorted(order) == list(range(8))


def test_interleaved_islands_are_found_by_connectivity_not_by_numbering():
    perf = [[0 if a == b else (0 if a % 2 == b % 2 else 2) for b in range(4)] for a in range(4)]
    px, py, order, kind = grid_from_topology(perf, 4096)
    assert (px, py) == (2, 2) and kind == "islands"
    assert {order[0], order[2]} in ({0, 2}, {1, 3}) and {order[1], order[3]} in ({0, 2}, {1, 3})


def test_irregular_topology_falls_back():
    perf = [[0, 0, 1, 1], [0, 0, 1, 1], [1, 1, 0, 1], [1, 1, 1, 0]]  # islands {0,1}, {2}, {3}: unequal
    px, py, order, kind = grid_from_topology(perf, 4096)
    assert (px, py) == (2, 2) and order == [0, 1, 2, 3] and kind == "fallback"


def test_single_device():
    assert grid_from_topology([[0]], 4096)[:2] == (1, 1)
