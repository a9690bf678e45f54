"""pytest configuration: the ``gpu`` marker and shared fixtures."""
import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")
    config.addinivalue_line("markers", "slow: long CPU oracle runs")


@pytest.fixture(scope="session")
def native():
    from poisson_ellipse_openmp_mpi_cuda_b200.ops import _native

    _native.lib()
    return _native


# iteration counts of the reference
ITERS_WEIGHTED = {(10, 10): 15, (20, 20): 26, (40, 40): 50, (80, 80): 96, (160, 160): 183, (400, 600): 546,
                  (800, 1200): 989}
ITERS_UNWEIGHTED = {(10, 10): 17, (20, 20): 31, (40, 40): 61}
# L2 (in D), L2 (all), max error, max w  -- BASELINE.md section 3
ERRORS = {
    (10, 10): (5.603706175e-3, 5.669960579e-3, 9.394889283e-3, 0.09758020841),
    (20, 20): (7.662608376e-3, 7.666506656e-3, 1.119665031e-2, 0.09409613374),
    (40, 40): (3.677324436e-3, 3.677820166e-3, 6.418248093e-3, 0.09797040155),
    (80, 80): (1.691136735e-3, 1.691201676e-3, 3.432152822e-3, 0.09881080759),
    (160, 160): (8.956752999e-4, 8.956829639e-4, 1.85682095e-3, 0.09943637621),
    (400, 600): (3.060682255e-4, 3.060688e-4, 6.215795962e-4, 0.09971655665),
    (800, 1200): (1.91571282e-4, 1.915713392e-4, 3.154020681e-4, 0.09972306141),
}
