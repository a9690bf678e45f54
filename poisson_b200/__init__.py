"""Short alias of :mod:`poisson_ellipse_openmp_mpi_cuda_b200` (``import poisson_b200``)."""
from poisson_ellipse_openmp_mpi_cuda_b200 import *  # noqa: F401,F403
from poisson_ellipse_openmp_mpi_cuda_b200 import PoissonSolver, clear_cache, get_solver, solve  # noqa: F401
