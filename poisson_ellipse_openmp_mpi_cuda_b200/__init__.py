"""poisson_ellipse_openmp_mpi_cuda_b200 -- B200-native distributed PCG solver for the fictitious-domain
Poisson problem on the ellipse x^2 + 4 y^2 < 1 (capabilities of mxy-kit/poisson-ellipse-openmp-mpi-cuda).

Layout:  ``models``  problem definition and the solver families (CPU stages, GPU fused, baselines)
         ``ops``     native library loader and the GPU solver object (hand-written sm_100a kernels)
         ``parallel`` decomposition, multi-GPU fabric (CUDA IPC peer memory), library-collective solvers
         ``utils``   reports in the reference's formats, solution dumps
"""
from .api import PoissonSolver, clear_cache, get_solver, solve  # noqa: F401

__version__ = "0.1.0"
