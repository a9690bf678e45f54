"""``python -m poisson_b200`` = ``python -m poisson_ellipse_openmp_mpi_cuda_b200``."""
import sys

from poisson_ellipse_openmp_mpi_cuda_b200.__main__ import main

sys.exit(main())
