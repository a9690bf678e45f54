"""Problem definition and solver families."""
from .problem import EllipseProblem  # noqa: F401
from .families import FAMILIES, REFERENCE_PROGRAMS, run_family, run_program  # noqa: F401
