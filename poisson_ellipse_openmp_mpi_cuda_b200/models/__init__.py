"""Problem definition and solver families."""
from .problem import EllipseProblem  # noqa: F401
from .families import FAMILIES, run_family  # noqa: F401
