"""Reports in the reference's stdout formats and solution dumps."""
from . import dump, report  # noqa: F401
