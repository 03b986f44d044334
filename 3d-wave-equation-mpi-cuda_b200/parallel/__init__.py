"""Block decomposition and inter-rank plumbing (peer memory wiring, staged halo transport)."""
from .decomp import Decomposition, dims_create, parse_dims  # noqa: F401
