"""Solvers: WaveSolver (one process drives all ranks) and DistributedWaveSolver
(one rank per process under torch.distributed)."""
from .wave import WaveSolver, SolveResult  # noqa: F401
from .distributed import DistributedWaveSolver  # noqa: F401
