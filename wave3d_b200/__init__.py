"""Import alias for the package directory ``3d-wave-equation-mpi-cuda_b200`` (whose
name is not a Python identifier): ``import wave3d_b200`` exposes its sub-packages
``models``, ``ops``, ``parallel`` and ``utils`` and the main entry points."""
import os as _os

_root = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
__path__.insert(0, _os.path.join(_root, "3d-wave-equation-mpi-cuda_b200"))

from . import _lib  # noqa: E402
from ._lib import build, cuda_available  # noqa: E402,F401
from .utils.spec import Problem, parse_args  # noqa: E402,F401
from .parallel.decomp import Decomposition, dims_create  # noqa: E402,F401
from .models.wave import WaveSolver, SolveResult  # noqa: E402,F401
from .models.distributed import DistributedWaveSolver  # noqa: E402,F401

__all__ = ["build", "cuda_available", "Problem", "parse_args", "Decomposition", "dims_create", "WaveSolver",
           "SolveResult", "DistributedWaveSolver"]
