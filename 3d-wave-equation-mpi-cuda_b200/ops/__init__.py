"""Kernel-level ops on torch tensors (``kernels``) and their plain-PyTorch fp64
references (``reference``) used by the numerics tests."""
from . import kernels, reference  # noqa: F401
