"""Decomposition / topology, peer-memory runtime, multi-GPU solvers."""
from .decomp import Decomposition, TileSpec, split, DIRS, OPPOSITE  # noqa: F401
