"""Kernel bindings (see include/heat2d.h for the C ABI they wrap)."""
from . import kernels  # noqa: F401
