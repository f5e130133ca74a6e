"""Solvers ("model families" of this framework): the GPU heat solver, the CPU
oracle solver, and the presets that mirror the four reference programs."""
