"""gpu_fluid_simulation_b200 — B200-native (sm_100a CUDA) backend for the per-frame fluid step of
rookieCookies/gpu-fluid-simulation, behind the reference's own FluidSimulation API.

The compute path is libfluid_b200.so (hand-written CUDA, C ABI in include/fluid_b200.h). This package is
the Python host mirror of the reference interface; it never falls back to a CPU implementation.
"""
from ._lib import PARTICLE_DTYPE, FluidError, load  # noqa: F401
from .simulation import FluidSimulation, SimulationSettings, TickSettings  # noqa: F401
from . import field, jfa, scenes  # noqa: F401
