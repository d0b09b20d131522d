"""grayscott.jl_b200 -- B200 (sm_100a) backend for GrayScott.jl's per-step hot path.

Contents: csrc/ (hand-written CUDA kernels + the C ABI of include/gsb200.h, built into
libgsb200.so next to this file) and simulation.py (host-side mirror of the reference's
Simulation API for backend = "B200").  The directory name contains a dot, so load it with
`__graft_entry__.load_package()` (registers it as module `grayscott_jl_b200`).
"""
from . import _lib, simulation  # noqa: F401
from ._lib import GsbError, LIB_PATH, SYMBOLS  # noqa: F401
from .simulation import (Fields, MPICartDomain, Settings, SettingsKeys, calculate, exchange,  # noqa: F401
                         get_fields, get_settings, init_domain, init_fields, initialization, iterate,
                         parse_settings_toml)
