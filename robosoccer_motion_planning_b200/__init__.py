"""B200-native kinodynamic RRT* hot path (sm_100a CUDA behind the C ABI in include/krrt.h).

Only what the hot path needs lives here: csrc/ (kernels + C ABI), capi.py (ctypes host mirror),
scenes.py (synthetic soccer-field inputs for tests and bench), sweep.py (multi-GPU query sharding).
"""
from .capi import (KrrtConfig, KrrtError, Planner, RESULT_DTYPE, build_library, default_config, lib,  # noqa: F401
                   LIB_PATH)
