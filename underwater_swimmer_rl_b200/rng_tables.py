"""Host-seeded MT19937 word tables shared with the reference.

The reference draws food positions from the process-global ``random`` module
(src/salp/environments/salp_snake_env.py:101-104,239-242) and thrust jitter from the global
``np.random`` (scripts/utilities/salp_robot.py:311).  Seeding both with ``s`` and reading the
raw 32-bit generator outputs gives two word streams; row ``r`` of the tables built here holds
the first ``len`` words of the streams for seed ``seed0 + env_id0 + r``.  Generation is done by
the C++ generator inside libsalp_b200.so (multi-threaded), checked against CPython / numpy in
tests/test_rng_tables.py.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi

FOOD_STREAM, JITTER_STREAM = 0, 1


def fill_words(rows: int, length: int, seed0: int, first_row: int, kind: int, threads: int = 0) -> np.ndarray:
    """uint32 [rows, length] of raw MT19937 outputs (kind 0: random.seed, 1: np.random.seed)."""
    if length <= 0 or length & (length - 1):
        raise ValueError("table length must be a power of two")
    out = np.empty((rows, length), dtype=np.uint32)
    L = _capi.load()
    _capi.check(L.salp_fill_rng_words(out.ctypes.data_as(C.c_void_p), rows, length, seed0, first_row, kind, threads))
    return out
