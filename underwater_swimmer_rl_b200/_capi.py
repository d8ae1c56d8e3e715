"""ctypes binding of include/salp_b200.h (libsalp_b200.so, built in-tree by build.py).

There is no fallback: if the shared library is missing or a symbol is absent, importing a
compute entry point raises.  The library itself refuses to run without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
_TAG = os.environ.get("SALP_LIB_TAG", "")  # developer A/B builds (build.py)
LIB_PATH = os.path.join(_PKG, "libsalp_b200" + ("_" + _TAG if _TAG else "") + ".so")

SALP_ABI_VERSION = 5
SALP_OK, SALP_EINVAL, SALP_ECUDA, SALP_EUNBOUND, SALP_EUNSUPPORTED = 0, -1, -2, -3, -4
SALP_F32, SALP_F64 = 0, 1
SALP_AUTORESET, SALP_RNG_PHILOX, SALP_GENERIC_KERNEL = 1, 2, 4
EV_TERMINATED, EV_TRUNCATED, EV_COLLISION, EV_FOOD = 1, 2, 4, 8

# every symbol include/salp_b200.h declares (tests/test_capi_symbols.py parses the header too)
EXPORTED_SYMBOLS = (
    "salp_abi_version", "salp_last_error", "salp_query_layout", "salp_create", "salp_destroy", "salp_bind",
    "salp_reset", "salp_refresh_rng", "salp_observe", "salp_set_num_food_items", "salp_set_rng_seed", "salp_forced_cycle", "salp_step", "salp_step_slot", "salp_step_host", "salp_step_host_noalloc", "salp_fill_rng_words", "salp_launch_count", "salp_stats", "salp_replay_append", "salp_actor_forward",
    "salp_cycle_reset", "salp_cycle_step",
)
SALP_CYCLE_AUTORESET = 1
SALP_NOALLOC_STAGING_BYTES = 1 << 20
SALP_CYCLE_OBS_DIM, SALP_CYCLE_ACT_DIM, SALP_CYCLE_MAX_SUBSTEPS = 6, 3, 4096


class SalpParams(C.Structure):
    _fields_ = [
        ("width", C.c_double), ("height", C.c_double),
        ("num_food_items", C.c_int32), ("max_observed_food", C.c_int32),
        ("forced_breathing", C.c_int32), ("random_food_count", C.c_int32),
        ("respawn_food", C.c_int32), ("max_steps_without_food", C.c_int32),
        ("food_reward", C.c_double), ("collision_penalty", C.c_double), ("time_penalty", C.c_double),
        ("efficiency_bonus", C.c_double), ("proximity_reward_weight", C.c_double),
        ("flags", C.c_int32), ("reserved_", C.c_int32),
    ]


class SalpLayout(C.Structure):
    _fields_ = [(n, C.c_int32) for n in
                ("real_bytes", "obs_dim", "act_dim", "food_slots", "body_reals", "ctrl_ints", "stats_len", "reserved_")]


class SalpBuffers(C.Structure):
    _fields_ = [
        ("body", C.c_void_p), ("ctrl", C.c_void_p), ("ep_return", C.c_void_p), ("food", C.c_void_p),
        ("obs", C.c_void_p), ("reward", C.c_void_p),
        ("terminated", C.c_void_p), ("truncated", C.c_void_p), ("events", C.c_void_p),
        ("final_obs", C.c_void_p), ("final_info", C.c_void_p), ("stats", C.c_void_p),
        ("food_words", C.c_void_p), ("jit_words", C.c_void_p),
        ("table_rows", C.c_int64), ("food_len", C.c_int64), ("jit_len", C.c_int64),
        ("max_substeps", C.c_int32), ("reserved_", C.c_int32),
    ]


class SalpReplayAppend(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "obs_prev", "actions", "reward", "reward_aux", "terminated", "truncated", "obs_next", "final_obs",
        "rb_obs", "rb_act", "rb_rew", "rb_next", "rb_done", "head", "size", "ticket")] + [
        ("n", C.c_int64), ("capacity", C.c_int64), ("obs_dim", C.c_int32), ("act_dim", C.c_int32),
        ("dtype", C.c_int32), ("w_env", C.c_float), ("w_aux", C.c_float), ("reserved_", C.c_int32)]


class SalpActorForward(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("obs", "w1_packed", "b1", "w2_packed", "b2", "w3", "b3", "noise", "actions",
                                          "extra", "rng_state")] + [
        ("n", C.c_int64), ("obs_dim", C.c_int32), ("act_dim", C.c_int32), ("hidden", C.c_int32), ("dtype", C.c_int32),
        ("extra_dim", C.c_int32), ("mode", C.c_int32), ("act_rescale", C.c_int32), ("reserved_", C.c_int32),
        ("act_low", C.c_float * 4), ("act_high", C.c_float * 4)]


class SalpCycleParams(C.Structure):
    _fields_ = [(n, C.c_double) for n in (
        "dry_mass", "init_length", "init_width", "max_contraction", "density",
        "nozzle_length1", "nozzle_length2", "nozzle_length3", "nozzle_area", "nozzle_mass", "width", "height")] + [
        ("max_cycles", C.c_int32), ("max_substeps", C.c_int32), ("flags", C.c_int32), ("reserved_", C.c_int32)]


class SalpCycleBuffers(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "state", "ctrl", "obs", "reward", "terminated", "truncated", "n_sub", "final_obs", "stats", "words")] + [
        ("table_rows", C.c_int64), ("words_len", C.c_int64), ("n", C.c_int64), ("env_id0", C.c_int64),
        ("dtype", C.c_int32), ("reserved_", C.c_int32)]


class SalpError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libsalp_b200 error {code}: {msg}")
        self.code = code


_lib = None


def load():
    """Load libsalp_b200.so and declare prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -m underwater_swimmer_rl_b200.build` "
            "(nvcc, sm_100a).  There is no CPU or PyTorch fallback for the SALP env step.")
    L = C.CDLL(LIB_PATH)
    for s in EXPORTED_SYMBOLS:
        getattr(L, s)  # AttributeError if the library does not export it
    L.salp_abi_version.restype = C.c_int
    L.salp_last_error.restype = C.c_char_p
    L.salp_query_layout.argtypes = [C.POINTER(SalpParams), C.c_int, C.POINTER(SalpLayout)]
    L.salp_create.argtypes = [C.POINTER(SalpParams), C.c_int64, C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    L.salp_destroy.argtypes = [C.c_void_p]
    L.salp_bind.argtypes = [C.c_void_p, C.POINTER(SalpBuffers)]
    L.salp_reset.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.salp_refresh_rng.argtypes = [C.c_void_p, C.c_void_p]
    L.salp_observe.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    L.salp_set_num_food_items.argtypes = [C.c_void_p, C.c_int32]
    L.salp_set_rng_seed.argtypes = [C.c_void_p, C.c_uint64]
    L.salp_forced_cycle.argtypes = [C.POINTER(SalpParams)] + [C.c_void_p] * 5
    L.salp_step.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    L.salp_step_slot.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    L.salp_step_host.argtypes = [C.c_void_p] + [C.c_void_p] * 7
    L.salp_step_host_noalloc.argtypes = [C.c_void_p] + [C.c_void_p] * 6
    L.salp_fill_rng_words.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_uint64, C.c_int64, C.c_int, C.c_int]
    L.salp_launch_count.argtypes = [C.c_void_p]
    L.salp_stats.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.salp_actor_forward.argtypes = [C.POINTER(SalpActorForward), C.c_void_p]
    L.salp_replay_append.argtypes = [C.POINTER(SalpReplayAppend), C.c_void_p]
    L.salp_launch_count.restype = C.c_int64
    L.salp_cycle_reset.argtypes = [C.POINTER(SalpCycleParams), C.POINTER(SalpCycleBuffers), C.c_void_p, C.c_int, C.c_void_p]
    L.salp_cycle_step.argtypes = [C.POINTER(SalpCycleParams), C.POINTER(SalpCycleBuffers), C.c_void_p, C.c_void_p]
    if L.salp_abi_version() != SALP_ABI_VERSION:
        raise ImportError("libsalp_b200.so ABI version mismatch; rebuild it")
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != SALP_OK:
        raise SalpError(rc, load().salp_last_error().decode("utf-8", "replace"))
