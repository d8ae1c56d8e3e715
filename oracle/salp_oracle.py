"""ctypes front-end for oracle/salp_oracle.c (the CPU restatement of the reference path).

TEST INFRASTRUCTURE ONLY -- see the header of salp_oracle.c.  Allowed importers:
tests/, __graft_entry__.smoke(), bench.py (cpu_baseline / --impl reference).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "salp_oracle.c")
_LIB = os.path.join(_HERE, "libsalp_oracle.so")
MAX_FOOD = 32

PARAM_DEFAULTS = dict(  # SalpSnakeEnv.__init__ defaults, salp_snake_env.py:29-33
    width=800, height=600, num_food_items=5, food_reward=10.0, collision_penalty=-50.0,
    time_penalty=-0.1, efficiency_bonus=1.0, forced_breathing=True, max_observed_food=3,
    random_food_count=False, respawn_food=True, proximity_reward_weight=0.0,
    max_steps_without_food=1500,
)


class OracleParams(C.Structure):
    _fields_ = [
        ("width", C.c_double), ("height", C.c_double),
        ("num_food_items", C.c_int32), ("max_observed_food", C.c_int32),
        ("forced_breathing", C.c_int32), ("random_food_count", C.c_int32),
        ("respawn_food", C.c_int32), ("max_steps_without_food", C.c_int32),
        ("food_reward", C.c_double), ("collision_penalty", C.c_double),
        ("time_penalty", C.c_double), ("efficiency_bonus", C.c_double),
        ("proximity_reward_weight", C.c_double),
    ]


class OracleEnv(C.Structure):
    _fields_ = [
        ("px", C.c_double), ("py", C.c_double), ("vx", C.c_double), ("vy", C.c_double),
        ("angle", C.c_double), ("omega", C.c_double),
        ("nozzle", C.c_double), ("target_nozzle", C.c_double),
        ("wv", C.c_double), ("a", C.c_double), ("b", C.c_double), ("score", C.c_double),
        ("food", (C.c_double * 2) * MAX_FOOD), ("alive", C.c_int32 * MAX_FOOD),
        ("n_food", C.c_int32), ("phase", C.c_int32), ("timer", C.c_int32),
        ("exhale_dur", C.c_int32), ("is_inhaling", C.c_int32),
        ("food_collected", C.c_int32), ("steps_since_food", C.c_int32),
        ("shape_kind", C.c_int32), ("shape_timer", C.c_int32), ("pad_", C.c_int32),
        ("cur_food", C.c_uint64), ("cur_jit", C.c_uint64),
    ]


class OracleTables(C.Structure):
    _fields_ = [
        ("food_words", C.c_void_p), ("jit_words", C.c_void_p),
        ("rows", C.c_int64), ("food_len", C.c_int64), ("jit_len", C.c_int64),
    ]


ENV_DTYPE = np.dtype(dict(
    names=[f[0] for f in OracleEnv._fields_],
    formats=["f8"] * 12 + [("f8", (MAX_FOOD, 2)), ("i4", (MAX_FOOD,))] + ["i4"] * 10 + ["u8", "u8"],
    offsets=[getattr(OracleEnv, f[0]).offset for f in OracleEnv._fields_],
    itemsize=C.sizeof(OracleEnv),
))


def build(force: bool = False, openmp: bool = True) -> str:
    """Compile the C restatement (gcc, strict IEEE: no contraction, no fast-math)."""
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(_SRC):
        cmd = ["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-std=c11"]
        if openmp:
            cmd.append("-fopenmp")
        cmd += ["-o", _LIB, _SRC, "-lm"]
        subprocess.check_call(cmd)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        assert L.salp_oracle_env_size() == C.sizeof(OracleEnv)
        L.salp_oracle_rollout.restype = C.c_int64
        _lib = L
    return _lib


def fill_words(rows: int, length: int, seed0: int, first_row: int, kind: int) -> np.ndarray:
    """uint32 [rows, length]: the raw MT19937 outputs of random.seed(s) (kind 0) / np.random.seed(s) (kind 1) for
    s = seed0 + first_row + r, generated inside the oracle library (no dependency on the product package)."""
    out = np.empty((rows, length), dtype=np.uint32)
    lib().salp_oracle_fill_rng_words(_ptr(out), C.c_int64(rows), C.c_int64(length), C.c_uint64(seed0), C.c_int64(first_row), C.c_int(kind))
    return out


def make_params(**kw) -> OracleParams:
    d = dict(PARAM_DEFAULTS)
    unknown = set(kw) - set(d)
    if unknown:
        raise TypeError(f"unknown SalpSnakeEnv parameter(s): {sorted(unknown)}")
    d.update(kw)
    d["num_food_items"] = max(0, int(d["num_food_items"]))
    if d["num_food_items"] > MAX_FOOD:
        raise ValueError("oracle supports at most 32 food items")
    p = OracleParams()
    for k, _ in OracleParams._fields_:
        setattr(p, k, type(getattr(p, k))(d[k]))
    return p


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Oracle:
    """A batch of independent oracle envs with global ids env_id0 .. env_id0+n-1."""

    def __init__(self, n: int, food_words: np.ndarray, jit_words: np.ndarray, env_id0: int = 0, **params):
        self.p = make_params(**params)
        self.n = int(n)
        self.env_id0 = int(env_id0)
        self.food_words = np.ascontiguousarray(food_words, dtype=np.uint32)
        self.jit_words = np.ascontiguousarray(jit_words, dtype=np.uint32)
        assert self.food_words.ndim == 2 and self.jit_words.ndim == 2
        assert self.food_words.shape[0] == self.jit_words.shape[0]
        for L in (self.food_words.shape[1], self.jit_words.shape[1]):
            assert L & (L - 1) == 0, "table lengths must be powers of two"
        self.t = OracleTables(self.food_words.ctypes.data, self.jit_words.ctypes.data,
                              self.food_words.shape[0], self.food_words.shape[1], self.jit_words.shape[1])
        self.envs = np.zeros(self.n, dtype=ENV_DTYPE)
        self.obs_dim = 10 + 4 * self.p.max_observed_food + 2
        self.act_dim = 1 if self.p.forced_breathing else 2

    def reset(self) -> np.ndarray:
        obs = np.zeros((self.n, self.obs_dim), np.float32)
        lib().salp_oracle_reset(C.byref(self.p), _ptr(self.envs), C.c_int64(self.n), C.c_int64(self.env_id0),
                                C.byref(self.t), _ptr(obs))
        return obs

    def reset_one(self, i: int) -> np.ndarray:
        obs = np.zeros((1, self.obs_dim), np.float32)
        sub = self.envs[i:i + 1]
        lib().salp_oracle_reset(C.byref(self.p), _ptr(sub), C.c_int64(1), C.c_int64(self.env_id0 + i),
                                C.byref(self.t), _ptr(obs))
        return obs[0]

    def observe(self) -> np.ndarray:
        """_get_extended_observation() of the current (possibly poked) state, without stepping."""
        obs = np.zeros((self.n, self.obs_dim), np.float32)
        lib().salp_oracle_observe(C.byref(self.p), _ptr(self.envs), C.c_int64(self.n), _ptr(obs))
        return obs

    def step(self, actions: np.ndarray):
        a = np.ascontiguousarray(np.asarray(actions, dtype=np.float64).reshape(self.n, self.act_dim))
        obs = np.zeros((self.n, self.obs_dim), np.float32)
        rew = np.zeros(self.n, np.float64)
        flags = np.zeros(self.n, np.uint8)
        lib().salp_oracle_step(C.byref(self.p), _ptr(self.envs), C.c_int64(self.n), C.c_int64(self.env_id0),
                               C.byref(self.t), _ptr(a), _ptr(obs), _ptr(rew), _ptr(flags))
        return obs, rew, flags

    def rollout(self, actions: np.ndarray, autoreset: bool = True, record=True):
        """actions [T, n, act_dim] float32.  Returns dict (obs/reset_obs/reward/flags if record;
        record="flags": only the per-step event flags and rewards -- long free-running comparisons)."""
        a = np.ascontiguousarray(actions, dtype=np.float32)
        T = a.shape[0]
        assert a.shape[1] == self.n and a.reshape(T, self.n, -1).shape[2] == self.act_dim
        obs = reset_obs = rew = flags = None
        if record == "flags":
            rew = np.zeros((T, self.n), np.float64)
            flags = np.zeros((T, self.n), np.uint8)
        elif record:
            obs = np.zeros((T, self.n, self.obs_dim), np.float32)
            reset_obs = np.zeros((T, self.n, self.obs_dim), np.float32)
            rew = np.zeros((T, self.n), np.float64)
            flags = np.zeros((T, self.n), np.uint8)
        stats = np.zeros(6, np.int64)
        rsum = C.c_double(0.0)
        lib().salp_oracle_rollout(C.byref(self.p), _ptr(self.envs), C.c_int64(self.n), C.c_int64(self.env_id0),
                                  C.byref(self.t), C.c_int64(T), _ptr(a), C.c_int(1 if autoreset else 0),
                                  _ptr(obs), _ptr(reset_obs), _ptr(rew), _ptr(flags), _ptr(stats), C.byref(rsum))
        return dict(obs=obs, reset_obs=reset_obs, reward=rew, flags=flags, stats=stats, return_sum=rsum.value)
