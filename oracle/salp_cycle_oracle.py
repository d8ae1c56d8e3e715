"""ctypes front-end for oracle/salp_cycle_oracle.c (CPU restatement of the reference's model (B) cycle env).

TEST INFRASTRUCTURE ONLY -- see the header of salp_cycle_oracle.c.  Allowed importers:
tests/, __graft_entry__.smoke(), bench.py (cpu_baseline / --impl reference).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "salp_cycle_oracle.c")
_LIB = os.path.join(_HERE, "libsalp_cycle_oracle.so")

# the construction every reference script uses (salp_robot_env.py:1305-1310, train_robot.py:12-17)
PARAM_DEFAULTS = dict(
    dry_mass=1.0, init_length=0.3, init_width=0.15, max_contraction=0.06, density=1000.0,
    nozzle_length1=0.05, nozzle_length2=0.05, nozzle_length3=0.05, nozzle_area=0.00016, nozzle_mass=1.0,
    width=900, height=700, max_cycles=500, max_substeps=0,
)
OBS_DIM, ACT_DIM = 6, 3


class CycleParams(C.Structure):
    _fields_ = [(n, C.c_double) for n in (
        "dry_mass", "init_length", "init_width", "max_contraction", "density",
        "noz_l1", "noz_l2", "noz_l3", "noz_area", "noz_mass", "width", "height")] + [
        ("max_cycles", C.c_int32), ("max_substeps", C.c_int32)]


class CycleEnv(C.Structure):
    _fields_ = [
        ("pos", C.c_double * 3), ("vel", C.c_double * 3), ("w", C.c_double * 3), ("euler", C.c_double * 3),
        ("prev_I", C.c_double * 3), ("vel_world", C.c_double * 3),
        ("volume", C.c_double), ("prev_dist", C.c_double), ("ep_return", C.c_double), ("time", C.c_double),
        ("length", C.c_double), ("width_b", C.c_double),
        ("target", C.c_float * 2),
        ("cycle", C.c_int32), ("n_sub", C.c_int32), ("phase", C.c_int32), ("pad_", C.c_int32),
        ("cur", C.c_uint64),
    ]


class CycleTables(C.Structure):
    _fields_ = [("words", C.c_void_p), ("rows", C.c_int64), ("len", C.c_int64)]


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
        L.salp_cycle_oracle_env_size.restype = C.c_int64
        assert L.salp_cycle_oracle_env_size() == C.sizeof(CycleEnv)
        L.salp_cycle_oracle_step.restype = C.c_int64
        _lib = L
    return _lib


def make_params(**kw) -> CycleParams:
    d = dict(PARAM_DEFAULTS)
    unknown = set(kw) - set(d)
    if unknown:
        raise TypeError(f"unknown parameters: {sorted(unknown)}")
    d.update(kw)
    return CycleParams(d["dry_mass"], d["init_length"], d["init_width"], d["max_contraction"], d["density"],
                       d["nozzle_length1"], d["nozzle_length2"], d["nozzle_length3"], d["nozzle_area"],
                       d["nozzle_mass"], float(d["width"]), float(d["height"]), int(d["max_cycles"]),
                       int(d["max_substeps"]))


STATE_FIELDS = ("pos", "vel", "w", "euler", "vel_world")  # each [.., 3]


class CycleOracle:
    """N independent model (B) envs; env i consumes row i of the np.random word table."""

    def __init__(self, num_envs: int, words: np.ndarray, **params):
        self.n = int(num_envs)
        self.params = make_params(**params)
        self.words = np.ascontiguousarray(words, dtype=np.uint32)
        assert self.words.ndim == 2 and self.words.shape[0] >= self.n
        self.tables = CycleTables(self.words.ctypes.data, self.words.shape[0], self.words.shape[1])
        self.envs = (CycleEnv * self.n)()
        self.L = lib()

    def reset(self, mask=None, rewind_rng: bool = True) -> np.ndarray:
        obs = np.zeros((self.n, OBS_DIM), np.float32)
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        self.L.salp_cycle_oracle_reset(C.byref(self.params), self.envs, C.c_int64(self.n), C.byref(self.tables),
                                       C.c_int64(0), None if m is None else m.ctypes.data_as(C.c_void_p),
                                       C.c_int(int(rewind_rng)), obs.ctypes.data_as(C.c_void_p))
        return obs

    def step(self, actions: np.ndarray, autoreset: bool = True) -> dict:
        a = np.ascontiguousarray(actions, np.float64).reshape(self.n, ACT_DIM)
        out = dict(obs=np.zeros((self.n, OBS_DIM), np.float32), reward=np.zeros(self.n, np.float64),
                   terminated=np.zeros(self.n, np.uint8), truncated=np.zeros(self.n, np.uint8),
                   n_sub=np.zeros(self.n, np.int32), final_obs=np.zeros((self.n, OBS_DIM), np.float32))
        p = lambda x: x.ctypes.data_as(C.c_void_p)
        out["substeps"] = int(self.L.salp_cycle_oracle_step(
            C.byref(self.params), self.envs, C.c_int64(self.n), C.byref(self.tables), C.c_int64(0), p(a),
            p(out["obs"]), p(out["reward"]), p(out["terminated"]), p(out["truncated"]), p(out["n_sub"]),
            p(out["final_obs"]), C.c_int(int(autoreset))))
        return out

    def state(self) -> dict:
        """Continuous + discrete state after the last step (before any auto-reset it is the reset state)."""
        d = {k: np.array([list(getattr(e, k)) for e in self.envs], np.float64) for k in STATE_FIELDS}
        d["target"] = np.array([list(e.target) for e in self.envs], np.float32)
        d["cycle"] = np.array([e.cycle for e in self.envs], np.int32)
        d["phase"] = np.array([e.phase for e in self.envs], np.int32)
        d["prev_dist"] = np.array([e.prev_dist for e in self.envs], np.float64)
        return d

    def rollout(self, actions: np.ndarray, autoreset: bool = True) -> dict:
        """actions [T, N, 3] -> per-step stacks of step() outputs, plus the pre-reset state of every step
        (cont [T, N, 5, 3] in STATE_FIELDS order, cycle, phase, target)."""
        T = actions.shape[0]
        keys = ("obs", "reward", "terminated", "truncated", "n_sub", "final_obs")
        acc = {k: [] for k in keys + ("cont", "cycle", "phase", "target")}
        for t in range(T):
            o = self.step(actions[t], autoreset=False)
            st = self.state()
            acc["cont"].append(np.stack([st[k] for k in STATE_FIELDS], axis=1))
            for k in ("cycle", "phase", "target"):
                acc[k].append(st[k])
            done = (o["terminated"] | o["truncated"]).astype(np.uint8)
            if autoreset and done.any():
                ro = self.reset(mask=done, rewind_rng=False)
                o["obs"][done != 0] = ro[done != 0]
            for k in keys:
                acc[k].append(o[k])
        return {k: np.stack(v) for k, v in acc.items()}
