"""Pin oracle/salp_cycle_oracle.c against the UNMODIFIED reference model (B) (build container only).

    python -m oracle.validate_cycle_against_reference [--quick]

For every case: same np.random seeds, same float64 action sequences; DISCRETE outputs (substeps per cycle,
phase, cycle count, terminated, truncated) must be bit-identical, continuous state / reward / observation
within 1e-6 (abs + rel).  Why not tighter: the reference's nozzle IK (robot.py:74-78) evaluates
arcsin(c / sqrt(a^2 + b^2)) where the argument is analytically -sign(yaw) for EVERY yaw, so angle1 carries
+-sqrt(2 ulp) ~ 2e-8 of rounding noise that depends on the libm in use (numpy's SIMD sin/cos vs glibc vs
CUDA); trajectories whose IK roundings happen to coincide agree to 1e-14.
"""
from __future__ import annotations

import sys
import time

import numpy as np

from . import ref_harness as rh
from . import ref_harness_cycle as rc
from .salp_cycle_oracle import CycleOracle

CASES = {
    # name: (params, T, N, action style, first seed)
    "uniform": (dict(), 40, 3, "uniform", 21_000),
    "short_cycles": (dict(), 120, 2, "short", 22_000),
    "seek": (dict(), 60, 3, "seek", 23_000),
    # the three ways an episode ends (salp_robot_env.py:178-188): 500 cycles, target reached, > 5 m from the target
    "max_cycles": (dict(), 1010, 2, "tiny", 24_000),
    "tiny_tank_reached": (dict(width=102, height=102), 40, 3, "short", 26_000),
    "big_tank_out_of_range": (dict(width=3000, height=2400), 24, 3, "uniform", 27_000),
    "small_tank_thin_nozzle": (dict(width=500, height=400, nozzle_area=0.00009, nozzle_length1=0.01,
                                    nozzle_length2=0.01, nozzle_length3=0.01), 40, 2, "uniform", 25_000),
}


def make_cycle_actions(T: int, N: int, style: str, seed: int) -> np.ndarray:
    """float64 actions inside the env's Box ([0,1], [0,1], [-1,1]; salp_robot_env.py:51-55)."""
    g = np.random.default_rng(seed)
    a = np.zeros((T, N, 3))
    if style == "uniform":
        a[..., 0] = g.uniform(0, 1, (T, N)); a[..., 1] = g.uniform(0, 1, (T, N)); a[..., 2] = g.uniform(-1, 1, (T, N))
    elif style == "short":  # many short cycles incl. zero-length ones and exact end-of-phase boundaries
        a[..., 0] = g.choice([0.0, 0.25, 0.5, 1.0], (T, N)); a[..., 1] = g.choice([0.0, 0.001, 0.05, 0.1], (T, N))
        a[..., 2] = g.choice([-1.0, -0.5, 0.0, 0.5, 1.0], (T, N))
    elif style == "seek":  # strong strokes, little coasting, slowly varying yaw
        a[..., 0] = g.uniform(0.6, 1.0, (T, N)); a[..., 1] = g.uniform(0, 0.15, (T, N))
        a[..., 2] = np.clip(np.cumsum(g.normal(0, 0.15, (T, N)), axis=0), -1, 1)
    elif style == "tiny":  # ~25-substep cycles: reaches the 500-cycle truncation cheaply
        a[..., 0] = g.uniform(0.02, 0.08, (T, N)); a[..., 1] = g.choice([0.0, 0.002], (T, N)); a[..., 2] = g.uniform(-1, 1, (T, N))
    else:
        raise ValueError(style)
    return a


def compare(ref: dict, orc: dict, tol: float = 1e-6) -> dict:
    for k in ("n_sub", "terminated", "truncated", "cycle", "phase"):
        if not np.array_equal(ref[k], orc[k]):
            bad = np.argwhere(ref[k] != orc[k])[0]
            raise AssertionError(f"discrete field {k} differs first at (t, env) = {tuple(bad)}: "
                                 f"ref {ref[k][tuple(bad)]} oracle {orc[k][tuple(bad)]}")
    assert np.array_equal(ref["target"], orc["target"]), "target points differ"
    err = {}
    for k in ("cont", "reward"):
        d = np.abs(ref[k] - orc[k]) / (1.0 + np.abs(ref[k]))
        err[k] = float(d.max())
        assert err[k] <= tol, (k, err[k])
    for k in ("obs", "final_obs"):
        d = np.abs(ref[k].astype(np.float64) - orc[k]) / (1.0 + np.abs(ref[k]))
        err[k] = float(d.max())
        assert err[k] <= tol, (k, err[k])
    return err


def run_case(name: str, quick: bool = False) -> dict:
    params, T, N, style, seed0 = CASES[name]
    if quick:
        T, N = max(8, T // 4), min(N, 2)
    seeds = [seed0 + e for e in range(N)]
    acts = make_cycle_actions(T, N, style, seed=777 + seed0)
    ref = rc.rollout_reference_cycle(params, seeds, acts, autoreset=True)
    words = np.stack([rh.jitter_words(s, 256) for s in seeds])
    orc = CycleOracle(N, words, **params)
    o0 = orc.reset()
    assert np.array_equal(o0, ref["obs0"]), "reset observation differs"
    out = orc.rollout(acts, autoreset=True)
    err = compare(ref, out)
    err.update(substeps=int(ref["n_sub"].sum()), episodes=int((ref["terminated"] | ref["truncated"]).sum()))
    return err


def main():
    quick = "--quick" in sys.argv
    for name in CASES:
        t0 = time.time()
        err = run_case(name, quick)
        print(f"{name:26s} OK  {err}  ({time.time() - t0:.1f} s)")


if __name__ == "__main__":
    main()
