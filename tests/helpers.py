"""Shared helpers for the parity tests (oracle side + GPU side)."""
from __future__ import annotations

import glob
import json
import os
import random

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")

# environment.params of the reference YAMLs (configs/{single_food,single_food_long_horizon,defaults,sac_gail}.yaml)
CONFIGS = {
    "single_food": dict(num_food_items=1, food_reward=1000.0, collision_penalty=-10.0, time_penalty=-0.1,
                        proximity_reward_weight=5.0, respawn_food=True, forced_breathing=True,
                        max_steps_without_food=1500, efficiency_bonus=0.0),
    "single_food_long_horizon": dict(num_food_items=1, food_reward=1000.0, collision_penalty=-500.0,
                                     time_penalty=-0.5, proximity_reward_weight=3.0, respawn_food=True,
                                     forced_breathing=True, max_steps_without_food=2000, efficiency_bonus=0.0),
    "defaults": dict(num_food_items=12, food_reward=15.0, collision_penalty=-30.0, time_penalty=-0.05,
                     efficiency_bonus=2.0, forced_breathing=True),
    "sac_gail": dict(num_food_items=12, food_reward=15.0, collision_penalty=-30.0, time_penalty=-0.05,
                     efficiency_bonus=2.0, forced_breathing=True),
}


def golden_cases():
    return sorted(os.path.basename(p)[5:-4] for p in glob.glob(os.path.join(GOLDEN, "traj_*.npz")))


def load_golden(name: str) -> dict:
    z = np.load(os.path.join(GOLDEN, f"traj_{name}.npz"))
    d = {k: z[k] for k in z.files}
    d["params"] = json.loads(str(d["params"]))
    return d


def py_food_words(seed: int, n: int) -> np.ndarray:
    """Raw MT19937 outputs of CPython's random.seed(seed) stream (ground truth for the tables)."""
    r = random.Random(int(seed))
    return np.array([r.getrandbits(32) for _ in range(n)], dtype=np.uint32)


def py_jitter_words(seed: int, n: int) -> np.ndarray:
    """Raw MT19937 outputs of numpy's legacy np.random.seed(seed) stream."""
    st = np.random.RandomState(int(seed)).get_state()
    r = random.Random()
    r.setstate((3, tuple(int(k) for k in st[1]) + (int(st[2]),), None))
    return np.array([r.getrandbits(32) for _ in range(n)], dtype=np.uint32)


def py_tables(seeds, food_len: int, jit_len: int):
    return (np.stack([py_food_words(s, food_len) for s in seeds]),
            np.stack([py_jitter_words(s, jit_len) for s in seeds]))


def synth_actions(T: int, N: int, act_dim: int, seed: int = 1234, style: str = "mixed") -> np.ndarray:
    """Seeded float32 actions [T, N, act_dim]; 'mixed' = uniform noise on top of held steering
    segments so that walls, food and timeouts all occur."""
    g = np.random.default_rng(seed)
    seg = 60
    base = np.repeat(g.uniform(-1.0, 1.0, ((T + seg - 1) // seg, N, 1)), seg, axis=0)[:T]
    noise = g.uniform(-1.0, 1.0, (T, N, 1))
    w = g.uniform(0.0, 1.0, (1, N, 1))
    steer = np.clip(w * base + (1 - w) * noise, -1, 1)
    if style == "uniform":
        steer = noise
    if act_dim == 1:
        return steer.astype(np.float32)
    seg2 = 45
    hold = np.repeat(g.uniform(0.0, 1.0, ((T + seg2 - 1) // seg2, N, 1)), seg2, axis=0)[:T]
    return np.concatenate([hold, steer], axis=2).astype(np.float32)


def bits_equal(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b).astype(a.dtype)
    if a.dtype.kind == "f":
        ia = a.view(f"u{a.dtype.itemsize}")
        ib = b.view(f"u{a.dtype.itemsize}")
        return (ia == ib) | ((a == 0) & (b == 0))  # +0 == -0
    return a == b
