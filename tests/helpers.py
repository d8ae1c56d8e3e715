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


def push_oracle_state(env, orc) -> None:
    """Teacher forcing: overwrite the device state of ``env`` (SalpVectorEnv, any build) with the
    fp64 state of the oracle batch ``orc`` (tests only; the kernel then advances exactly one step
    from the same state the oracle advances from)."""
    import torch
    e = orc.envs
    F = env.food_slots
    np_real = np.float64 if env.real == torch.float64 else np.float32
    body = np.stack([np.stack([e["px"], e["py"], e["vx"], e["vy"]], 1),
                     np.stack([e["angle"], e["omega"], e["nozzle"], e["wv"]], 1)])
    env.body.copy_(torch.from_numpy(body.astype(np_real)))
    ctrl = env.ctrl.cpu().numpy()
    ctrl[0, :, 0] = e["phase"] | (e["timer"] << 2) | (e["exhale_dur"] << 10) | (e["shape_kind"] << 18) | (e["shape_timer"] << 20)
    ctrl[0, :, 1] = e["steps_since_food"]
    ctrl[0, :, 2] = e["food_collected"]
    mask = (e["alive"][:, :F].astype(np.int64) << np.arange(F)).sum(1)
    ctrl[0, :, 3] = np.where(mask >= 2 ** 31, mask - 2 ** 32, mask)
    ctrl[1, :, 0] = e["cur_food"]
    ctrl[1, :, 1] = e["cur_jit"]
    ctrl[1, :, 2] = (ctrl[1, :, 2] & 0x3FFFFFF) | (e["n_food"].astype(np.int64) << 26).astype(np.int32)
    env.ctrl.copy_(torch.from_numpy(ctrl))
    env.food.copy_(torch.from_numpy(np.ascontiguousarray(e["food"][:, :F].transpose(1, 0, 2)).astype(np_real)))
    env.refresh_rng()  # the cursors were overwritten from outside


def steering_actions(obs: np.ndarray, t: int, g: np.random.Generator, act_dim: int = 1) -> np.ndarray:
    """Closed-loop float32 actions computed from the ORACLE's observation so that every kind of event
    occurs: a third of the lanes steer towards the nearest food (column 13 = relative bearing / pi, the
    recipe of oracle.validate_against_reference.seeker_actions), a third steer away from it (they end
    on a wall), the rest hold random steering segments (they time out or drift into a wall)."""
    n = obs.shape[0]
    lane = np.arange(n) % 3
    noise = 0.15 * g.standard_normal(n)
    seek = np.clip(-0.4 * obs[:, 13] * 3.0 + noise, -1, 1)
    flee = np.clip(+0.4 * obs[:, 13] * 3.0 + noise, -1, 1)
    seg = np.random.default_rng(1000 + t // 60).uniform(-1, 1, n)
    a = np.where(lane == 0, seek, np.where(lane == 1, flee, np.clip(0.7 * seg + 2 * noise, -1, 1))).astype(np.float32)
    if act_dim == 1:
        return a[:, None]
    hold = (np.random.default_rng(2000 + t // 45).uniform(0, 1, n)).astype(np.float32)
    return np.stack([hold, a], axis=1)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Philox4x32-10 (Salmon et al., SC'11) on numpy uint32 arrays; independent restatement of the kernel's generator."""
    M0, M1, W0, W1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), 0x9E3779B9, 0xBB67AE85
    c = [np.asarray(x, dtype=np.uint32).copy() for x in np.broadcast_arrays(c0, c1, c2, c3)]
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = c[0].astype(np.uint64) * M0
        p1 = c[2].astype(np.uint64) * M1
        hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
        hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
        c = [hi1 ^ c[1] ^ np.uint32(k0), lo1, hi0 ^ c[3] ^ np.uint32(k1), lo0]
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return c


def philox_tables(seed: int, env_ids, food_len: int, jit_len: int):
    """The word streams of SALP_RNG_PHILOX (include/salp_b200.h) as tables: word w of stream s (0 food, 1 jitter) of the
    env with global id g = philox4x32_10(counter = (w >> 2, s, g_lo, g_hi), key = (seed_lo, seed_hi))[w & 3]."""
    g = np.asarray(list(env_ids), dtype=np.uint64)
    out = []
    for s, n in ((0, food_len), (1, jit_len)):
        blocks = np.arange((n + 3) // 4, dtype=np.uint32)
        r = philox4x32_10(blocks[None, :], np.uint32(s), (g & np.uint64(0xFFFFFFFF)).astype(np.uint32)[:, None],
                          (g >> np.uint64(32)).astype(np.uint32)[:, None], seed & 0xFFFFFFFF, seed >> 32)
        out.append(np.stack(r, axis=-1).reshape(len(g), -1)[:, :n].astype(np.uint32))
    return out[0], out[1]
