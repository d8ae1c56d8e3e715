"""Pin oracle/salp_oracle.c against the UNMODIFIED Python reference (build container only).

Runs the reference SalpSnakeEnv (oracle/ref_harness.py) and the C restatement on the same
seeds, RNG word tables and float32 action sequences, and requires BIT-EXACT agreement of
every continuous state variable, the float32 observation, the float64 reward, all flags and
counters, the food layout and the number of RNG words consumed from each stream.

    python -m oracle.validate_against_reference            # all configs, prints a table

TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import sys

import numpy as np

from . import ref_harness as rh
from .salp_oracle import Oracle

# environment.params of the four reference YAMLs (configs/*.yaml) -- read as data
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

STATE_MAP = dict(px="px", py="py", vx="vx", vy="vy", angle="angle", omega="omega", nozzle="nozzle",
                 wv="wv", ellipse_a="a", ellipse_b="b")


def make_actions(T: int, N: int, act_dim: int, seed: int = 1234, style: str = "uniform") -> np.ndarray:
    """Synthetic float32 actions shared by reference, oracle and kernel."""
    g = np.random.default_rng(seed)
    if style == "uniform":
        a = g.uniform(-1.0, 1.0, (T, N, act_dim))
    elif style == "hold":  # piecewise-constant steering: reaches walls/food far more often
        seg = 40
        base = g.uniform(-1.0, 1.0, ((T + seg - 1) // seg, N, act_dim))
        a = np.repeat(base, seg, axis=0)[:T]
    else:
        raise ValueError(style)
    a = a.astype(np.float32)
    if act_dim == 2:  # manual breathing: column 0 is inhale control in [0,1], held in runs
        seg = 25
        hold = g.uniform(0.0, 1.0, ((T + seg - 1) // seg, N))
        a[:, :, 0] = np.repeat(hold, seg, axis=0)[:T].astype(np.float32)
    return a


def tables_for(seeds, food_len: int, jit_len: int):
    fw = np.stack([rh.food_words(s, food_len) for s in seeds])
    jw = np.stack([rh.jitter_words(s, jit_len) for s in seeds])
    return fw, jw


def compare(params: dict, seeds, actions: np.ndarray, food_len: int = 4096, jit_len: int = 4096):
    """Return (n_mismatch_total, report dict, reference rollout, tables)."""
    T, N = actions.shape[:2]
    ref = rh.rollout_reference(params, seeds, actions, autoreset=True)
    fw, jw = tables_for(seeds, food_len, jit_len)
    orc = Oracle(N, fw, jw, **params)
    rep = {}

    def chk(name, a, b):
        a = np.asarray(a)
        b = np.asarray(b)
        bad = int((a.view(np.uint8) != b.astype(a.dtype).view(np.uint8)).reshape(a.shape + (-1,)).any(-1).sum()) \
            if a.dtype.kind == "f" else int((a != b).sum())
        rep[name] = rep.get(name, 0) + bad

    o0 = orc.reset()
    chk("reset_obs0", ref["reset_obs0"], o0)
    F = ref["food"].shape[2]
    chk("food0", ref["food0"] * ref["alive0"][..., None], orc.envs["food"][:, :F] * orc.envs["alive"][:, :F, None])
    for t in range(T):
        obs, rew, flags = orc.step(actions[t].astype(np.float64))
        e = orc.envs
        chk("obs", ref["obs"][t], obs)
        chk("reward", ref["reward"][t], rew)
        chk("terminated", ref["terminated"][t], flags & 1)
        chk("truncated", ref["truncated"][t], (flags >> 1) & 1)
        chk("collision", ref["collision"][t], (flags >> 2) & 1)
        for rk, ok in STATE_MAP.items():
            chk(rk, ref[rk][t], e[ok])
        for k in ("phase", "timer", "steps_since_food", "food_collected", "n_food"):
            chk(k, ref[k][t], e[k])
        chk("score", ref["score"][t], e["score"])
        chk("alive", ref["alive"][t], e["alive"][:, :F])
        chk("food", ref["food"][t] * ref["alive"][t][..., None], e["food"][:, :F] * e["alive"][:, :F, None])
        for i in np.nonzero(flags & 3)[0]:
            ro = orc.reset_one(int(i))
            chk("reset_obs", ref["reset_obs"][t, i], ro)
    chk("cur_food", ref["cur_food"], orc.envs["cur_food"].astype(np.int64))
    chk("cur_jit", ref["cur_jit"], orc.envs["cur_jit"].astype(np.int64))
    assert int(orc.envs["cur_food"].max()) <= food_len and int(orc.envs["cur_jit"].max()) <= jit_len, "table wrapped"
    return sum(rep.values()), rep, ref, (fw, jw)


def main(argv=None):
    T, N = 3000, 6
    total = 0
    print(f"{'case':44s} {'events (term/trunc/food)':26s} mismatches")
    cases = []
    for name, p in CONFIGS.items():
        cases.append((name + "/forced/uniform", p, "uniform"))
        cases.append((name + "/forced/hold", p, "hold"))
        pm = dict(p, forced_breathing=False)
        cases.append((name + "/manual/hold", pm, "hold"))
    cases.append(("no_respawn_5food/forced/hold", dict(num_food_items=5, respawn_food=False), "hold"))
    cases.append(("random_food_count/forced/hold", dict(num_food_items=7, random_food_count=True), "hold"))
    cases.append(("max_observed_food=5/forced/hold", dict(num_food_items=8, max_observed_food=5,
                                                           proximity_reward_weight=2.0), "hold"))
    for ci, (name, p, style) in enumerate(cases):
        ad = 1 if p.get("forced_breathing", True) else 2
        acts = make_actions(T, N, ad, seed=1234 + ci, style=style)
        seeds = [10_000 + 100 * ci + e for e in range(N)]
        bad, rep, ref, _ = compare(p, seeds, acts)
        ev = f"{int(ref['terminated'].sum())}/{int(ref['truncated'].sum())}/{int(ref['food_collected'].max(0).sum())}"
        print(f"{name:44s} {ev:26s} {bad}  {[k for k, v in rep.items() if v]}")
        total += bad
    print("TOTAL mismatches:", total)
    return 0 if total == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
