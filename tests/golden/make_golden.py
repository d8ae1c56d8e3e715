"""Generate the committed golden fixtures from the UNMODIFIED reference (build container only).

    python tests/golden/make_golden.py

Writes, next to this file:
  traj_<case>.npz   reference trajectories (SalpSnakeEnv over the legacy SalpRobotEnv,
                    /root/reference/src/salp/environments/salp_snake_env.py +
                    /root/reference/scripts/utilities/salp_robot.py) on seeded RNG streams:
                    every step  -> obs f32[24], reward f64, flags, discrete state;
                    every 25th  -> full float64 continuous state + food layout.
  human_demos.npz   the reference's own recorded human demonstrations
                    (/root/reference/data/expert_demos/human/*.pkl): the f64 action
                    sequences and the robot-state observation columns 0..9.
The RNG word tables are NOT stored: tests rebuild them from the seeds with the Python
stdlib / numpy MT19937 generators (oracle.ref_harness.food_words / jitter_words).
"""
from __future__ import annotations

import glob
import json
import os
import pickle
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_harness as rh  # noqa: E402
from oracle.validate_against_reference import CONFIGS, make_actions, seeker_actions  # noqa: E402

SUB = 25
CASES = [
    # name, params, T, N, action style, first seed
    ("single_food_uniform", CONFIGS["single_food"], 1200, 3, "uniform", 10_000),
    ("single_food_seek", CONFIGS["single_food"], 2400, 3, "seek", 11_000),
    ("single_food_flee", CONFIGS["single_food"], 1800, 2, "flee", 11_500),
    ("single_food_manual", dict(CONFIGS["single_food"], forced_breathing=False), 2000, 2, "seek", 12_000),
    ("long_horizon_flee", CONFIGS["single_food_long_horizon"], 2600, 2, "flee", 13_000),
    ("defaults_seek", CONFIGS["defaults"], 1800, 3, "seek", 14_000),
    ("defaults_manual", dict(CONFIGS["defaults"], forced_breathing=False), 1500, 2, "seek", 15_000),
    ("no_respawn", dict(num_food_items=3, respawn_food=False, proximity_reward_weight=1.0), 2400, 2, "seek", 16_000),
    ("random_food_count", dict(num_food_items=7, random_food_count=True), 1500, 2, "seek", 17_000),
    ("observe5", dict(num_food_items=8, max_observed_food=5, proximity_reward_weight=2.0), 1000, 2, "seek", 18_000),
]


def main():
    for ci, (name, params, T, N, style, seed0) in enumerate(CASES):
        ad = 1 if params.get("forced_breathing", True) else 2
        seeds = [seed0 + e for e in range(N)]
        if style in ("seek", "flee"):
            acts = seeker_actions(params, seeds, T, seed=4321 + ci, gain=0.4 if style == "seek" else -0.4)
        else:
            acts = make_actions(T, N, ad, seed=4321 + ci, style=style)
        ref = rh.rollout_reference(params, seeds, acts, autoreset=True)
        flags = (ref["terminated"] | (ref["truncated"] << 1) | (ref["collision"] << 2)).astype(np.uint8)
        sub = np.arange(0, T, SUB)
        cont = np.stack([ref[k][sub] for k in rh.STATE_FIELDS], axis=-1)  # [T/SUB, N, 10] f64
        out = dict(
            params=json.dumps(params), seeds=np.array(seeds, np.int64), actions=acts,
            obs=ref["obs"], reset_obs=ref["reset_obs"], reset_obs0=ref["reset_obs0"],
            reward=ref["reward"], flags=flags,
            phase=ref["phase"].astype(np.uint8), timer=ref["timer"].astype(np.uint8),
            steps_since_food=ref["steps_since_food"], food_collected=ref["food_collected"],
            n_food=ref["n_food"].astype(np.uint8), alive=ref["alive"],
            sub_steps=sub, sub_state=cont, sub_food=ref["food"][sub], sub_score=ref["score"][sub],
            food0=ref["food0"], alive0=ref["alive0"],
            cur_food=ref["cur_food"], cur_jit=ref["cur_jit"],
        )
        # reset_obs rows are only meaningful on done steps: zero elsewhere compresses away
        path = os.path.join(HERE, f"traj_{name}.npz")
        np.savez_compressed(path, **out)
        ev = (int(ref["terminated"].sum()), int(ref["truncated"].sum()), int(ref["food_collected"].max(0).sum()))
        print(f"{name:24s} T={T} N={N} events term/trunc/food={ev} words food/jit={ref['cur_food'].max()}/"
              f"{ref['cur_jit'].max()}  {os.path.getsize(path) / 1024:.0f} KiB")

    demos = {}
    for f in sorted(glob.glob(os.path.join(rh.REFERENCE_ROOT, "data/expert_demos/human/*.pkl"))):
        d = pickle.load(open(f, "rb"))
        key = os.path.basename(f).replace("human_demo_", "").replace(".pkl", "")
        demos[key + "_actions"] = np.asarray(d["actions"], np.float64)
        demos[key + "_obs10"] = np.asarray(d["next_observations"], np.float32)[:, :10]
        demos[key + "_obs10_first"] = np.asarray(d["observations"], np.float32)[0, :10]
        demos[key + "_dones"] = np.asarray(d["dones"], np.uint8)
    path = os.path.join(HERE, "human_demos.npz")
    np.savez_compressed(path, **demos)
    print(f"human_demos.npz: {len(demos) // 4} demos, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
