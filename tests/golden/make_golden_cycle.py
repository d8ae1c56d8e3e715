"""Generate the committed model (B) fixtures from the UNMODIFIED reference (build container only).

    python tests/golden/make_golden_cycle.py

Writes cycle_<case>.npz next to this file: reference trajectories of SalpRobotEnv over Robot / Nozzle
(/root/reference/src/salp/environments/salp_robot_env.py + robot.py) on seeded np.random streams with float64
action sequences: per env step the float32 observation, reward, terminated / truncated, the number of
integrator substeps, cycle count, phase, the target point and the float64 continuous state
(position, velocity, angular velocity, Euler angles, world-frame velocity).
The RNG word tables are not stored: tests rebuild them from the seeds.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_harness_cycle as rc  # noqa: E402
from oracle.validate_cycle_against_reference import CASES, make_cycle_actions  # noqa: E402


def main():
    for name, (params, T, N, style, seed0) in CASES.items():
        seeds = [seed0 + e for e in range(N)]
        acts = make_cycle_actions(T, N, style, seed=777 + seed0)
        ref = rc.rollout_reference_cycle(params, seeds, acts, autoreset=True)
        meta = dict(case=name, params=params, seeds=seeds, style=style,
                    source="SalpRobotEnv over Robot/Nozzle, /root/reference/src/salp/environments")
        path = os.path.join(HERE, f"cycle_{name}.npz")
        np.savez_compressed(path, meta=json.dumps(meta), actions=acts, obs0=ref["obs0"], obs=ref["obs"],
                            reward=ref["reward"], terminated=ref["terminated"], truncated=ref["truncated"],
                            n_sub=ref["n_sub"], final_obs=ref["final_obs"], cycle=ref["cycle"].astype(np.int16),
                            phase=ref["phase"].astype(np.int8), target=ref["target"], cont=ref["cont"])
        ep = int((ref["terminated"] | ref["truncated"]).sum())
        print(f"{name:26s} T={T} N={N} substeps={int(ref['n_sub'].sum())} episodes={ep} -> {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()
