"""Run the step kernel for ncu captures (builder tool): python tools/run_fast_once.py CFG N MODE [steps]
MODE = fast/table | fast/philox | generic/table ...; primes 2048 steps, then runs `steps` (default 8) eager launches."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import DEFAULTS, SINGLE_FOOD  # noqa: E402
from underwater_swimmer_rl_b200 import SalpVectorEnv  # noqa: E402

cfg, N, mode = sys.argv[1], int(sys.argv[2]), sys.argv[3]
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 8
k, rng = mode.split("/")
env = SalpVectorEnv(N, dtype="float32", seed=10_000, food_table_len=128, jitter_table_len=512, autoreset=True,
                    generic_kernel=(k == "generic"), rng=rng, **{"single_food": SINGLE_FOOD, "defaults": DEFAULTS}[cfg])
env.reset()
g = torch.Generator(device="cuda")
g.manual_seed(1)
pool = [torch.rand(N, 1, device="cuda", generator=g) * 2 - 1 for _ in range(32)]
for i in range(2048 + steps):
    env.step(pool[i % 32])
torch.cuda.synchronize()
print("done", env.episode_stats()["mean_length"])
