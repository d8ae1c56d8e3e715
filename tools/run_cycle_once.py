"""A few steps of the model (B) cycle env for ncu captures (tools/: not part of the product)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
env = SalpCycleVectorEnv(N, table_rows=1 << 16, table_len=64)
env.reset()
g = torch.Generator(device="cuda").manual_seed(1)
for i in range(4):
    a = torch.rand(N, 3, device="cuda", generator=g)
    a[:, 2] = a[:, 2] * 2 - 1
    env.episode_stats(reset=True)
    env.step(a)
torch.cuda.synchronize()
print(env.episode_stats())
