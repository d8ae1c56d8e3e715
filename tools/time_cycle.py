"""Time the model (B) cycle kernel (tools/: tuning aid)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
env = SalpCycleVectorEnv(N, table_rows=1 << 16, table_len=64)
env.reset()
g = torch.Generator(device="cuda").manual_seed(1)
acts = []
for i in range(8):
    a = torch.rand(N, 3, device="cuda", generator=g)
    a[:, 2] = a[:, 2] * 2 - 1
    acts.append(a)
for i in range(4):
    env.step(acts[i])
torch.cuda.synchronize()
env.episode_stats(reset=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(40):
    env.step(acts[i % 8])
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 40
st = env.episode_stats()
print(f"N={N}: {ms:.3f} ms/step, {st['substeps'] / 40 / ms / 1e6:.1f} G substeps/s, "
      f"{N / ms / 1e3:.1f} M env-steps/s")
