"""Time the host-buffer step (salp_step_host) alone: python tools/time_e2e.py [N] [steps]  (builder tool)"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import SINGLE_FOOD  # noqa: E402
from underwater_swimmer_rl_b200 import SalpVectorEnv  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
env = SalpVectorEnv(N, dtype="float32", rng="philox", autoreset=True, **SINGLE_FOOD)
env.reset()
a_h = (torch.rand(N, 1) * 2 - 1).pin_memory()
o_h = torch.empty(N, 24).pin_memory()
r_h = torch.empty(N).pin_memory()
te_h = torch.empty(N, dtype=torch.bool).pin_memory()
tr_h = torch.empty(N, dtype=torch.bool).pin_memory()
for _ in range(5):
    env.step_host(a_h, o_h, r_h, te_h, tr_h)
best = 1e9
for rep in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        env.step_host(a_h, o_h, r_h, te_h, tr_h)
    torch.cuda.synchronize()
    best = min(best, (time.perf_counter() - t0) / steps)
d2h = N * 102
print(f"N={N}: {best * 1e6:.1f} us per host-buffer step, {N / best:.4g} env-steps/s, D2H {d2h / best / 1e9:.2f} GB/s")
