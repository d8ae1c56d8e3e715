"""Time the forced-breathing kernel against the generic one on the same box (builder tool, not a bench):
    python tools/time_fast.py [--sizes 1048576,131072] [--cfgs single_food,defaults] [--steps 300]
Prints us/step for {fast, generic} x {table, philox} per (config, slab size), steady state (primed), CUDA events around
graph replays of 32 steps (launch overhead excluded) and around eager launches."""
import argparse
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import DEFAULTS, SINGLE_FOOD  # noqa: E402
from underwater_swimmer_rl_b200 import SalpVectorEnv  # noqa: E402

CFG = {"single_food": SINGLE_FOOD, "defaults": DEFAULTS}


def time_one(cfg, N, generic, rng, steps, prime, graph, rows=0):
    env = SalpVectorEnv(N, dtype="float32", seed=10_000, food_table_len=128, jitter_table_len=512, autoreset=True,
                        generic_kernel=generic, rng=rng, table_rows=rows or None, **CFG[cfg])
    env.reset()
    g = torch.Generator(device="cuda")
    g.manual_seed(1)
    pool = [torch.rand(N, 1, device="cuda", generator=g) * 2 - 1 for _ in range(32)]
    if graph:
        rep = env.capture_graph(pool)
        run = lambda n: [rep() for _ in range(n // 32)]  # noqa: E731
    else:
        run = lambda n: [env.step(pool[i % 32]) for i in range(n)]  # noqa: E731
    run(prime)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(steps)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / (steps // 32 * 32 if graph else steps) * 1e3)
    st = env.episode_stats()
    env.close()
    return best, st


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1048576,131072,32768")
    ap.add_argument("--cfgs", default="single_food,defaults")
    ap.add_argument("--steps", type=int, default=320)
    ap.add_argument("--prime", type=int, default=2048)
    ap.add_argument("--modes", default="fast/table,generic/table,fast/philox")
    ap.add_argument("--rows", type=int, default=0, help="rows of the RNG tables (default: one per env)")
    ap.add_argument("--graph-only", action="store_true")
    a = ap.parse_args()
    for cfg in a.cfgs.split(","):
        for N in [int(x) for x in a.sizes.split(",")]:
            for mode in a.modes.split(","):
                k, rng = mode.split("/")
                for graph in ((True,) if a.graph_only else (True, False)):
                    us, st = time_one(cfg, N, k == "generic", rng, a.steps, a.prime, graph, a.rows)
                    nf = CFG[cfg]["num_food_items"]
                    bpe = 250 if nf == 1 else 426
                    print(f"{cfg:12s} N={N:8d} {mode:14s} {'graph' if graph else 'eager'}: {us:8.2f} us/step  "
                          f"{N / us * 1e6:.3e} env-steps/s  alg {bpe * N / us / 1e3:7.1f} GB/s  mean_len {st['mean_length']:.0f}", flush=True)


if __name__ == "__main__":
    main()
