"""Multi-GPU layout of the env step: contiguous env-index slabs, one process per GPU, nothing but
an 8-double statistics vector ever crosses ranks.

Pure host logic (works with any torch.distributed backend: NCCL on the GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist

STATS_KEYS = ("episodes", "return_sum", "length_sum", "food_sum", "collisions", "truncations", "nonfinite", "env_steps")
# model (B) cycle env (salp_cycle_step, include/salp_b200.h)
CYCLE_STATS_KEYS = ("episodes", "return_sum", "cycle_sum", "targets_reached", "truncations", "nonfinite", "env_steps", "substeps")


def slab_for_rank(total_envs: int, rank: int, world: int) -> Tuple[int, int]:
    """(env_id0, num_envs) of rank's slab: envs [r*N/R, (r+1)*N/R) with the remainder spread over
    the first ranks.  RNG streams are indexed by GLOBAL env id, so results do not depend on R."""
    if not 0 <= rank < world:
        raise ValueError("rank outside world")
    base, rem = divmod(int(total_envs), int(world))
    n = base + (1 if rank < rem else 0)
    start = rank * base + min(rank, rem)
    return start, n


def all_reduce_stats(stats: torch.Tensor, group: Optional[dist.ProcessGroup] = None, keys=STATS_KEYS) -> dict:
    """SUM the per-rank statistics vector over ranks (no-op without an initialised process group)
    and derive the per-episode means."""
    v = stats.detach().clone().to(torch.float64)
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(v, op=dist.ReduceOp.SUM, group=group)
    out = dict(zip(keys, v.tolist()))
    ep = max(out["episodes"], 1.0)
    out["mean_return"] = out["return_sum"] / ep
    if "length_sum" in out:
        out.update(mean_length=out["length_sum"] / ep, mean_food=out["food_sum"] / ep)
    if "cycle_sum" in out:
        out.update(mean_cycles=out["cycle_sum"] / ep, substeps_per_env_step=out["substeps"] / max(out["env_steps"], 1.0))
    return out
