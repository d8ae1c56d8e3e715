"""CPU, world_size 2, gloo: the N>1 host path -- slab ownership by global env id and the ONE
collective on this path, the all-reduce (SUM) of the 8-double episode-statistics vector."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, total, out_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from underwater_swimmer_rl_b200.sharding import all_reduce_stats, slab_for_rank
    start, n = slab_for_rank(total, rank, world)
    # per-rank statistics as the kernel would leave them: episodes, sum return, sum length, ...
    stats = torch.tensor([n, -1.5 * n, 10.0 * n, 0.5 * n, n - 1, 1, 0, 100.0 * n], dtype=torch.float64)
    red = all_reduce_stats(stats)
    # model (B) vector: episodes, sum return, sum cycles, reached, truncations, non-finite, env steps, substeps
    from underwater_swimmer_rl_b200.sharding import CYCLE_STATS_KEYS
    cyc = all_reduce_stats(torch.tensor([n, 2.0 * n, 500.0 * n, 1, n - 1, 0, 4.0 * n, 2900.0 * n], dtype=torch.float64),
                           keys=CYCLE_STATS_KEYS)
    red["cycle"] = cyc
    out_q.put((rank, start, n, red))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_stats_all_reduce():
    world, total = 2, 1001
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, s0, n0, red0), (r1, s1, n1, red1) = res
    assert (s0, n0, s1, n1) == (0, 501, 501, 500)
    assert red0 == red1
    assert red0["episodes"] == total and red0["env_steps"] == 100.0 * total
    assert red0["collisions"] == total - 2 and red0["truncations"] == 2
    assert abs(red0["mean_return"] + 1.5) < 1e-12 and abs(red0["mean_length"] - 10.0) < 1e-12
    cyc = red0["cycle"]
    assert cyc["episodes"] == total and cyc["targets_reached"] == 2 and cyc["truncations"] == total - 2
    assert cyc["mean_cycles"] == 500.0 and cyc["substeps_per_env_step"] == 725.0 and cyc["mean_return"] == 2.0
