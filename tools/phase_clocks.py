"""Where does one tile's latency go?  (builder tool for the forced-breathing kernel, needs the instrumented build)

    SALP_LIB_TAG=clk SALP_STEP_DEFS=-DSALP_LAB_CLOCK python -m underwater_swimmer_rl_b200.build
    SALP_LIB_TAG=clk python tools/phase_clocks.py defaults 32768

The instrumented kernel stores clock64() of every warp's FIRST tile at fixed points; this prints, per phase, the median and
the maximum over warps (cycles), and the same split for the warp that finished last -- the one that sets a small slab's time.
"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import DEFAULTS, SINGLE_FOOD  # noqa: E402
from underwater_swimmer_rl_b200 import SalpVectorEnv, _capi  # noqa: E402

PHASES = [("entry -> dependency wait done", 0, 1), ("issue loads -> loop top", 1, 2), ("wait core records", 2, 3),
          ("physics", 3, 4), ("wait food records", 4, 5), ("food loop", 5, 6), ("reward / counters / flags", 6, 7),
          ("obs pass 0 (+ wait obs buffer)", 7, 8), ("stats / reset / obs pass 1", 8, 9), ("publish obs + state", 9, 10),
          ("exit", 10, 11), ("TOTAL", 0, 11)]
RESET_PHASES = [("  episode statistics", 8, 16), ("  final_info / final_obs stores", 16, 17), ("  food layout (layout_warp)", 17, 18),
                ("  state reset + reselect", 18, 19), ("  obs pass 1", 19, 9)]


def main():
    cfg, n = sys.argv[1], int(sys.argv[2])
    rng = sys.argv[3] if len(sys.argv) > 3 else "philox"
    params = {"defaults": DEFAULTS, "single_food": SINGLE_FOOD}[cfg]
    env = SalpVectorEnv(n, dtype="float32", seed=10_000, food_table_len=128, jitter_table_len=512, autoreset=True, rng=rng, **params)
    lib = _capi.load()
    lib.salp_lab_read_clocks.argtypes = [C.c_void_p, C.c_int, C.c_int]
    env.reset()
    g = torch.Generator(device="cuda")
    g.manual_seed(1)
    pool = [torch.rand(n, 1, device="cuda", generator=g) * 2 - 1 for _ in range(32)]
    # back-to-back launches from a CUDA graph, as in the bench (an eager launch runs alone on a cold instruction cache);
    # every launch overwrites the clocks, so a read-out holds the LAST of the 32 launches of a replay
    eager = os.environ.get("SALP_LAB_EAGER") == "1"
    rep = None if eager else env.capture_graph(pool)
    for i in range(2048 if eager else 64):
        env.step(pool[i % 32]) if eager else rep()
    torch.cuda.synchronize()
    warps = min(8192, (n + 31) // 32)
    buf = np.zeros((8192, 24), dtype=np.int64)
    rows = []
    for i in range(24):
        assert lib.salp_lab_read_clocks(buf.ctypes.data, buf.size, 1) == 0
        env.step(pool[i % 32]) if eager else rep()
        torch.cuda.synchronize()
        assert lib.salp_lab_read_clocks(buf.ctypes.data, buf.size, 0) == 0
        rows.append(buf[:warps].copy())
    lay = np.zeros(8, dtype=np.uint64)
    lib.salp_lab_read_layout.argtypes = [C.c_void_p, C.c_int]
    if lib.salp_lab_read_layout(lay.ctypes.data, 0) == 0 and lay[0] > 0:
        c = float(lay[0])
        print(f"layout_warp: {int(lay[0])} calls, {lay[5] / c:.2f} windows per call; cycles per call: total {lay[4] / c:.0f} = "
              f"words+uniform {lay[1] / c:.0f} + placed-food checks {lay[2] / c:.0f} + conflict mask {lay[3] / c:.0f} + "
              f"placements/rest {(float(lay[4]) - float(lay[1]) - float(lay[2]) - float(lay[3])) / c:.0f}")
    ghz = 1.0
    try:
        import pynvml
        pynvml.nvmlInit()
        ghz = pynvml.nvmlDeviceGetClockInfo(pynvml.nvmlDeviceGetHandleByIndex(0), pynvml.NVML_CLOCK_SM) / 1000.0
    except Exception:
        pass
    print(f"{cfg} N={n} rng={rng}: {warps} warps with a first tile, {len(rows)} launches, SM clock ~{ghz:.2f} GHz")
    allr = np.concatenate(rows)
    ok = allr[:, 10] > 0
    allr = allr[ok]
    has_reset = allr[:, 12] > 0
    print(f"warps whose tile had a finished episode: {has_reset.mean() * 100:.1f} %")
    for name, sel in (("tiles without a reset", ~has_reset), ("tiles WITH a reset", has_reset)):
        r = allr[sel]
        if len(r) == 0:
            continue
        print(f"--- {name} ({len(r)})")
        for label, i0, i1 in PHASES:
            d = r[:, i1] - r[:, i0]
            print(f"  {label:34s} median {np.median(d):8.0f}  p90 {np.percentile(d, 90):8.0f}  max {d.max():8.0f} cycles")
        if sel is has_reset:
            one = r[r[:, 12] == 1]
            for label, i0, i1 in RESET_PHASES:
                d = one[:, i1] - one[:, i0]
                print(f"  {label:34s} median {np.median(d):8.0f}  p90 {np.percentile(d, 90):8.0f}  max {d.max():8.0f} cycles (tiles with ONE reset)")
    # the last warp to finish on each launch (per-SM clocks are not comparable across SMs: use per-warp totals instead)
    print("--- slowest warp of each launch (total cycles, resets in tile, per-phase)")
    for r in rows[:8]:
        r = r[r[:, 10] > 0]
        tot = r[:, 11] - r[:, 0]
        w = int(np.argmax(tot))
        print(f"  total {tot[w]:7d}  resets {r[w, 12]}  sm {r[w, 14]:3d}  " + " ".join(f"{r[w, i1] - r[w, i0]:6d}" for _, i0, i1 in PHASES[:-1]))
    # warps sharing an SM: spread of their start and end (same clock)
    r = rows[-1]
    r = r[r[:, 10] > 0]
    spans = []
    for sm in np.unique(r[:, 14]):
        q = r[r[:, 14] == sm]
        spans.append((q[:, 11].max() - q[:, 0].min(), len(q)))
    spans = np.array(spans)
    print(f"--- per SM (last launch): first entry -> last exit: median {np.median(spans[:, 0]):.0f}  max {spans[:, 0].max():.0f} cycles; "
          f"warps per SM median {np.median(spans[:, 1]):.0f} max {spans[:, 1].max()}")


if __name__ == "__main__":
    main()
