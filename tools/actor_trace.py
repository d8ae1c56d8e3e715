"""Per-phase clock trace of the fused actor kernel (developer tool).

Build with `SALP_ACTOR_DEFS=-DSALP_ACTOR_TRACE python -m underwater_swimmer_rl_b200.build --force`, run this on the GPU, then rebuild
without the variable.  Prints, for CTA 0 and warp 0 and the warp that issues the MMAs, SM clocks between the trace points of each
loop iteration."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from underwater_swimmer_rl_b200 import _capi
if os.environ.get("SALP_LAB_LIB"):
    _capi.LIB_PATH = os.path.abspath(os.environ["SALP_LAB_LIB"])
from underwater_swimmer_rl_b200.rollout import FusedActor, TanhGaussianActor

N = 1 << 18
fused = FusedActor(TanhGaussianActor(24, 1).cuda())
obs = torch.rand(N, 24, device="cuda"); act = torch.zeros(N, 1, device="cuda"); noise = torch.randn(N, 1, device="cuda")
for _ in range(5):
    fused(obs, act, noise=noise)
torch.cuda.synchronize()
IT, PT = 12, 8
buf = (C.c_longlong * (2 * IT * PT + 8))()
L = _capi.load()
assert L.salp_actor_trace_read(buf) == 0
names = ["top", "waits done", "own epi2", "epi1 done", "meet5+issue", "epi2 sums", "meet6+store"]
for w, wname in enumerate(["warp 0 (column group 0: stores)", "issuing warp"]):
    print(wname, " ".join(f"{n:>11s}" for n in names), "  iter_total")
    base = None
    for it in range(IT):
        row = [buf[(w * IT + it) * PT + k] for k in range(len(names))]
        if base is None:
            base = row[0]
        d = [row[0] - base] + [row[k] - row[k - 1] for k in range(1, len(names))]
        nxt = buf[(w * IT + it + 1) * PT] if it + 1 < IT else row[-1]
        print(f"  it {it:2d}", " ".join(f"{x:11d}" for x in d), f"  {nxt - row[0]:8d}")

t = [buf[2 * IT * PT + k] for k in range(6)]
print(f"entry -> loop start: {t[2] - t[0]} clocks, {t[3] - t[1]} ns; loop start -> end: {t[4] - t[2]} clocks, {t[5] - t[3]} ns "
      f"=> {(t[4] - t[2]) / max(1, t[5] - t[3]):.3f} clocks/ns")
