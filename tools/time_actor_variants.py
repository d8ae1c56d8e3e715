"""Time the fused actor for every library build in tools/_lab/ (developer tool), each in its own process, interleaved
and repeated so that box-to-box and clock drift show up as spread rather than as a difference between variants."""
import glob, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
libs = [l for l in sorted(glob.glob(os.path.join(root, "tools", "_lab", "libsalp_*.so"))) if "libsalp_t_" not in l]  # t_ = trace builds
code = r'''
import os, sys, torch
sys.path.insert(0, %r)
from underwater_swimmer_rl_b200 import _capi
_capi.LIB_PATH = os.environ["SALP_LAB_LIB"]
from underwater_swimmer_rl_b200.rollout import FusedActor, TanhGaussianActor
N = 1 << 18
fused = FusedActor(TanhGaussianActor(24, 1).cuda())
obs = torch.rand(N, 24, device="cuda") * 2 - 1; act = torch.zeros(N, 1, device="cuda"); noise = torch.randn(N, 1, device="cuda")
for _ in range(20): fused(obs, act, noise=noise)
torch.cuda.synchronize()
# 20 launches per CUDA graph: the Python/ctypes launch path (tens of microseconds) must not be what is timed
g = torch.cuda.CUDAGraph()
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    with torch.cuda.graph(g, stream=side):
        for _ in range(20): fused(obs, act, noise=noise)
torch.cuda.synchronize()
best = 1e9
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): g.replay()
    e1.record(); torch.cuda.synchronize()
    best = min(best, 1e3 * e0.elapsed_time(e1) / 200)
print(f"{best:.1f}")
''' % root
for rnd in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    for lib in libs:
        r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, SALP_LAB_LIB=lib), capture_output=True, text=True, timeout=60)
        print(f"round {rnd} {os.path.basename(lib):40s} {r.stdout.strip() or r.stderr.strip()[-300:]} us", flush=True)
