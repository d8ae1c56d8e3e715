"""Time bench.py workloads for every library build in tools/_lab/ (developer tool): each build in its own process, same box.
usage: python tools/time_step_variants.py workload [workload ...]"""
import glob, json, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
libs = [l for l in sorted(glob.glob(os.path.join(root, "tools", "_lab", "libsalp_*.so"))) if "libsalp_t_" not in l]
code = r'''
import os, sys
sys.path.insert(0, %r)
from underwater_swimmer_rl_b200 import _capi
_capi.LIB_PATH = os.environ["SALP_LAB_LIB"]
import bench
sys.argv = ["bench.py", "--workload", os.environ["SALP_LAB_WORKLOAD"], "--steps", "600", "--warmup", "20", "--no-cpu-baseline", "--e2e-steps", "1"]
bench.main()
''' % root
for wl in sys.argv[1:] or ["single_food_1m"]:
    for lib in libs:
        r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, SALP_LAB_LIB=lib, SALP_LAB_WORKLOAD=wl),
                           capture_output=True, text=True, timeout=200)
        try:
            d = json.loads(r.stdout.strip().splitlines()[-1])
            print(f"{wl:20s} {os.path.basename(lib):30s} {1e3 * d['ms_per_step']:8.2f} us/step  frac {d['roofline']['frac']:.3f}", flush=True)
        except Exception:
            print(wl, os.path.basename(lib), "FAILED", r.stderr.strip()[-400:], flush=True)
