"""Build libsalp_b200.so (the C-ABI shared library) in-tree with nvcc for sm_100a.

    python -m underwater_swimmer_rl_b200.build [--force] [--verbose]

Nine translation units, compiled in parallel:
  salp_kernels.cu  -DSALP_BUILD_F64=0               float   kernels (throughput build)
  salp_kernels.cu  -DSALP_BUILD_F64=1 -fmad=false   double  kernels (validation build: no FMA
                                                    contraction, reference operation order)
  salp_kernels_fast.cu                              forced-breathing f32 kernel: persistent warps, TMA in/out, cycle table
  salp_capi.cu                                      extern "C" boundary (include/salp_b200.h)
  salp_replay.cu                                    rollout -> replay ring-buffer append kernel
  salp_actor.cu                                     fused actor MLP (tcgen05 / TMEM) of the rollout loop
  salp_cycle_kernels.cu  (twice, like salp_kernels.cu)  model (B) breathing-cycle kernels, float / double
  salp_cycle_capi.cu                                extern "C" boundary of the cycle env (salp_cycle_*)
nvcc cross-compiles without a GPU; the built .so is git-ignored but travels with gpurun.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.environ.get("SALP_CSRC") or os.path.join(PKG, "csrc")  # SALP_CSRC: developer A/B builds from another source tree
# developer A/B builds: SALP_LIB_TAG=x builds libsalp_b200_x.so from objects in _build_x/ (and _capi loads that one)
TAG = os.environ.get("SALP_LIB_TAG", "")
OBJ = os.path.join(PKG, "_build" + ("_" + TAG if TAG else ""))
LIB = os.path.join(PKG, "libsalp_b200" + ("_" + TAG if TAG else "") + ".so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC"] + ARCH


def _sources():
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hdrs.append(os.path.join(os.path.dirname(PKG), "include", "salp_b200.h"))
    return hdrs


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _sources()
    kern = os.path.join(CSRC, "salp_kernels.cu")
    fast = os.path.join(CSRC, "salp_kernels_fast.cu")
    capi = os.path.join(CSRC, "salp_capi.cu")
    replay = os.path.join(CSRC, "salp_replay.cu")
    actor = os.path.join(CSRC, "salp_actor.cu")
    cyck = os.path.join(CSRC, "salp_cycle_kernels.cu")
    cycc = os.path.join(CSRC, "salp_cycle_capi.cu")
    vflag = ["-Xptxas", "-v"] if verbose else []
    step_defs = os.environ.get("SALP_STEP_DEFS", "").split()  # developer builds of the step kernels, e.g. -DSALP_L2_HINTS=2
    actor_defs = os.environ.get("SALP_ACTOR_DEFS", "").split()  # developer builds, e.g. SALP_ACTOR_DEFS=-DSALP_ACTOR_TRACE
    jobs = [
        (os.path.join(OBJ, "salp_kernels_f32.o"), [NVCC, "-c", kern, "-DSALP_BUILD_F64=0"] + COMMON + vflag + step_defs, [kern] + hdrs),
        (os.path.join(OBJ, "salp_kernels_fast.o"), [NVCC, "-c", fast] + COMMON + vflag + step_defs, [fast] + hdrs),
        (os.path.join(OBJ, "salp_kernels_f64.o"), [NVCC, "-c", kern, "-DSALP_BUILD_F64=1", "-fmad=false"] + COMMON + vflag + os.environ.get("SALP_F64_DEFS", "").split(), [kern] + hdrs),
        (os.path.join(OBJ, "salp_capi.o"), [NVCC, "-c", capi] + COMMON, [capi] + hdrs),
        (os.path.join(OBJ, "salp_replay.o"), [NVCC, "-c", replay] + COMMON, [replay] + hdrs),
        (os.path.join(OBJ, "salp_actor.o"), [NVCC, "-c", actor] + COMMON + vflag + actor_defs, [actor] + hdrs),
        (os.path.join(OBJ, "salp_cycle_f32.o"), [NVCC, "-c", cyck, "-DSALP_BUILD_F64=0"] + COMMON + vflag, [cyck] + hdrs),
        (os.path.join(OBJ, "salp_cycle_f64.o"), [NVCC, "-c", cyck, "-DSALP_BUILD_F64=1", "-fmad=false"] + COMMON + vflag, [cyck] + hdrs),
        (os.path.join(OBJ, "salp_cycle_capi.o"), [NVCC, "-c", cycc] + COMMON, [cycc] + hdrs),
    ]
    todo = [(o, c + ["-o", o]) for o, c, d in jobs if force or _stale(o, d)]

    def run(job):
        o, cmd = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {o}:\n{r.stdout}\n{r.stderr}")
        return o, r.stderr

    if todo:
        with ThreadPoolExecutor(max_workers=8) as ex:
            for o, log in ex.map(run, todo):
                if verbose and log:
                    print(f"---- {os.path.basename(o)}\n{log}")
    objs = [o for o, _, _ in jobs]
    if todo or force or _stale(LIB, objs):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ARCH + ["-Xcompiler", "-fPIC", "-lcudart_static", "-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
