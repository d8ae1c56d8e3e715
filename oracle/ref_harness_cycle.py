"""Run the UNMODIFIED reference model (B): SalpRobotEnv over Robot/Nozzle.

TEST INFRASTRUCTURE ONLY; works only where ``/root/reference`` exists (the build container).  Used to pin
``oracle/salp_cycle_oracle.c`` and to generate ``tests/golden/cycle_*.npz``.

Composition: ``src/salp/environments/salp_robot_env.py`` does ``from robot import Robot, Nozzle`` and imports
gymnasium, pygame, matplotlib.pyplot and PIL.Image at module level; none of the four carries arithmetic of
the step path (render / GIF only), so empty stub modules stand in for the absent ones.  ``robot.py`` needs
numpy only.  The env draws its target point from the process-global ``np.random`` (salp_robot_env.py:279-282):
``np.random.seed(s)`` is called AFTER construction (the constructor's own reset() draws too) and immediately
BEFORE ``reset()``; the oracle / GPU consume the same MT19937 stream from a word table
(oracle.ref_harness.jitter_words).
"""
from __future__ import annotations

import sys
import types

import numpy as np

from . import ref_harness as rh

_CACHE: dict = {}


def load_reference_cycle():
    if "mods" in _CACHE:
        return _CACHE["mods"]
    if not rh.reference_available():
        raise RuntimeError(f"reference tree not found at {rh.REFERENCE_ROOT}")
    rh._install_stubs()
    for name in ("matplotlib", "matplotlib.pyplot", "PIL", "PIL.Image"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["PIL"].Image = sys.modules["PIL.Image"]
    robot = rh._load("robot", "src/salp/environments/robot.py")
    env = rh._load("salp_robot_env_model_b", "src/salp/environments/salp_robot_env.py")
    _CACHE["mods"] = (robot, env)
    return robot, env


def make_reference_cycle_env(**params):
    """Build Robot + SalpRobotEnv the way salp_robot_env.py:1305-1310 / train_robot.py:12-17 do."""
    from .salp_cycle_oracle import PARAM_DEFAULTS
    d = dict(PARAM_DEFAULTS)
    d.update(params)
    robot, envm = load_reference_cycle()
    nz = robot.Nozzle(length1=d["nozzle_length1"], length2=d["nozzle_length2"], length3=d["nozzle_length3"],
                      area=d["nozzle_area"], mass=d["nozzle_mass"])
    rb = robot.Robot(dry_mass=d["dry_mass"], init_length=d["init_length"], init_width=d["init_width"],
                     max_contraction=d["max_contraction"], nozzle=nz)
    rb.nozzle.set_angles(angle1=0.0, angle2=0.0)
    rb.set_environment(density=d["density"])
    return envm.SalpRobotEnv(render_mode=None, width=d["width"], height=d["height"], robot=rb)


CONT_FIELDS = ("position", "velocity", "angular_velocity", "euler_angle", "velocity_world")


def rollout_reference_cycle(params: dict, seeds, actions: np.ndarray, autoreset: bool = True) -> dict:
    """actions [T, N, 3] float64.  Env e is rolled out on its own np.random stream (seed seeds[e])."""
    T, N, _ = actions.shape
    out = dict(
        obs0=np.zeros((N, 6), np.float32), obs=np.zeros((T, N, 6), np.float32), reward=np.zeros((T, N)),
        terminated=np.zeros((T, N), np.uint8), truncated=np.zeros((T, N), np.uint8),
        n_sub=np.zeros((T, N), np.int32), final_obs=np.zeros((T, N, 6), np.float32),
        cycle=np.zeros((T, N), np.int32), phase=np.zeros((T, N), np.int32),
        cont=np.zeros((T, N, 5, 3)), target=np.zeros((T, N, 2), np.float32),
    )
    max_cycles = int(params.get("max_cycles", 500))
    assert max_cycles == 500, "the reference hard-codes 500 cycles per episode (salp_robot_env.py:187)"
    for e in range(N):
        env = make_reference_cycle_env(**{k: v for k, v in params.items() if k not in ("max_cycles", "max_substeps")})
        np.random.seed(int(seeds[e]))
        o, _ = env.reset()
        out["obs0"][e] = o
        for t in range(T):
            o, r, term, trunc, info = env.step(actions[t, e].copy())
            rb = env.robot
            out["reward"][t, e] = r
            out["terminated"][t, e] = term
            out["truncated"][t, e] = trunc
            out["n_sub"][t, e] = len(rb.position_history) - 1
            out["cycle"][t, e] = rb.cycle
            out["phase"][t, e] = rb.state.value
            out["target"][t, e] = env.target_point
            for j, f in enumerate(CONT_FIELDS):
                out["cont"][t, e, j] = getattr(rb, f)
            if term or trunc:
                out["final_obs"][t, e] = o
                if autoreset:
                    o, _ = env.reset()
            out["obs"][t, e] = o
    return out
