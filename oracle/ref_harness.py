"""Run the UNMODIFIED reference SalpSnakeEnv (over the legacy SalpRobotEnv integrator).

TEST INFRASTRUCTURE ONLY.  This module works only where ``/root/reference`` exists
(the build container).  It is used to (1) pin the C restatement in
``oracle/salp_oracle.c`` and (2) generate the golden fixtures under ``tests/golden/``.
Nothing in the product package imports it and nothing in the ``-m gpu`` tests,
``smoke()`` or ``bench.py`` needs it at run time.

How the reference is composed:

* ``gymnasium`` and ``pygame`` are not installed.  Neither carries hot-path arithmetic
  (``gym.Env.__init__/reset(seed)`` + ``spaces.Box`` metadata; pygame only in render()).
  Two in-memory stub modules stand in for them.
* ``src/salp/environments/salp_snake_env.py`` imports ``SalpRobotEnv`` from
  ``salp.environments.salp_robot_env`` (salp_snake_env.py:14), which at HEAD is the new
  model (B).  The snake env is written against the legacy model (A) in
  ``scripts/utilities/salp_robot.py``; that file is loaded under the module name
  ``salp.environments.salp_robot_env``.
* RNG discipline: the env draws food from the process-global ``random``
  and thrust jitter from the process-global ``np.random``.  ``seed_reference(s)`` seeds
  both AFTER construction and immediately BEFORE ``reset()``; the oracle/GPU consume the
  same two MT19937 word streams from tables.
"""
from __future__ import annotations

import importlib.util
import os
import random
import sys
import types

import numpy as np

REFERENCE_ROOT = os.environ.get("SALP_REFERENCE_ROOT", "/root/reference")
_ROBOT_FILE = "scripts/utilities/salp_robot.py"
_SNAKE_FILE = "src/salp/environments/salp_snake_env.py"

PHASE_CODE = {"rest": 0, "inhaling": 1, "exhaling": 2}


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, _ROBOT_FILE)) and os.path.isfile(
        os.path.join(REFERENCE_ROOT, _SNAKE_FILE)
    )


def _install_stubs() -> None:
    if "gymnasium" not in sys.modules:
        gym = types.ModuleType("gymnasium")

        class Env:  # the slice of gym.Env the reference touches
            metadata: dict = {}

            def reset(self, seed=None, options=None):
                if seed is not None:  # real gymnasium seeds self.np_random; nothing reads it
                    self._np_random = np.random.default_rng(seed)
                return None

            def close(self):
                pass

        spaces = types.ModuleType("gymnasium.spaces")

        class Box:
            def __init__(self, low, high, shape=None, dtype=np.float32):
                low = np.asarray(low, dtype=dtype)
                high = np.asarray(high, dtype=dtype)
                if shape is None:
                    shape = low.shape
                self.low = np.broadcast_to(low, shape).copy()
                self.high = np.broadcast_to(high, shape).copy()
                self.shape = tuple(shape)
                self.dtype = np.dtype(dtype)

        spaces.Box = Box
        gym.Env = Env
        gym.spaces = spaces
        sys.modules["gymnasium"] = gym
        sys.modules["gymnasium.spaces"] = spaces
    if "pygame" not in sys.modules:
        sys.modules["pygame"] = types.ModuleType("pygame")


def _load(name: str, relpath: str):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REFERENCE_ROOT, relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


_CACHE: dict = {}


def load_reference():
    """Return the reference ``SalpSnakeEnv`` class (composed over legacy model A)."""
    if "cls" in _CACHE:
        return _CACHE["cls"]
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    _install_stubs()
    for pkg in ("salp", "salp.environments"):
        if pkg not in sys.modules:
            m = types.ModuleType(pkg)
            m.__path__ = []  # bare package: skips src/salp/environments/__init__.py (model B)
            sys.modules[pkg] = m
    _load("salp.environments.salp_robot_env", _ROBOT_FILE)
    snake = _load("salp.environments.salp_snake_env", _SNAKE_FILE)
    _CACHE["cls"] = snake.SalpSnakeEnv
    return snake.SalpSnakeEnv


def make_reference_env(**params):
    return load_reference()(render_mode=None, **params)


def seed_reference(seed: int) -> None:
    """Seed BOTH global streams the reference consumes (food: random, jitter: np.random)."""
    random.seed(int(seed))
    np.random.seed(int(seed))


# ---------------------------------------------------------------------------------------
# MT19937 word streams, as the GPU/oracle tables hold them
# ---------------------------------------------------------------------------------------

def food_words(seed: int, n_words: int) -> np.ndarray:
    """First n_words 32-bit outputs of random.Random(seed) (== global random after random.seed)."""
    r = random.Random(int(seed))
    return np.array([r.getrandbits(32) for _ in range(n_words)], dtype=np.uint32)


def jitter_words(seed: int, n_words: int) -> np.ndarray:
    """First n_words 32-bit outputs of np.random.RandomState(seed) (== np.random after seed)."""
    return _raw_words_numpy(np.random.RandomState(int(seed)), n_words)


def _raw_words_numpy(rs: np.random.RandomState, n_words: int) -> np.ndarray:
    # Re-run the generator from its key: state tuple is ('MT19937', key[624], pos, ...).
    # tomaxint / randint paths mask or reject; replaying the key with random.Random's
    # core (same MT19937 recurrence) gives the raw outputs without guessing.
    st = rs.get_state()
    key, pos = st[1], int(st[2])
    r = random.Random()
    r.setstate((3, tuple(int(k) for k in key) + (pos,), None))
    return np.array([r.getrandbits(32) for _ in range(n_words)], dtype=np.uint32)


def words_consumed_food(seed: int, max_words: int = 1 << 20) -> int:
    """How many words of random.Random(seed) the global ``random`` has consumed so far."""
    target = random.getstate()
    r = random.Random(int(seed))
    n = 0
    while r.getstate() != target:
        r.getrandbits(32)
        n += 1
        if n > max_words:
            raise RuntimeError("global random is not on the seeded stream")
    return n


def words_consumed_jitter(seed: int, max_words: int = 1 << 20) -> int:
    tgt = np.random.get_state()
    rs = np.random.RandomState(int(seed))
    n = 0
    while True:
        st = rs.get_state()
        if int(st[2]) == int(tgt[2]) and np.array_equal(st[1], tgt[1]):
            return n
        rs.random_sample()
        n += 2
        if n > max_words:
            raise RuntimeError("global np.random is not on the seeded stream")


# ---------------------------------------------------------------------------------------
# Trajectory recording
# ---------------------------------------------------------------------------------------

STATE_FIELDS = (
    "px", "py", "vx", "vy", "angle", "omega", "nozzle", "wv", "ellipse_a", "ellipse_b",
)


def _snap(env, fmax: int) -> dict:
    food = np.zeros((fmax, 2), dtype=np.float64)
    alive = np.zeros((fmax,), dtype=np.uint8)
    for i, fp in enumerate(env.food_positions):
        if fp is not None:
            food[i] = fp
            alive[i] = 1
    return dict(
        px=float(env.robot_pos[0]), py=float(env.robot_pos[1]),
        vx=float(env.robot_velocity[0]), vy=float(env.robot_velocity[1]),
        angle=float(env.robot_angle), omega=float(env.robot_angular_velocity),
        nozzle=float(env.nozzle_angle), wv=float(env.water_volume),
        ellipse_a=float(env.ellipse_a), ellipse_b=float(env.ellipse_b),
        phase=PHASE_CODE[env.breathing_phase], timer=int(env.breathing_timer),
        steps_since_food=int(env.steps_since_food), food_collected=int(env.food_collected),
        score=float(env.score), n_food=len(env.food_positions), food=food, alive=alive,
    )


def rollout_reference(params: dict, seeds, actions: np.ndarray, autoreset: bool = True,
                      fmax: int | None = None) -> dict:
    """Roll the reference for T steps on each seed.

    actions: float array [T, N, act_dim] (same values fed to oracle and kernel).
    Returns arrays shaped [T, N, ...] with the post-step state (BEFORE any auto-reset),
    the step outputs, and for done/truncated steps the reset observation that the
    caller's ``env.reset()`` produced.  ``reset_obs0`` is the initial reset observation.
    Envs are run one after the other because the reference's RNGs are process-global.
    """
    T, N = actions.shape[0], actions.shape[1]
    seeds = list(seeds)
    assert len(seeds) == N
    base_f = int(params.get("num_food_items", 5))
    fmax = fmax or max(1, base_f)
    out = dict(
        obs=np.zeros((T, N, 0), np.float32), reward=np.zeros((T, N), np.float64),
        terminated=np.zeros((T, N), np.uint8), truncated=np.zeros((T, N), np.uint8),
        collision=np.zeros((T, N), np.uint8),
        phase=np.zeros((T, N), np.int32), timer=np.zeros((T, N), np.int32),
        steps_since_food=np.zeros((T, N), np.int32), food_collected=np.zeros((T, N), np.int32),
        score=np.zeros((T, N), np.float64), n_food=np.zeros((T, N), np.int32),
        food=np.zeros((T, N, fmax, 2), np.float64), alive=np.zeros((T, N, fmax), np.uint8),
        cur_food=np.zeros((N,), np.int64), cur_jit=np.zeros((N,), np.int64),
    )
    for k in STATE_FIELDS:
        out[k] = np.zeros((T, N), np.float64)
    obs_dim = None
    for e, s in enumerate(seeds):
        env = make_reference_env(**params)
        seed_reference(s)
        o0, _ = env.reset()
        if obs_dim is None:
            obs_dim = o0.shape[0]
            out["obs"] = np.zeros((T, N, obs_dim), np.float32)
            out["reset_obs"] = np.zeros((T, N, obs_dim), np.float32)
            out["reset_obs0"] = np.zeros((N, obs_dim), np.float32)
            out["food0"] = np.zeros((N, fmax, 2), np.float64)
            out["alive0"] = np.zeros((N, fmax), np.uint8)
        out["reset_obs0"][e] = o0
        s0 = _snap(env, fmax)
        out["food0"][e], out["alive0"][e] = s0["food"], s0["alive"]
        for t in range(T):
            o, r, d, tr, info = env.step(actions[t, e])
            out["obs"][t, e] = o
            out["reward"][t, e] = r
            out["terminated"][t, e] = d
            out["truncated"][t, e] = tr
            out["collision"][t, e] = info["collision"]
            sn = _snap(env, fmax)
            for k in STATE_FIELDS + ("phase", "timer", "steps_since_food", "food_collected",
                                     "score", "n_food", "food", "alive"):
                out[k][t, e] = sn[k]
            if autoreset and (d or tr):
                ro, _ = env.reset()
                out["reset_obs"][t, e] = ro
        out["cur_food"][e] = words_consumed_food(s)
        out["cur_jit"][e] = words_consumed_jitter(s)
    return out
