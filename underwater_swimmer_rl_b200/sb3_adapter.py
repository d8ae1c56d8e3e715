"""Stable-Baselines3 ``VecEnv`` protocol over SalpVectorEnv (callers).

The reference trains with ``stable_baselines3.SAC(env=SalpSnakeEnv(...))`` (train.py:60-70), which
wraps the env in a DummyVecEnv: ``reset() -> obs``, ``step_async/step_wait -> (obs, rewards, dones,
infos: list[dict])`` with ``terminal_observation`` and ``TimeLimit.truncated`` in the info of lanes
that finished.  This adapter speaks that protocol with numpy arrays so SB3 runs unchanged on top of
the CUDA env.  It materialises Python dicts per env and is meant for small ``num_envs``; at scale
use the tensor API of SalpVectorEnv.  stable_baselines3 is not installed on this image; when it is
importable the class registers as a virtual subclass of its VecEnv.
"""
from __future__ import annotations

from typing import Any, List, Optional, Sequence

import numpy as np

from .vector_env import SalpVectorEnv


class SalpSB3VecEnv:
    def __init__(self, venv: SalpVectorEnv):
        if not venv.autoreset:
            raise ValueError("the SB3 VecEnv protocol needs autoreset=True")
        self.venv = venv
        self.num_envs = venv.num_envs
        self.observation_space = venv.single_observation_space
        self.action_space = venv.single_action_space
        self.render_mode = None
        self._actions = None
        self.reset_infos: List[dict] = [{} for _ in range(self.num_envs)]

    def reset(self) -> np.ndarray:
        obs, _ = self.venv.reset()
        return obs.cpu().numpy().copy()

    def step_async(self, actions: np.ndarray) -> None:
        self._actions = np.asarray(actions, dtype=np.float32).reshape(self.num_envs, -1)

    def step_wait(self):
        obs, rew, term, trunc, infos = self.venv.step(self._actions)
        o = obs.cpu().numpy().copy()
        r = rew.cpu().numpy().astype(np.float32)
        te, tr = term.cpu().numpy(), trunc.cpu().numpy()
        dones = te | tr
        ev = infos["events"].cpu().numpy()
        fc = self.venv.food_collected.cpu().numpy()
        ssf = self.venv.steps_since_food.cpu().numpy()
        out: List[dict] = []
        fo = fi = None
        if dones.any():
            fo = infos["final_observation"].cpu().numpy()
            fi = infos["final_info"].cpu().numpy()
        for i in range(self.num_envs):
            d = {"collision": bool(ev[i] & 4), "food_collected": int(fc[i]), "steps_since_food": int(ssf[i]),
                 "score": float(fc[i] * self.venv.params.food_reward)}
            if dones[i]:
                d["terminal_observation"] = fo[i].copy()
                d["TimeLimit.truncated"] = bool(tr[i] and not te[i])
                d["episode"] = {"r": float(fi[i, 0]), "l": int(fi[i, 1])}
                d["food_collected"], d["steps_since_food"] = int(fi[i, 2]), int(fi[i, 3])
                d["score"] = float(fi[i, 2] * self.venv.params.food_reward)
            out.append(d)
        return o, r, dones, out

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self) -> None:
        self.venv.close()

    def seed(self, seed: Optional[int] = None) -> Sequence[Optional[int]]:
        if seed is not None:
            self.venv.reset(seed=seed)
        return [None if seed is None else seed + i for i in range(self.num_envs)]

    def _idx(self, indices):
        if indices is None:
            return list(range(self.num_envs))
        return [indices] if isinstance(indices, int) else list(indices)

    def get_attr(self, attr_name: str, indices=None) -> List[Any]:
        v = self.venv.get_attr(attr_name)
        if hasattr(v, "cpu"):
            v = v.cpu().numpy()
            return [v[i] for i in self._idx(indices)]
        return [v for _ in self._idx(indices)]

    def set_attr(self, attr_name: str, value: Any, indices=None) -> None:
        cur = self.venv.get_attr(attr_name)
        for i in self._idx(indices):
            cur[i] = value

    def env_method(self, method_name: str, *args, indices=None, **kwargs) -> List[Any]:
        return [getattr(self.venv, method_name)(*args, **kwargs) for _ in self._idx(indices)]

    def env_is_wrapped(self, wrapper_class, indices=None) -> List[bool]:
        return [False for _ in self._idx(indices)]

    def get_images(self):
        return [None] * self.num_envs

    def render(self, mode: Optional[str] = None):
        return None


try:  # pragma: no cover - stable_baselines3 is absent on this image
    from stable_baselines3.common.vec_env import VecEnv as _SB3VecEnv
    _SB3VecEnv.register(SalpSB3VecEnv)
except Exception:
    pass


class SalpCycleSB3VecEnv:
    """The same protocol over the model (B) env -- what ``make_vec_env(make_env, n_envs=4, vec_env_cls=SubprocVecEnv)``
    gives SB3 in the reference's src/salp/environments/train_robot.py:10-25.  Episode returns / lengths (cycles) come
    from the kernel's per-env accumulators; the info dict of salp_robot_env.py:194-198 (per-substep histories for
    rendering) is not produced."""

    def __init__(self, venv):
        if not venv.autoreset:
            raise ValueError("the SB3 VecEnv protocol needs autoreset=True")
        self.venv = venv
        self.num_envs = venv.num_envs
        self.observation_space = venv.single_observation_space
        self.action_space = venv.single_action_space
        self.render_mode = None
        self._actions = None
        self._ret = np.zeros(self.num_envs, np.float64)
        self._len = np.zeros(self.num_envs, np.int64)
        self.reset_infos: List[dict] = [{} for _ in range(self.num_envs)]

    def reset(self) -> np.ndarray:
        obs, _ = self.venv.reset()
        self._ret[:] = 0.0
        self._len[:] = 0
        return obs.cpu().numpy().copy()

    def step_async(self, actions: np.ndarray) -> None:
        self._actions = np.asarray(actions, dtype=np.float32).reshape(self.num_envs, 3)

    def step_wait(self):
        obs, rew, term, trunc, infos = self.venv.step(self._actions)
        o = obs.cpu().numpy().copy()
        r = rew.cpu().numpy().astype(np.float32)
        te, tr = term.cpu().numpy().astype(bool), trunc.cpu().numpy().astype(bool)
        dones = te | tr
        n_sub = infos["n_sub"].cpu().numpy()
        self._ret += r
        self._len += 1
        fo = infos["final_obs"].cpu().numpy() if dones.any() else None
        out: List[dict] = []
        for i in range(self.num_envs):
            d = {"n_sub": int(n_sub[i])}
            if dones[i]:
                d["terminal_observation"] = fo[i].copy()
                d["TimeLimit.truncated"] = bool(tr[i] and not te[i])
                d["episode"] = {"r": float(self._ret[i]), "l": int(self._len[i])}
                self._ret[i], self._len[i] = 0.0, 0
            out.append(d)
        return o, r, dones, out

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self) -> None:
        self.venv.close()

    def seed(self, seed: Optional[int] = None) -> Sequence[Optional[int]]:
        if seed is not None:
            self.venv.reset(seed=seed)
        return [None if seed is None else seed + i for i in range(self.num_envs)]

    def get_attr(self, attr_name: str, indices=None) -> List[Any]:
        v = getattr(self.venv, attr_name)
        idx = list(range(self.num_envs)) if indices is None else ([indices] if isinstance(indices, int) else list(indices))
        if hasattr(v, "cpu"):
            v = v.cpu().numpy()
            return [v[i] for i in idx]
        return [v for _ in idx]

    def env_is_wrapped(self, wrapper_class, indices=None) -> List[bool]:
        return [False] * self.num_envs

    def render(self, mode: Optional[str] = None):
        return None


try:  # pragma: no cover
    _SB3VecEnv.register(SalpCycleSB3VecEnv)
except Exception:
    pass
