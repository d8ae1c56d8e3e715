"""SalpSnakeEnv -- single-environment facade with the reference's exact call shape.

    env = SalpSnakeEnv(render_mode=None, **config.environment['params'])   # train.py:50-52
    obs, info = env.reset()
    obs, reward, done, truncated, info = env.step(action)                  # salp_snake_env.py:157

Backed by a 1-env ``SalpVectorEnv`` (validation build, fp64, no auto-reset -- the reference never
auto-resets), so every step is the same sm_100a kernel.  Returns numpy / Python scalars and the
reference's 14-key info dict (salp_robot.py:390-403 + salp_snake_env.py:195-200).  This is a
compatibility shim for single-env callers (train.py, eval scripts); anything that cares about
throughput uses SalpVectorEnv directly.
"""
from __future__ import annotations

import random
from typing import Optional

import numpy as np
import torch

from ._gym_compat import Env
from .vector_env import PHASE_NAMES, SalpVectorEnv


BACKEND = SalpVectorEnv


class SalpSnakeEnv(Env):
    metadata = {"render_modes": [], "render_fps": 60}

    def __init__(self, render_mode: Optional[str] = None, width: int = 800, height: int = 600, *,
                 device="cuda:0", dtype="float64", seed: Optional[int] = None, **kwargs):
        if seed is None:  # the reference draws from unseeded global generators; pick a stream at random
            seed = random.SystemRandom().randrange(1, 2 ** 31)
        # BACKEND is the batched env class; the CPU test of the reference's train.py call sequence swaps in an oracle-backed
        # stand-in (tests/fake_backend.py) because the build container has no GPU.  The product has exactly one backend.
        self._v = BACKEND(1, device=device, dtype=dtype, seed=seed, autoreset=False, food_table_len=4096,
                          jitter_table_len=16384, render_mode=render_mode, width=width, height=height, **kwargs)
        self.render_mode = None
        self.width, self.height = width, height
        self.tank_margin, self.base_radius, self.food_radius, self.min_food_distance = 50, 30, 15, 80
        self.action_space = self._v.single_action_space
        self.observation_space = self._v.single_observation_space
        for k in ("food_reward", "collision_penalty", "time_penalty", "efficiency_bonus", "forced_breathing",
                  "max_observed_food", "random_food_count", "respawn_food", "proximity_reward_weight",
                  "max_steps_without_food"):
            setattr(self, k, getattr(self._v, k))
        self.reset()  # the reference constructor resets (salp_robot.py:93)

    # ------------------------------------------------------------------ gym API
    def reset(self, seed: Optional[int] = None, options: Optional[dict] = None):
        obs, _ = self._v.reset(seed=seed)
        return obs[0].cpu().numpy().copy(), {}

    def step(self, action):
        a = np.asarray(action, dtype=np.float64).reshape(1, self._v.act_dim)
        obs, rew, term, trunc, infos = self._v.step(a)
        o = obs[0].cpu().numpy().copy()
        ev = int(infos["events"][0].item())
        info = self._get_info()
        info.update(score=self.score, food_collected=self.food_collected, steps_since_food=self.steps_since_food,
                    collision=bool(ev & 4))
        return o, float(rew[0].item()), bool(term[0].item()), bool(trunc[0].item()), info

    def _get_extended_observation(self):
        """salp_snake_env.py:366-428: the observation of the current (possibly poked) state, no step
        (eval/collect_navigation_data.py:89, eval/watch_navigation.py:63) -- the salp_observe kernel."""
        return self._v.observe()[0].cpu().numpy().copy()

    @property
    def base_num_food_items(self): return self._v.base_num_food_items
    @base_num_food_items.setter
    def base_num_food_items(self, n): self._v.base_num_food_items = n  # continuous_trainer.py:408-411

    def _get_info(self) -> dict:
        a, b = self._v.ellipse_axes()
        return {
            "robot_position": tuple(self.robot_pos.tolist()), "robot_velocity": tuple(self.robot_velocity.tolist()),
            "robot_angle": self.robot_angle, "nozzle_angle": self.nozzle_angle,
            "ellipse_a": float(a[0].item()), "ellipse_b": float(b[0].item()), "body_radius": self.base_radius,
            "breathing_phase": self.breathing_phase, "water_volume": self.water_volume,
            "breathing_timer": self.breathing_timer,
        }

    def close(self):
        self._v.close()

    def render(self):
        return None

    # ------------------------------------------------------------------ reference attributes
    def _f(self, t):
        return float(t[0].item())

    @property
    def robot_pos(self): return self._v.robot_pos[0].cpu().numpy().astype(float)
    @robot_pos.setter
    def robot_pos(self, v): self._v.robot_pos[0] = torch.as_tensor(np.asarray(v, dtype=np.float64))
    @property
    def robot_velocity(self): return self._v.robot_velocity[0].cpu().numpy().astype(float)
    @robot_velocity.setter
    def robot_velocity(self, v): self._v.robot_velocity[0] = torch.as_tensor(np.asarray(v, dtype=np.float64))
    @property
    def robot_angle(self): return self._f(self._v.robot_angle)
    @robot_angle.setter
    def robot_angle(self, v): self._v.robot_angle[0] = float(v)
    @property
    def robot_angular_velocity(self): return self._f(self._v.robot_angular_velocity)
    @robot_angular_velocity.setter
    def robot_angular_velocity(self, v): self._v.robot_angular_velocity[0] = float(v)
    @property
    def nozzle_angle(self): return self._f(self._v.nozzle_angle)
    @property
    def water_volume(self): return self._f(self._v.water_volume)
    @property
    def breathing_phase(self): return PHASE_NAMES[int(self._v.breathing_phase[0].item())]
    @property
    def breathing_timer(self): return int(self._v.breathing_timer[0].item())
    @property
    def is_inhaling(self):  # salp_robot.py:158-167 for forced mode; last command otherwise unknown
        return self.breathing_timer < 120 if self.forced_breathing else self.breathing_phase == "inhaling"
    @property
    def ellipse_a(self): return float(self._v.ellipse_axes()[0][0].item())
    @property
    def ellipse_b(self): return float(self._v.ellipse_axes()[1][0].item())
    @property
    def steps_since_food(self): return int(self._v.steps_since_food[0].item())
    @steps_since_food.setter
    def steps_since_food(self, v): self._v.steps_since_food[0] = int(v)
    @property
    def food_collected(self): return int(self._v.food_collected[0].item())
    @property
    def score(self): return float(self.food_collected * self.food_reward)
    @property
    def num_food_items(self): return int(self._v.num_food_items_per_env[0].item())

    @property
    def food_positions(self):
        pos, alive = self._v.food_positions
        n = self.num_food_items
        p, a = pos[0].cpu().numpy(), alive[0].cpu().numpy()
        return [[float(p[i, 0]), float(p[i, 1])] if a[i] else None for i in range(n)]

    @food_positions.setter
    def food_positions(self, items):
        items = list(items)
        pos = np.zeros((1, len(items), 2))
        alive = np.zeros((1, len(items)), dtype=bool)
        for i, it in enumerate(items):
            if it is not None:
                pos[0, i], alive[0, i] = it, True
        self._v.set_attr("food_positions", (pos, alive))
