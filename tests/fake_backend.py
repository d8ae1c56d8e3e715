"""CPU stand-in for SalpVectorEnv behind the 1-env facade (tests only).

The build container has no GPU and the GPU box has no /root/reference, so the test that runs the reference's unmodified
train.py on top of the facade (tests/test_train_py_runs_unchanged.py) needs the facade without CUDA.  This class offers the
slice of the SalpVectorEnv surface that snake_env.SalpSnakeEnv touches, computed by the CPU oracle; tensors are CPU torch
tensors refreshed from the oracle's state after every call, and writes to them are pushed back before the next call."""
import numpy as np
import torch

from oracle import salp_oracle
from underwater_swimmer_rl_b200._gym_compat import Box
from underwater_swimmer_rl_b200.vector_env import ENV_DEFAULTS
import helpers


class _Params:
    pass


class OracleVectorBackend:
    def __init__(self, num_envs, device=None, dtype="float64", seed=10_000, autoreset=False, food_table_len=4096,
                 jitter_table_len=16384, render_mode=None, **env_kwargs):
        assert num_envs == 1 and render_mode is None and not autoreset
        unknown = set(env_kwargs) - set(ENV_DEFAULTS)
        if unknown:
            raise TypeError(f"SalpSnakeEnv got unexpected keyword argument(s) {sorted(unknown)}")
        kw = dict(ENV_DEFAULTS, **env_kwargs)
        for k, v in kw.items():
            setattr(self, k, v)
        self.seed_base = int(seed)
        self._tables(self.seed_base)
        self.orc = salp_oracle.Oracle(1, self.fw, self.jw, **kw)
        self.act_dim, self.obs_dim = self.orc.act_dim, self.orc.obs_dim
        self.base_num_food_items = max(0, int(kw["num_food_items"]))
        self.single_action_space = Box(-1.0, 1.0, (1,), np.float32) if self.act_dim == 1 else Box(np.array([0.0, -1.0]), np.array([1.0, 1.0]), dtype=np.float32)
        self.single_observation_space = Box(-np.inf, np.inf, (self.obs_dim,), np.float32)
        self._pull()

    def _tables(self, seed):
        self.fw, self.jw = helpers.py_tables([seed], 4096, 16384)

    # -- state mirror ------------------------------------------------------------------------------------------------
    def _pull(self):
        e = self.orc.envs
        self.robot_pos = torch.tensor([[e["px"][0], e["py"][0]]], dtype=torch.float64)
        self.robot_velocity = torch.tensor([[e["vx"][0], e["vy"][0]]], dtype=torch.float64)
        self.robot_angle = torch.tensor([e["angle"][0]], dtype=torch.float64)
        self.robot_angular_velocity = torch.tensor([e["omega"][0]], dtype=torch.float64)
        self.nozzle_angle = torch.tensor([e["nozzle"][0]])
        self.water_volume = torch.tensor([e["wv"][0]])
        self.breathing_phase = torch.tensor([int(e["phase"][0])])
        self.breathing_timer = torch.tensor([int(e["timer"][0])])
        self.steps_since_food = torch.tensor([int(e["steps_since_food"][0])])
        self.food_collected = torch.tensor([int(e["food_collected"][0])])
        self.num_food_items_per_env = torch.tensor([int(e["n_food"][0])])

    def _push(self):
        e = self.orc.envs
        e["px"][0], e["py"][0] = self.robot_pos[0].tolist()
        e["vx"][0], e["vy"][0] = self.robot_velocity[0].tolist()
        e["angle"][0], e["omega"][0] = float(self.robot_angle[0]), float(self.robot_angular_velocity[0])
        e["steps_since_food"][0] = int(self.steps_since_food[0])

    def ellipse_axes(self):
        e = self.orc.envs
        return torch.tensor([e["a"][0]]), torch.tensor([e["b"][0]])

    @property
    def food_positions(self):
        e = self.orc.envs
        return torch.from_numpy(e["food"][:, :].copy()), torch.from_numpy(e["alive"][:, :].astype(bool))

    def set_attr(self, name, values):
        assert name == "food_positions"
        pos, alive = values
        e = self.orc.envs
        n = pos.shape[1]
        e["food"][0, :n] = pos[0]
        e["alive"][0, :] = 0
        e["alive"][0, :n] = alive[0]
        e["n_food"][0] = n

    # -- gym surface ----------------------------------------------------------------------------------------------------
    def reset(self, *, seed=None, options=None, mask=None):
        if seed is not None:
            self.seed_base = int(seed)
            self._tables(self.seed_base)
            self.orc = salp_oracle.Oracle(1, self.fw, self.jw, **{k: getattr(self, k) for k in ENV_DEFAULTS})
        self.orc.p.num_food_items = self.base_num_food_items
        obs = self.orc.reset()
        self._pull()
        return torch.from_numpy(obs), {}

    def step(self, actions, slot=0):
        self._push()
        obs, rew, flags = self.orc.step(np.asarray(actions, dtype=np.float64))
        self._pull()
        return (torch.from_numpy(obs), torch.from_numpy(rew), torch.from_numpy((flags & 1).astype(bool)),
                torch.from_numpy(((flags >> 1) & 1).astype(bool)), {"events": torch.from_numpy(flags)})

    def observe(self, slot=0):
        self._push()
        return torch.from_numpy(self.orc.observe())

    def close(self, **kw):
        pass
