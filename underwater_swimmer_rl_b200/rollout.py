"""GPU-resident rollout collection: policy MLP -> env-step kernel -> replay ring, nothing leaves HBM.

The step either side of the env kernel (BASELINE.json configs[4]):

* ``TanhGaussianActor``    the reference's SAC actor shape (record_model.py:27-53): ReLU MLP
                           [obs -> 256 -> 256 -> 2*act], ``log_std`` clamped to [-20, 2], tanh-squashed
                           reparameterised sample.  A torch module (library GEMMs) -- it is the consumer of the
                           observation tensor, not part of the env path; it reads ``obs`` in place and writes
                           straight into the tensor whose pointer the env kernel reads.
* ``GailDiscriminator``    D(s, a) = sigmoid(MLP(cat(s, a))) and the imitation reward -log(1 - D + 1e-8)
                           (src/salp/agents/discriminator.py:43-85).
* ``DeviceReplayBuffer``   ring buffer of (obs, action, reward, next_obs, done) in HBM, filled by the
                           ``salp_replay_append`` kernel (csrc/salp_replay.cu) -- the batched, on-device form of
                           ``replay_buffer.add(...)`` in src/salp/training/continuous_trainer.py:296-303; rewards
                           optionally mixed ``w_env * r_env + w_gail * r_gail`` (base_config.py:77-78: 0.3 / 0.7).
* ``DeviceExpertBuffer``   expert (obs, action) pairs on the device (format of data/expert_demos/README.md:57-73 /
                           src/salp/training/expert_buffer.py:13-259), uniform sampling.
* ``RolloutCollector``     the loop itself (continuous_trainer.py:284-338 without the learner update): two output
                           rows of the env are used alternately so the previous observation stays alive without a
                           copy; optionally the whole step (policy + env + append) is replayed from a CUDA graph.
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Dict, Optional, Sequence

import numpy as np
import torch
import torch.nn as nn

from . import _capi
from .vector_env import SalpVectorEnv


class TanhGaussianActor(nn.Module):
    """record_model.py:27-53 (CustomActor): hidden ReLU layers, last layer -> (mean, log_std)."""

    def __init__(self, obs_dim: int, action_dim: int, hidden_sizes: Sequence[int] = (256, 256)):
        super().__init__()
        sizes = [obs_dim, *hidden_sizes]
        self.layers = nn.ModuleList([nn.Linear(a, b) for a, b in zip(sizes[:-1], sizes[1:])]
                                    + [nn.Linear(sizes[-1], 2 * action_dim)])
        self.action_dim = action_dim

    def forward(self, obs: torch.Tensor, deterministic: bool = False) -> torch.Tensor:
        x = obs.to(self.layers[0].weight.dtype)
        for layer in self.layers[:-1]:
            x = torch.relu(layer(x))
        mean, log_std = self.layers[-1](x).float().chunk(2, dim=-1)
        if deterministic:
            return torch.tanh(mean)
        std = torch.exp(torch.clamp(log_std, -20, 2))
        return torch.tanh(mean + std * torch.randn_like(mean))  # Normal(mean, std).rsample()


class GailDiscriminator(nn.Module):
    """src/salp/agents/discriminator.py:16-85: MLP on cat(obs, action) -> expert probability."""

    def __init__(self, obs_dim: int, action_dim: int, hidden_sizes: Sequence[int] = (256, 256)):
        super().__init__()
        sizes = [obs_dim + action_dim, *hidden_sizes, 1]
        self.layers = nn.ModuleList([nn.Linear(a, b) for a, b in zip(sizes[:-1], sizes[1:])])

    def forward(self, obs: torch.Tensor, action: torch.Tensor) -> torch.Tensor:
        x = torch.cat([obs, action.to(obs.dtype)], dim=-1).to(self.layers[0].weight.dtype)
        for layer in self.layers[:-1]:
            x = torch.relu(layer(x))
        return torch.sigmoid(self.layers[-1](x).float())

    @torch.no_grad()
    def predict_reward(self, obs: torch.Tensor, action: torch.Tensor) -> torch.Tensor:
        return -torch.log(1 - self.forward(obs, action) + 1e-8)  # discriminator.py:82


class DeviceReplayBuffer:
    """(obs, action, reward, next_obs, done) ring in HBM; head/size are device scalars."""

    def __init__(self, capacity: int, obs_dim: int, act_dim: int, device):
        self.capacity, self.obs_dim, self.act_dim = int(capacity), int(obs_dim), int(act_dim)
        self.device = torch.device(device)
        z = lambda *s, dtype=torch.float32: torch.zeros(*s, dtype=dtype, device=self.device)  # noqa: E731
        self.obs, self.next_obs = z(capacity, obs_dim), z(capacity, obs_dim)
        self.act, self.rew = z(capacity, act_dim), z(capacity)
        self.done = z(capacity, dtype=torch.bool)
        self.head, self.size = z(1, dtype=torch.int64), z(1, dtype=torch.int64)
        self._L = _capi.load()

    def append(self, obs_prev, actions, reward, terminated, truncated, obs_next, final_obs=None,
               reward_aux=None, w_env: float = 1.0, w_aux: float = 0.0) -> None:
        """Append one step of n = obs_prev.shape[0] transitions (one kernel + a one-thread head bump)."""
        n = obs_prev.shape[0]
        p = _capi.SalpReplayAppend()
        for name, t in (("obs_prev", obs_prev), ("actions", actions), ("reward", reward), ("terminated", terminated),
                        ("truncated", truncated), ("obs_next", obs_next), ("rb_obs", self.obs), ("rb_act", self.act),
                        ("rb_rew", self.rew), ("rb_next", self.next_obs), ("rb_done", self.done), ("head", self.head),
                        ("size", self.size)):
            assert t.is_cuda and t.is_contiguous(), name
            setattr(p, name, t.data_ptr())
        assert obs_prev.dtype == torch.float32 and obs_next.dtype == torch.float32
        assert actions.dtype == reward.dtype and actions.dtype in (torch.float32, torch.float64)
        p.final_obs = final_obs.data_ptr() if final_obs is not None else None
        p.reward_aux = reward_aux.data_ptr() if reward_aux is not None else None
        if reward_aux is not None:
            assert reward_aux.dtype == torch.float32 and reward_aux.is_contiguous() and reward_aux.numel() == n
        p.n, p.capacity, p.obs_dim, p.act_dim = n, self.capacity, self.obs_dim, self.act_dim
        p.dtype = _capi.SALP_F64 if actions.dtype == torch.float64 else _capi.SALP_F32
        p.w_env, p.w_aux = float(w_env), float(w_aux)
        with torch.cuda.device(self.device):
            _capi.check(self._L.salp_replay_append(C.byref(p), C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))

    def __len__(self) -> int:
        return int(self.size.item())

    def sample(self, batch_size: int, generator: Optional[torch.Generator] = None) -> Dict[str, torch.Tensor]:
        """Uniform batch over the filled rows (continuous_trainer.py:312 ``replay_buffer.sample``); the batch
        indices are drawn on the device, the only host read is the fill level."""
        size = len(self)
        if size == 0:
            raise ValueError("replay buffer is empty")
        idx = torch.randint(0, size, (batch_size,), device=self.device, generator=generator)
        return {"observations": self.obs[idx], "actions": self.act[idx], "rewards": self.rew[idx],
                "next_observations": self.next_obs[idx], "dones": self.done[idx]}


class DeviceExpertBuffer:
    """Expert transitions on the device.  ``episodes``: dicts with 'observations' [T, obs_dim] and
    'actions' [T, act_dim] (the pickle format of data/expert_demos/README.md:57-73)."""

    def __init__(self, episodes, device):
        obs = np.concatenate([np.asarray(e["observations"], dtype=np.float32) for e in episodes])
        act = np.concatenate([np.asarray(e["actions"], dtype=np.float32).reshape(len(e["observations"]), -1) for e in episodes])
        self.observations = torch.from_numpy(obs).to(device)
        self.actions = torch.from_numpy(act).to(device)
        self.num_transitions = obs.shape[0]

    def sample(self, batch_size: int, generator: Optional[torch.Generator] = None) -> Dict[str, torch.Tensor]:
        idx = torch.randint(0, self.num_transitions, (batch_size,), device=self.observations.device, generator=generator)
        return {"observations": self.observations[idx], "actions": self.actions[idx]}


class RolloutCollector:
    """Steps ``env`` with ``actor`` in the loop and appends every transition to ``buffer``.

    ``env`` must have been built with ``max_substeps >= 2`` (two output rows, used alternately) and
    auto-reset on.  With ``discriminator`` the stored reward is ``w_env*r_env + w_gail*r_gail``.
    """

    def __init__(self, env: SalpVectorEnv, actor: nn.Module, buffer: Optional[DeviceReplayBuffer] = None,
                 discriminator: Optional[GailDiscriminator] = None, w_env: float = 0.3, w_gail: float = 0.7,
                 deterministic: bool = False, autocast_dtype: Optional[torch.dtype] = None):
        if env.max_substeps < 2:
            raise ValueError("RolloutCollector needs an env with max_substeps >= 2 (ping-pong output rows)")
        self.env, self.actor, self.buffer, self.disc = env, actor, buffer, discriminator
        self.w_env, self.w_gail, self.deterministic = w_env, w_gail, deterministic
        self.autocast_dtype = autocast_dtype
        self.cur = 0          # output row holding the current observation
        self.act = torch.zeros(env.num_envs, env.act_dim, dtype=env.real, device=env.device)  # the kernel reads this
        self.steps = 0
        self._graph = None

    def reset(self):
        obs, info = self.env.reset()
        self.cur = 0
        return obs, info

    @torch.no_grad()
    def _one_step(self, cur: int) -> int:
        env = self.env
        obs = env.obs[cur]
        a = self.actor(obs, deterministic=self.deterministic)
        self.act.copy_(a.reshape(env.num_envs, env.act_dim))  # cast to the build's real type, in place
        nxt = 1 - cur
        env.step(self.act, slot=nxt)
        if self.buffer is not None:
            aux = self.disc.predict_reward(obs, self.act).reshape(-1).contiguous() if self.disc is not None else None
            self.buffer.append(obs, self.act, env.reward[nxt], env.terminated[nxt], env.truncated[nxt], env.obs[nxt],
                               final_obs=env.final_obs[nxt] if env.autoreset else None, reward_aux=aux,
                               w_env=self.w_env if aux is not None else 1.0, w_aux=self.w_gail if aux is not None else 0.0)
        return nxt

    def collect(self, n_steps: int) -> torch.Tensor:
        """Run ``n_steps`` env steps (eager launches).  Returns the current observation tensor."""
        for _ in range(n_steps):
            self.cur = self._one_step(self.cur)
        self.steps += n_steps
        return self.env.obs[self.cur]

    def capture(self) -> Callable[[], None]:
        """Capture TWO steps (row 0 -> 1 -> 0) into a CUDA graph; each replay advances the rollout by two env
        steps with a single graph launch.  Must be called with the current observation in row 0."""
        if self.cur != 0:
            self.collect(1)
        dev = self.env.device
        s = torch.cuda.Stream(dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):  # warm-up outside capture (cuBLAS workspaces, lazy module loads)
            for _ in range(2):
                self.cur = self._one_step(self.cur)
        torch.cuda.current_stream(dev).wait_stream(s)
        self.steps += 2
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            c = self._one_step(0)
            c = self._one_step(c)
        assert c == 0
        self._graph = g

        def replay():
            g.replay()
            self.steps += 2
        return replay
