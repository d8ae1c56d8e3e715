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
    """record_model.py:27-53 (CustomActor): hidden ReLU layers, last layer -> (mean, log_std).

    ``action_low`` / ``action_high``: bounds of the env's action Box.  When given, the squashed action is mapped
    from [-1, 1] to the Box the way SB3's SAC does before stepping the env (``unscale_action``:
    low + 0.5 * (a + 1) * (high - low)) -- model (B)'s Box is [0, 1] x [0, 1] x [-1, 1] (salp_robot_env.py:51-56)."""

    def __init__(self, obs_dim: int, action_dim: int, hidden_sizes: Sequence[int] = (256, 256),
                 action_low: Optional[Sequence[float]] = None, action_high: Optional[Sequence[float]] = None):
        super().__init__()
        if (action_low is None) != (action_high is None):
            raise ValueError("give both action_low and action_high, or neither")
        self.rescale = action_low is not None
        if self.rescale:
            lo, hi = torch.as_tensor(action_low, dtype=torch.float32), torch.as_tensor(action_high, dtype=torch.float32)
            if lo.shape != (action_dim,) or hi.shape != (action_dim,):
                raise ValueError("action_low / action_high must have action_dim entries")
            self.register_buffer("action_low", lo)
            self.register_buffer("action_high", hi)
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
            return self.to_box(torch.tanh(mean))
        std = torch.exp(torch.clamp(log_std, -20, 2))
        return self.to_box(torch.tanh(mean + std * torch.randn_like(mean)))  # Normal(mean, std).rsample()

    def to_box(self, squashed: torch.Tensor) -> torch.Tensor:
        if not self.rescale:
            return squashed
        return squashed * (0.5 * (self.action_high - self.action_low)) + 0.5 * (self.action_high + self.action_low)


def pack_weight(w: torch.Tensor, k_pad: int) -> torch.Tensor:
    """torch Linear weight [out, in] -> bf16 [k_pad/8][out][8]: the tcgen05 K-major, no-swizzle operand layout
    (8x8 core matrices, include/salp_b200.h SalpActorForward); ``in`` is zero-padded to ``k_pad``."""
    out_f, in_f = w.shape
    wp = torch.zeros(out_f, k_pad, dtype=torch.bfloat16, device=w.device)
    wp[:, :in_f] = w.detach().to(torch.bfloat16)
    return wp.view(out_f, k_pad // 8, 8).permute(1, 0, 2).contiguous()


class FusedActor:
    """The same function as ``TanhGaussianActor.forward`` computed by ONE sm_100a kernel
    (csrc/salp_actor.cu: bf16 tcgen05 MMAs, accumulator in TMEM, weights resident in shared memory).
    Built from a ``TanhGaussianActor`` with hidden sizes (256, 256); ``refresh()`` re-packs the weights after
    a learner update.  ``__call__`` writes the actions straight into ``out`` (the tensor the env kernel reads)."""

    def __init__(self, actor: TanhGaussianActor, seed: int = 0):
        self.actor = actor
        lin = list(actor.layers)
        if len(lin) != 3 or lin[0].out_features != 256 or lin[1].out_features != 256:
            raise ValueError("FusedActor needs hidden_sizes == (256, 256)")
        self.obs_dim, self.act_dim = lin[0].in_features, actor.action_dim
        if self.obs_dim > 32 or not 1 <= self.act_dim <= 4:
            raise ValueError("FusedActor supports obs_dim <= 32 and act_dim in 1..4")
        self._L = _capi.load()
        # {seed, draw counter, ticket, pad}: the kernel draws its own normals (Philox4x32-10) and bumps the counter
        self.rng_state = torch.tensor([int(seed), 0, 0, 0], dtype=torch.int64, device=lin[0].weight.device)
        self.refresh()

    def refresh(self) -> None:
        l1, l2, l3 = self.actor.layers
        self.w1p, self.w2p = pack_weight(l1.weight, 32), pack_weight(l2.weight, 256)
        f32 = lambda t: t.detach().float().contiguous()  # noqa: E731
        # W3 goes through bf16 like the other layers of a bf16 actor would, then is held in f32
        self.b1, self.b2, self.b3 = f32(l1.bias), f32(l2.bias), f32(l3.bias)
        self.w3 = f32(l3.weight.detach().to(torch.bfloat16))
        if self.actor.rescale:  # host copies: the bounds travel by value in SalpActorForward
            self._low, self._high = self.actor.action_low.tolist(), self.actor.action_high.tolist()

    def __call__(self, obs: torch.Tensor, out: torch.Tensor, deterministic: bool = False,
                 noise: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``noise``: explicit standard-normal draws [n, act]; without it (and not deterministic) the kernel
        samples in place from its own counter-based generator -- no extra launch, CUDA-graph replay safe."""
        n = obs.shape[0]
        assert obs.is_cuda and obs.dtype == torch.float32 and obs.is_contiguous() and obs.shape[1] == self.obs_dim
        assert out.is_contiguous() and out.numel() == n * self.act_dim and out.dtype in (torch.float32, torch.float64)
        p = _capi.SalpActorForward()
        p.obs, p.w1_packed, p.b1, p.w2_packed, p.b2 = (obs.data_ptr(), self.w1p.data_ptr(), self.b1.data_ptr(),
                                                       self.w2p.data_ptr(), self.b2.data_ptr())
        p.w3, p.b3, p.actions = self.w3.data_ptr(), self.b3.data_ptr(), out.data_ptr()
        p.noise = None if (deterministic or noise is None) else noise.data_ptr()
        p.rng_state = None if (deterministic or noise is not None) else self.rng_state.data_ptr()
        p.extra, p.extra_dim, p.mode = None, 0, 0
        p.act_rescale = 1 if self.actor.rescale else 0
        if self.actor.rescale:
            for d in range(self.act_dim):
                p.act_low[d], p.act_high[d] = self._low[d], self._high[d]
        p.n, p.obs_dim, p.act_dim, p.hidden = n, self.obs_dim, self.act_dim, 256
        p.dtype = _capi.SALP_F64 if out.dtype == torch.float64 else _capi.SALP_F32
        with torch.cuda.device(obs.device):
            _capi.check(self._L.salp_actor_forward(C.byref(p), C.c_void_p(torch.cuda.current_stream(obs.device).cuda_stream)))
        return out


class FusedGailReward:
    """``GailDiscriminator.predict_reward`` -- -log(1 - sigmoid(MLP(cat(obs, action))) + 1e-8),
    src/salp/agents/discriminator.py:43-85 -- by the same fused tcgen05 kernel as ``FusedActor`` (mode 1: the
    action columns are appended to the observation row on the way into shared memory).  Built from a
    ``GailDiscriminator`` with hidden sizes (256, 256); ``refresh()`` after a discriminator update."""

    def __init__(self, disc: "GailDiscriminator", obs_dim: int, act_dim: int):
        self.disc, self.obs_dim, self.act_dim = disc, int(obs_dim), int(act_dim)
        lin = list(disc.layers)
        if len(lin) != 3 or lin[0].out_features != 256 or lin[1].out_features != 256 or lin[2].out_features != 1:
            raise ValueError("FusedGailReward needs a discriminator obs+act -> 256 -> 256 -> 1")
        if lin[0].in_features != obs_dim + act_dim or obs_dim + act_dim > 32:
            raise ValueError("FusedGailReward supports obs_dim + act_dim <= 32")
        self._L = _capi.load()
        self.refresh()

    def refresh(self) -> None:
        l1, l2, l3 = self.disc.layers
        self.w1p, self.w2p = pack_weight(l1.weight, 32), pack_weight(l2.weight, 256)
        f32 = lambda t: t.detach().float().contiguous()  # noqa: E731
        self.b1, self.b2, self.b3 = f32(l1.bias), f32(l2.bias), f32(l3.bias)
        self.w3 = f32(l3.weight.detach().to(torch.bfloat16))

    def __call__(self, obs: torch.Tensor, action: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        n = obs.shape[0]
        assert obs.is_cuda and obs.dtype == torch.float32 and obs.is_contiguous() and obs.shape[1] == self.obs_dim
        assert action.is_contiguous() and action.numel() == n * self.act_dim and action.dtype in (torch.float32, torch.float64)
        assert out.dtype == torch.float32 and out.is_contiguous() and out.numel() == n
        p = _capi.SalpActorForward()
        p.obs, p.w1_packed, p.b1, p.w2_packed, p.b2 = (obs.data_ptr(), self.w1p.data_ptr(), self.b1.data_ptr(),
                                                       self.w2p.data_ptr(), self.b2.data_ptr())
        p.w3, p.b3, p.actions = self.w3.data_ptr(), self.b3.data_ptr(), out.data_ptr()
        p.noise, p.rng_state = None, None
        p.extra, p.extra_dim, p.mode = action.data_ptr(), self.act_dim, 1
        p.n, p.obs_dim, p.act_dim, p.hidden = n, self.obs_dim, self.act_dim, 256
        p.dtype = _capi.SALP_F64 if action.dtype == torch.float64 else _capi.SALP_F32
        with torch.cuda.device(obs.device):
            _capi.check(self._L.salp_actor_forward(C.byref(p), C.c_void_p(torch.cuda.current_stream(obs.device).cuda_stream)))
        return out


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
        self._ticket = z(1, dtype=torch.int64)
        self._L = _capi.load()

    def append(self, obs_prev, actions, reward, terminated, truncated, obs_next, final_obs=None,
               reward_aux=None, w_env: float = 1.0, w_aux: float = 0.0) -> None:
        """Append one step of n = obs_prev.shape[0] transitions (one kernel + a one-thread head bump)."""
        n = obs_prev.shape[0]
        p = _capi.SalpReplayAppend()
        for name, t in (("obs_prev", obs_prev), ("actions", actions), ("reward", reward), ("terminated", terminated),
                        ("truncated", truncated), ("obs_next", obs_next), ("rb_obs", self.obs), ("rb_act", self.act),
                        ("rb_rew", self.rew), ("rb_next", self.next_obs), ("rb_done", self.done), ("head", self.head),
                        ("size", self.size), ("ticket", self._ticket)):
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


class _OutputRows:
    """The [K, N, ...] output tensors of either env: ``SalpVectorEnv`` (model A) holds them under the plain names,
    ``SalpCycleVectorEnv`` (model B, built with ``rows=K``) under ``*_rows``."""

    def __init__(self, env):
        cyc = hasattr(env, "obs_rows")
        self.K = env.output_rows
        for name in ("obs", "reward", "terminated", "truncated", "final_obs"):
            setattr(self, name, getattr(env, name + "_rows") if cyc else getattr(env, name))


class StepRingReplay:
    """Zero-copy replay storage: the env's own ``[K, N, ...]`` output rows ARE the buffer.

    ``salp_step_slot`` writes step j's observation, reward, flags and terminal observation into row ``j % K`` of the
    tensors the env already owns, and the actor writes its action into row ``j % K`` of ``self.act`` -- so storing a
    transition costs no kernel and no extra HBM traffic (the transition-major ``DeviceReplayBuffer`` re-reads and
    re-writes ~410 B per transition in ``salp_replay_append``).  Transition j is
        (obs[(j-1) % K], act[j % K], reward[j % K], final_obs[j % K] if done else obs[j % K], done[j % K])
    and stays intact until row ``(j-1) % K`` is overwritten by step ``j-1+K``: the ring holds the last K-1 steps,
    i.e. ``(K-1) * N`` transitions (K = 5 at 262 144 envs = 1 048 576, the reference's buffer_size).
    With a GAIL discriminator the imitation reward of each step is kept in ``self.aux`` and mixed when sampling
    (``w_env * r_env + w_aux * r_gail``, src/salp/config/base_config.py:77-78)."""

    def __init__(self, env):
        if env.output_rows < 3:
            raise ValueError("StepRingReplay needs an env with >= 3 output rows (max_substeps of SalpVectorEnv, "
                             "rows of SalpCycleVectorEnv)")
        if not env.autoreset:
            raise ValueError("StepRingReplay needs autoreset=True (terminal observations come from final_obs)")
        self.env, self.K, self.N = env, env.output_rows, env.num_envs
        self.out = _OutputRows(env)
        self.device = env.device
        self.act = torch.zeros(self.K, self.N, env.act_dim, dtype=env.real, device=env.device)
        self.aux = torch.zeros(self.K, self.N, dtype=torch.float32, device=env.device)
        self.w_env, self.w_aux = 1.0, 0.0
        self.steps = 0                 # env steps written so far (host counter; the rollout is launched from the host)

    @property
    def capacity(self) -> int:
        return (self.K - 1) * self.N

    def __len__(self) -> int:
        return min(self.steps, self.K - 1) * self.N

    def note_steps(self, n: int) -> None:
        self.steps += int(n)

    def sample(self, batch_size: int, generator: Optional[torch.Generator] = None) -> Dict[str, torch.Tensor]:
        """Uniform batch over the stored transitions (continuous_trainer.py:312 ``replay_buffer.sample``)."""
        have = min(self.steps, self.K - 1)
        if have == 0:
            raise ValueError("replay buffer is empty")
        env = self.out
        back = torch.randint(0, have, (batch_size,), device=self.device, generator=generator)   # 0 = newest step
        e = torch.randint(0, self.N, (batch_size,), device=self.device, generator=generator)
        row = (self.steps - back) % self.K           # row of step j = steps - back (steps are numbered from 1)
        prev = (row - 1) % self.K
        done = (env.terminated[row, e] | env.truncated[row, e]).bool()
        nxt = torch.where(done[:, None], env.final_obs[row, e], env.obs[row, e])
        rew = env.reward[row, e].float()
        if self.w_aux != 0.0:
            rew = self.w_env * rew + self.w_aux * self.aux[row, e]
        return {"observations": env.obs[prev, e], "actions": self.act[row, e].float(), "rewards": rew,
                "next_observations": nxt, "dones": done}


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

    ``env``: a ``SalpVectorEnv`` built with ``max_substeps >= 2`` or a ``SalpCycleVectorEnv`` built with
    ``rows >= 2`` (two output rows, used alternately; all rows with a ``StepRingReplay``) and auto-reset on.
    With ``discriminator`` the stored reward is ``w_env*r_env + w_gail*r_gail``.
    """

    def __init__(self, env, actor, buffer: Optional[DeviceReplayBuffer] = None,
                 discriminator=None, w_env: float = 0.3, w_gail: float = 0.7,
                 deterministic: bool = False, autocast_dtype: Optional[torch.dtype] = None):
        if env.output_rows < 2:
            raise ValueError("RolloutCollector needs an env with >= 2 output rows (ping-pong)")
        self.env, self.actor, self.buffer, self.disc = env, actor, buffer, discriminator
        self.out = _OutputRows(env)
        self.w_env, self.w_gail, self.deterministic = w_env, w_gail, deterministic
        self.autocast_dtype = autocast_dtype
        self.cur = 0          # output row holding the current observation
        self.act = torch.zeros(env.num_envs, env.act_dim, dtype=env.real, device=env.device)  # the kernel reads this
        self.steps = 0
        self._graph = None
        self._aux = torch.zeros(env.num_envs, dtype=torch.float32, device=env.device)  # GAIL reward of the step
        self.ring = buffer if isinstance(buffer, StepRingReplay) else None
        self.rows = env.output_rows if self.ring is not None else 2   # output rows the rollout cycles through
        if self.ring is not None and discriminator is not None:
            self.ring.w_env, self.ring.w_aux = float(w_env), float(w_gail)

    def reset(self):
        obs, info = self.env.reset()
        self.cur = 0
        if self.ring is not None:
            self.ring.steps = 0    # rows are numbered from the reset observation in row 0
        return obs, info

    @torch.no_grad()
    def _one_step(self, cur: int) -> int:
        env, out = self.env, self.out
        obs = out.obs[cur]
        nxt = (cur + 1) % self.rows
        act = self.ring.act[nxt] if self.ring is not None else self.act   # ring: the action row is the storage
        if isinstance(self.actor, FusedActor):
            self.actor(obs, act, deterministic=self.deterministic)  # one kernel, writes the env's action tensor
        else:
            a = self.actor(obs, deterministic=self.deterministic)
            act.copy_(a.reshape(env.num_envs, env.act_dim))  # cast to the build's real type, in place
        env.step(act, slot=nxt)
        if self.ring is not None:
            # zero-copy storage: everything of the transition already sits in row `nxt`; only the imitation reward
            # (when there is a discriminator) is computed, straight into its ring row
            if isinstance(self.disc, FusedGailReward):
                self.disc(obs, act, self.ring.aux[nxt])
            elif self.disc is not None:
                self.ring.aux[nxt].copy_(self.disc.predict_reward(obs, act).reshape(-1))
        elif self.buffer is not None:
            if isinstance(self.disc, FusedGailReward):
                aux = self.disc(obs, self.act, self._aux)
            else:
                aux = self.disc.predict_reward(obs, self.act).reshape(-1).contiguous() if self.disc is not None else None
            self.buffer.append(obs, self.act, out.reward[nxt], out.terminated[nxt], out.truncated[nxt], out.obs[nxt],
                               final_obs=out.final_obs[nxt] if env.autoreset else None, reward_aux=aux,
                               w_env=self.w_env if aux is not None else 1.0, w_aux=self.w_gail if aux is not None else 0.0)
        return nxt

    def collect(self, n_steps: int) -> torch.Tensor:
        """Run ``n_steps`` env steps (eager launches).  Returns the current observation tensor."""
        for _ in range(n_steps):
            self.cur = self._one_step(self.cur)
        self.steps += n_steps
        if self.ring is not None:
            self.ring.note_steps(n_steps)
        return self.out.obs[self.cur]

    def capture(self) -> Callable[[], None]:
        """Capture one trip through the output rows (2 steps with ping-pong rows, K with a StepRingReplay) into a
        CUDA graph; each replay advances the rollout by that many env steps with a single graph launch.
        ``steps_per_replay`` holds the count."""
        R = self.rows
        while self.cur != 0:
            self.collect(1)
        dev = self.env.device
        s = torch.cuda.Stream(dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):  # warm-up outside capture (cuBLAS workspaces, lazy module loads)
            for _ in range(R):
                self.cur = self._one_step(self.cur)
        torch.cuda.current_stream(dev).wait_stream(s)
        self.steps += R
        if self.ring is not None:
            self.ring.note_steps(R)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            c = 0
            for _ in range(R):
                c = self._one_step(c)
        assert c == 0
        self._graph = g
        self.steps_per_replay = R

        def replay():
            g.replay()
            self.steps += R
            if self.ring is not None:
                self.ring.note_steps(R)
        return replay
