"""GPU: no kernel writes outside the buffers it was given.  compute-sanitizer is not available on the GPU pool, so every
output / state tensor of the env is swapped for a view into canary-padded storage (re-bound through salp_bind) and the
bands are checked after resets, single steps, fused K-step launches and slot-addressed steps of odd-sized slabs --
the observation tiles leave the SM through TMA bulk stores, whose last partial tile is the interesting case."""
import numpy as np
import pytest
import torch

from helpers import CONFIGS, synth_actions

pytestmark = pytest.mark.gpu
PAD = 8192  # bytes of canary on both sides of every buffer (a multiple of every alignment the ABI asks for)


def _guard(t: torch.Tensor):
    nbytes = t.numel() * t.element_size()
    body = (nbytes + 255) // 256 * 256
    raw = torch.full((PAD + body + PAD,), 0xA5, dtype=torch.uint8, device=t.device)
    view = raw[PAD:PAD + nbytes].view(t.dtype).reshape(t.shape)
    view.copy_(t)
    return raw, view, nbytes


def _check(guards):
    for name, (raw, _, nbytes) in guards.items():
        assert bool((raw[:PAD] == 0xA5).all()), f"{name}: bytes before the buffer were written"
        assert bool((raw[PAD + nbytes:] == 0xA5).all()), f"{name}: bytes after the buffer were written"


ENV_BUFFERS = ("body", "ctrl", "ep_return", "food", "obs", "reward", "terminated", "truncated", "events", "final_obs",
               "final_info", "stats")


@pytest.mark.parametrize("cfg,extra,dtype", [("single_food", {}, "float32"), ("defaults", {}, "float32"),
                                             ("defaults", dict(num_food_items=5), "float32"),
                                             ("defaults", dict(num_food_items=20, forced_breathing=False), "float32"),
                                             ("single_food", dict(max_steps_without_food=9), "float64"),
                                             ("defaults", dict(max_steps_without_food=9), "float64")])
@pytest.mark.parametrize("N", [1003, 77])
def test_env_kernels_stay_inside_their_buffers(salp_lib, cfg, extra, dtype, N):
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    K = 3
    p = dict(CONFIGS[cfg], **extra)
    env = SalpVectorEnv(N, dtype=dtype, max_substeps=K, seed=123, **p)
    ref = SalpVectorEnv(N, dtype=dtype, max_substeps=K, seed=123, **p)
    guards = {}
    for name in ENV_BUFFERS:
        raw, view, nbytes = _guard(getattr(env, name))
        setattr(env, name, view)
        guards[name] = (raw, view, nbytes)
    env._bind()
    env.reset(); ref.reset()
    acts = torch.from_numpy(synth_actions(40, N, env.act_dim, seed=8)).cuda().to(env.real)
    for t in range(12):
        env.step(acts[t]); ref.step(acts[t])
    env.step_many(acts[12:15].contiguous(), K); ref.step_many(acts[12:15].contiguous(), K)
    for t in range(15, 21):
        env.step(acts[t], slot=t % K); ref.step(acts[t], slot=t % K)
    m = torch.zeros(N, dtype=torch.bool, device="cuda"); m[::3] = True
    env.reset(mask=m); ref.reset(mask=m)
    env.step(acts[21]); ref.step(acts[21])
    torch.cuda.synchronize()
    _check(guards)
    for name in ("body", "ctrl", "food", "obs", "reward", "final_obs"):
        assert torch.equal(getattr(env, name), getattr(ref, name)), name


def test_replay_append_and_actor_stay_inside_their_buffers(salp_lib):
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    from underwater_swimmer_rl_b200.rollout import DeviceReplayBuffer, FusedActor, FusedGailReward, GailDiscriminator, TanhGaussianActor
    torch.manual_seed(0)
    N = 1003
    env = SalpVectorEnv(N, max_substeps=2, seed=5, **CONFIGS["sac_gail"])
    env.reset()
    buf = DeviceReplayBuffer(2 * N + 17, env.obs_dim, env.act_dim, "cuda:0")   # rows wrap inside a step
    guards = {}
    for name in ("obs", "next_obs", "act", "rew", "done"):
        raw, view, nbytes = _guard(getattr(buf, name))
        setattr(buf, name, view)
        guards["rb_" + name] = (raw, view, nbytes)
    actor = FusedActor(TanhGaussianActor(env.obs_dim, env.act_dim).cuda())
    disc = FusedGailReward(GailDiscriminator(env.obs_dim, env.act_dim).cuda(), env.obs_dim, env.act_dim)
    raw_a, act, nb_a = _guard(torch.zeros(N, env.act_dim, device="cuda"))
    raw_r, aux, nb_r = _guard(torch.zeros(N, device="cuda"))
    guards["actions"], guards["gail_reward"] = (raw_a, act, nb_a), (raw_r, aux, nb_r)
    cur = 0
    for t in range(5):
        obs = env.obs[cur]
        actor(obs, act)
        disc(obs, act, aux)
        nxt = 1 - cur
        env.step(act, slot=nxt)
        buf.append(obs, act, env.reward[nxt], env.terminated[nxt], env.truncated[nxt], env.obs[nxt], final_obs=env.final_obs[nxt],
                   reward_aux=aux, w_env=0.3, w_aux=0.7)
        cur = nxt
    torch.cuda.synchronize()
    _check(guards)
    assert len(buf) == 2 * N + 17 and torch.isfinite(buf.rew).all() and (act.abs() <= 1).all()
