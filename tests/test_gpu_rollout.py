"""GPU: rollout -> replay path (rollout.py over salp_step_slot + salp_replay_append), SURVEY 8(f) N2/N3.
The ring contents are checked against the same transitions assembled with plain torch indexing."""
import numpy as np
import pytest
import torch

from helpers import CONFIGS

pytestmark = pytest.mark.gpu


def _mk(n, cap, disc=False, dtype="float32", seed=3):
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    from underwater_swimmer_rl_b200.rollout import DeviceReplayBuffer, GailDiscriminator, RolloutCollector, TanhGaussianActor
    torch.manual_seed(seed)
    env = SalpVectorEnv(n, device="cuda:0", dtype=dtype, max_substeps=2, seed=555, **CONFIGS["sac_gail"])
    actor = TanhGaussianActor(env.obs_dim, env.act_dim).cuda()
    buf = DeviceReplayBuffer(cap, env.obs_dim, env.act_dim, "cuda:0")
    d = GailDiscriminator(env.obs_dim, env.act_dim).cuda() if disc else None
    return env, actor, buf, RolloutCollector(env, actor, buf, discriminator=d, deterministic=True)


@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_ring_holds_the_transitions_the_reference_loop_would_store(salp_lib, dtype):
    n, cap, T = 1000, 300 * 1000 + 500, 900  # capacity is not a multiple of n: rows wrap inside a step
    env, actor, buf, col = _mk(n, cap, dtype=dtype)
    col.reset()
    # force early episode ends so that final_obs rows are exercised inside the window kept by the ring
    pos = torch.stack([torch.linspace(450.0, 705.0, n), torch.full((n,), 300.0)], 1)
    env.set_attr("robot_pos", pos.cuda())
    exp = {k: [] for k in ("obs", "act", "rew", "next", "done")}
    for t in range(T):
        cur = col.cur
        obs_prev = env.obs[cur].clone()
        col.collect(1)
        nxt = col.cur
        done = env.terminated[nxt] | env.truncated[nxt]
        exp["obs"].append(obs_prev)
        exp["act"].append(col.act.float().clone())
        exp["rew"].append(env.reward[nxt].float().clone())
        exp["next"].append(torch.where(done[:, None], env.final_obs[nxt], env.obs[nxt]).clone())
        exp["done"].append(done.clone())
    assert len(buf) == cap and int(buf.head.item()) == (n * T) % cap
    flat = {k: torch.cat(v) for k, v in exp.items()}
    rows = torch.arange(n * T - cap, n * T, device="cuda")  # the last `cap` transitions survive, at row (i mod cap)
    assert int(flat["done"][rows].sum()) > n // 4  # episodes ended inside the window the ring keeps
    ring = rows % cap
    assert torch.equal(buf.obs[ring], flat["obs"][rows])
    assert torch.equal(buf.next_obs[ring], flat["next"][rows])
    assert torch.equal(buf.act[ring], flat["act"][rows])
    assert torch.equal(buf.rew[ring], flat["rew"][rows])
    assert torch.equal(buf.done[ring], flat["done"][rows])
    # next_obs of a non-final transition is the obs of the same env's next transition (n rows later)
    r0 = rows[:-n][~flat["done"][rows[:-n]]]
    assert torch.equal(flat["next"][r0], flat["obs"][r0 + n])
    batch = buf.sample(256)
    assert batch["observations"].shape == (256, env.obs_dim) and batch["dones"].dtype == torch.bool


def test_gail_reward_mixing(salp_lib):
    n = 512
    env, actor, buf, col = _mk(n, 4 * n, disc=True)
    col.reset()
    obs_prev = env.obs[0].clone()
    col.collect(1)
    r_env = env.reward[col.cur].float()
    r_gail = col.disc.predict_reward(obs_prev, col.act).reshape(-1)
    assert torch.allclose(buf.rew[:n], 0.3 * r_env + 0.7 * r_gail, rtol=1e-5, atol=1e-6)
    assert (r_gail > 0).all()  # -log(1 - D) with D in (0, 1)


def test_graph_replay_equals_eager_rollout(salp_lib):
    n, steps = 2048, 40
    outs = []
    for use_graph in (False, True):
        env, actor, buf, col = _mk(n, n * 8, seed=11)
        col.reset()
        if use_graph:
            replay = col.capture()       # runs 2 warm-up steps eagerly
            for _ in range((steps - 2) // 2):
                replay()
        else:
            col.collect(steps)
        torch.cuda.synchronize()
        assert col.steps == steps
        outs.append((env.body.clone(), env.ctrl.clone(), buf.obs.clone(), buf.rew.clone(), buf.head.clone()))
    for x, y in zip(*outs):
        assert torch.equal(x, y)


def test_expert_buffer_and_bad_arguments(salp_lib):
    from underwater_swimmer_rl_b200 import _capi
    from underwater_swimmer_rl_b200.rollout import DeviceExpertBuffer
    z = np.load(__import__("os").path.join(__import__("os").path.dirname(__file__), "golden", "human_demos.npz"))
    keys = sorted(k for k in z.files if k.endswith("_actions"))
    eps = [{"observations": np.zeros((len(z[k]), 24), np.float32), "actions": z[k]} for k in keys]
    eb = DeviceExpertBuffer(eps, "cuda:0")
    b = eb.sample(128)
    assert b["observations"].shape == (128, 24) and b["actions"].shape == (128, 1)
    assert eb.num_transitions == sum(len(z[k]) for k in keys)
    import ctypes as C
    p = _capi.SalpReplayAppend()
    assert salp_lib.salp_replay_append(C.byref(p), None) == _capi.SALP_EINVAL


@pytest.mark.parametrize("dtype,fused,disc", [("float32", True, True), ("float64", False, False)])
def test_step_ring_replay_is_zero_copy_and_exact(salp_lib, dtype, fused, disc):
    """StepRingReplay: the env's [K, N] output rows are the replay storage.  Every transition still inside the ring
    must equal the one assembled the way the reference loop stores it (continuous_trainer.py:296-303)."""
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    from underwater_swimmer_rl_b200.rollout import (FusedActor, FusedGailReward, GailDiscriminator, RolloutCollector,
                                                     StepRingReplay, TanhGaussianActor)
    torch.manual_seed(1)
    n, K, T = 768, 5, 16
    # episodes are cut after 7 steps without food so that terminal observations fall inside the ring window
    env = SalpVectorEnv(n, device="cuda:0", dtype=dtype, max_substeps=K, seed=99,
                        **dict(CONFIGS["sac_gail"], max_steps_without_food=7))
    actor = TanhGaussianActor(env.obs_dim, env.act_dim).cuda()
    d = GailDiscriminator(env.obs_dim, env.act_dim).cuda() if disc else None
    ring = StepRingReplay(env)
    col = RolloutCollector(env, FusedActor(actor) if fused else actor,
                           ring, discriminator=FusedGailReward(d, env.obs_dim, env.act_dim) if disc else None,
                           w_env=0.3, w_gail=0.7, deterministic=True)
    col.reset()
    hist = []
    for t in range(T):
        prev = env.obs[col.cur].clone()
        col.collect(1)
        r = col.cur
        assert r == (t + 1) % K
        done = (env.terminated[r] | env.truncated[r]).bool()
        rew = env.reward[r].float().clone()
        if disc:
            with torch.no_grad():
                rg = d.predict_reward(prev, ring.act[r].float()).reshape(-1)
            rew_mixed = 0.3 * rew + 0.7 * rg
        else:
            rew_mixed = rew
        hist.append(dict(obs=prev, act=ring.act[r].float().clone(), rew=rew_mixed,
                         next=torch.where(done[:, None], env.final_obs[r], env.obs[r]).clone(), done=done.clone()))
    assert len(ring) == (K - 1) * n and ring.capacity == (K - 1) * n
    assert any(h["done"].any() for h in hist[-(K - 1):]), "no episode ended inside the ring window"
    g = torch.Generator(device="cuda").manual_seed(7)
    # the sampler draws (steps back, env) pairs: replay its draws and compare every field
    g2 = torch.Generator(device="cuda").manual_seed(7)
    batch = ring.sample(4096, generator=g)
    back = torch.randint(0, K - 1, (4096,), device="cuda", generator=g2)
    e = torch.randint(0, n, (4096,), device="cuda", generator=g2)
    for b in range(K - 1):
        m = back == b
        h = hist[T - 1 - b]
        assert torch.equal(batch["observations"][m], h["obs"][e[m]])
        assert torch.equal(batch["next_observations"][m], h["next"][e[m]])
        assert torch.equal(batch["actions"][m], h["act"][e[m]])
        assert torch.equal(batch["dones"][m], h["done"][e[m]])
        assert torch.allclose(batch["rewards"][m], h["rew"][e[m]], rtol=2e-2 if disc else 0, atol=2e-2 if disc else 0)
    # graph replay of one trip through the ring continues the same rollout
    replay = col.capture()
    assert col.steps_per_replay == K
    before = ring.steps
    replay(); replay()
    torch.cuda.synchronize()
    assert ring.steps == before + 2 * K and len(ring) == (K - 1) * n
    s = ring.sample(256)
    assert s["observations"].shape == (256, env.obs_dim) and torch.isfinite(s["rewards"]).all()


@pytest.mark.parametrize("dtype,fused", [("float64", True), ("float32", True), ("float32", False)])
def test_model_b_rollout_with_step_ring(salp_lib, dtype, fused):
    """Model (B) in the same loop: SAC actor obs 6 -> 256 -> 256 -> 2*3 with SB3's Box rescale (train_robot.py:36-49,
    salp_robot_env.py:51-56) -> salp_cycle_step into ring row j % K.  A second env stepped by hand with the ring's
    actions must produce the same rows bit for bit, and the sampler must return exactly those transitions."""
    from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv
    from underwater_swimmer_rl_b200.rollout import FusedActor, RolloutCollector, StepRingReplay, TanhGaussianActor
    torch.manual_seed(2)
    n, K, T = 512, 4, 11
    env = SalpCycleVectorEnv(n, device="cuda:0", dtype=dtype, rows=K, seed=5)
    shadow = SalpCycleVectorEnv(n, device="cuda:0", dtype=dtype, seed=5)
    env.params.max_cycles = shadow.params.max_cycles = 5      # episodes end inside the ring window
    actor = TanhGaussianActor(6, 3, action_low=[0.0, 0.0, -1.0], action_high=[1.0, 1.0, 1.0]).cuda()
    ring = StepRingReplay(env)
    col = RolloutCollector(env, FusedActor(actor) if fused else actor, ring, deterministic=True)
    obs0, _ = col.reset()
    s_obs, _ = shadow.reset()
    assert torch.equal(obs0, s_obs)
    lo, hi = torch.tensor([0.0, 0.0, -1.0], device="cuda"), torch.tensor([1.0, 1.0, 1.0], device="cuda")
    hist = []
    for t in range(T):
        prev = env.obs_rows[col.cur].clone()
        col.collect(1)
        r = col.cur
        assert r == (t + 1) % K
        a = ring.act[r]
        assert (a.float() >= lo).all() and (a.float() <= hi).all()
        o, rew, term, trunc, info = shadow.step(a)
        assert torch.equal(o, env.obs_rows[r]) and torch.equal(rew, env.reward_rows[r])
        assert torch.equal(term, env.terminated_rows[r]) and torch.equal(trunc, env.truncated_rows[r])
        assert torch.equal(info["n_sub"], env.n_sub_rows[r])
        done = (term | trunc).bool()
        assert torch.equal(info["final_obs"][done], env.final_obs_rows[r][done])
        hist.append(dict(obs=prev, act=a.float().clone(), rew=rew.float().clone(), done=done.clone(),
                         next=torch.where(done[:, None], env.final_obs_rows[r], env.obs_rows[r]).clone()))
    assert any(h["done"].any() for h in hist[-(K - 1):]), "no episode ended inside the ring window"
    assert len(ring) == (K - 1) * n
    g, g2 = torch.Generator(device="cuda").manual_seed(3), torch.Generator(device="cuda").manual_seed(3)
    batch = ring.sample(2048, generator=g)
    back = torch.randint(0, K - 1, (2048,), device="cuda", generator=g2)
    e = torch.randint(0, n, (2048,), device="cuda", generator=g2)
    for b in range(K - 1):
        m, h = back == b, hist[T - 1 - b]
        for key, name in (("observations", "obs"), ("next_observations", "next"), ("actions", "act"),
                          ("rewards", "rew"), ("dones", "done")):
            assert torch.equal(batch[key][m], h[name][e[m]]), key
    # one CUDA graph = one trip through the ring; the shadow env follows it step by step through the ring's actions
    replay = col.capture()
    assert col.steps_per_replay == K and col.cur == 0
    shadow.state.copy_(env.state); shadow.ctrl.copy_(env.ctrl)
    replay()
    torch.cuda.synchronize()
    for r in list(range(1, K)) + [0]:
        o, rew, term, trunc, _ = shadow.step(ring.act[r])
        assert torch.equal(o, env.obs_rows[r]) and torch.equal(rew, env.reward_rows[r])
        assert torch.equal(term, env.terminated_rows[r]) and torch.equal(trunc, env.truncated_rows[r])
    st = env.episode_stats()
    assert st["episodes"] > 0
