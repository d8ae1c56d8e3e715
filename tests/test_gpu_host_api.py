"""GPU: the host-side API around the kernel -- single-env facade, SB3 VecEnv adapter, CUDA-graph
replay, state poking, checkpoint/restore, size-independent properties at BASELINE full size."""
import numpy as np
import pytest
import torch

from helpers import CONFIGS, py_tables, synth_actions

pytestmark = pytest.mark.gpu


def test_single_env_facade_matches_oracle(oracle_lib, salp_lib):
    from underwater_swimmer_rl_b200 import SalpSnakeEnv
    params = CONFIGS["single_food"]
    env = SalpSnakeEnv(render_mode=None, seed=4321, **params)
    fw, jw = py_tables([4321], 4096, 16384)
    orc = oracle_lib.Oracle(1, fw, jw, **params)
    orc.reset()            # the reference constructor consumes one food layout ...
    o_ref = orc.reset()    # ... and the caller's reset() another
    obs, info = env.reset()
    assert info == {} and obs.dtype == np.float32 and np.abs(obs - o_ref[0]).max() <= 2e-6
    assert env.action_space.shape == (1,) and env.observation_space.shape == (24,)
    acts = synth_actions(1200, 1, 1, seed=8)
    for t in range(1200):
        o, r, d, tr, info = env.step(acts[t, 0])
        o_ref, r_ref, f_ref = orc.step(acts[t].astype(np.float64))
        assert np.abs(o - o_ref[0]).max() <= 2e-6 and abs(r - r_ref[0]) <= 1e-9
        assert (d, tr, info["collision"]) == (bool(f_ref[0] & 1), bool(f_ref[0] & 2), bool(f_ref[0] & 4))
        assert isinstance(r, float) and isinstance(d, bool)
        if d or tr:
            break
    assert set(info) == {"robot_position", "robot_velocity", "robot_angle", "nozzle_angle", "ellipse_a", "ellipse_b",
                         "body_radius", "breathing_phase", "water_volume", "breathing_timer", "score",
                         "food_collected", "steps_since_food", "collision"}
    assert info["breathing_phase"] in ("rest", "inhaling", "exhaling")
    e = orc.envs[0]
    assert abs(info["ellipse_a"] - e["a"]) < 1e-12 and abs(env.robot_pos[0] - e["px"]) < 1e-9
    # state injection the way eval/collect_navigation_data.py:76-89 does it
    env.reset()
    env.robot_pos = np.array([200.0, 150.0]); env.robot_angle = 1.0; env.food_positions = [[600.0, 450.0]]
    o, *_ = env.step(np.array([0.0]))
    assert abs(o[0] - 200.0 / 800) < 1e-6 and abs(o[4] - 1.0 / np.pi) < 1e-6 and abs(o[10] - 0.5) < 1e-6
    env.close()


def test_manual_breathing_facade_shapes(salp_lib):
    from underwater_swimmer_rl_b200 import SalpSnakeEnv
    env = SalpSnakeEnv(forced_breathing=False, num_food_items=5, seed=1)
    assert env.action_space.shape == (2,) and len(env.food_positions) == 5
    o, r, d, tr, info = env.step(np.array([1.0, 0.3], dtype=np.float32))
    assert info["breathing_phase"] == "inhaling" and o.shape == (24,)
    env.close()


def test_sb3_vecenv_protocol(salp_lib):
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    from underwater_swimmer_rl_b200.sb3_adapter import SalpSB3VecEnv
    N = 8
    venv = SalpSB3VecEnv(SalpVectorEnv(N, dtype="float32", seed=3, **CONFIGS["single_food"]))
    obs = venv.reset()
    assert obs.shape == (N, 24) and obs.dtype == np.float32
    acts = synth_actions(1600, N, 1, seed=2)
    n_done = 0
    for t in range(1600):
        venv.step_async(acts[t])
        obs, rews, dones, infos = venv.step_wait()
        assert rews.dtype == np.float32 and dones.dtype == bool and len(infos) == N
        for i in np.nonzero(dones)[0]:
            n_done += 1
            assert infos[i]["terminal_observation"].shape == (24,) and "TimeLimit.truncated" in infos[i]
            assert infos[i]["episode"]["l"] > 0
            assert obs[i, 0] == 0.5 and obs[i, 6] == 1.0  # auto-reset: first observation of the new episode
    assert n_done >= N
    assert len(venv.get_attr("steps_since_food")) == N
    venv.close()


@pytest.mark.parametrize("N,T", [(4096, 64), (200_000, 24)])   # the larger slab launches with programmatic dependent launch
def test_cuda_graph_replay_equals_eager(salp_lib, N, T):
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    p = CONFIGS["single_food_long_horizon"]
    acts = torch.from_numpy(synth_actions(T + 1, N, 1, seed=4)).cuda()
    e1, e2 = SalpVectorEnv(N, seed=9, **p), SalpVectorEnv(N, seed=9, **p)
    e1.reset(); e2.reset()
    buf = acts[0].clone()
    replay = e2.capture_graph(buf)   # performs one real step (warm-up) with buf == acts[0]
    e1.step(acts[0])
    for t in range(1, T + 1):
        e1.step(acts[t])
        buf.copy_(acts[t]); replay()
    torch.cuda.synchronize()
    assert torch.equal(e1.obs, e2.obs) and torch.equal(e1.body, e2.body) and torch.equal(e1.ctrl, e2.ctrl)


def test_state_dict_roundtrip(salp_lib):
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    N = 1000
    acts = torch.from_numpy(synth_actions(60, N, 1, seed=6)).cuda()
    env = SalpVectorEnv(N, seed=5, **CONFIGS["defaults"])
    env.reset()
    for t in range(30):
        env.step(acts[t])
    sd = env.state_dict()
    ref = [tuple(x.clone() for x in env.step(acts[t])[:4]) for t in range(30, 60)]
    env.load_state_dict(sd)
    for t, r in zip(range(30, 60), ref):
        out = env.step(acts[t])[:4]
        assert all(torch.equal(a, b) for a, b in zip(out, r))


def test_full_size_properties_1m_envs(salp_lib):
    """BASELINE full size (1,048,576 envs, fp32): properties that need no oracle --
    determinism, bounds, flag consistency, statistics bookkeeping, partial reset."""
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    N, T = 1 << 20, 1300
    p = CONFIGS["single_food"]
    gen = torch.Generator(device="cuda"); gen.manual_seed(0)
    pool = [torch.rand(N, 1, device="cuda", generator=gen) * 2 - 1 for _ in range(16)]
    finals = []
    for rep in range(2):
        env = SalpVectorEnv(N, seed=123, table_rows=1 << 16, food_table_len=128, jitter_table_len=1024, **p)
        env.reset()
        n_term = torch.zeros((), dtype=torch.int64, device="cuda"); n_trunc = torch.zeros_like(n_term)
        for t in range(T):
            obs, rew, term, trunc, info = env.step(pool[t % 16])
            n_term += term.sum(); n_trunc += trunc.sum()
            if t % 100 == 99:
                assert not (term & trunc).any()                     # never both (S:182-189)
                ev = info["events"]
                assert torch.equal((ev & 1) != 0, term) and torch.equal((ev & 2) != 0, trunc)
                assert torch.equal((ev & 4) != 0, term)             # respawn mode: terminated <=> collision
                assert torch.isfinite(obs).all() and torch.isfinite(rew).all()
                pos = env.robot_pos
                assert (pos[:, 0] >= 80 - 1e-3).all() and (pos[:, 0] <= 720 + 1e-3).all()   # 50 + min radius 30
                assert (pos[:, 1] >= 80 - 1e-3).all() and (pos[:, 1] <= 520 + 1e-3).all()
                assert (obs[:, 6] >= 1.0).all() and (obs[:, 6] <= 1.3 + 1e-6).all()
                assert (env.nozzle_angle.abs() <= np.pi / 3 + 1e-6).all()
                assert (env.breathing_phase <= 2).all() and (env.steps_since_food <= 1501).all()
        s = env.episode_stats()
        assert s["episodes"] == float(n_term + n_trunc) and s["collisions"] == float(n_term) and s["truncations"] == float(n_trunc)
        assert s["env_steps"] == float(N) * T and s["nonfinite"] == 0 and s["episodes"] > 0.5 * N
        assert abs(s["length_sum"] + float(env.episode_length.sum()) - float(N) * T) < 0.5   # every step is in exactly one episode
        finals.append((env.body.clone(), env.ctrl.clone(), env.obs.clone()))
    assert all(torch.equal(a, b) for a, b in zip(*finals)), "two identical runs must be bit-identical"
    # partial reset: only masked lanes change
    mask = torch.zeros(N, dtype=torch.bool, device="cuda"); mask[::3] = True
    before = env.body.clone()
    env.reset(mask=mask)
    assert torch.equal(env.body[:, ~mask], before[:, ~mask])
    assert (env.robot_pos[mask] == torch.tensor([400.0, 300.0], device="cuda")).all() and (env.episode_length[mask] == 0).all()


def test_salp_stats_entry_point_matches_the_bound_vector(salp_lib):
    """salp_stats: the eight accumulators copied to host memory through the C ABI equal the bound device vector;
    reset zeroes it; a handle without a stats buffer is refused."""
    import ctypes as C
    from underwater_swimmer_rl_b200 import SalpVectorEnv, _capi
    env = SalpVectorEnv(4096, device="cuda:0", dtype="float32", rng="philox", max_steps_without_food=40, **{
        k: v for k, v in CONFIGS["single_food"].items() if k != "max_steps_without_food"})
    env.reset()
    g = torch.Generator(device="cuda")
    g.manual_seed(3)
    for _ in range(120):
        env.step(torch.rand(4096, 1, device="cuda", generator=g) * 2 - 1)
    want = env.stats.cpu().numpy().copy()
    got = env.stats_host()
    assert want[0] > 0 and want[7] == 4096 * 120            # episodes ended (timeouts after 40 steps); env-steps counted
    assert list(got.values()) == want.tolist()
    again = env.stats_host(reset=True)
    assert list(again.values()) == want.tolist()
    assert float(env.stats.abs().sum()) == 0.0
    env.step(torch.zeros(4096, 1, device="cuda"))
    assert env.stats_host()["env_steps"] == 4096.0
    # no stats buffer bound -> SALP_EUNBOUND
    env._bind(omit=("stats",))
    out = (C.c_double * 8)()
    assert salp_lib.salp_stats(env._handle, out, 0, None) == _capi.SALP_EUNBOUND
    assert "stats" in salp_lib.salp_last_error().decode()
    env.step(torch.zeros(4096, 1, device="cuda"))     # stepping without a stats buffer is fine
    env._bind()
    assert env.stats_host()["env_steps"] == 4096.0   # the unbound step was not counted
