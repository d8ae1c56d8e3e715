"""GPU: the model (B) breathing-cycle kernel (csrc/salp_cycle.cuh, salp_cycle_step) against the CPU oracle
(oracle/salp_cycle_oracle.c) on the same seeds and action sequences, and against the committed reference
trajectories (tests/golden/cycle_*.npz, generated from the unmodified Python reference).

Bar (BASELINE.json north_star): discrete state bit-exact in the fp64 validation build, continuous state within
1e-6 (fp64) / 2e-3 (fp32).  The fp64 floor is the reference's own: its nozzle IK carries ~2e-8 of libm-dependent
rounding noise (see oracle/salp_cycle_oracle.c)."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _words(seeds, n=256):
    from underwater_swimmer_rl_b200 import rng_tables
    assert list(seeds) == list(range(seeds[0], seeds[0] + len(seeds)))
    return rng_tables.fill_words(len(seeds), n, seeds[0], 0, rng_tables.JITTER_STREAM)


def _gpu_rollout(dtype, words, acts, **kw):
    from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv
    T, N, _ = acts.shape
    env = SalpCycleVectorEnv(N, dtype=dtype, **kw)
    env.set_rng_table(words)
    obs0, _ = env.reset()
    out = dict(obs0=obs0.cpu().numpy().copy(), obs=[], reward=[], terminated=[], truncated=[], n_sub=[], final_obs=[])
    a = torch.from_numpy(acts).cuda()
    for t in range(T):
        o, r, te, tr, info = env.step(a[t])
        out["obs"].append(o.cpu().numpy().copy()); out["reward"].append(r.cpu().numpy().astype(np.float64))
        out["terminated"].append(te.cpu().numpy().copy()); out["truncated"].append(tr.cpu().numpy().copy())
        out["n_sub"].append(info["n_sub"].cpu().numpy().copy()); out["final_obs"].append(info["final_obs"].cpu().numpy().copy())
    res = {k: (np.stack(v) if isinstance(v, list) else v) for k, v in out.items()}
    res["env"] = env
    return res


def _oracle_rollout(words, acts, **params):
    from oracle.salp_cycle_oracle import CycleOracle
    orc = CycleOracle(acts.shape[1], words, **params)
    o0 = orc.reset()
    out = orc.rollout(acts, autoreset=True)
    out["obs0"] = o0
    return out


def _relerr(a, b):
    return float((np.abs(a.astype(np.float64) - b) / (1.0 + np.abs(b))).max())


FP64_CASES = [
    # style, T, N, env keywords, which way episodes must end
    ("uniform", 10, 512, {}, None),
    ("short", 40, 300, {}, None),
    ("seek", 16, 256, {}, None),
    ("tiny", 1005, 96, {}, "truncated"),                                  # 500-cycle truncation, twice
    ("short", 30, 200, dict(width=102, height=102), "terminated"),        # target within 1 cm: reached
    ("uniform", 12, 200, dict(width=3000, height=2400), "truncated"),     # targets > 5 m away: out of range
    ("uniform", 25, 4096, {}, None),                                      # 7e7 integrator substeps
    ("seek", 20, 300, dict(width=500, height=400, robot=dict(nozzle_area=0.00009, nozzle_length1=0.01, nozzle_length2=0.01,
                                                             nozzle_length3=0.01)), None),
    ("uniform", 12, 300, dict(robot=dict(dry_mass=2.5, init_length=0.4, init_width=0.12, max_contraction=0.05, density=1025.0,
                                         nozzle_mass=0.3, nozzle_area=0.0003)), None),
]


def _split_kw(kw):
    """Env keywords of a case -> (oracle parameters, SalpCycleVectorEnv keywords)."""
    from underwater_swimmer_rl_b200.cycle_env import Nozzle, Robot
    kw = dict(kw)
    rp = kw.pop("robot", None)
    okw, gkw = dict(kw), dict(kw)
    if rp:
        okw.update(rp)
        nz = Nozzle(rp.get("nozzle_length1", 0.05), rp.get("nozzle_length2", 0.05), rp.get("nozzle_length3", 0.05),
                    rp.get("nozzle_area", 0.00016), rp.get("nozzle_mass", 1.0))
        rb = Robot(dry_mass=rp.get("dry_mass", 1.0), init_length=rp.get("init_length", 0.3), init_width=rp.get("init_width", 0.15),
                   max_contraction=rp.get("max_contraction", 0.06), nozzle=nz)
        rb.set_environment(rp.get("density", 1000.0))
        gkw["robot"] = rb
    return okw, gkw


@pytest.mark.parametrize("style,T,N,kw,ends", FP64_CASES)
def test_fp64_vs_oracle(salp_lib, style, T, N, kw, ends):
    from oracle.validate_cycle_against_reference import make_cycle_actions
    acts = make_cycle_actions(T, N, style, seed=99)
    seeds = list(range(31_000, 31_000 + N))
    words = _words(seeds)
    okw, gkw = _split_kw(kw)
    ref = _oracle_rollout(words, acts, **okw)
    got = _gpu_rollout("float64", words, acts, **gkw)
    assert np.array_equal(got["obs0"], ref["obs0"])
    for k in ("n_sub", "terminated", "truncated"):
        assert np.array_equal(got[k], ref[k]), k
    assert _relerr(got["obs"], ref["obs"]) <= 1e-6
    assert _relerr(got["reward"], ref["reward"]) <= 1e-6
    done = (ref["terminated"] | ref["truncated"]) != 0
    if ends:
        assert ref[ends].any(), f"this case must end episodes by {ends}"
    if done.any():
        assert _relerr(got["final_obs"][done], ref["final_obs"][done]) <= 1e-6
    st = got["env"].episode_stats()
    assert st["env_steps"] == T * N and st["substeps"] == float(ref["n_sub"].sum())
    assert st["episodes"] == float(done.sum()) and st["nonfinite"] == 0


@pytest.mark.parametrize("style,T,N", [("uniform", 10, 512), ("seek", 16, 256)])
def test_fp32_vs_oracle(salp_lib, style, T, N):
    from oracle.validate_cycle_against_reference import make_cycle_actions
    acts = make_cycle_actions(T, N, style, seed=7).astype(np.float32)
    seeds = list(range(32_000, 32_000 + N))
    words = _words(seeds)
    ref = _oracle_rollout(words, acts.astype(np.float64))
    got = _gpu_rollout("float32", words, acts)
    assert np.array_equal(got["obs0"], ref["obs0"])
    assert np.array_equal(got["n_sub"], ref["n_sub"]), "substep counts must be exact in the fp32 build too"
    assert np.array_equal(got["terminated"], ref["terminated"]) and np.array_equal(got["truncated"], ref["truncated"])
    err = _relerr(got["obs"], ref["obs"])
    print(f"fp32 {style}: max obs err over {T} cycles {err:.2e}")
    assert err <= 2e-3
    assert _relerr(got["reward"], ref["reward"]) <= 2e-2   # 100 x (difference of two distances) in float


def test_golden_reference_trajectories(salp_lib):
    """fp64 build against trajectories recorded from the unmodified Python reference."""
    files = sorted(glob.glob(os.path.join(GOLDEN, "cycle_*.npz")))
    assert files, "tests/golden/cycle_*.npz missing"
    from underwater_swimmer_rl_b200.cycle_env import Nozzle, Robot
    for f in files:
        g = np.load(f, allow_pickle=False)
        import json
        meta = json.loads(str(g["meta"]))
        pr = meta["params"]
        robot = Robot(nozzle=Nozzle(pr.get("nozzle_length1", 0.05), pr.get("nozzle_length2", 0.05), pr.get("nozzle_length3", 0.05),
                                    pr.get("nozzle_area", 0.00016), 1.0))
        kw = dict(robot=robot, width=pr.get("width", 900), height=pr.get("height", 700))
        seeds = meta["seeds"]
        got = _gpu_rollout("float64", _words(seeds), g["actions"], **kw)
        assert np.array_equal(got["obs0"], g["obs0"]), f
        for k in ("n_sub", "terminated", "truncated"):
            assert np.array_equal(got[k], g[k]), (f, k)
        assert _relerr(got["obs"], g["obs"]) <= 1e-6, f
        assert _relerr(got["reward"], g["reward"]) <= 1e-6, f


def test_masked_reset_and_no_autoreset(salp_lib):
    from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv
    N = 64
    env = SalpCycleVectorEnv(N, dtype="float64", autoreset=False, seed=500)
    env.reset()
    a = torch.tensor([[1.0, 0.0, 0.0]], device="cuda", dtype=torch.float64).repeat(N, 1)
    for _ in range(3):
        env.step(a)
    pos = env.position.clone()
    assert (env.cycle == 3).all() and pos[:, 0].abs().min() > 1e-3
    mask = torch.zeros(N, dtype=torch.uint8, device="cuda"); mask[::2] = 1
    tgt = env.target_point.clone()
    env.reset(mask=mask)
    assert (env.cycle[::2] == 0).all() and (env.cycle[1::2] == 3).all()
    assert torch.equal(env.position[1::2], pos[1::2]) and (env.position[::2] == 0).all()
    assert torch.equal(env.target_point[1::2], tgt[1::2]) and not torch.equal(env.target_point[::2], tgt[::2])


def test_bad_arguments_are_reported(salp_lib):
    import ctypes as C
    from underwater_swimmer_rl_b200 import _capi
    from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv
    env = SalpCycleVectorEnv(8)
    with pytest.raises(ValueError):
        env.step(torch.zeros(8, 2, device="cuda"))
    rc = salp_lib.salp_cycle_step(C.byref(env.params), C.byref(env._buffers), None, None)
    assert rc == _capi.SALP_EINVAL and b"actions" in salp_lib.salp_last_error()
    p = _capi.SalpCycleParams.from_buffer_copy(env.params); p.nozzle_area = 0.0
    assert salp_lib.salp_cycle_reset(C.byref(p), C.byref(env._buffers), None, 1, None) == _capi.SALP_EINVAL


def test_throughput_print(salp_lib):
    """Not a benchmark (bench.py --workload robot_cycle_256k is): prints the substep rate of both builds."""
    from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv
    for dtype, N in (("float32", 262144), ("float64", 65536)):
        env = SalpCycleVectorEnv(N, dtype=dtype, table_rows=4096)
        env.reset()
        g = torch.Generator(device="cuda").manual_seed(1)
        a = torch.rand(N, 3, device="cuda", generator=g, dtype=env.real)
        a[:, 2] = a[:, 2] * 2 - 1
        env.step(a); torch.cuda.synchronize()
        env.episode_stats(reset=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); env.step(a); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1); st = env.episode_stats()
        print(f"{dtype}: {N} envs, {st['substeps'] / N:.0f} substeps/env, {ms:.2f} ms/step, "
              f"{st['substeps'] / ms / 1e6:.2f} G substeps/s, {N / ms / 1e3:.2f} M env-steps/s")


def test_slab_sharding_is_rank_invariant(salp_lib):
    """Two slabs (env_id0 = 0 and 96) reproduce one 200-env run: RNG rows are indexed by global env id."""
    from oracle.validate_cycle_against_reference import make_cycle_actions
    from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv
    N, T, split = 200, 6, 96
    acts = torch.from_numpy(make_cycle_actions(T, N, "short", seed=3)).cuda()
    whole = SalpCycleVectorEnv(N, dtype="float64", seed=777, width=3000, height=2400)   # frequent out-of-range resets
    parts = [SalpCycleVectorEnv(split, dtype="float64", seed=777, env_id0=0, width=3000, height=2400),
             SalpCycleVectorEnv(N - split, dtype="float64", seed=777, env_id0=split, width=3000, height=2400)]
    whole.reset()
    for p in parts:
        p.reset()
    n_reset = 0
    for t in range(T):
        o, r, te, tr, _ = whole.step(acts[t])
        o0, r0, te0, tr0, _ = parts[0].step(acts[t, :split])
        o1, r1, te1, tr1, _ = parts[1].step(acts[t, split:])
        assert torch.equal(o, torch.cat([o0, o1])) and torch.equal(r, torch.cat([r0, r1]))
        assert torch.equal(te, torch.cat([te0, te1])) and torch.equal(tr, torch.cat([tr0, tr1]))
        n_reset += int((te | tr).sum())
    assert n_reset > 0
    assert torch.equal(whole.target_point, torch.cat([parts[0].target_point, parts[1].target_point]))


def test_reset_seed_replays_the_stream(salp_lib):
    from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv
    env = SalpCycleVectorEnv(32, dtype="float32", seed=1)
    o1, _ = env.reset(seed=42)
    t1 = env.target_point.clone()
    a = torch.rand(32, 3, device="cuda")
    env.step(a)
    o2, _ = env.reset(seed=42)
    assert torch.equal(t1, env.target_point) and torch.equal(o1, o2)
    env.reset(seed=43)
    assert not torch.equal(t1, env.target_point)
    # target points are np.random.uniform draws inside the tank (salp_robot_env.py:274-282): check against numpy
    rs = np.random.RandomState(43)
    assert np.allclose(env.target_point[0].cpu().numpy(), [np.float32(rs.uniform(-2.0, 2.0)), np.float32(rs.uniform(-1.5, 1.5))], rtol=0, atol=0)
    assert env.single_action_space.shape == (3,) and env.single_observation_space.shape == (6,)


def test_sb3_vecenv_protocol_model_b(salp_lib):
    """train_robot.py:20-25 hands SB3 a 4-env VecEnv of SalpRobotEnv: same protocol over the CUDA env."""
    from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv
    from underwater_swimmer_rl_b200.sb3_adapter import SalpCycleSB3VecEnv
    N = 4
    venv = SalpCycleSB3VecEnv(SalpCycleVectorEnv(N, seed=5, width=3000, height=2400))   # far targets: episodes end quickly
    obs = venv.reset()
    assert obs.shape == (N, 6) and obs.dtype == np.float32
    g = np.random.default_rng(0)
    n_done = 0
    for t in range(12):
        a = np.stack([g.uniform(0, 1, N), g.uniform(0, 0.2, N), g.uniform(-1, 1, N)], axis=1).astype(np.float32)
        obs, rews, dones, infos = venv.step(a)
        assert rews.dtype == np.float32 and dones.dtype == bool and len(infos) == N
        for i in np.nonzero(dones)[0]:
            n_done += 1
            assert infos[i]["terminal_observation"].shape == (6,) and infos[i]["episode"]["l"] >= 1
            assert obs[i, 2] == 0.0 and obs[i, 4] == 0.0      # auto-reset: robot at rest in the new episode
    assert n_done > 0 and len(venv.get_attr("cycle")) == N
    venv.close()


def test_no_write_outside_the_bound_buffers(salp_lib):
    """compute-sanitizer is not available on the GPU pool: every buffer handed to salp_cycle_reset / salp_cycle_step sits
    between two canary bands; after resets and steps of an odd-sized slab the bands must be untouched and the results must
    equal those of the regular env."""
    import ctypes as C
    from underwater_swimmer_rl_b200 import _capi, rng_tables
    from underwater_swimmer_rl_b200.cycle_env import SalpCycleVectorEnv, default_robot, make_cycle_params
    N, PAD = 1003, 4096                       # bytes of canary on both sides (multiple of every alignment)
    dev = torch.device("cuda:0")
    words = rng_tables.fill_words(N, 64, 900, 0, rng_tables.JITTER_STREAM)

    def guarded(nbytes):
        total = PAD + ((nbytes + 255) // 256) * 256 + PAD
        raw = torch.full((total,), 0xA5, dtype=torch.uint8, device=dev)
        return raw, raw[PAD:PAD + nbytes]

    for dtype, code, rb in ((torch.float32, _capi.SALP_F32, 4), (torch.float64, _capi.SALP_F64, 8)):
        sizes = dict(state=6 * N * 4 * rb, ctrl=N * 16, obs=N * 24, reward=N * rb, terminated=N, truncated=N, n_sub=N * 4,
                     final_obs=N * 24, stats=64, actions=N * 3 * rb)
        bufs = {k: guarded(v) for k, v in sizes.items()}
        for k in ("state", "ctrl", "stats"):
            bufs[k][1].zero_()
        w = torch.from_numpy(words.view(np.int32)).to(dev)
        b = _capi.SalpCycleBuffers()
        for k in ("state", "ctrl", "obs", "reward", "terminated", "truncated", "n_sub", "final_obs", "stats"):
            setattr(b, k, bufs[k][1].data_ptr())
        b.words, b.table_rows, b.words_len, b.n, b.env_id0, b.dtype = w.data_ptr(), N, 64, N, 0, code
        p = make_cycle_params(default_robot(), 3000, 2400, True)     # far targets: auto-resets happen inside the steps
        ref = SalpCycleVectorEnv(N, dtype=dtype, width=3000, height=2400)
        ref.set_rng_table(words)
        ref.reset()
        _capi.check(salp_lib.salp_cycle_reset(C.byref(p), C.byref(b), None, 1, None))
        g = torch.Generator(device="cuda").manual_seed(3)
        for t in range(5):
            a = torch.rand(N, 3, device=dev, dtype=dtype, generator=g)
            a[:, 2] = a[:, 2] * 2 - 1
            bufs["actions"][1].copy_(a.contiguous().view(torch.uint8).reshape(-1))
            _capi.check(salp_lib.salp_cycle_step(C.byref(p), C.byref(b), C.c_void_p(bufs["actions"][1].data_ptr()), None))
            ref.step(a)
        torch.cuda.synchronize()
        for k, (raw, view) in bufs.items():
            n = view.numel()
            assert bool((raw[:PAD] == 0xA5).all()) and bool((raw[PAD + n:] == 0xA5).all()), f"{k}: canary overwritten ({dtype})"
        assert torch.equal(bufs["obs"][1].view(torch.float32).reshape(N, 6), ref.obs)
        assert torch.equal(bufs["state"][1].view(dtype).reshape(6, N, 4), ref.state)
        assert torch.equal(bufs["n_sub"][1].view(torch.int32), ref.n_sub)
        assert int(ref.terminated.sum() + ref.truncated.sum()) >= 0
