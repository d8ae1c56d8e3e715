"""GPU parity tests: the CUDA path (through the C ABI, via SalpVectorEnv) against
 (1) the committed reference trajectories (tests/golden/), and
 (2) the C oracle on the same seeded inputs, at BASELINE.json configs[1] size (4096 envs, 1000 steps).

Bars (BASELINE.json north_star):
  fp64 validation build: discrete state (flags, breathing phase/timer, counters, food liveness,
      RNG cursors) BIT-EXACT; continuous state within 1e-9 abs (stated bar: 1e-5) -- the only
      source of difference is CUDA libdevice vs glibc sin/cos/atan2 (<= 2 ulp) and pow(x,2) vs x*x.
  observations (f32): |err| <= 2e-6 (one f32 rounding of near-identical doubles).
  fp32 throughput build: per-step error from synced states <= 1e-3 (normalised obs units).
"""
import numpy as np
import pytest
import torch

from helpers import CONFIGS, bits_equal, golden_cases, load_golden, py_tables, synth_actions

pytestmark = pytest.mark.gpu

OBS_ATOL_F64 = 2e-6
STATE_ATOL_F64 = 1e-9


def _env(n, params, dtype="float64", **kw):
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    return SalpVectorEnv(n, device="cuda:0", dtype=dtype, **kw, **params)


def _unpack_state(env):
    body = env.body.cpu().numpy()
    ctrl = env.ctrl.cpu().numpy()
    a, b = env.ellipse_axes()
    return dict(px=body[0, :, 0], py=body[0, :, 1], vx=body[0, :, 2], vy=body[0, :, 3], angle=body[1, :, 0],
                omega=body[1, :, 1], nozzle=body[1, :, 2], wv=body[1, :, 3], a=a.cpu().numpy(), b=b.cpu().numpy(),
                phase=ctrl[0, :, 0] & 3, timer=(ctrl[0, :, 0] >> 2) & 0xFF, steps_since_food=ctrl[0, :, 1],
                food_collected=ctrl[0, :, 2], mask=ctrl[0, :, 3].astype(np.uint32), cur_food=ctrl[1, :, 0],
                cur_jit=ctrl[1, :, 1], n_food=(ctrl[1, :, 2] >> 26) & 0x3F)


@pytest.mark.parametrize("case", golden_cases())
def test_fp64_kernel_reproduces_reference_trajectory(case):
    g = load_golden(case)
    T, N = g["actions"].shape[:2]
    fw, jw = py_tables(g["seeds"], 2048, 2048)
    env = _env(N, g["params"], autoreset=True)
    env.set_rng_tables(fw, jw)
    obs0, _ = env.reset()
    np.testing.assert_allclose(obs0.cpu().numpy(), g["reset_obs0"], atol=OBS_ATOL_F64, rtol=0)
    F = g["alive"].shape[2]
    sub = {int(s): i for i, s in enumerate(g["sub_steps"])}
    acts = torch.from_numpy(g["actions"].astype(np.float64)).cuda()
    worst_obs = worst_state = 0.0
    for t in range(T):
        obs, rew, term, trunc, info = env.step(acts[t])
        ev = info["events"].cpu().numpy()
        assert ((ev & 7) == g["flags"][t]).all(), f"{case}: flags differ at step {t}"
        assert (term.cpu().numpy() == (g["flags"][t] & 1)).all() and (trunc.cpu().numpy() == ((g["flags"][t] >> 1) & 1)).all()
        done = (g["flags"][t] & 3) != 0
        o = obs.cpu().numpy()
        fo = info["final_observation"].cpu().numpy()
        term_obs = np.where(done[:, None], fo, o)
        worst_obs = max(worst_obs, float(np.abs(term_obs - g["obs"][t]).max()))
        if done.any():
            worst_obs = max(worst_obs, float(np.abs(o[done] - g["reset_obs"][t][done]).max()))
        np.testing.assert_allclose(rew.cpu().numpy(), g["reward"][t], atol=1e-9, rtol=1e-12)
        if t in sub or t == T - 1:
            st = _unpack_state(env)
            nd = ~done  # auto-reset lanes hold the NEW episode's state
            assert (st["phase"][nd] == g["phase"][t][nd]).all() and (st["timer"][nd] == g["timer"][t][nd]).all()
            assert (st["steps_since_food"][nd] == g["steps_since_food"][t][nd]).all()
            assert (st["food_collected"][nd] == g["food_collected"][t][nd]).all()
            assert (st["n_food"][nd] == g["n_food"][t][nd]).all()
            alive = (st["mask"][:, None] >> np.arange(F)[None, :]) & 1
            assert (alive[nd] == g["alive"][t][nd]).all()
            if t in sub:
                i = sub[t]
                for c, k in enumerate(("px", "py", "vx", "vy", "angle", "omega", "nozzle", "wv", "a", "b")):
                    worst_state = max(worst_state, float(np.abs(st[k][nd] - g["sub_state"][i, :, c][nd]).max()))
                food = env.food.cpu().numpy().transpose(1, 0, 2)[:, :F]
                al = g["alive"][t][..., None] * nd[:, None, None]
                assert bits_equal(g["sub_food"][i] * al, food * al).all(), "food positions must be bit-exact"
    st = _unpack_state(env)
    assert (st["cur_food"] == g["cur_food"]).all() and (st["cur_jit"] == g["cur_jit"]).all(), "RNG consumption"
    assert worst_obs <= OBS_ATOL_F64, worst_obs
    assert worst_state <= STATE_ATOL_F64, worst_state


def _oracle_rollout(oracle_lib, params, fw, jw, acts):
    orc = oracle_lib.Oracle(acts.shape[1], fw, jw, **params)
    o0 = orc.reset()
    out = orc.rollout(acts, autoreset=True)
    return orc, o0, out


@pytest.mark.parametrize("cfg,forced", [("single_food", True), ("defaults", True), ("single_food", False)])
def test_fp64_4096_envs_1000_steps_vs_oracle(cfg, forced, oracle_lib, salp_lib):
    """BASELINE.json configs[1]: 4096 envs, fp64 parity build, 1000 steps, K-fused launches."""
    from underwater_swimmer_rl_b200 import rng_tables
    N, T, K = 4096, 1000, 50
    params = dict(CONFIGS[cfg], forced_breathing=forced)
    acts = synth_actions(T, N, 1 if forced else 2, seed=99)
    fw = rng_tables.fill_words(N, 512, 10_000, 0, 0)
    jw = rng_tables.fill_words(N, 1024, 10_000, 0, 1)
    orc, o0, ref = _oracle_rollout(oracle_lib, params, fw, jw, acts)
    env = _env(N, params, autoreset=True, max_substeps=K)
    env.set_rng_tables(fw, jw)
    obs0, _ = env.reset()
    np.testing.assert_allclose(obs0.cpu().numpy(), o0, atol=OBS_ATOL_F64, rtol=0)
    a_dev = torch.from_numpy(acts.astype(np.float64)).cuda()
    worst_obs = worst_rew = 0.0
    n_done = 0
    # float32 observation values that are bit-identical to the oracle's, over the columns that involve no libm call
    # (position, velocity, radius, phase, water volume, nozzle, food offsets and distances, count, mean distance): the fp64
    # build scales them with host reciprocals instead of dividing (<= 1 ulp(double) before the cast, csrc/salp_step.cuh divc)
    plain_cols = [c for c in range(o0.shape[1]) if c not in (4, 13, 17, 21)]
    same_bits = total_vals = 0
    for t0 in range(0, T, K):
        obs, rew, term, trunc, info = env.step_many(a_dev[t0:t0 + K])
        ev = info["events"].cpu().numpy()
        rf = ref["flags"][t0:t0 + K]
        assert (ev == rf).all(), f"discrete events differ in steps {t0}..{t0 + K}: {int((ev != rf).sum())} lanes"
        done = (rf & 3) != 0
        n_done += int(done.sum())
        o = obs.cpu().numpy()
        fo = info["final_observation"].cpu().numpy()
        term_obs = np.where(done[..., None], fo, o)
        worst_obs = max(worst_obs, float(np.abs(term_obs - ref["obs"][t0:t0 + K]).max()))
        if t0 < 4 * K:  # early steps: the continuous state itself is still bit-identical (libm differences have not spread)
            a_ = np.ascontiguousarray(term_obs[..., plain_cols]).view(np.uint32)
            b_ = np.ascontiguousarray(ref["obs"][t0:t0 + K][..., plain_cols].astype(np.float32)).view(np.uint32)
            same_bits += int((a_ == b_).sum())
            total_vals += a_.size
        if done.any():
            worst_obs = max(worst_obs, float(np.abs(o[done] - ref["reset_obs"][t0:t0 + K][done]).max()))
        worst_rew = max(worst_rew, float(np.abs(rew.cpu().numpy() - ref["reward"][t0:t0 + K]).max()))
    st = _unpack_state(env)
    e = orc.envs
    for k in ("phase", "timer", "steps_since_food", "food_collected", "n_food"):
        assert (st[k] == e[k]).all(), k
    assert (st["cur_food"] == e["cur_food"].astype(np.int64)).all() and (st["cur_jit"] == e["cur_jit"].astype(np.int64)).all()
    F = env.food_slots
    alive = (st["mask"][:, None] >> np.arange(F)[None, :]) & 1
    assert (alive == e["alive"][:, :F]).all()
    food = env.food.cpu().numpy().transpose(1, 0, 2)
    assert bits_equal(e["food"][:, :F] * alive[..., None], food * alive[..., None]).all()
    worst_state = max(float(np.abs(st[k] - e[k]).max()) for k in ("px", "py", "vx", "vy", "angle", "omega", "nozzle", "wv", "a", "b"))
    assert n_done > 100, "the action sequence must actually end episodes"
    assert worst_obs <= OBS_ATOL_F64 and worst_rew <= 1e-8 and worst_state <= STATE_ATOL_F64, (worst_obs, worst_rew, worst_state)
    print(f"\n[fp64 parity] {cfg} forced={forced}: worst obs {worst_obs:.2e}, reward {worst_rew:.2e}, state {worst_state:.2e}; "
          f"float32 obs values bit-identical to the oracle's in the first {4 * K} steps: {same_bits}/{total_vals} = {same_bits / total_vals:.6f}")
    assert same_bits / total_vals >= 0.999, (same_bits, total_vals)
    # episode statistics accumulated on the device == the oracle's
    s = env.episode_stats()
    assert s["episodes"] == ref["stats"][0] and s["collisions"] == ref["stats"][2] and s["truncations"] == ref["stats"][3]
    assert s["env_steps"] == N * T and s["nonfinite"] == 0
    assert abs(s["return_sum"] - float(ref["reward"].sum() - _open_episode_returns(env))) <= 1e-6 * max(1.0, abs(s["return_sum"]))


def _open_episode_returns(env):
    return float(env.ep_return.sum().item())


def test_k_fusion_is_bitwise_identical_to_single_steps(salp_lib):
    """K fused substeps in one launch == K launches of one step (same arithmetic, f32 build)."""
    from underwater_swimmer_rl_b200 import rng_tables
    N, T, K = 3000, 240, 8  # N not a multiple of the CTA size on purpose
    params = CONFIGS["defaults"]
    acts = torch.from_numpy(synth_actions(T, N, 1, seed=5)).cuda()
    outs = []
    for k in (1, K):
        env = _env(N, params, dtype="float32", max_substeps=k, seed=77)
        env.reset()
        rec = []
        for t0 in range(0, T, k):
            obs, rew, term, trunc, info = env.step_many(acts[t0:t0 + k])
            rec.append((obs.clone(), rew.clone(), info["events"].clone(), info["final_observation"].clone()))
        outs.append([torch.cat([r[i] for r in rec]) for i in range(4)] + [env.body.clone(), env.ctrl.clone(), env.food.clone()])
    ev = outs[0][2]
    done = (ev & 3) != 0
    for i, (x, y) in enumerate(zip(*outs)):
        if i == 3:  # final_obs rows are only defined where an episode ended
            x, y = x[done], y[done]
        assert torch.equal(x, y), i


def test_autoreset_off_keeps_stepping_like_the_reference(oracle_lib, salp_lib):
    """autoreset=False: done lanes are NOT reset; the next step continues from the terminal state
    exactly as the reference does when its caller ignores `done`."""
    N, T = 64, 1300
    params = CONFIGS["single_food"]
    acts = synth_actions(T, N, 1, seed=3)
    fw, jw = py_tables(range(500, 500 + N), 256, 1024)
    orc = oracle_lib.Oracle(N, fw, jw, **params)
    orc.reset()
    ref = orc.rollout(acts, autoreset=False)
    env = _env(N, params, autoreset=False)
    env.set_rng_tables(fw, jw)
    env.reset()
    a = torch.from_numpy(acts.astype(np.float64)).cuda()
    for t in range(T):
        obs, rew, term, trunc, info = env.step(a[t])
        assert (info["events"].cpu().numpy() == ref["flags"][t]).all(), t
        assert float(np.abs(obs.cpu().numpy() - ref["obs"][t]).max()) <= OBS_ATOL_F64
    assert ((ref["flags"] & 3) != 0).sum() > 0


def test_fp32_per_step_error_from_synced_states(oracle_lib, salp_lib):
    """fp32 throughput build: load the oracle's fp64 state every step, advance one step on the GPU,
    compare -> single-step rounding error only.  Bar: 1e-3 (north_star); measured ~1e-6."""
    N, T = 512, 600
    params = CONFIGS["defaults"]
    acts = synth_actions(T, N, 1, seed=11)
    fw, jw = py_tables(range(9000, 9000 + N), 512, 1024)
    orc = oracle_lib.Oracle(N, fw, jw, **params)
    orc.reset()
    env = _env(N, params, dtype="float32", autoreset=False)
    env.set_rng_tables(fw, jw)
    env.reset()
    F = env.food_slots
    worst = 0.0
    flips = 0
    for t in range(T):
        e = orc.envs
        body = np.stack([np.stack([e["px"], e["py"], e["vx"], e["vy"]], 1),
                         np.stack([e["angle"], e["omega"], e["nozzle"], e["wv"]], 1)])
        env.body.copy_(torch.from_numpy(body.astype(np.float32)))
        ctrl = env.ctrl.cpu().numpy()
        fsm = e["phase"] | (e["timer"] << 2) | (e["exhale_dur"] << 10) | (e["shape_kind"] << 18) | (e["shape_timer"] << 20)
        ctrl[0, :, 0] = fsm
        ctrl[0, :, 1] = e["steps_since_food"]
        ctrl[0, :, 2] = e["food_collected"]
        mask = (e["alive"][:, :F].astype(np.int64) << np.arange(F)).sum(1)
        ctrl[0, :, 3] = np.where(mask >= 2 ** 31, mask - 2 ** 32, mask)
        ctrl[1, :, 0] = e["cur_food"]
        ctrl[1, :, 1] = e["cur_jit"]
        ctrl[1, :, 2] = (ctrl[1, :, 2] & 0x3FFFFFF) | (e["n_food"].astype(np.int64) << 26).astype(np.int32)
        env.ctrl.copy_(torch.from_numpy(ctrl))
        env.refresh_rng()  # cursors were overwritten from outside
        env.food.copy_(torch.from_numpy(e["food"][:, :F].transpose(1, 0, 2).astype(np.float32)))
        obs_ref, rew_ref, flags_ref = orc.step(acts[t].astype(np.float64))
        obs, rew, term, trunc, info = env.step(torch.from_numpy(acts[t]).cuda())
        same = info["events"].cpu().numpy() == flags_ref
        err = np.abs(obs.cpu().numpy() - obs_ref).max(axis=1)
        # a discrete divergence is either an event flag flip or two foods whose distances tie to
        # within fp32 rounding being listed in the other order (whole 4-float entries swap)
        swapped = same & (err > 1e-3)
        flips += int((~same).sum()) + int(swapped.sum())
        ok = same & ~swapped
        worst = max(worst, float(err[ok].max()))
        for i in np.nonzero(flags_ref & 3)[0]:
            orc.reset_one(int(i))
    assert worst <= 1e-3, worst
    assert flips <= 3, f"{flips} discrete flips (event flags or near-tie food order) in {N * T} fp32 steps from synced states"


def test_step_host_equals_device_path(salp_lib):
    """salp_step_host (pinned host buffers, chunked copy/compute overlap) == the device path."""
    N, T = 5000, 40
    params = CONFIGS["single_food"]
    acts = synth_actions(T, N, 1, seed=21)
    env_d = _env(N, params, dtype="float32", seed=5)
    env_h = _env(N, params, dtype="float32", seed=5)
    env_d.reset()
    env_h.reset()
    a_h = torch.empty(N, 1, dtype=torch.float32).pin_memory()
    o_h = torch.empty(N, env_h.obs_dim, dtype=torch.float32).pin_memory()
    r_h = torch.empty(N, dtype=torch.float32).pin_memory()
    te_h = torch.empty(N, dtype=torch.bool).pin_memory()
    tr_h = torch.empty(N, dtype=torch.bool).pin_memory()
    for t in range(T):
        a_h.copy_(torch.from_numpy(acts[t]))
        env_h.step_host(a_h, o_h, r_h, te_h, tr_h)
        obs, rew, term, trunc, _ = env_d.step(torch.from_numpy(acts[t]).cuda())
        assert torch.equal(o_h, obs.cpu()) and torch.equal(r_h, rew.cpu())
        assert torch.equal(te_h, term.cpu()) and torch.equal(tr_h, trunc.cpu())
    assert torch.equal(env_d.body, env_h.body) and torch.equal(env_d.ctrl, env_h.ctrl)


def test_slab_sharding_is_rank_invariant(salp_lib):
    """RNG streams are indexed by GLOBAL env id (seed + env_id0 + e), so an env's
    trajectory does not depend on which rank's slab it lives in: one 2048-env slab == two
    1024-env slabs with env_id0 = 0 / 1024, bit for bit (f32 throughput build, auto-reset on)."""
    N, T = 2048, 1400  # long enough for wall collisions, resets and food-table draws
    params = CONFIGS["single_food"]
    acts = torch.from_numpy(synth_actions(T, N, 1, seed=9)).cuda()
    whole = _env(N, params, dtype="float32", seed=4242)
    halves = [_env(N // 2, params, dtype="float32", seed=4242, env_id0=r * (N // 2)) for r in range(2)]
    o_w, _ = whole.reset()
    o_h = torch.cat([h.reset()[0] for h in halves])
    assert torch.equal(o_w, o_h)
    n_done = 0
    for t in range(T):
        ow, rw, tw, uw, _ = whole.step(acts[t])
        parts = [h.step(acts[t, r * (N // 2):(r + 1) * (N // 2)].contiguous()) for r, h in enumerate(halves)]
        assert torch.equal(ow, torch.cat([p[0] for p in parts])), t
        assert torch.equal(rw, torch.cat([p[1] for p in parts])), t
        assert torch.equal(tw, torch.cat([p[2] for p in parts])), t
        n_done += int((tw | uw).sum().item())
    assert n_done > N // 4  # episodes did end (and were reset from the tables) inside the window
    assert torch.equal(whole.body, torch.cat([h.body for h in halves], dim=1))
    assert torch.equal(whole.ctrl, torch.cat([h.ctrl for h in halves], dim=1))
    s = sum(h.stats for h in halves)
    idx = torch.tensor([0, 2, 3, 4, 5], device=s.device)  # integer-valued sums are exact in f64
    assert torch.equal(whole.stats[idx], s[idx])
    assert torch.allclose(whole.stats[1], s[1], rtol=1e-12)  # return sum: atomic order differs


EDGE_CASES = {
    # every food-slot instantiation of the kernel (1, 4, 8, 12, 16, 32) and the generic obs-row path (observed != 3)
    "two_food": dict(num_food_items=2, max_steps_without_food=400),
    "five_food_default_ctor": dict(max_steps_without_food=300),                                   # SalpSnakeEnv() defaults: 5 foods, eff. bonus 1.0
    "sixteen_food": dict(num_food_items=16, max_steps_without_food=500),
    "thirty_two_food": dict(num_food_items=32, max_steps_without_food=500, max_observed_food=8),
    "thirteen_food_observe_1": dict(num_food_items=13, max_observed_food=1, max_steps_without_food=300),
    "observe_more_than_food": dict(num_food_items=2, max_observed_food=6, max_steps_without_food=300),
    "no_food_at_all": dict(num_food_items=0, max_steps_without_food=150),
    "small_tank": dict(width=400, height=300, num_food_items=3, max_steps_without_food=300, proximity_reward_weight=2.0),
    "big_tank_manual": dict(width=1600, height=900, num_food_items=8, forced_breathing=False, max_steps_without_food=600),
    "crowded_no_respawn_random_count": dict(num_food_items=12, respawn_food=False, random_food_count=True,
                                            max_steps_without_food=500, efficiency_bonus=0.5, proximity_reward_weight=1.0),
    "timeout_zero": dict(num_food_items=1, max_steps_without_food=0),
    # twelve foods cannot keep 80 px apart in a 460 x 380 tank: the reset layout runs into the 100-attempt cap (S:100-117) and
    # spans several 32-candidate windows, respawns reject most candidates and some take the unconditional 51st draw (S:232-276)
    "crowded_tank_twelve": dict(width=460, height=380, num_food_items=12, max_steps_without_food=200, efficiency_bonus=0.5),
}


@pytest.mark.parametrize("name", sorted(EDGE_CASES))
def test_fp64_edge_cases_vs_oracle(name, oracle_lib, salp_lib):
    """Parameter corners the reference accepts (SalpSnakeEnv.__init__ keywords, S:29-47): food counts that select
    every kernel instantiation, more observed slots than foods, zero foods, other tank sizes, manual breathing,
    no respawn + random food count, immediate timeout.  fp64 build vs the oracle: events bit-exact, obs <= 2e-6."""
    from underwater_swimmer_rl_b200 import rng_tables
    params = EDGE_CASES[name]
    N, T, K = 384, 700, 35
    act_dim = 1 if params.get("forced_breathing", True) else 2
    acts = synth_actions(T, N, act_dim, seed=len(name))
    fw = rng_tables.fill_words(N, 4096, 31_000, 0, 0)
    jw = rng_tables.fill_words(N, 1024, 31_000, 0, 1)
    orc, o0, ref = _oracle_rollout(oracle_lib, params, fw, jw, acts)
    env = _env(N, params, autoreset=True, max_substeps=K)
    env.set_rng_tables(fw, jw)
    obs0, _ = env.reset()
    np.testing.assert_allclose(obs0.cpu().numpy(), o0, atol=OBS_ATOL_F64, rtol=0)
    a_dev = torch.from_numpy(acts.astype(np.float64)).cuda()
    worst = 0.0
    for t0 in range(0, T, K):
        obs, rew, term, trunc, info = env.step_many(a_dev[t0:t0 + K])
        ev = info["events"].cpu().numpy()
        rf = ref["flags"][t0:t0 + K]
        assert (ev == rf).all(), f"{name}: events differ in steps {t0}..{t0 + K}"
        done = (rf & 3) != 0
        o = obs.cpu().numpy()
        term_obs = np.where(done[..., None], info["final_observation"].cpu().numpy(), o)
        worst = max(worst, float(np.abs(term_obs - ref["obs"][t0:t0 + K]).max()))
        if done.any():
            worst = max(worst, float(np.abs(o[done] - ref["reset_obs"][t0:t0 + K][done]).max()))
        assert float(np.abs(rew.cpu().numpy() - ref["reward"][t0:t0 + K]).max()) <= 1e-8
    st, e = _unpack_state(env), orc.envs
    for k in ("phase", "timer", "steps_since_food", "food_collected", "n_food"):
        assert (st[k] == e[k]).all(), k
    assert (st["cur_food"] == e["cur_food"].astype(np.int64)).all() and (st["cur_jit"] == e["cur_jit"].astype(np.int64)).all()
    assert worst <= OBS_ATOL_F64, worst
    assert int(((ref["flags"] & 3) != 0).sum()) > 0, "the case must end episodes"


def test_back_to_back_launches_of_large_slabs_match_the_oracle(salp_lib, oracle_lib):
    """Slabs of at least one resident wave are launched with programmatic dependent launch (csrc/salp_launch.cuh): the
    CTAs of step t+1 become resident while step t drains.  200 000 envs x 12 back-to-back steps, fp32 build, against
    the CPU oracle on the same seeds (discrete events bit-exact)."""
    import torch
    from underwater_swimmer_rl_b200 import SalpVectorEnv, rng_tables
    N, T = 200_000, 12
    p = CONFIGS["single_food"]
    fw = rng_tables.fill_words(N, 64, 40_000, 0, rng_tables.FOOD_STREAM)
    jw = rng_tables.fill_words(N, 64, 40_000, 0, rng_tables.JITTER_STREAM)
    acts = synth_actions(T, N, 1, seed=11)
    orc = oracle_lib.Oracle(N, fw, jw, **p)
    orc.reset()
    ref = orc.rollout(acts, autoreset=True)
    env = SalpVectorEnv(N, dtype="float32", **p)
    env.set_rng_tables(fw, jw)
    env.reset()
    a = torch.from_numpy(acts).cuda()
    ev = []
    for t in range(T):           # no synchronisation between the launches
        _, _, _, _, info = env.step(a[t])
        ev.append(info["events"].clone())
    torch.cuda.synchronize()
    got = torch.stack(ev).cpu().numpy()
    assert (got == ref["flags"]).all()
    assert np.abs(env.obs[0].cpu().numpy() - ref["obs"][-1]).max() <= 1e-4
