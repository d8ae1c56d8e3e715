"""CPU: known-answer tests of the breathing FSM (SURVEY.md section 8(c) invariants, probed on the
unmodified reference) and of the integer restatement of its thresholds that the kernel uses."""
import numpy as np

from helpers import CONFIGS, py_tables


def _run(oracle_lib, T, action=0.0, **kw):
    fw, jw = py_tables([7], 256, 4096)
    orc = oracle_lib.Oracle(1, fw, jw, **dict(CONFIGS["single_food"], **kw))
    orc.reset()
    rec = []
    for t in range(T):
        cj = int(orc.envs["cur_jit"][0])
        obs, rew, flags = orc.step(np.array([[action]]))
        e = orc.envs[0]
        rec.append((int(e["phase"]), int(e["timer"]), float(e["wv"]), int(orc.envs["cur_jit"][0]) - cj, int(flags[0]), obs[0].copy()))
        if flags[0] & 3:
            break
    return rec, orc


def test_forced_breathing_period_273_and_61_thrust_steps(oracle_lib):
    rec, _ = _run(oracle_lib, 900)
    phases = [r[0] for r in rec]
    # cycle: 1 rest step, 121 steps in 'inhaling' (120 inhale + the transition step), 151 in 'exhaling'
    runs, cur, n = [], phases[0], 0
    for p in phases:
        if p == cur:
            n += 1
        else:
            runs.append((cur, n)); cur, n = p, 1
    assert runs[:6] == [(1, 121), (2, 151), (0, 1), (1, 121), (2, 151), (0, 1)]
    thrust = [r[3] == 2 for r in rec[:273]]
    assert sum(thrust) == 61
    # thrust exactly on exhale timers 15..75 (0.1 <= t/150 <= 0.5)
    assert [r[1] for r, th in zip(rec[:273], thrust) if th] == list(range(15, 76))


def test_water_volume_known_answers(oracle_lib):
    rec, _ = _run(oracle_lib, 300)
    ex = {r[1]: r[2] for r in rec[:273] if r[0] == 2 and r[1] > 0}
    assert ex[15] == 0.4364539654671127
    assert ex[75] == 7.151666631483946e-11
    assert 0 < ex[149] < 1e-60 and ex[150] == 0.0


def test_first_step_body_size_sequence(oracle_lib):
    """obs[6]: 1.0 at reset -> 1.3 (rest shape) -> 1.2983333 (first inhale step)."""
    fw, jw = py_tables([7], 256, 4096)
    orc = oracle_lib.Oracle(1, fw, jw, **CONFIGS["single_food"])
    assert orc.reset()[0, 6] == np.float32(1.0)
    assert orc.step(np.zeros((1, 1)))[0][0, 6] == np.float32(1.3)
    assert orc.step(np.zeros((1, 1)))[0][0, 6] == np.float32(1.2983333)


def test_zero_action_episode_hits_right_wall_at_step_1019(oracle_lib):
    rec, orc = _run(oracle_lib, 1200)
    assert len(rec) == 1019 and rec[-1][4] & 1 and rec[-1][4] & 4
    assert abs(float(orc.envs["px"][0]) - 713.88) < 0.01


def test_exhale_duration_lut_matches_python_expression():
    """int(150 * max(t/120, 0.3)) is NOT floor(5t/4): 54 @ t=44, 109 @ t=88."""
    lut = [int(150 * max(t / 120, 0.3)) for t in range(121)]
    assert lut[44] == 54 and lut[88] == 109 and lut[120] == 150 and min(lut) == 45
    assert [t for t in range(36, 121) if lut[t] != (5 * t) // 4] == [44, 88]


def test_integer_thresholds_equal_the_float_comparisons():
    """The kernel evaluates 0.1 <= t/D <= 0.5, t/D <= 1 and wv > 0.05 on integers; check the
    equivalence against IEEE double AND float for every reachable (t, D)."""
    for D in range(45, 151):
        for t in range(1, D + 2):
            for dt in (np.float64, np.float32):
                p = dt(t) / dt(D)
                assert (dt(0.1) <= p) == (10 * t >= D), (t, D, dt)
                assert (p <= dt(0.5)) == (2 * t <= D), (t, D, dt)
                assert (p <= dt(1.0)) == (t <= D), (t, D, dt)
    for t in range(0, 121):
        for dt in (np.float64, np.float32):
            assert (dt(t) / dt(120) > dt(0.05)) == (t >= 7)
