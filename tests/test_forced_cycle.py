"""The forced-breathing cycle table (salp_forced_cycle, include/salp_b200.h) against the oracle's recurrences (CPU).

Forced breathing (`forced_breathing=True`, every reference YAML; scripts/utilities/salp_robot.py:158-167) makes phase, timer,
ellipse axes, water volume and the thrust window functions of the position in a 273-step cycle.  The f32
forced-breathing kernel reads them from a host-built table; here the table's double sources are checked BIT FOR BIT against
the oracle, which runs the reference's update (salp_robot.py:184-259) step by step.  tests/test_gpu_forced_cycle.py checks
the same table against the fp64 validation kernel on the GPU.
"""
import ctypes as C

import numpy as np

from helpers import CONFIGS, py_tables


def forced_cycle(salp_lib, **params):
    from underwater_swimmer_rl_b200 import vector_env
    p = vector_env.make_params(True, **params)
    wv, a, b = (np.zeros(274) for _ in range(3))
    fsm, touch = np.zeros(274, np.int32), np.zeros(274, np.int32)
    rc = salp_lib.salp_forced_cycle(C.byref(p), *(x.ctypes.data_as(C.c_void_p) for x in (wv, a, b, fsm, touch)))
    assert rc == 0
    return dict(wv=wv, a=a, b=b, fsm=fsm, touch=touch, phase=fsm & 3, timer=(fsm >> 2) & 0xFF, exd=(fsm >> 10) & 0xFF)


def test_cycle_table_equals_the_oracle_recurrences_bit_for_bit(salp_lib, oracle_lib):
    t = forced_cycle(salp_lib, **CONFIGS["single_food"])
    fw, jw = py_tables([123], 64, 2048)
    orc = oracle_lib.Oracle(1, fw, jw, **dict(CONFIGS["single_food"], max_steps_without_food=10 ** 6))
    orc.reset()
    e = orc.envs
    # index 0 = fresh reset: a = b = 30 (salp_robot.py:111-113), rest, water 0
    assert (t["a"][0], t["b"][0], t["wv"][0], t["phase"][0], t["timer"][0]) == (30.0, 30.0, 0.0, 0, 0)
    thrust_steps = 0
    for step in range(1, 3 * 273 + 1):
        cur0 = int(e["cur_jit"][0])
        orc.step(np.zeros((1, 1)))
        i = (step - 1) % 273 + 1
        assert e["phase"][0] == t["phase"][i] and e["timer"][0] == t["timer"][i], (step, i)
        for k, ok in (("wv", "wv"), ("a", "a"), ("b", "b")):
            assert e[ok][0].tobytes() == t[k][i].tobytes(), (step, k, e[ok][0], t[k][i])
        if step > 273:  # steady state: the exhale duration of the previous cycle stays in the fsm word
            assert t["exd"][i] == 150
        thrust_steps += int(e["cur_jit"][0]) != cur0
        # re-centre the robot so that no wall ends the episode
        e["px"][0], e["py"][0], e["vx"][0], e["vy"][0] = 400.0, 300.0, 0.0, 0.0
    assert thrust_steps == 3 * 61


def test_known_answers(salp_lib):
    """wv after exhale step 15 / 75 / 150; period 273 = 1 + 121 + 151; D-LUT irrelevant in forced mode (150)."""
    t = forced_cycle(salp_lib, **CONFIGS["single_food"])
    assert t["wv"][122 + 15] == 0.4364539654671127
    assert t["wv"][122 + 75] == 7.151666631483946e-11
    assert t["wv"][122 + 150] == 0.0 and t["wv"][273] == 0.0
    phases = t["phase"][1:]
    assert (phases == 1).sum() == 121 and (phases == 2).sum() == 151 and (phases == 0).sum() == 1
    assert t["phase"][273] == 0 and t["phase"][1] == 1 and t["timer"][1] == 0


def test_wall_touch_bits_are_the_reference_rounding(salp_lib):
    """after a bounce the robot sits on 50 + r; salp_snake_env.py:225-228 then tests pos - r <= 50 (and the mirror
    on the far walls) in doubles.  The table's bits must equal that expression evaluated in numpy float64."""
    for W, H in ((800, 600), (400, 300), (1600, 900)):
        t = forced_cycle(salp_lib, **dict(CONFIGS["single_food"], width=W, height=H))
        r = np.maximum(t["a"], t["b"])
        m = 50.0 + r
        lo = (m - r) <= 50.0
        right = ((W - m) + r) >= (W - 50.0)
        bottom = ((H - m) + r) >= (H - 50.0)
        assert ((t["touch"] & 1) != 0).tolist() == lo.tolist()
        assert ((t["touch"] & 2) != 0).tolist() == right.tolist()
        assert ((t["touch"] & 4) != 0).tolist() == bottom.tolist()
        assert 0 < lo.sum() < 274  # both outcomes occur (~5 % of the touches bounce and survive)


def test_sqrt_free_min_food_distance_threshold():
    """csrc/salp_step.cuh closer_than_min_food_dist<double>: sqrt(x) < 80 == x < 6400 - 2^-40 for every double (correctly rounded
    sqrt), so the fp64 build's pair tests of the food layout need no square root and keep the reference's verdicts."""
    t = float.fromhex("0x1.8ffffffffffffp+12")
    assert t == np.nextafter(6400.0, 0.0)
    x = np.float64(6400.0)
    near = [x]
    for _ in range(64):
        x = np.nextafter(x, 0.0)
        near.append(x)
    x = np.float64(6400.0)
    for _ in range(64):
        x = np.nextafter(x, np.inf)
        near.append(x)
    rng = np.random.default_rng(7)
    v = np.concatenate([np.array(near), rng.uniform(0.0, 1.0e6, 1_000_000), 6400.0 + rng.uniform(-1e-8, 1e-8, 1_000_000),
                        np.array([0.0, 5e-324, 1e-300, 1e300])])
    assert ((np.sqrt(v) < 80.0) == (v < t)).all()
