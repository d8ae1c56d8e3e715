"""CPU: replay the reference's own recorded human demonstrations
(/root/reference/data/expert_demos/human/*.pkl -> tests/golden/human_demos.npz) through the oracle.
They were recorded from SalpSnakeEnv over the legacy integrator with 12 food / forced breathing
and f64 actions.  The thrust jitter came from an UN-seeded np.random, so:
  obs columns 6..9 (FSM-derived: radius, phase, water volume, nozzle)  -> exact in f32
  obs columns 4..5 (angle, angular velocity: jitter-free)               -> exact in f32
  obs columns 0..3 (position, velocity: jitter enters velocity)         -> |err| <= 1e-4
"""
import os

import numpy as np
import pytest

from helpers import GOLDEN, py_tables


def _demos():
    z = np.load(os.path.join(GOLDEN, "human_demos.npz"))
    keys = sorted(k[: -len("_actions")] for k in z.files if k.endswith("_actions"))
    return z, keys


@pytest.mark.parametrize("idx", range(5))
def test_demo_replay(idx, oracle_lib):
    z, keys = _demos()
    key = keys[idx]
    acts, obs10 = z[key + "_actions"], z[key + "_obs10"]
    T = acts.shape[0]
    fw, jw = py_tables([4242], 1024, 8192)
    orc = oracle_lib.Oracle(1, fw, jw, num_food_items=12, forced_breathing=True)
    o0 = orc.reset()
    assert np.array_equal(o0[0, :10], z[key + "_obs10_first"])
    worst = 0.0
    for t in range(T):
        obs, _, flags = orc.step(acts[t].reshape(1, 1))
        assert np.array_equal(obs[0, 4:10], obs10[t, 4:10]), f"{key} step {t}: FSM/angle columns differ"
        worst = max(worst, float(np.abs(obs[0, :4] - obs10[t, :4]).max()))
    assert worst <= 1e-4, worst
