"""CPU: host-side logic that needs no GPU -- config loader (mirror of the reference's
src/salp/config/config_loader.py), spaces, slab partitioning, algorithmic-byte accounting."""
import os

import numpy as np
import pytest

from helpers import CONFIGS

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("name", ["single_food", "single_food_long_horizon", "defaults", "sac_gail"])
def test_load_config_presets(name):
    from underwater_swimmer_rl_b200.config import env_kwargs, load_config
    cfg = load_config(name, config_dir=os.path.join(ROOT, "configs"))
    assert cfg.environment["type"] == "salp_snake"
    assert env_kwargs(cfg) == CONFIGS[name]
    kw = env_kwargs(cfg, with_size=True)
    assert kw["width"] == 800 and kw["height"] == 600
    assert cfg.get("agent", "hidden_sizes") == [256, 256]


def test_load_config_overrides_and_fallback(tmp_path):
    from underwater_swimmer_rl_b200.config import load_config
    cfgdir = os.path.join(ROOT, "configs")
    cfg = load_config("single_food", config_dir=cfgdir, environment={"params": {"num_food_items": 3}})
    assert cfg.environment["params"]["num_food_items"] == 3 and cfg.environment["params"]["food_reward"] == 1000.0
    assert load_config("no_such_preset", config_dir=cfgdir).environment["name"] == "salp_snake"  # falls back to defaults
    p = tmp_path / "x.yaml"
    p.write_text("environment:\n  params:\n    num_food_items: 2\n")
    assert load_config(str(p)).environment["params"]["num_food_items"] == 2
    with pytest.raises(FileNotFoundError):
        from underwater_swimmer_rl_b200.config import Config
        Config.from_yaml(str(tmp_path / "missing.yaml"))


def test_spaces_match_reference_shapes(salp_lib):
    import ctypes as C

    from underwater_swimmer_rl_b200 import _capi
    from underwater_swimmer_rl_b200.vector_env import make_params, single_spaces
    for forced, adim in ((True, 1), (False, 2)):
        p = make_params(True, forced_breathing=forced)
        lay = _capi.SalpLayout()
        assert salp_lib.salp_query_layout(C.byref(p), _capi.SALP_F32, C.byref(lay)) == 0
        obs, act = single_spaces(p, lay)
        assert obs.shape == (24,) and obs.dtype == np.float32 and act.shape == (adim,)
        assert obs.low[4] == np.float32(-np.pi) and obs.high[6] == 2.0 and obs.low[10] == -1
    assert act.low.tolist() == [0.0, -1.0] and act.high.tolist() == [1.0, 1.0]


def test_slab_partition_covers_every_env_once():
    from underwater_swimmer_rl_b200.sharding import slab_for_rank
    for total, world in ((1 << 20, 8), (1000, 3), (7, 8), (262144, 4)):
        slabs = [slab_for_rank(total, r, world) for r in range(world)]
        assert slabs[0][0] == 0 and sum(n for _, n in slabs) == total
        for (s0, n0), (s1, _) in zip(slabs, slabs[1:]):
            assert s0 + n0 == s1
    with pytest.raises(ValueError):
        slab_for_rank(10, 3, 3)


def test_algorithmic_bytes_match_survey():
    import bench
    assert bench.algorithmic_bytes_per_env_step(1, 4, 1) == 250.0
    assert bench.algorithmic_bytes_per_env_step(1, 4, 8) == 124.0
    assert bench.algorithmic_bytes_per_env_step(12, 4, 1) == 426.0
    assert bench.algorithmic_bytes_per_env_step(12, 4, 8) == 146.0
