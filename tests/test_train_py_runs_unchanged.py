"""SURVEY 8(f) N1, as far as this image allows: the reference's UNMODIFIED train.py (`--sb3` path, train.py:42-85,165-177)
runs on top of this repo's SalpSnakeEnv facade and YAML configs.

What is real: /root/reference/train.py (imported from where it lies, not copied), this repo's `load_config`, configs/*.yaml
and `snake_env.SalpSnakeEnv`.  What is a stand-in: `stable_baselines3` and `gymnasium` (not installed, no network) -- a stub SAC
that checks the constructor keywords train.py passes and drives the env the way SB3's DummyVecEnv + collect_rollouts do
(reset, step with actions sampled from action_space, reset on done, info keys read) -- and the CUDA handle (no GPU in the
build container): tests/fake_backend.py computes the same calls with the CPU oracle.  tests/test_gpu_host_api.py runs the
facade's own behaviour on the real kernel.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import pytest

REF = "/root/reference"
pytestmark = [pytest.mark.reference, pytest.mark.skipif(not os.path.isfile(os.path.join(REF, "train.py")), reason="needs the reference tree")]

INFO_KEYS = {"robot_position", "robot_velocity", "robot_angle", "nozzle_angle", "ellipse_a", "ellipse_b", "body_radius",
             "breathing_phase", "water_volume", "breathing_timer", "score", "food_collected", "steps_since_food", "collision"}


def _stub_sb3(record):
    sb3 = types.ModuleType("stable_baselines3")
    common = types.ModuleType("stable_baselines3.common")
    callbacks = types.ModuleType("stable_baselines3.common.callbacks")

    class BaseCallback:
        def __init__(self, verbose=0):
            self.n_calls, self.model = 0, None

        def _on_step(self):
            return True

    class CheckpointCallback(BaseCallback):
        def __init__(self, save_freq, save_path, name_prefix="rl_model"):
            super().__init__()
            record["checkpoint_cb"] = dict(save_freq=save_freq, save_path=save_path, name_prefix=name_prefix)

    class SAC:
        def __init__(self, policy, env, learning_rate=3e-4, buffer_size=1_000_000, batch_size=256, tau=0.005, gamma=0.99,
                     policy_kwargs=None, verbose=0, **kw):
            assert not kw, f"train.py passed keywords SB3's SAC does not have: {sorted(kw)}"
            record["sac_kwargs"] = dict(policy=policy, learning_rate=learning_rate, buffer_size=buffer_size, batch_size=batch_size,
                                        tau=tau, gamma=gamma, policy_kwargs=policy_kwargs, verbose=verbose)
            self.env = env
            # what SB3 checks when it wraps a gym.Env: spaces with shape / dtype / bounds
            assert env.observation_space.shape == (24,) and env.observation_space.dtype == np.float32
            assert env.action_space.shape == (1,) and float(env.action_space.low[0]) == -1.0 and float(env.action_space.high[0]) == 1.0

        def learn(self, total_timesteps, callback=None, log_interval=4):
            env, rng = self.env, np.random.default_rng(0)
            obs, info = env.reset()
            assert isinstance(obs, np.ndarray) and obs.shape == (24,) and obs.dtype == np.float32 and info == {}
            episodes = 0
            for t in range(total_timesteps):
                action = rng.uniform(env.action_space.low, env.action_space.high).astype(np.float32)   # SAC's warm-up sampling
                obs, reward, terminated, truncated, info = env.step(action)
                assert obs.shape == (24,) and obs.dtype == np.float32 and isinstance(reward, float)
                assert isinstance(terminated, bool) and isinstance(truncated, bool) and not (terminated and truncated)
                assert set(info) == INFO_KEYS, set(info) ^ INFO_KEYS
                record["steps"] = record.get("steps", 0) + 1
                if callback is not None:
                    callback.n_calls += 1
                    callback.model = self
                    callback._on_step()
                if terminated or truncated:   # DummyVecEnv.step_wait resets and stores terminal_observation
                    episodes += 1
                    obs, info = env.reset()
            record["episodes"] = episodes
            return self

        def save(self, path):
            record.setdefault("saved", []).append(path)

        @classmethod
        def load(cls, path, env=None):
            return cls("MlpPolicy", env)

    sb3.SAC = SAC
    callbacks.CheckpointCallback, callbacks.BaseCallback = CheckpointCallback, BaseCallback
    common.callbacks = callbacks
    sb3.common = common
    return {"stable_baselines3": sb3, "stable_baselines3.common": common, "stable_baselines3.common.callbacks": callbacks}


@pytest.mark.parametrize("preset", ["single_food", "defaults", "sac_gail", "single_food_long_horizon"])
def test_reference_train_py_sb3_path_on_the_facade(preset, monkeypatch, tmp_path, oracle_lib):
    import fake_backend
    from underwater_swimmer_rl_b200 import config as our_config
    from underwater_swimmer_rl_b200 import snake_env

    monkeypatch.setattr(snake_env, "BACKEND", fake_backend.OracleVectorBackend)
    record = {}
    mods = _stub_sb3(record)
    # the two imports train.py makes from the reference package (train.py:37-38) resolve to this repo's loader and facade --
    # the one-line change INTEGRATION.md section 3 shows a maintainer
    salp = types.ModuleType("salp")
    salp.__path__ = []
    salp_config, salp_env = types.ModuleType("salp.config"), types.ModuleType("salp.environments")
    salp_config.__path__, salp_env.__path__ = [], []
    loader = types.ModuleType("salp.config.config_loader")
    repo_cfg_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "configs")
    loader.load_config = lambda name=None, **ov: our_config.load_config(name, config_dir=repo_cfg_dir, **ov)
    envmod = types.ModuleType("salp.environments.salp_snake_env")
    envmod.SalpSnakeEnv = snake_env.SalpSnakeEnv
    mods.update({"salp": salp, "salp.config": salp_config, "salp.config.config_loader": loader, "salp.environments": salp_env,
                 "salp.environments.salp_snake_env": envmod})
    for k, v in mods.items():
        monkeypatch.setitem(sys.modules, k, v)
    spec = importlib.util.spec_from_file_location("reference_train", os.path.join(REF, "train.py"))
    train = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(train)   # the reference's own file, unmodified

    monkeypatch.chdir(tmp_path)    # train.py writes data/models/<experiment>_<timestamp>/ relative to the cwd
    monkeypatch.setattr(sys, "argv", ["train.py", "--sb3", "--config", preset, "--timesteps", "2600"])
    train.main()

    cfg = our_config.load_config(preset, config_dir=repo_cfg_dir)
    k = record["sac_kwargs"]
    assert k["policy"] == "MlpPolicy" and k["policy_kwargs"] == dict(net_arch=cfg.agent["hidden_sizes"])
    for key in ("learning_rate", "buffer_size", "batch_size", "tau", "gamma"):
        assert k[key] == cfg.agent[key], key          # read from the restored agent section of configs/*.yaml
    assert record["steps"] == 2600 and record["episodes"] >= 1
    assert record["checkpoint_cb"]["save_freq"] == 10000 and record["saved"] and record["saved"][0].endswith("final_model")
    assert os.path.isdir(os.path.dirname(record["saved"][0])) and cfg.training["experiment_name"] in record["saved"][0]
