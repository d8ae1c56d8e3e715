"""Pin the torch checkers of the fused actor / GAIL-reward kernel to the reference's own networks (build container only).

tests/test_gpu_actor.py compares the tcgen05 kernel with `rollout.TanhGaussianActor` / `rollout.GailDiscriminator`; those are
this repo's restatements.  Here they are checked against the classes the reference ships -- `CustomActor`
(/root/reference/record_model.py:27-53) and `Discriminator` (/root/reference/src/salp/agents/discriminator.py:16-85 over
src/salp/core/base_agent.py:12-74) -- loaded from where they lie, with the same state_dict: outputs must be identical.
"""
import ast
import importlib.util
import os
import sys
import types

import pytest
import torch

REF = "/root/reference"
pytestmark = [pytest.mark.reference, pytest.mark.skipif(not os.path.isfile(os.path.join(REF, "record_model.py")), reason="needs the reference tree")]


def _reference_custom_actor():
    """record_model.py imports cv2 / pygame-backed modules at the top; take the CustomActor class definition alone."""
    src = open(os.path.join(REF, "record_model.py")).read()
    node = next(n for n in ast.parse(src).body if isinstance(n, ast.ClassDef) and n.name == "CustomActor")
    ns = {"torch": torch, "nn": torch.nn, "Normal": torch.distributions.Normal}
    exec(compile(ast.Module(body=[node], type_ignores=[]), "record_model.py:CustomActor", "exec"), ns)
    return ns["CustomActor"]


def _reference_discriminator(monkeypatch):
    salp, core = types.ModuleType("salp"), types.ModuleType("salp.core")
    salp.__path__, core.__path__ = [], []
    monkeypatch.setitem(sys.modules, "salp", salp)
    monkeypatch.setitem(sys.modules, "salp.core", core)
    for name, rel in (("salp.core.base_agent", "src/salp/core/base_agent.py"), ("salp.agents.discriminator", "src/salp/agents/discriminator.py")):
        spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
        mod = importlib.util.module_from_spec(spec)
        monkeypatch.setitem(sys.modules, name, mod)
        spec.loader.exec_module(mod)
    return sys.modules["salp.agents.discriminator"].Discriminator


@pytest.mark.parametrize("obs_dim,act_dim", [(24, 1), (24, 2), (6, 3)])
def test_tanh_gaussian_actor_equals_reference_custom_actor(obs_dim, act_dim):
    from underwater_swimmer_rl_b200.rollout import TanhGaussianActor
    torch.manual_seed(3)
    ref = _reference_custom_actor()(obs_dim, act_dim, [256, 256])
    ours = TanhGaussianActor(obs_dim, act_dim, (256, 256))
    ours.load_state_dict(ref.state_dict())       # same parameter names: layers.{0,1,2}.{weight,bias}
    obs = torch.randn(512, obs_dim)
    with torch.no_grad():
        assert torch.equal(ours(obs, deterministic=True), ref(obs, deterministic=True))
        torch.manual_seed(11)
        a = ours(obs)
        torch.manual_seed(11)
        b = ref(obs)                              # Normal(mean, std).rsample() == mean + std * randn_like(mean)
        assert torch.allclose(a, b, atol=1e-6, rtol=0)


def test_gail_discriminator_equals_reference_discriminator(monkeypatch):
    from underwater_swimmer_rl_b200.rollout import GailDiscriminator
    torch.manual_seed(5)
    ref = _reference_discriminator(monkeypatch)(24, 1, [256, 256])
    ours = GailDiscriminator(24, 1, (256, 256))
    ours.load_state_dict({k: v for k, v in ref.state_dict().items() if k.startswith("layers.")})
    obs, act = torch.randn(400, 24), torch.rand(400, 1) * 2 - 1
    with torch.no_grad():
        assert torch.equal(ours(obs, act), ref(obs, act))
        assert torch.equal(ours.predict_reward(obs, act), ref.predict_reward(obs, act))
