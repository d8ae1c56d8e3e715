"""YAML experiment configs, same surface as the reference loader
(src/salp/config/config_loader.py:12-115): ``load_config(name_or_path, **overrides) -> Config`` with
``.environment / .agent / .training / .gail`` dicts; ``environment['params']`` are the SalpSnakeEnv
keyword arguments (train.py:50-52).  ``make_vector_env`` builds the batched env from a Config.
"""
from __future__ import annotations

from dataclasses import dataclass
from pathlib import Path
from typing import Any, Dict, Optional

import yaml


@dataclass
class Config:
    environment: Dict[str, Any]
    agent: Dict[str, Any]
    training: Dict[str, Any]
    gail: Optional[Dict[str, Any]] = None

    @classmethod
    def from_yaml(cls, config_path: str, overrides: Optional[Dict] = None) -> "Config":
        path = Path(config_path)
        if not path.exists():
            raise FileNotFoundError(f"Config file not found: {config_path}")
        with open(path, "r") as f:
            d = yaml.safe_load(f) or {}
        if overrides:
            d = _deep_update(d, overrides)
        return cls(environment=d.get("environment", {}), agent=d.get("agent", {}), training=d.get("training", {}),
                   gail=d.get("gail"))

    @classmethod
    def from_preset(cls, preset: str = "defaults", overrides: Optional[Dict] = None, config_dir: str = "configs") -> "Config":
        path = Path(config_dir) / f"{preset}.yaml"
        if not path.exists():  # the reference falls back to defaults.yaml (config_loader.py:60-62)
            path = Path(config_dir) / "defaults.yaml"
        return cls.from_yaml(str(path), overrides)

    def get(self, section: str, key: str, default: Any = None) -> Any:
        return getattr(self, section, {}).get(key, default)


def _deep_update(base: Dict, upd: Dict) -> Dict:
    out = base.copy()
    for k, v in upd.items():
        out[k] = _deep_update(out[k], v) if k in out and isinstance(out[k], dict) and isinstance(v, dict) else v
    return out


def load_config(config_file: Optional[str] = None, config_dir: str = "configs", **overrides) -> Config:
    if config_file is None:
        config_file = "defaults"
    path = Path(config_file)
    if path.suffix == ".yaml" and path.exists():
        return Config.from_yaml(str(path), overrides or None)
    return Config.from_preset(config_file, overrides or None, config_dir=config_dir)


def env_kwargs(config: Config, with_size: bool = False) -> Dict[str, Any]:
    """SalpSnakeEnv kwargs from a Config.  train.py:50-52 forwards only ``params`` (tank size stays at
    the 800x600 default); scripts/train.py:78-85 also forwards width/height -> ``with_size=True``."""
    kw = dict(config.environment.get("params", {}) or {})
    if with_size:
        for k in ("width", "height"):
            if k in config.environment:
                kw[k] = config.environment[k]
    return kw


def make_vector_env(config: Config, num_envs: int, **vec_kwargs):
    from .vector_env import SalpVectorEnv
    return SalpVectorEnv(num_envs, **vec_kwargs, **env_kwargs(config))
