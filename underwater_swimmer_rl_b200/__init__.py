"""underwater_swimmer_rl_b200 -- B200-native batched SALP swimmer environment step.

One hot path of SungRoboticsGroup/UNDERWATER-SWIMMER_RL (SalpSnakeEnv.step/reset over the legacy
SalpRobotEnv integrator) as a hand-written sm_100a CUDA kernel behind a C ABI
(include/salp_b200.h), with a Python host that keeps the reference's Gymnasium API.
"""
__all__ = ["SalpVectorEnv", "SalpSnakeEnv", "load_config", "build"]


def __getattr__(name):  # lazy: importing the package must not need torch / a built .so
    if name == "SalpVectorEnv":
        from .vector_env import SalpVectorEnv
        return SalpVectorEnv
    if name == "SalpSnakeEnv":
        from .snake_env import SalpSnakeEnv
        return SalpSnakeEnv
    if name == "load_config":
        from .config import load_config
        return load_config
    if name == "build":
        from .build import build
        return build
    raise AttributeError(name)
