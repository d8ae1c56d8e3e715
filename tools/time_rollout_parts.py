"""Time the pieces of the sac_rollout step separately (developer tool): fused actor, torch actor,
replay append, env kernel (graph replay).  CUDA events, 200 launches each."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from underwater_swimmer_rl_b200 import _capi
if os.environ.get("SALP_LAB_LIB"):  # time another build of the library (tools/_lab/*.so)
    _capi.LIB_PATH = os.path.abspath(os.environ["SALP_LAB_LIB"])
from underwater_swimmer_rl_b200 import SalpVectorEnv
from underwater_swimmer_rl_b200.rollout import DeviceReplayBuffer, FusedActor, TanhGaussianActor

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
P = dict(num_food_items=12, food_reward=15.0, collision_penalty=-30.0, time_penalty=-0.05, efficiency_bonus=2.0, forced_breathing=True)
env = SalpVectorEnv(N, device="cuda:0", max_substeps=2, food_table_len=128, jitter_table_len=512, **P)
env.reset()
actor = TanhGaussianActor(24, 1).cuda()
fused = FusedActor(actor)
actor_bf = TanhGaussianActor(24, 1).cuda().to(torch.bfloat16)
buf = DeviceReplayBuffer(4 * N, 24, 1, "cuda:0")
act = torch.zeros(N, 1, device="cuda")
noise = torch.randn(N, 1, device="cuda")
obs = env.obs[0]


def timeit(name, fn, n=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:34s} {1e3 * e0.elapsed_time(e1) / n:8.1f} us")


timeit("fused actor (noise given)", lambda: fused(obs, act, noise=noise))
timeit("fused actor (deterministic)", lambda: fused(obs, act, deterministic=True))
timeit("fused actor + torch.randn", lambda: fused(obs, act))
with torch.no_grad():
    timeit("torch bf16 actor", lambda: actor_bf(obs))
timeit("replay append", lambda: buf.append(obs, act, env.reward[0], env.terminated[0], env.truncated[0], env.obs[1], final_obs=env.final_obs[1]))
g = env.capture_graph(act)
timeit("env step (graph replay)", g)
