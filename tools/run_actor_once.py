import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from underwater_swimmer_rl_b200.rollout import FusedActor, TanhGaussianActor
N = 1 << 18
fused = FusedActor(TanhGaussianActor(24, 1).cuda())
obs = torch.rand(N, 24, device="cuda"); act = torch.zeros(N, 1, device="cuda"); noise = torch.randn(N, 1, device="cuda")
for _ in range(5):
    fused(obs, act, noise=noise)
torch.cuda.synchronize()
