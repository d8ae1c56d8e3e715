"""GPU: the fused tcgen05 actor kernel (csrc/salp_actor.cu) against the torch module it replaces
(rollout.TanhGaussianActor == record_model.py:27-53), evaluated in fp32 from the same bf16-rounded weights."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(actor, obs, noise):
    """fp32 evaluation with the roundings the kernel makes: bf16 weights, bf16 inputs to both hidden GEMMs."""
    bf = lambda t: t.to(torch.bfloat16).float()  # noqa: E731
    l1, l2, l3 = actor.layers
    h = torch.relu(bf(obs) @ bf(l1.weight).T + l1.bias.float())
    h = torch.relu(bf(h) @ bf(l2.weight).T + l2.bias.float())
    out = h @ bf(l3.weight).T + l3.bias.float()
    mean, log_std = out.chunk(2, dim=-1)
    if noise is None:
        return actor.to_box(torch.tanh(mean))
    return actor.to_box(torch.tanh(mean + torch.exp(torch.clamp(log_std, -20, 2)) * noise))


# sizes chosen for the kernel's software pipeline on 148 SMs: one tile per CTA (<= 18 944 rows), some CTAs with two tiles
# and some with one (21 504), three to four (70 001), 13-14 with a ragged last tile (262 221); act_dim 2 = 4 output rows
@pytest.mark.parametrize("n,obs_dim,act_dim", [(128, 24, 1), (1000, 24, 1), (21504, 24, 1), (70001, 24, 1), (262221, 24, 1),
                                               (4096, 24, 2), (40000, 24, 2), (513, 32, 2), (300, 16, 1)])
def test_fused_actor_matches_torch(salp_lib, n, obs_dim, act_dim):
    from underwater_swimmer_rl_b200.rollout import FusedActor, TanhGaussianActor
    torch.manual_seed(n + obs_dim)
    actor = TanhGaussianActor(obs_dim, act_dim).cuda()
    with torch.no_grad():
        for l in actor.layers:  # larger weights than the default init so that errors would show
            l.weight.mul_(2.0)
            l.bias.uniform_(-0.5, 0.5)
    fused = FusedActor(actor)
    obs = (torch.rand(n, obs_dim, device="cuda") * 2 - 1)
    noise = torch.randn(n, act_dim, device="cuda")
    for nz, out_dtype in ((None, torch.float32), (noise, torch.float32), (noise, torch.float64)):
        out = torch.full((n, act_dim), 7.0, device="cuda", dtype=out_dtype)
        fused(obs, out, deterministic=nz is None, noise=nz)
        torch.cuda.synchronize()
        with torch.no_grad():
            ref = _ref(actor, obs, nz)
        err = (out.float() - ref).abs().max().item()
        assert err <= 5e-3, (err, nz is None, out_dtype)
        assert out.abs().max().item() <= 1.0


@pytest.mark.parametrize("n,obs_dim,act_dim,box", [
    (4097, 6, 3, ([0.0, 0.0, -1.0], [1.0, 1.0, 1.0])),   # model (B): salp_robot_env.py:51-71
    (50001, 6, 3, ([0.0, 0.0, -1.0], [1.0, 1.0, 1.0])),  # ... with two to three tiles per CTA (8 rows of W3: sums meet in TMEM)
    (1000, 6, 3, None), (777, 7, 4, ([-2.0, 0.0, 0.5, -1.0], [2.0, 3.0, 1.5, 1.0])), (300, 31, 3, None), (129, 1, 1, None)])
def test_fused_actor_other_shapes_and_box_rescale(salp_lib, n, obs_dim, act_dim, box):
    """Rows that are not float4 groups, up to 4 action dims (8 rows of W3), and SB3's [-1, 1] -> Box mapping."""
    from underwater_swimmer_rl_b200.rollout import FusedActor, TanhGaussianActor
    torch.manual_seed(n + obs_dim)
    kw = {} if box is None else {"action_low": box[0], "action_high": box[1]}
    actor = TanhGaussianActor(obs_dim, act_dim, **kw).cuda()
    with torch.no_grad():
        for l in actor.layers:
            l.weight.mul_(2.0)
            l.bias.uniform_(-0.5, 0.5)
    fused = FusedActor(actor)
    base = torch.rand(n * obs_dim + 1, device="cuda") * 2 - 1
    noise = torch.randn(n, act_dim, device="cuda")
    for off in (0, 1):  # off = 1: the observation rows start 4 bytes off a 16-byte boundary
        obs = base[off:off + n * obs_dim].view(n, obs_dim)
        for nz, out_dtype in ((None, torch.float32), (noise, torch.float32), (noise, torch.float64)):
            out = torch.full((n, act_dim), 7.0, device="cuda", dtype=out_dtype)
            fused(obs, out, deterministic=nz is None, noise=nz)
            torch.cuda.synchronize()
            with torch.no_grad():
                ref = _ref(actor, obs, nz)
            err = (out.float() - ref).abs().max().item()
            assert err <= 5e-3 * (1.0 if box is None else 2.0), (err, nz is None, out_dtype, off)
            if box is not None:
                lo, hi = (torch.tensor(b, device="cuda") for b in box)
                assert (out.float() >= lo - 1e-6).all() and (out.float() <= hi + 1e-6).all()
    # in-kernel noise with 3-4 action dims: every dim gets its own draw
    if act_dim >= 3:
        f2 = FusedActor(TanhGaussianActor(obs_dim, act_dim).cuda())
        with torch.no_grad():
            f2.actor.layers[2].weight.zero_(); f2.actor.layers[2].bias.zero_()
        f2.refresh()
        out = torch.empty(1 << 16, act_dim, device="cuda", dtype=torch.float64)
        f2(torch.rand(1 << 16, obs_dim, device="cuda"), out)
        z = torch.atanh(out)
        assert abs(z.mean().item()) < 0.02 and abs(z.std().item() - 1.0) < 0.02
        c = torch.corrcoef(z.T)
        assert (c - torch.eye(act_dim, device="cuda", dtype=torch.float64)).abs().max().item() < 0.02


def test_fused_gail_reward_model_b_shape(salp_lib):
    """Discriminator on cat(obs 6, action 3): the appended columns start inside the first 8-feature chunk."""
    from underwater_swimmer_rl_b200.rollout import FusedGailReward, GailDiscriminator
    torch.manual_seed(3)
    n = 5000
    disc = GailDiscriminator(6, 3).cuda()
    with torch.no_grad():
        for l in disc.layers:
            l.weight.mul_(2.0)
            l.bias.uniform_(-0.5, 0.5)
    fused = FusedGailReward(disc, 6, 3)
    obs = torch.rand(n, 6, device="cuda") * 2 - 1
    act = torch.rand(n, 3, device="cuda") * 2 - 1
    out = torch.empty(n, device="cuda")
    fused(obs, act, out)
    bf = lambda t: t.to(torch.bfloat16).float()  # noqa: E731
    with torch.no_grad():
        l1, l2, l3 = disc.layers
        h = torch.relu(bf(torch.cat([obs, act], -1)) @ bf(l1.weight).T + l1.bias)
        h = torch.relu(bf(h) @ bf(l2.weight).T + l2.bias)
        ref = -torch.log(1 - torch.sigmoid((h @ bf(l3.weight).T + l3.bias).reshape(-1)) + 1e-8)
    assert (out - ref).abs().max().item() <= 5e-3 * max(1.0, ref.abs().max().item())


def test_fused_actor_in_the_rollout_loop(salp_lib):
    """Collector with the fused actor == collector with the torch actor up to the actor's bf16 rounding:
    same env trajectories for a while (deterministic policy), and the graph replay path works."""
    from helpers import CONFIGS
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    from underwater_swimmer_rl_b200.rollout import DeviceReplayBuffer, FusedActor, RolloutCollector, TanhGaussianActor
    n = 4096
    torch.manual_seed(5)
    actor = TanhGaussianActor(24, 1).cuda()
    env = SalpVectorEnv(n, device="cuda:0", max_substeps=2, seed=77, **CONFIGS["sac_gail"])
    buf = DeviceReplayBuffer(8 * n, 24, 1, "cuda:0")
    col = RolloutCollector(env, FusedActor(actor), buf, deterministic=True)
    col.reset()
    replay = col.capture()
    for _ in range(10):
        replay()
    torch.cuda.synchronize()
    assert col.steps == 22 and len(buf) == 8 * n
    with torch.no_grad():
        a_ref = _ref(actor, buf.obs[:n], None)
    assert (buf.act[:n] - a_ref).abs().max().item() <= 5e-3


def test_in_kernel_noise_is_standard_normal_and_advances(salp_lib):
    """Without a noise tensor the kernel draws its own normals (Philox4x32-10 + Box-Muller): with W3 = 0 the action
    is tanh(b_mean + exp(b_logstd) z), so z is recoverable; it must look N(0,1), differ between calls, and be
    reproducible from (seed, draw counter)."""
    from underwater_swimmer_rl_b200.rollout import FusedActor, TanhGaussianActor
    n = 1 << 18
    actor = TanhGaussianActor(24, 2).cuda()
    with torch.no_grad():
        actor.layers[2].weight.zero_()
        actor.layers[2].bias.copy_(torch.tensor([0.1, -0.2, -1.0, -1.5]))  # means 0.1/-0.2, log_std -1/-1.5
    obs = torch.rand(n, 24, device="cuda")
    draws = []
    for seed in (7, 7, 8):
        fused = FusedActor(actor, seed=seed)
        zs = []
        for _ in range(2):
            out = torch.empty(n, 2, device="cuda", dtype=torch.float64)
            fused(obs, out)
            z = (torch.atanh(out) - torch.tensor([0.1, -0.2], device="cuda", dtype=torch.float64)) / torch.exp(
                torch.tensor([-1.0, -1.5], device="cuda", dtype=torch.float64))
            zs.append(z)
        assert int(fused.rng_state[1].item()) == 2 and int(fused.rng_state[2].item()) == 0
        draws.append(zs)
    z = draws[0][0]
    assert abs(z.mean().item()) < 0.01 and abs(z.std().item() - 1.0) < 0.01
    assert abs((z ** 4).mean().item() - 3.0) < 0.1                       # kurtosis of a normal
    assert abs((z[:, 0] * z[:, 1]).mean().item()) < 0.01                 # the two action dims are independent
    assert abs((z[1:, 0] * z[:-1, 0]).mean().item()) < 0.01              # neighbouring rows too
    assert (draws[0][0] - draws[0][1]).abs().mean().item() > 0.5         # consecutive calls differ
    assert torch.equal(draws[0][0], draws[1][0]) and torch.equal(draws[0][1], draws[1][1])  # same seed -> same draws
    assert (draws[0][0] - draws[2][0]).abs().mean().item() > 0.5         # other seed differs


@pytest.mark.parametrize("n,act_dim,dtype", [(1000, 1, torch.float32), (70001, 2, torch.float64)])
def test_fused_gail_reward_matches_torch(salp_lib, n, act_dim, dtype):
    from underwater_swimmer_rl_b200.rollout import FusedGailReward, GailDiscriminator
    torch.manual_seed(n)
    disc = GailDiscriminator(24, act_dim).cuda()
    with torch.no_grad():
        for l in disc.layers:
            l.weight.mul_(2.0)
            l.bias.uniform_(-0.5, 0.5)
    fused = FusedGailReward(disc, 24, act_dim)
    obs = torch.rand(n, 24, device="cuda") * 2 - 1
    act = (torch.rand(n, act_dim, device="cuda") * 2 - 1).to(dtype)
    out = torch.empty(n, device="cuda")
    fused(obs, act, out)
    bf = lambda t: t.to(torch.bfloat16).float()  # noqa: E731
    with torch.no_grad():
        l1, l2, l3 = disc.layers
        x = torch.cat([obs, act.float()], -1)
        h = torch.relu(bf(x) @ bf(l1.weight).T + l1.bias)
        h = torch.relu(bf(h) @ bf(l2.weight).T + l2.bias)
        logit = (h @ bf(l3.weight).T + l3.bias).reshape(-1)
        ref = -torch.log(1 - torch.sigmoid(logit) + 1e-8)
    assert (out - ref).abs().max().item() <= 5e-3 * max(1.0, ref.abs().max().item())
    assert (out > 0).all()


def test_fused_actor_against_the_true_fp32_actor(salp_lib):
    """the kernel runs bf16 GEMMs where the reference actor (record_model.py:27-53) is fp32.  Bound,
    stated and tested: on observations produced by the env itself, with torch's default init AND with weights twice that
    scale, the deterministic action differs from the fp32 actor's by <= 2e-2 (worst) and <= 2e-3 (mean) -- bf16 keeps 8 bits
    of mantissa in the inputs of the two hidden GEMMs, accumulation and both epilogues are f32.  (tests/test_reference_networks.py
    pins TanhGaussianActor itself to the reference class.)"""
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    from underwater_swimmer_rl_b200.rollout import FusedActor, TanhGaussianActor
    env = SalpVectorEnv(32768, dtype="float32", rng="philox", seed=3, num_food_items=12, forced_breathing=True)
    env.reset()
    g = torch.Generator(device="cuda")
    g.manual_seed(0)
    for _ in range(400):   # observations of moving robots in all breathing phases
        obs, *_ = env.step(torch.rand(32768, 1, device="cuda", generator=g) * 2 - 1)
    obs = obs.clone()
    for scale in (1.0, 2.0):
        torch.manual_seed(int(scale * 10))
        actor = TanhGaussianActor(24, 1).cuda()
        with torch.no_grad():
            for l in actor.layers:
                l.weight.mul_(scale)
        out = torch.empty(32768, 1, device="cuda")
        FusedActor(actor)(obs, out, deterministic=True)
        with torch.no_grad():
            ref = actor(obs, deterministic=True)     # fp32 weights, fp32 activations
        err = (out - ref).abs()
        assert err.max().item() <= 2e-2 and err.mean().item() <= 2e-3, (scale, err.max().item(), err.mean().item())
