"""SalpVectorEnv -- N independent SalpSnakeEnv instances stepped by one sm_100a CUDA kernel.

Drop-in for the reference's ``SalpSnakeEnv`` (src/salp/environments/salp_snake_env.py:17-202 over
the legacy integrator scripts/utilities/salp_robot.py:15-403) behind the
``gymnasium.vector.VectorEnv`` API: same constructor keywords, same action / observation spaces,
same reward / terminated / truncated semantics, Gymnasium 0.29.1 vector auto-reset.

All state and all step outputs are torch tensors on the GPU, allocated here and handed to
``libsalp_b200.so`` as raw device pointers (include/salp_b200.h); nothing is copied to the host
unless the caller passes numpy actions or asks for numpy outputs.  There is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Any, Optional

import numpy as np
import torch

from . import _capi, rng_tables
from ._gym_compat import HAVE_GYMNASIUM, Box, VectorEnv, batch_box

# SalpSnakeEnv.__init__ defaults (salp_snake_env.py:29-33)
ENV_DEFAULTS = dict(
    width=800, height=600, num_food_items=5, food_reward=10.0, collision_penalty=-50.0, time_penalty=-0.1,
    efficiency_bonus=1.0, forced_breathing=True, max_observed_food=3, random_food_count=False,
    respawn_food=True, proximity_reward_weight=0.0, max_steps_without_food=1500,
)
PHASE_NAMES = ("rest", "inhaling", "exhaling")  # salp_robot.py:71,374
STATS_KEYS = ("episodes", "return_sum", "length_sum", "food_sum", "collisions", "truncations", "nonfinite", "env_steps")


def _torch_dtype(dtype) -> torch.dtype:
    if dtype in ("float32", "f32", torch.float32, np.float32):
        return torch.float32
    if dtype in ("float64", "f64", torch.float64, np.float64):
        return torch.float64
    raise ValueError(f"dtype must be float32 (throughput build) or float64 (validation build), got {dtype!r}")


def make_params(autoreset: bool, **kw) -> _capi.SalpParams:
    d = dict(ENV_DEFAULTS)
    unknown = set(kw) - set(d)
    if unknown:
        raise TypeError(f"SalpSnakeEnv got unexpected keyword argument(s) {sorted(unknown)}")
    d.update(kw)
    p = _capi.SalpParams()
    p.width, p.height = float(d["width"]), float(d["height"])
    p.num_food_items = max(0, int(d["num_food_items"]))  # salp_snake_env.py:36
    p.max_observed_food = int(d["max_observed_food"])
    p.forced_breathing = int(bool(d["forced_breathing"]))
    p.random_food_count = int(bool(d["random_food_count"]))
    p.respawn_food = int(bool(d["respawn_food"]))
    p.max_steps_without_food = int(d["max_steps_without_food"])
    for k in ("food_reward", "collision_penalty", "time_penalty", "efficiency_bonus", "proximity_reward_weight"):
        setattr(p, k, float(d[k]))
    p.flags = _capi.SALP_AUTORESET if autoreset else 0
    return p


def single_spaces(params: _capi.SalpParams, layout: _capi.SalpLayout):
    """action/observation spaces exactly as salp_snake_env.py:69-88 builds them."""
    if params.forced_breathing:
        act = Box(low=-1.0, high=1.0, shape=(1,), dtype=np.float32)
    else:
        act = Box(low=np.array([0.0, -1.0]), high=np.array([1.0, 1.0]), dtype=np.float32)
    food_obs = params.max_observed_food * 4 + 2
    obs = Box(low=np.array([0, 0, -10, -10, -math.pi, -0.1, 0.5, 0, 0, -1] + [-1] * food_obs),
              high=np.array([1, 1, 10, 10, math.pi, 0.1, 2.0, 2, 1, 1] + [1] * food_obs), dtype=np.float32)
    assert obs.shape == (layout.obs_dim,)
    return obs, act


class SalpVectorEnv(VectorEnv):
    """Batched SALP snake environment on one B200.

    Parameters beyond the reference keywords:
      num_envs, device           slab size and CUDA device of this process (one process per GPU)
      dtype                      "float32" throughput build | "float64" validation build
      seed, env_id0              env e uses the reference streams random.seed(seed+env_id0+e) /
                                 np.random.seed(seed+env_id0+e)
      table_rows                 rows of the RNG word tables (default num_envs; fewer rows are
                                 shared round-robin -- throughput runs only)
      food_table_len, jitter_table_len   words per row (powers of two; cursors wrap)
      rng                        "table": the two random streams are read from host-seeded MT19937 word tables (parity with
                                 the reference's `random` / `np.random`); "philox": computed in the kernel (Philox4x32-10
                                 keyed by seed, global env id and word index) -- unbounded streams, no memory, no parity
      max_substeps               K: env steps fused per kernel launch (``step_many``)
      autoreset                  Gymnasium vector auto-reset inside the kernel (default True)
    """

    metadata = {"render_modes": [], "autoreset_mode": "next_obs_in_obs"}

    def __init__(self, num_envs: int, device: Any = "cuda:0", dtype: Any = "float32", seed: int = 10_000,
                 env_id0: int = 0, table_rows: Optional[int] = None, food_table_len: int = 256,
                 jitter_table_len: int = 1024, max_substeps: int = 1, autoreset: bool = True,
                 render_mode: Optional[str] = None, rng: str = "table", generic_kernel: bool = False, **env_kwargs):
        if render_mode is not None:
            raise ValueError("render_mode must be None: there is no graphics stack on the GPU box (SURVEY.md F11)")
        if not torch.cuda.is_available():
            raise RuntimeError("SalpVectorEnv needs a CUDA device; there is no CPU fallback")
        self._L = _capi.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise ValueError("device must be a CUDA device")
        self.real = _torch_dtype(dtype)
        self._dtype_code = _capi.SALP_F64 if self.real == torch.float64 else _capi.SALP_F32
        self.num_envs = int(num_envs)
        self.env_id0 = int(env_id0)
        self.seed_base = int(seed)
        self.max_substeps = int(max_substeps)
        self.autoreset = bool(autoreset)
        self.render_mode = None
        if rng not in ("table", "philox"):
            raise ValueError("rng must be 'table' (host-seeded MT19937 word tables shared with the reference) or 'philox'")
        self.rng = rng
        self.params = make_params(self.autoreset, **env_kwargs)
        if rng == "philox":
            self.params.flags |= _capi.SALP_RNG_PHILOX
        if generic_kernel:  # diagnostics / tests: never select the specialised forced-breathing kernel
            self.params.flags |= _capi.SALP_GENERIC_KERNEL
        self.layout = _capi.SalpLayout()
        _capi.check(self._L.salp_query_layout(C.byref(self.params), self._dtype_code, C.byref(self.layout)))
        self.obs_dim, self.act_dim, self.food_slots = self.layout.obs_dim, self.layout.act_dim, self.layout.food_slots
        self.table_rows = int(table_rows) if table_rows else self.num_envs
        self.food_table_len, self.jitter_table_len = int(food_table_len), int(jitter_table_len)
        # reference-visible attributes
        for k in ENV_DEFAULTS:
            setattr(self, k, env_kwargs.get(k, ENV_DEFAULTS[k]))
        self._base_num_food_items = self.params.num_food_items
        obs_space, act_space = single_spaces(self.params, self.layout)
        if HAVE_GYMNASIUM:  # gymnasium 0.29.1: VectorEnv.__init__(num_envs, observation_space, action_space) batches the spaces
            try:
                VectorEnv.__init__(self, self.num_envs, obs_space, act_space)
            except TypeError:  # gymnasium >= 1.0: VectorEnv has no constructor arguments
                VectorEnv.__init__(self)
        self.single_observation_space, self.single_action_space = obs_space, act_space
        self.observation_space = batch_box(self.single_observation_space, self.num_envs)
        self.action_space = batch_box(self.single_action_space, self.num_envs)

        N, K, F, dev, real = self.num_envs, self.max_substeps, self.food_slots, self.device, self.real
        z = lambda *shape, dtype=real: torch.zeros(*shape, dtype=dtype, device=dev)  # noqa: E731
        self.body = z(2, N, 4)
        self.ctrl = z(2, N, 4, dtype=torch.int32)
        self.ep_return = z(N)
        self.food = z(F, N, 2)
        self.actions = z(K, N, self.act_dim)
        self.obs = z(K, N, self.obs_dim, dtype=torch.float32)
        self.reward = z(K, N)
        self.terminated = z(K, N, dtype=torch.bool)
        self.truncated = z(K, N, dtype=torch.bool)
        self.events = z(K, N, dtype=torch.uint8)
        self.final_obs = z(K, N, self.obs_dim, dtype=torch.float32)
        self.final_info = z(K, N, 4)
        self.stats = z(8, dtype=torch.float64)
        self._handle = C.c_void_p()
        _capi.check(self._L.salp_create(C.byref(self.params), N, self.env_id0, self.device.index or 0,
                                        self._dtype_code, C.byref(self._handle)))
        self._load_tables(self.seed_base)
        self._bind()
        self._needs_reset = True

    # ------------------------------------------------------------------ plumbing
    def _load_tables(self, seed: int) -> None:
        if self.rng == "philox":  # both streams are computed in the kernel: no tables, no memory
            _capi.check(self._L.salp_set_rng_seed(self._handle, int(seed)))
            self.food_words = self.jit_words = torch.zeros(8, dtype=torch.int32, device=self.device)
            return
        rows, first = self.table_rows, self.env_id0
        fw = rng_tables.fill_words(rows, self.food_table_len, seed, first, rng_tables.FOOD_STREAM)
        jw = rng_tables.fill_words(rows, self.jitter_table_len, seed, first, rng_tables.JITTER_STREAM)
        self.food_words = torch.from_numpy(fw.view(np.int32)).to(self.device)
        self.jit_words = torch.from_numpy(jw.view(np.int32)).to(self.device)

    def set_rng_tables(self, food_words: np.ndarray, jitter_words: np.ndarray) -> None:
        """Install caller-built word tables (uint32 [rows, len]); used by the parity tests."""
        fw = np.ascontiguousarray(food_words, dtype=np.uint32)
        jw = np.ascontiguousarray(jitter_words, dtype=np.uint32)
        assert fw.shape[0] == jw.shape[0]
        self.table_rows, self.food_table_len, self.jitter_table_len = fw.shape[0], fw.shape[1], jw.shape[1]
        self.food_words = torch.from_numpy(fw.view(np.int32)).to(self.device)
        self.jit_words = torch.from_numpy(jw.view(np.int32)).to(self.device)
        self._bind()

    def _bind(self, omit=()) -> None:
        """(Re-)attach the tensors to the handle; ``omit`` names optional buffers to leave unbound (NULL) -- tests."""
        b = _capi.SalpBuffers()
        for name in ("body", "ctrl", "ep_return", "food", "obs", "reward", "terminated", "truncated",
                     "events", "final_obs", "final_info", "stats", "food_words", "jit_words"):
            setattr(b, name, None if name in omit else getattr(self, name).data_ptr())
        b.table_rows, b.food_len, b.jit_len = self.table_rows, self.food_table_len, self.jitter_table_len
        b.max_substeps = self.max_substeps
        _capi.check(self._L.salp_bind(self._handle, C.byref(b)))
        self.refresh_rng()

    def refresh_rng(self) -> None:
        """Re-read the cached jitter word (ctrl rec1.w) after the tables or the cursors in ``ctrl`` changed."""
        with torch.cuda.device(self.device):
            _capi.check(self._L.salp_refresh_rng(self._handle, self._stream()))

    def _stream(self) -> C.c_void_p:
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    @property
    def launch_count(self) -> int:
        return int(self._L.salp_launch_count(self._handle))

    # ------------------------------------------------------------------ gymnasium API
    @property
    def output_rows(self) -> int:
        """Rows of the [max_substeps, N, ...] output tensors (``step(..., slot=r)`` writes row r)."""
        return self.max_substeps

    def reset(self, *, seed: Optional[int] = None, options: Optional[dict] = None, mask: Optional[torch.Tensor] = None):
        """Reset all envs (or those where ``mask`` is set).  ``seed`` re-seeds the RNG word tables
        (rows seed+env_id0+r) and rewinds the cursors; without it the streams continue, exactly as
        the reference keeps consuming its global generators across episodes."""
        rewind = 0
        if seed is not None:
            self.seed_base = int(seed)
            self._load_tables(self.seed_base)
            self._bind()
            rewind = 1
        mptr = None
        if mask is not None:
            mask = mask.to(device=self.device, dtype=torch.uint8).contiguous()
            mptr = C.c_void_p(mask.data_ptr())
        with torch.cuda.device(self.device):
            _capi.check(self._L.salp_reset(self._handle, mptr, rewind, self._stream()))
        self._needs_reset = False
        return self.obs[0], {}

    def _stage_actions(self, actions, k_steps: int) -> torch.Tensor:
        """Return a contiguous device tensor [k, N, act_dim] of this build's real type.  A CUDA
        tensor that already has that form is used IN PLACE (its pointer goes to the kernel);
        anything else is copied into the env's own staging buffer."""
        shape = (k_steps, self.num_envs, self.act_dim)
        if (isinstance(actions, torch.Tensor) and actions.device == self.device and actions.dtype == self.real
                and actions.is_contiguous() and actions.numel() == k_steps * self.num_envs * self.act_dim):
            return actions
        buf = self.actions[:k_steps]
        if isinstance(actions, torch.Tensor):
            buf.copy_(actions.reshape(shape), non_blocking=True)
        else:
            buf.copy_(torch.as_tensor(np.asarray(actions)).reshape(shape), non_blocking=True)
        return buf

    def step(self, actions, slot: int = 0):
        """One env step for all N envs.  ``actions``: a contiguous [N, act_dim] tensor of this
        build's dtype on this device is consumed in place (zero copy: its pointer is passed to
        the kernel); anything else (numpy, other dtype/device) is staged through a copy.
        Returns views of the bound output buffers (overwritten by the next step)."""
        if self._needs_reset:
            raise RuntimeError("call reset() before step()")
        a = self._stage_actions(actions, 1)
        with torch.cuda.device(self.device):
            if slot == 0:
                _capi.check(self._L.salp_step(self._handle, 1, a.data_ptr(), self._stream()))
            else:  # outputs go to row `slot` of the [max_substeps, N, ...] buffers (ping-pong, see rollout.py)
                _capi.check(self._L.salp_step_slot(self._handle, int(slot), a.data_ptr(), self._stream()))
        return self.obs[slot], self.reward[slot], self.terminated[slot], self.truncated[slot], StepInfos(self, slot)

    def observe(self, slot: int = 0) -> torch.Tensor:
        """Observation of the CURRENT state without stepping: the reference's ``_get_extended_observation()``
        (salp_snake_env.py:366-428) as its callers use it after poking robot_pos / robot_angle / food_positions
        (eval/collect_navigation_data.py:76-89).  Writes and returns ``obs[slot]``."""
        with torch.cuda.device(self.device):
            _capi.check(self._L.salp_observe(self._handle, int(slot), self._stream()))
        return self.obs[slot]

    @property
    def base_num_food_items(self) -> int:
        return self._base_num_food_items

    @base_num_food_items.setter
    def base_num_food_items(self, n: int) -> None:
        """``env.base_num_food_items = n`` (continuous_trainer.py:408-411): episodes that start after this use n foods."""
        _capi.check(self._L.salp_set_num_food_items(self._handle, int(n)))
        self._base_num_food_items = max(0, int(n))

    def step_many(self, actions, k_steps: Optional[int] = None):
        """K fused env steps in ONE launch (open-loop actions [K, N, act_dim]): state stays in
        registers across the K substeps.  Returns [K, N, ...] views."""
        k_steps = int(k_steps or (actions.shape[0] if hasattr(actions, "shape") else self.max_substeps))
        if not 1 <= k_steps <= self.max_substeps:
            raise ValueError(f"k_steps={k_steps} outside 1..max_substeps={self.max_substeps}")
        if self._needs_reset:
            raise RuntimeError("call reset() before step()")
        a = self._stage_actions(actions, k_steps)
        with torch.cuda.device(self.device):
            _capi.check(self._L.salp_step(self._handle, k_steps, a.data_ptr(), self._stream()))
        return (self.obs[:k_steps], self.reward[:k_steps], self.terminated[:k_steps], self.truncated[:k_steps],
                {"events": self.events[:k_steps], "final_observation": self.final_obs[:k_steps],
                 "final_info": self.final_info[:k_steps]})

    def step_host(self, actions_host: torch.Tensor, obs_host: torch.Tensor, reward_host: torch.Tensor,
                  terminated_host: torch.Tensor, truncated_host: torch.Tensor) -> None:
        """Reference-style call with HOST buffers (pinned CPU tensors): H2D actions, one step,
        D2H obs/reward/terminated/truncated, pipelined in chunks inside the library; returns when
        the results are in host memory."""
        if self._needs_reset:
            raise RuntimeError("call reset() before step()")
        for t in (actions_host, obs_host, reward_host, terminated_host, truncated_host):
            assert t.device.type == "cpu" and t.is_contiguous()
        assert actions_host.dtype == self.real and reward_host.dtype == self.real and obs_host.dtype == torch.float32
        _capi.check(self._L.salp_step_host(self._handle, self.actions.data_ptr(), actions_host.data_ptr(), obs_host.data_ptr(),
                                           reward_host.data_ptr(), terminated_host.data_ptr(),
                                           truncated_host.data_ptr(), self._stream()))

    def step_host_noalloc(self, actions_host: torch.Tensor, obs_host: torch.Tensor, reward_host: torch.Tensor,
                          terminated_host: torch.Tensor, truncated_host: torch.Tensor) -> None:
        """step_host for callers that may not allocate device memory: the action staging lives in a 1 MiB static
        device array inside the library (salp_step_host_noalloc; num_envs * act_dim * itemsize must fit), the work
        runs on the current stream alone.  Calls that overlap on one GPU race on that array."""
        if self._needs_reset:
            raise RuntimeError("call reset() before step()")
        for t in (actions_host, obs_host, reward_host, terminated_host, truncated_host):
            assert t.device.type == "cpu" and t.is_contiguous()
        assert actions_host.dtype == self.real and reward_host.dtype == self.real and obs_host.dtype == torch.float32
        _capi.check(self._L.salp_step_host_noalloc(self._handle, actions_host.data_ptr(), obs_host.data_ptr(),
                                                   reward_host.data_ptr(), terminated_host.data_ptr(),
                                                   truncated_host.data_ptr(), self._stream()))

    def capture_graph(self, actions, k_steps: int = 1):
        """Capture step launches into a CUDA graph and return a zero-argument replay function.
        ``actions``: one tensor (one launch per replay, read in place) or a list of tensors (one launch per
        tensor, in order, per replay).  For small slabs (<= ~256k envs) a step is a few microseconds of device
        time, less than the Python/ctypes launch path; replaying the graph removes that overhead
        (SURVEY.md section 7 step 6).  The caller refills the action tensors between replays."""
        acts = list(actions) if isinstance(actions, (list, tuple)) else [actions]
        staged = [self._stage_actions(a, k_steps) for a in acts]
        if any(s.data_ptr() != a.data_ptr() for s, a in zip(staged, acts)):
            raise ValueError("capture_graph needs contiguous device tensors of this build's dtype")
        s = torch.cuda.Stream(self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):  # warm-up launch outside capture (module load, smem attribute)
            _capi.check(self._L.salp_step(self._handle, k_steps, staged[0].data_ptr(), C.c_void_p(s.cuda_stream)))
        torch.cuda.current_stream(self.device).wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for a in staged:
                _capi.check(self._L.salp_step(self._handle, k_steps, a.data_ptr(), self._stream()))
        return g.replay

    def step_async(self, actions):
        self._pending = actions

    def step_wait(self, **kwargs):
        return self.step(self._pending)

    @property
    def action_buffer(self) -> torch.Tensor:
        """[N, act_dim] device tensor the kernel reads; write actions here to avoid a copy."""
        return self.actions[0]

    def close_extras(self, **kwargs):
        if getattr(self, "_handle", None) and self._handle.value:
            self._L.salp_destroy(self._handle)
            self._handle = C.c_void_p()

    def close(self, **kwargs):
        self.close_extras()
        self.closed = True

    def __del__(self):
        try:
            self.close_extras()
        except Exception:
            pass

    # ------------------------------------------------------------------ episode statistics
    def episode_stats(self, reduce_group=None, reset: bool = False) -> dict:
        """Accumulated on-device statistics.  With ``reduce_group`` (or an initialised default
        process group) the 8-double vector is all-reduced (SUM) over ranks with NCCL -- the only
        collective on this path."""
        from .sharding import all_reduce_stats
        out = all_reduce_stats(self.stats, reduce_group)
        if reset:
            self.stats.zero_()
        return out

    def stats_host(self, reset: bool = False) -> dict:
        """The same eight accumulators through the C ABI (salp_stats): copied to host memory after the work queued on the
        current stream; this rank only (episode_stats all-reduces over ranks)."""
        import numpy as np
        from .sharding import STATS_KEYS
        out = np.zeros(8, dtype=np.float64)
        with torch.cuda.device(self.device):
            _capi.check(self._L.salp_stats(self._handle, out.ctypes.data, 1 if reset else 0, self._stream()))
        return dict(zip(STATS_KEYS, out.tolist()))

    # ------------------------------------------------------------------ state access (tensor views)
    # The reference's callers read and poke these attributes directly
    # (eval/collect_navigation_data.py:76-89, src/salp/training/continuous_trainer.py:380,409-411).
    @property
    def robot_pos(self): return self.body[0, :, 0:2]
    @property
    def robot_velocity(self): return self.body[0, :, 2:4]
    @property
    def robot_angle(self): return self.body[1, :, 0]
    @property
    def robot_angular_velocity(self): return self.body[1, :, 1]
    @property
    def nozzle_angle(self): return self.body[1, :, 2]
    @property
    def water_volume(self): return self.body[1, :, 3]
    @property
    def steps_since_food(self): return self.ctrl[0, :, 1]
    @property
    def food_collected(self): return self.ctrl[0, :, 2]
    @property
    def food_mask(self): return self.ctrl[0, :, 3]
    @property
    def breathing_phase(self): return self.ctrl[0, :, 0] & 3
    @property
    def breathing_timer(self): return (self.ctrl[0, :, 0] >> 2) & 0xFF
    @property
    def current_exhale_duration(self): return (self.ctrl[0, :, 0] >> 10) & 0xFF
    @property
    def num_food_items_per_env(self): return (self.ctrl[1, :, 2] >> 26) & 0x3F
    @property
    def rng_cursors(self): return self.ctrl[1, :, 0:2]
    @property
    def episode_length(self): return self.ctrl[1, :, 2] & 0x3FFFFFF
    @property
    def score(self): return self.ctrl[0, :, 2].to(torch.float64) * float(self.params.food_reward)

    @property
    def food_positions(self):
        """([N, F, 2] positions, [N, F] bool alive) -- the reference's list with None for eaten food."""
        pos = self.food.permute(1, 0, 2)
        bits = torch.arange(self.food_slots, device=self.device, dtype=torch.int32)
        alive = ((self.ctrl[0, :, 3].unsqueeze(1) >> bits) & 1).bool()
        return pos, alive

    def ellipse_axes(self):
        """(ellipse_a, ellipse_b) [N] float64, recomputed from the fsm shape fields with the
        reference's expressions (salp_robot.py:190-191,207-213,240-246)."""
        w = self.ctrl[0, :, 0]
        kind, st, exd = (w >> 18) & 3, ((w >> 20) & 0xFF).double(), ((w >> 10) & 0xFF).double()
        R = 30.0
        a_rest, b_rest, a_full, b_full = R * 1.3, R * 0.8, R * 1.1, R * 1.1
        inh = kind == 2
        denom = torch.where(inh, torch.full_like(st, 120.0), torch.clamp(exd, min=1.0))
        p = st / denom
        a_m = torch.where(inh, a_rest + (a_full - a_rest) * p, a_full + (a_rest - a_full) * p)
        b_m = torch.where(inh, b_rest + (b_full - b_rest) * p, b_full + (b_rest - b_full) * p)
        a = torch.where(kind == 0, torch.full_like(p, R), torch.where(kind == 1, torch.full_like(p, a_rest), a_m))
        b = torch.where(kind == 0, torch.full_like(p, R), torch.where(kind == 1, torch.full_like(p, b_rest), b_m))
        return a, b

    def get_attr(self, name: str):
        if name == "ellipse_a":
            return self.ellipse_axes()[0]
        if name == "ellipse_b":
            return self.ellipse_axes()[1]
        if name == "body_radius":
            return torch.full((self.num_envs,), 30.0, device=self.device, dtype=torch.float64)
        return getattr(self, name)

    def set_attr(self, name: str, values) -> None:
        """Overwrite per-env state (eval/collect_navigation_data.py:76-89 pattern)."""
        if name == "food_positions":  # ([N, F', 2] positions, optional [N, F'] alive)
            pos, alive = values if isinstance(values, (tuple, list)) else (values, None)
            pos = torch.as_tensor(pos, dtype=self.real, device=self.device)
            nf = pos.shape[1]
            if nf > self.food_slots:
                raise ValueError("more food than food_slots")
            self.food[:nf] = pos.permute(1, 0, 2)
            if alive is None:
                alive = torch.ones(pos.shape[:2], dtype=torch.bool, device=self.device)
            alive = torch.as_tensor(alive, device=self.device).to(torch.int64)
            bits = (alive << torch.arange(nf, device=self.device)).sum(1)
            self.ctrl[0, :, 3] = torch.where(bits >= 2 ** 31, bits - 2 ** 32, bits).to(torch.int32)
            w = (self.ctrl[1, :, 2].to(torch.int64) & 0x3FFFFFF) | (int(nf) << 26)
            self.ctrl[1, :, 2] = torch.where(w >= 2 ** 31, w - 2 ** 32, w).to(torch.int32)
            return
        if name == "base_num_food_items":
            self.base_num_food_items = int(values)
            return
        target = getattr(self, name)
        if not isinstance(target, torch.Tensor):
            raise AttributeError(f"{name!r} is not per-env state; pass it to the constructor")
        target.copy_(torch.as_tensor(values, device=self.device).to(target.dtype).expand_as(target))

    def call(self, name: str, *args, **kwargs):
        return getattr(self, name)(*args, **kwargs)

    def state_dict(self) -> dict:
        """Env-state checkpoint: the SoA tensors, cloned."""
        return {k: getattr(self, k).clone() for k in ("body", "ctrl", "ep_return", "food", "stats")} | {
            "seed_base": self.seed_base}

    def load_state_dict(self, sd: dict) -> None:
        for k in ("body", "ctrl", "ep_return", "food", "stats"):
            getattr(self, k).copy_(sd[k])
        if sd.get("seed_base", self.seed_base) != self.seed_base:
            self.seed_base = int(sd["seed_base"])
            self._load_tables(self.seed_base)
            self._bind()
        self.refresh_rng()
        self._needs_reset = False


class StepInfos(dict):
    """Gymnasium-0.29-style vector ``infos`` whose entries are computed on first access so a step
    costs no extra kernels unless the caller looks.  Keys (salp_snake_env.py:195-200 plus the
    vector auto-reset keys): final_observation, _final_observation, final_info, events, collision,
    food_collected, steps_since_food, score."""

    _KEYS = ("final_observation", "_final_observation", "final_info", "_final_info", "events", "collision",
             "_collision", "food_eaten", "food_collected", "steps_since_food", "score")

    def __init__(self, env: SalpVectorEnv, k: int):
        super().__init__()
        self._env, self._k = env, k

    def __missing__(self, key):
        e, k = self._env, self._k
        if key == "final_observation":
            v = e.final_obs[k]
        elif key in ("_final_observation", "_final_info"):
            v = e.terminated[k] | e.truncated[k]
        elif key == "final_info":  # columns: episode return, length, food_collected, steps_since_food
            v = e.final_info[k]
        elif key == "events":
            v = e.events[k]
        elif key == "collision":
            v = (e.events[k] & _capi.EV_COLLISION) != 0
        elif key == "_collision":
            v = torch.ones(e.num_envs, dtype=torch.bool, device=e.device)
        elif key == "food_eaten":
            v = (e.events[k] & _capi.EV_FOOD) != 0
        elif key == "food_collected":
            v = e.food_collected
        elif key == "steps_since_food":
            v = e.steps_since_food
        elif key == "score":
            v = e.score
        else:
            raise KeyError(key)
        self[key] = v
        return v

    def keys(self):
        return list(self._KEYS)

    def __contains__(self, key):
        return key in self._KEYS
