"""SalpCycleVectorEnv -- N independent model (B) ``SalpRobotEnv`` instances, one breathing cycle per step.

Drop-in for the reference's HEAD ``SalpRobotEnv`` (src/salp/environments/salp_robot_env.py:19-201) over
``Robot`` / ``Nozzle`` (src/salp/environments/robot.py) behind the ``gymnasium.vector.VectorEnv`` API:
3-D action (contraction, coast time, nozzle yaw) in Box([0,0,-1],[1,1,1]), 6-D observation, the reference's
reward / terminated / truncated rules, Gymnasium vector auto-reset.  ``Robot`` and ``Nozzle`` here are plain
parameter holders with the reference's constructor keywords; the dynamics run in ``libsalp_b200.so``
(salp_cycle_step, include/salp_b200.h).  All tensors live on the GPU; there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Any, Optional

import numpy as np
import torch

from . import _capi, rng_tables
from ._gym_compat import Box, VectorEnv, batch_box
from .vector_env import _torch_dtype

from .sharding import CYCLE_STATS_KEYS  # noqa: E402,F401  (re-exported)
PHASE_NAMES = ("REFILL", "JET", "COAST", "REST")  # Robot.Phase, robot.py:188-192


@dataclass
class Nozzle:
    """Parameters of the steerable nozzle (robot.py:19-25)."""
    length1: float = 0.0
    length2: float = 0.0
    length3: float = 0.0
    area: float = 0.0
    mass: float = 0.0


@dataclass
class Robot:
    """Parameters of the jet-propelled robot (robot.py:196-222; density via set_environment :279-285)."""
    dry_mass: float = 1.0
    init_length: float = 0.3
    init_width: float = 0.15
    max_contraction: float = 0.06
    nozzle: Nozzle = field(default_factory=lambda: Nozzle(0.05, 0.05, 0.05, 0.00016, 1.0))
    density: float = 1000.0

    def set_environment(self, density: float) -> None:
        self.density = float(density)


def default_robot() -> Robot:
    """The construction every reference script uses (salp_robot_env.py:1305-1310, train_robot.py:12-17)."""
    return Robot(dry_mass=1.0, init_length=0.3, init_width=0.15, max_contraction=0.06,
                 nozzle=Nozzle(length1=0.05, length2=0.05, length3=0.05, area=0.00016, mass=1.0))


def make_cycle_params(robot: Robot, width: float, height: float, autoreset: bool, max_substeps: int = 0) -> _capi.SalpCycleParams:
    p = _capi.SalpCycleParams()
    p.dry_mass, p.init_length, p.init_width = float(robot.dry_mass), float(robot.init_length), float(robot.init_width)
    p.max_contraction, p.density = float(robot.max_contraction), float(robot.density)
    nz = robot.nozzle
    p.nozzle_length1, p.nozzle_length2, p.nozzle_length3 = float(nz.length1), float(nz.length2), float(nz.length3)
    p.nozzle_area, p.nozzle_mass = float(nz.area), float(nz.mass)
    p.width, p.height = float(width), float(height)
    p.max_cycles = 500                      # salp_robot_env.py:187
    p.max_substeps = int(max_substeps)
    p.flags = _capi.SALP_CYCLE_AUTORESET if autoreset else 0
    return p


class SalpCycleVectorEnv(VectorEnv):
    """Batched model (B) environment on one B200.

    Reference keywords: ``render_mode`` (must be None), ``width``, ``height``, ``robot``.
    Beyond them: num_envs, device, dtype ("float32" throughput | "float64" validation build), seed / env_id0
    (env e draws its target points from np.random.seed(seed + env_id0 + e)), table_rows / table_len (RNG word
    table shape; rows are shared round-robin when fewer than num_envs), autoreset, max_substeps (safety cap on
    the cycle loop for actions outside the Box; 0 = 4096).
    """

    metadata = {"render_modes": [], "autoreset_mode": "next_obs_in_obs"}

    def __init__(self, num_envs: int, device: Any = "cuda:0", dtype: Any = "float32", seed: int = 10_000,
                 env_id0: int = 0, table_rows: Optional[int] = None, table_len: int = 256, autoreset: bool = True,
                 max_substeps: int = 0, render_mode: Optional[str] = None, width: int = 900, height: int = 700,
                 robot: Optional[Robot] = None, rows: int = 1):
        if render_mode is not None:
            raise ValueError("render_mode must be None: there is no graphics stack on the GPU box (SURVEY.md F11)")
        if not torch.cuda.is_available():
            raise RuntimeError("SalpCycleVectorEnv needs a CUDA device; there is no CPU fallback")
        self._L = _capi.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise ValueError("device must be a CUDA device")
        self.real = _torch_dtype(dtype)
        self._dtype_code = _capi.SALP_F64 if self.real == torch.float64 else _capi.SALP_F32
        self.num_envs, self.env_id0, self.seed_base = int(num_envs), int(env_id0), int(seed)
        self.autoreset = bool(autoreset)
        self.render_mode = None
        self.width, self.height = width, height
        self.robot = robot if robot is not None else default_robot()
        self.params = make_cycle_params(self.robot, width, height, self.autoreset, max_substeps)
        self.obs_dim, self.act_dim = _capi.SALP_CYCLE_OBS_DIM, _capi.SALP_CYCLE_ACT_DIM
        # spaces exactly as salp_robot_env.py:51-71 builds them
        self.single_action_space = Box(low=np.array([0.0, 0.0, -1.0]), high=np.array([1.0, 1.0, 1.0]), dtype=np.float32)
        self.single_observation_space = Box(low=np.full(6, -np.inf), high=np.full(6, np.inf), dtype=np.float32)
        self.observation_space = batch_box(self.single_observation_space, self.num_envs)
        self.action_space = batch_box(self.single_action_space, self.num_envs)
        self.table_rows = int(table_rows) if table_rows else self.num_envs
        self.table_len = int(table_len)
        if rows < 1:
            raise ValueError("rows must be >= 1")
        self.rows = int(rows)

        N, R, dev = self.num_envs, self.rows, self.device
        with torch.cuda.device(dev):
            self.state = torch.zeros((6, N, 4), dtype=self.real, device=dev)
            self.ctrl = torch.zeros((N, 4), dtype=torch.int32, device=dev)
            # outputs: `rows` rows each; step(actions, slot=r) writes row r (rollout.StepRingReplay keeps the last
            # rows - 1 steps in place).  obs / reward / ... are row 0, the only row of the default rows = 1.
            self.obs_rows = torch.zeros((R, N, 6), dtype=torch.float32, device=dev)
            self.reward_rows = torch.zeros((R, N), dtype=self.real, device=dev)
            self.terminated_rows = torch.zeros((R, N), dtype=torch.uint8, device=dev)
            self.truncated_rows = torch.zeros((R, N), dtype=torch.uint8, device=dev)
            self.n_sub_rows = torch.zeros((R, N), dtype=torch.int32, device=dev)
            self.final_obs_rows = torch.zeros((R, N, 6), dtype=torch.float32, device=dev)
            self.obs, self.reward = self.obs_rows[0], self.reward_rows[0]
            self.terminated, self.truncated = self.terminated_rows[0], self.truncated_rows[0]
            self.n_sub, self.final_obs = self.n_sub_rows[0], self.final_obs_rows[0]
            self.stats = torch.zeros((8,), dtype=torch.float64, device=dev)
            self.words = None
        self._buffers = None
        self.launch_count = 0
        self.set_rng_table(rng_tables.fill_words(self.table_rows, self.table_len, self.seed_base,
                                                 self.env_id0, rng_tables.JITTER_STREAM))

    # ------------------------------------------------------------------ plumbing
    def set_rng_table(self, words: np.ndarray) -> None:
        """uint32 [rows, len] of np.random MT19937 outputs (row r = stream of seed + env_id0 + r)."""
        w = np.ascontiguousarray(words, dtype=np.uint32)
        if w.ndim != 2 or w.shape[1] & (w.shape[1] - 1):
            raise ValueError("words must be [rows, power-of-two length]")
        self.table_rows, self.table_len = int(w.shape[0]), int(w.shape[1])
        self.words = torch.from_numpy(w.view(np.int32)).to(self.device)
        self._slot_buffers = [self._make_buffers(r) for r in range(self.rows)]
        self._buffers = self._slot_buffers[0]

    def _make_buffers(self, r: int) -> _capi.SalpCycleBuffers:
        """The caller-owned buffer set with the outputs of output row ``r`` (the ABI is stateless: a row is just
        another set of pointers)."""
        b = _capi.SalpCycleBuffers()
        b.state, b.ctrl, b.stats, b.words = self.state.data_ptr(), self.ctrl.data_ptr(), self.stats.data_ptr(), self.words.data_ptr()
        b.obs, b.reward = self.obs_rows[r].data_ptr(), self.reward_rows[r].data_ptr()
        b.terminated, b.truncated = self.terminated_rows[r].data_ptr(), self.truncated_rows[r].data_ptr()
        b.n_sub, b.final_obs = self.n_sub_rows[r].data_ptr(), self.final_obs_rows[r].data_ptr()
        b.table_rows, b.words_len, b.n = self.table_rows, self.table_len, self.num_envs
        # rows are indexed (env_id0 + e) % table_rows inside the kernel; the table passed here starts at this slab's row 0
        b.env_id0 = 0
        b.dtype = self._dtype_code
        return b

    @property
    def output_rows(self) -> int:
        return self.rows

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ------------------------------------------------------------------ Gymnasium vector API
    def reset(self, *, seed: Optional[int] = None, options: Optional[dict] = None, mask=None):
        """Reset all (or masked) envs.  ``seed`` rebuilds the RNG table from that base seed; every full reset
        rewinds the cursors so that env e replays np.random.seed(seed + env_id0 + e)."""
        if seed is not None:
            self.seed_base = int(seed)
            self.set_rng_table(rng_tables.fill_words(self.table_rows, self.table_len, self.seed_base,
                                                     self.env_id0, rng_tables.JITTER_STREAM))
        mptr = None
        if mask is not None:
            m = torch.as_tensor(mask, device=self.device).to(torch.uint8).contiguous()
            mptr = C.c_void_p(m.data_ptr())
        with torch.cuda.device(self.device):
            _capi.check(self._L.salp_cycle_reset(C.byref(self.params), C.byref(self._buffers), mptr,
                                                 1 if mask is None else 0, self._stream()))
        self.launch_count += 1
        return self.obs, {}

    def step(self, actions, slot: int = 0):
        """One breathing cycle per env.  actions [N, 3] (torch on the device, or numpy); the outputs go to output
        row ``slot`` (0 unless the env was built with ``rows`` > 1)."""
        if not 0 <= slot < self.rows:
            raise ValueError(f"slot={slot} outside 0..rows-1={self.rows - 1}")
        a = actions
        if not torch.is_tensor(a):
            a = torch.as_tensor(np.asarray(a), device=self.device)
        a = a.to(device=self.device, dtype=self.real).contiguous()
        if a.shape != (self.num_envs, 3):
            raise ValueError(f"actions must have shape ({self.num_envs}, 3), got {tuple(a.shape)}")
        with torch.cuda.device(self.device):
            _capi.check(self._L.salp_cycle_step(C.byref(self.params), C.byref(self._slot_buffers[slot]),
                                                C.c_void_p(a.data_ptr()), self._stream()))
        self.launch_count += 1
        info = {"n_sub": self.n_sub_rows[slot], "final_obs": self.final_obs_rows[slot]}
        return (self.obs_rows[slot], self.reward_rows[slot], self.terminated_rows[slot], self.truncated_rows[slot], info)

    def step_host(self, actions_h, obs_h, reward_h, terminated_h, truncated_h):
        """The reference-facing call for callers holding host arrays (what SB3's DummyVecEnv does with one env):
        pinned host actions -> device, one cycle per env, results -> pinned host buffers; returns when they are
        filled.  All five arguments are CPU tensors (pin them for full PCIe rate)."""
        if not hasattr(self, "_a_dev"):
            self._a_dev = torch.empty((self.num_envs, 3), dtype=self.real, device=self.device)
        self._a_dev.copy_(actions_h, non_blocking=True)
        self.step(self._a_dev)
        obs_h.copy_(self.obs, non_blocking=True)
        reward_h.copy_(self.reward, non_blocking=True)
        terminated_h.copy_(self.terminated.view(torch.bool) if terminated_h.dtype == torch.bool else self.terminated, non_blocking=True)
        truncated_h.copy_(self.truncated.view(torch.bool) if truncated_h.dtype == torch.bool else self.truncated, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()

    def close(self, **kwargs):
        self._buffers = None

    # ------------------------------------------------------------------ reference-visible state (views)
    @property
    def position(self): return self.state[0, :, :3]          # Robot.position (robot.py:247)
    @property
    def velocity(self): return self.state[1, :, :3]          # body frame (robot.py:248)
    @property
    def angular_velocity(self): return self.state[2, :, :3]
    @property
    def euler_angle(self): return self.state[3, :, :3]
    @property
    def velocity_world(self): return self.state[5, :, :3]
    @property
    def target_point(self): return self.ctrl[:, 2:4].view(torch.float32)
    @property
    def cycle(self): return self.ctrl[:, 0] & 0xFFFF
    @property
    def phase(self): return self.ctrl[:, 0] >> 16
    @property
    def prev_dist(self): return self.state[0, :, 3]
    @property
    def episode_return(self): return self.state[2, :, 3]

    def episode_stats(self, reduce_group=None, reset: bool = False) -> dict:
        """Counters accumulated by the kernel since the last reset=True call.  With an initialised process group the
        8-double vector is all-reduced (SUM) over ranks -- the path's only collective."""
        from .sharding import CYCLE_STATS_KEYS as keys, all_reduce_stats
        out = all_reduce_stats(self.stats, reduce_group, keys)
        if reset:
            self.stats.zero_()
        return out
