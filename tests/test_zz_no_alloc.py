"""salp_step_host_noalloc (include/salp_b200.h): the host-buffer step whose device staging is a static array inside
the library, called through the C ABI (ctypes) and compared bit for bit with the existing entry point salp_step_host
on twin environments -- both run the same step kernel on the same actions, so the bar is equality, the bar of
test_gpu_parity.py::test_step_host_equals_device_path.

Kept in a file of its own that pytest collects last.  Every test builds its own
environments, so the file passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import pytest
import torch

from helpers import CONFIGS, synth_actions

pytestmark = pytest.mark.gpu


def _env(n, params, dtype, **kw):
    from underwater_swimmer_rl_b200 import SalpVectorEnv
    return SalpVectorEnv(n, device="cuda:0", dtype=dtype, **kw, **params)


class _HostBuffers:
    def __init__(self, env):
        n = env.num_envs
        self.a = torch.empty(n, env.act_dim, dtype=env.real).pin_memory()
        self.o = torch.full((n, env.obs_dim), float("nan"), dtype=torch.float32).pin_memory()
        self.r = torch.full((n,), float("nan"), dtype=env.real).pin_memory()
        self.te = torch.zeros(n, dtype=torch.bool).pin_memory()
        self.tr = torch.zeros(n, dtype=torch.bool).pin_memory()

    def outputs(self):
        return self.o, self.r, self.te, self.tr


def _stream(env):
    return C.c_void_p(torch.cuda.current_stream(env.device).cuda_stream)


def _step_existing(lib, env, hb):
    """the existing entry point: caller-owned device staging (env.actions)"""
    rc = lib.salp_step_host(env._handle, env.actions.data_ptr(), hb.a.data_ptr(), hb.o.data_ptr(), hb.r.data_ptr(),
                            hb.te.data_ptr(), hb.tr.data_ptr(), _stream(env))
    assert rc == 0, lib.salp_last_error().decode()


def _step_noalloc(lib, env, hb):
    """the new entry point: no staging argument, nothing allocated"""
    rc = lib.salp_step_host_noalloc(env._handle, hb.a.data_ptr(), hb.o.data_ptr(), hb.r.data_ptr(), hb.te.data_ptr(),
                                    hb.tr.data_ptr(), _stream(env))
    assert rc == 0, lib.salp_last_error().decode()


def _assert_same(hb_new, hb_old, what):
    for name, x, y in zip(("obs", "reward", "terminated", "truncated"), hb_new.outputs(), hb_old.outputs()):
        assert torch.equal(x, y), f"{what}: {name} differs from salp_step_host"
    assert torch.isfinite(hb_new.o).all() and torch.isfinite(hb_new.r).all(), f"{what}: outputs were not written"


CASES = [
    ("single_food", "float32", 5000, dict()),                       # forced-breathing kernel, table RNG
    ("defaults", "float32", 4096, dict(rng="philox")),              # twelve foods, in-kernel Philox
    ("single_food", "float64", 1000, dict()),                       # validation build
    ("manual", "float64", 777, dict()),                             # two action columns, generic kernel, ragged N
]


@pytest.mark.parametrize("config,dtype,n,kw", CASES, ids=[f"{c[0]}-{c[1]}-{c[2]}" for c in CASES])
def test_noalloc_entry_point_equals_step_host(salp_lib, config, dtype, n, kw):
    params = dict(CONFIGS["single_food"], forced_breathing=False) if config == "manual" else CONFIGS[config]
    T = 24
    new, old = _env(n, params, dtype, seed=77, **kw), _env(n, params, dtype, seed=77, **kw)
    new.reset()
    old.reset()
    torch.cuda.synchronize()
    hb_new, hb_old = _HostBuffers(new), _HostBuffers(old)
    acts = synth_actions(T, n, new.act_dim, seed=5)
    free_after_first = None
    for t in range(T):  # "two calls in a row" and then some: every call is checked
        a = torch.from_numpy(acts[t]).to(new.real).reshape(n, new.act_dim)
        hb_new.a.copy_(a)
        hb_old.a.copy_(a)
        _step_noalloc(salp_lib, new, hb_new)
        _step_existing(salp_lib, old, hb_old)
        _assert_same(hb_new, hb_old, f"{config}/{dtype} step {t}")
        if t == 1:
            free_after_first = torch.cuda.mem_get_info(new.device)[0]
    # device memory in use did not move from the second call on (the first call of a process may load modules)
    assert torch.cuda.mem_get_info(new.device)[0] == free_after_first
    assert torch.equal(new.body, old.body) and torch.equal(new.ctrl, old.ctrl) and torch.equal(new.food, old.food)


def test_noalloc_calls_share_the_array_without_leaking_between_handles(salp_lib):
    """Two handles of different sizes and dtypes take turns on the one static array: a call reads only the records
    it has just copied in, so what the other handle's (larger) call left behind does not matter."""
    big = [_env(8192, CONFIGS["single_food"], "float64", seed=3) for _ in range(2)]
    small = [_env(1000, CONFIGS["defaults"], "float32", seed=4) for _ in range(2)]
    for e in big + small:
        e.reset()
    torch.cuda.synchronize()
    hb = {id(e): _HostBuffers(e) for e in big + small}
    a_big = synth_actions(6, 8192, 1, seed=8)
    a_small = synth_actions(6, 1000, 1, seed=9)
    for t in range(6):
        for pair, acts in ((big, a_big), (small, a_small)):
            for e in pair:
                hb[id(e)].a.copy_(torch.from_numpy(acts[t]).to(e.real).reshape(e.num_envs, 1))
            _step_noalloc(salp_lib, pair[0], hb[id(pair[0])])
            _step_existing(salp_lib, pair[1], hb[id(pair[1])])
            _assert_same(hb[id(pair[0])], hb[id(pair[1])], f"interleaved step {t}")


def test_noalloc_rejects_sizes_beyond_the_static_array(salp_lib):
    from underwater_swimmer_rl_b200 import _capi
    limit = _capi.SALP_NOALLOC_STAGING_BYTES
    assert limit == 1 << 20
    # exactly at the limit: 262144 f32 action records
    fits = _env(limit // 4, CONFIGS["single_food"], "float32", rng="philox")
    fits.reset()
    hb = _HostBuffers(fits)
    hb.a.zero_()
    _step_noalloc(salp_lib, fits, hb)
    assert torch.isfinite(hb.o).all()
    # one tile more: refused on the host, nothing launched, outputs untouched
    over = _env(limit // 4 + 32, CONFIGS["single_food"], "float32", rng="philox")
    over.reset()
    torch.cuda.synchronize()
    hb = _HostBuffers(over)
    hb.a.zero_()
    launches = salp_lib.salp_launch_count(over._handle)
    rc = salp_lib.salp_step_host_noalloc(over._handle, hb.a.data_ptr(), hb.o.data_ptr(), hb.r.data_ptr(), hb.te.data_ptr(),
                                         hb.tr.data_ptr(), _stream(over))
    assert rc == _capi.SALP_EUNSUPPORTED
    msg = salp_lib.salp_last_error().decode()
    assert "SALP_NOALLOC_STAGING_BYTES" in msg and str(limit) in msg and "salp_step_host" in msg, msg
    assert salp_lib.salp_launch_count(over._handle) == launches
    assert torch.isnan(hb.o).all()
    # f64 with two action columns: 16 bytes per env, 65536 envs fit, 65537 do not
    over64 = _env(limit // 16 + 1, dict(CONFIGS["single_food"], forced_breathing=False), "float64", rng="philox")
    over64.reset()
    hb = _HostBuffers(over64)
    hb.a.zero_()
    rc = salp_lib.salp_step_host_noalloc(over64._handle, hb.a.data_ptr(), hb.o.data_ptr(), hb.r.data_ptr(), hb.te.data_ptr(),
                                         hb.tr.data_ptr(), _stream(over64))
    assert rc == _capi.SALP_EUNSUPPORTED
    # and the python mirror raises
    with pytest.raises(_capi.SalpError):
        over64.step_host_noalloc(hb.a, hb.o, hb.r, hb.te, hb.tr)
    # the allocating entry point still serves that size
    _step_existing(salp_lib, over64, hb)
    assert torch.isfinite(hb.o).all()
