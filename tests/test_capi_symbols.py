"""CPU: libsalp_b200.so builds, loads, exports every function include/salp_b200.h declares, and
its host-only entry points behave.  No compute call is made (there is no GPU here); creating a
handle must FAIL loudly without a CUDA device -- there is no CPU fallback."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "salp_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(salp_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(salp_lib):
    from underwater_swimmer_rl_b200 import _capi
    names = _declared()
    assert set(names) == set(_capi.EXPORTED_SYMBOLS)
    for n in names:
        assert getattr(salp_lib, n) is not None
    assert salp_lib.salp_abi_version() == _capi.SALP_ABI_VERSION


def test_struct_sizes_match_header(salp_lib):
    from underwater_swimmer_rl_b200 import _capi
    assert C.sizeof(_capi.SalpParams) == 2 * 8 + 6 * 4 + 5 * 8 + 2 * 4
    assert C.sizeof(_capi.SalpLayout) == 32
    assert C.sizeof(_capi.SalpBuffers) == 14 * 8 + 3 * 8 + 2 * 4
    assert C.sizeof(_capi.SalpReplayAppend) == 16 * 8 + 2 * 8 + 6 * 4
    assert C.sizeof(_capi.SalpActorForward) == 11 * 8 + 8 + 8 * 4 + 8 * 4
    assert C.sizeof(_capi.SalpCycleParams) == 12 * 8 + 4 * 4
    assert C.sizeof(_capi.SalpCycleBuffers) == 10 * 8 + 4 * 8 + 2 * 4


@pytest.mark.parametrize("nfood,slots", [(0, 1), (1, 1), (3, 4), (5, 8), (12, 12), (13, 16), (32, 32)])
def test_query_layout(salp_lib, nfood, slots):
    from underwater_swimmer_rl_b200 import _capi
    from underwater_swimmer_rl_b200.vector_env import make_params
    lay = _capi.SalpLayout()
    p = make_params(True, num_food_items=nfood, forced_breathing=False, max_observed_food=5)
    assert salp_lib.salp_query_layout(C.byref(p), _capi.SALP_F64, C.byref(lay)) == 0
    assert (lay.real_bytes, lay.obs_dim, lay.act_dim, lay.food_slots) == (8, 32, 2, slots)


def test_invalid_params_are_rejected_with_message(salp_lib):
    from underwater_swimmer_rl_b200 import _capi
    from underwater_swimmer_rl_b200.vector_env import make_params
    lay = _capi.SalpLayout()
    p = make_params(True, num_food_items=33)
    assert salp_lib.salp_query_layout(C.byref(p), _capi.SALP_F32, C.byref(lay)) == _capi.SALP_EUNSUPPORTED
    assert b"num_food_items" in salp_lib.salp_last_error()
    p = make_params(True)
    assert salp_lib.salp_query_layout(C.byref(p), 7, C.byref(lay)) == _capi.SALP_EINVAL
    with pytest.raises(TypeError):
        make_params(True, not_a_reference_kwarg=1)


def test_no_cpu_fallback(salp_lib):
    """Without a CUDA device salp_create must fail (SALP_ECUDA), never silently run on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the failure path is for CPU-only machines")
    from underwater_swimmer_rl_b200 import _capi
    from underwater_swimmer_rl_b200.vector_env import SalpVectorEnv, make_params
    h = C.c_void_p()
    p = make_params(True)
    rc = salp_lib.salp_create(C.byref(p), 16, 0, 0, _capi.SALP_F32, C.byref(h))
    assert rc == _capi.SALP_ECUDA and not h.value
    assert b"no CPU fallback" in salp_lib.salp_last_error()
    with pytest.raises(RuntimeError):
        SalpVectorEnv(16)
