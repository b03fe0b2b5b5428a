"""CPU checks of the drop-in boundary: the shared library loads and exports every symbol that
include/blitzdg_b200.h declares (no compute calls without a GPU)."""
import ctypes
import os
import re

import pytest

from blitzdg_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "blitzdg_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bdg_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert _declared() == sorted(capi.EXPORTS)


def test_library_exports_every_declared_symbol():
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__

        __graft_entry__.build()
    lib = capi.load()
    for name in _declared():
        assert hasattr(lib, name), name
    assert lib.bdg_version() >= 100


def test_create_fails_loudly_without_gpu_or_with_bad_context():
    """No CPU fallback: on a box without CUDA, create returns BDG_E_CUDA; NULL context is BDG_E_INVALID."""
    lib = capi.load()
    h = ctypes.c_void_p()
    assert lib.bdg_sw2d_create(ctypes.byref(h), None, None, None, 0) == capi.BDG_E_INVALID
    assert b"NULL" in lib.bdg_last_error(None)


def test_struct_layout_matches_c():
    """sizeof of the ctypes mirrors equals the C structs (x86-64 SysV): 5*4 pad-> 24, then pointers."""
    assert ctypes.sizeof(capi.bdg_sw2d_params) == 4 * 8 + 4 * 4 + 3 * 8
    assert ctypes.sizeof(capi.bdg_dgcontext2d) == 24 + 13 * 8 + 9 * 8 + 9 * 4 + 4
    assert ctypes.sizeof(capi.bdg_halo) == 16 + 4 * 8 + 8
    assert ctypes.sizeof(capi.bdg_curved2d) == 8 + 10 * 8 + 8 + 9 * 8  # 2 ints, 10 pointers, 2 ints, 9 pointers


def test_null_handle_is_rejected_by_every_entry_point():
    """The reference-facing calls validate their handle before touching CUDA (BDG_E_INVALID, no crash, no GPU needed)."""
    lib = capi.load()
    nul = ctypes.c_void_p()
    dp = ctypes.POINTER(ctypes.c_double)()
    calls = {
        "bdg_sw2d_set_bathymetry": (nul, dp, dp, dp), "bdg_sw2d_set_state": (nul, dp, dp, dp, 0.0),
        "bdg_sw2d_get_state": (nul, dp, dp, dp, dp), "bdg_sw2d_rhs": (nul, dp, dp, dp), "bdg_sw2d_timestep": (nul, dp),
        "bdg_sw2d_step": (nul, 1, 0.1), "bdg_sw2d_step_ssprk2": (nul, 1, 0.1), "bdg_sw2d_sync": (nul,), "bdg_sw2d_check": (nul, dp),
        "bdg_sw2d_set_sponge": (nul, dp), "bdg_sw2d_set_output": (nul, dp, 3), "bdg_sw2d_output_begin": (nul, dp, dp, dp),
        "bdg_sw2d_output_wait": (nul,), "bdg_sw2d_limit_positivity": (nul,), "bdg_sw2d_stage_state": (nul, dp, dp, dp, 0.0),
        "bdg_sw2d_set_curved": (nul, None), "bdg_sw2d_rhs_curved": (nul, dp, dp, dp, dp, dp), "bdg_sw2d_step_curved": (nul, 1, 0.1, dp),
        "bdg_sw2d_commit_state": (nul,), "bdg_sw2d_get_state_async": (nul, dp, dp, dp), "bdg_sw2d_get_state_wait": (nul,),
        "bdg_sw2d_set_variant": (nul, 1), "bdg_sw2d_comm_init": (nul, nul), "bdg_sw2d_halo_pack_to_host": (nul, dp),
        "bdg_sw2d_halo_set_ghosts": (nul, dp), "bdg_sw2d_set_tracer": (nul, dp), "bdg_sw2d_get_tracer": (nul, dp),
        "bdg_sw2d_rhs_tracer": (nul, dp),
    }
    for name, args in calls.items():
        assert getattr(lib, name)(*args) == capi.BDG_E_INVALID, name
    assert lib.bdg_sw2d_launch_count(nul) == 0 and lib.bdg_sw2d_get_variant(nul) == -1
    lib.bdg_sw2d_destroy(nul)  # no-op
