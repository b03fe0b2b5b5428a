"""The C++ adapter of the drop-in boundary (include/blitzdg_b200_adapter.hpp) compiles against the
REFERENCE'S OWN headers (DGContext2D.hpp, BCtypes.hpp, Types.hpp, found under /root/reference/include) and links against
libblitzdg_b200.so with nothing left undefined.  blitz++ itself is absent from this image: oracle/blitz_shim stands in for
<blitz/array.h> (test infrastructure, see oracle/make_ref.py).  Skipped where the reference tree is absent (the GPU box)."""
import os
import shutil
import subprocess

import pytest

from blitzdg_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("BLITZDG_REFERENCE", "/root/reference")


def test_adapter_compiles_against_reference_headers_and_links(tmp_path):
    if not os.path.exists(os.path.join(REF, "include", "DGContext2D.hpp")):
        pytest.skip("reference tree not present")
    if not shutil.which("g++"):
        pytest.skip("g++ not available")
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__

        __graft_entry__.build()
    obj, so = str(tmp_path / "adapter_check.o"), str(tmp_path / "libadapter_check.so")
    inc = ["-I" + os.path.join(ROOT, "oracle", "blitz_shim"), "-I" + os.path.join(REF, "include"), "-I" + os.path.join(ROOT, "include")]
    r = subprocess.run(["g++", "-std=c++14", "-Wall", "-Werror=return-type", "-fPIC", "-c", os.path.join(ROOT, "tests", "adapter", "adapter_check.cpp"), "-o", obj] + inc,
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    nm = subprocess.run(["nm", "-C", obj], capture_output=True, text=True).stdout
    used = sorted({ln.split()[-1] for ln in nm.splitlines() if " U bdg_" in ln})
    assert {"bdg_sw2d_create", "bdg_sw2d_step", "bdg_sw2d_step_ssprk2", "bdg_sw2d_rhs", "bdg_sw2d_timestep", "bdg_sw2d_check"} <= set(used), used
    assert set(used) <= set(capi.EXPORTS)
    libdir = os.path.dirname(capi.LIB_PATH)
    r = subprocess.run(["g++", "-shared", "-o", so, obj, "-L" + libdir, "-l:" + os.path.basename(capi.LIB_PATH), "-Wl,--no-undefined",
                        "-Wl,--allow-shlib-undefined"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
