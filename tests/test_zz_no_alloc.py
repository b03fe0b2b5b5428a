"""bdg_sw2d_timestep_noalloc (include/blitzdg_b200.h): the allocation-free time-step reduction against the handle-based
bdg_sw2d_timestep, both through the C ABI.  The entry point has no counterpart in the reference; it was added
for callers that cannot call cudaMalloc.

Collected last on purpose.  The GPU test must pass alone in a fresh process and after
every other test: it makes two pairs of calls in a row on DIFFERENT states, the second with the smaller wave speed, so a
maximum left in the static scratch array by the first call would show.
"""
import ctypes as C

import numpy as np
import pytest

from blitzdg_b200 import capi

TOL = 1e-12  # relative, the RHS tolerance of the other GPU tests; the two reductions evaluate the same expression


def _call(lib, device, ctx, g, CFL, ptrs, K=None, stream=None):
    err = C.create_string_buffer(512)
    rc = lib.bdg_sw2d_timestep_noalloc(device, ctx["N"], ctx["Np"], ctx["Nfp"], ctx["Nfaces"], ctx["K"] if K is None else K, g, CFL,
                                       *ptrs, stream, err, len(err))
    return rc, err.value.decode()


def test_sizes_beyond_the_static_scratch_are_rejected_with_a_clear_error():
    """Argument checks come before any CUDA call: no GPU needed."""
    lib = capi.load()
    dims = dict(N=4, Np=15, Nfp=5, Nfaces=3)
    fake = [C.c_void_p(256)] * 6
    K = capi.BDG_NOALLOC_MAX_FACE_NODES // 15 + 1
    rc, msg = _call(lib, 0, dict(dims, K=K), 9.81, 0.25, fake)
    assert rc == capi.BDG_E_UNSUPPORTED
    assert "BDG_NOALLOC_MAX_FACE_NODES" in msg and str(15 * K) in msg and str(capi.BDG_NOALLOC_MAX_FACE_NODES) in msg
    rc, msg = _call(lib, 0, dict(dims, K=100), 9.81, 0.25, [C.c_void_p(256)] * 5 + [None])
    assert rc == capi.BDG_E_INVALID and "NULL" in msg
    rc, msg = _call(lib, 0, dict(dims, K=0), 9.81, 0.25, fake)
    assert rc == capi.BDG_E_INVALID and "K=0" in msg
    rc, msg = _call(lib, 64, dict(dims, K=100), 9.81, 0.25, fake)
    assert rc == capi.BDG_E_INVALID and "device" in msg
    assert lib.bdg_sw2d_timestep_noalloc(0, 4, 15, 5, 3, K, 9.81, 0.25, *fake, None, None, 0) == capi.BDG_E_UNSUPPORTED  # err may be NULL


@pytest.mark.gpu
@pytest.mark.parametrize("shape,N,nx,ny", [("tri", 4, 96, 80), ("quad", 3, 61, 47)], ids=["tri-N4", "quad-N3"])
def test_no_alloc_timestep_matches_the_handle_entry_point_twice_in_a_row(shape, N, nx, ny):
    import torch

    from blitzdg_b200.initial_conditions import gaussian_hump
    from blitzdg_b200.sw2d import SW2dSolver
    from blitzdg_b200.synthetic import provision, structured_mesh

    assert torch.cuda.is_available(), "needs the B200"
    lib = capi.load()
    dg = provision(structured_mesh(shape, nx, ny, jitter=0.2), N).get_DGContext()
    Nfn = dg.Nfaces * dg.Nfp
    dims = dict(N=dg.N, Np=dg.Np, Nfp=dg.Nfp, Nfaces=dg.Nfaces, K=dg.K)
    assert Nfn * dg.K > 3 * 4096  # several thread blocks and a ragged last one
    H, h0, _, _ = gaussian_hump(dg.x, dg.y)
    rng = np.random.default_rng(7)
    fast = (h0, 3.0 * rng.standard_normal(h0.shape), 2.0 * rng.standard_normal(h0.shape))
    slow = (0.25 * h0, 0.01 * rng.standard_normal(h0.shape), 0.01 * rng.standard_normal(h0.shape))
    sol = SW2dSolver(dg)
    g, CFL = sol.phys.g, sol.num.CFL
    dev = torch.device("cuda", 0)
    fmask = np.asarray(dg.vmapM, dtype=np.int64).reshape(-1)[:Nfn]  # element 0: vmapM = Fmask + Np*k
    d_fmask = torch.from_numpy(fmask.astype(np.int32)).to(dev)
    d_Fscale = torch.from_numpy(np.ascontiguousarray(dg.Fscale, dtype=np.float64)).to(dev)
    d_q = [torch.empty(h0.shape, dtype=torch.float64, device=dev) for _ in range(3)]
    d_dt = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
    ptrs = [C.c_void_p(t.data_ptr()) for t in (*d_q, d_Fscale, d_fmask, d_dt)]

    def both(state):
        st = [np.ascontiguousarray(a, dtype=np.float64) for a in state]
        sol.set_state(*st)
        want = sol.computeTimeStep()  # bdg_sw2d_timestep: the existing entry point
        for d, a in zip(d_q, st):
            d.copy_(torch.from_numpy(a))
        d_dt.fill_(-1.0)
        torch.cuda.synchronize()
        got = []
        for _ in range(2):  # two calls in a row on the same inputs, nothing cleared by the caller in between
            rc, msg = _call(lib, 0, dims, g, CFL, ptrs)
            assert rc == capi.BDG_OK, msg
            torch.cuda.synchronize()
            got.append(float(d_dt.cpu()[0]))
        return want, got

    want_fast, got_fast = both(fast)
    want_slow, got_slow = both(slow)  # smaller maximum: a partial or a ticket left over from the calls above would show
    sol.close()
    assert 0 < want_fast < want_slow
    for want, got in ((want_fast, got_fast), (want_slow, got_slow)):
        for v in got:
            assert abs(v - want) <= TOL * want, (want, got)
    # a pointer that is not device memory is refused before any launch
    host = np.zeros(4)
    bad = list(ptrs)
    bad[0] = C.c_void_p(host.ctypes.data)
    rc, msg = _call(lib, 0, dims, g, CFL, bad)
    assert rc == capi.BDG_E_INVALID and "d_h" in msg
    # and the limit holds on the GPU box as well
    rc, msg = _call(lib, 0, dims, g, CFL, ptrs, K=capi.BDG_NOALLOC_MAX_FACE_NODES // Nfn + 1)
    assert rc == capi.BDG_E_UNSUPPORTED and "BDG_NOALLOC_MAX_FACE_NODES" in msg
