"""Over-integrated RHS on curved triangles: sw2dComputeRHS_curved, reference swhelpers/rhs.py:6-176.

  * the host mirrors of the cubature volume mesh and the Gauss face nodes (blitzdg_b200/curved.py) hold what the
    reference's builders guarantee (src/TriangleNodesProvisioner.cpp:81-381);
  * the numpy oracle (oracle/curved_numpy.py) is pinned to outputs of the reference's OWN function, executed unmodified by
    tests/golden/make_curved_reference_vectors.py (tests/golden/curved_reference_rhs.npz);
  * the element functions of the CUDA kernel, compiled for the host with one thread per element
    (-DBDG_CURVED_HOST_EMULATION), reproduce the fixture without a GPU;
  * on the GPU the kernel is called through the C ABI (bdg_sw2d_set_curved / bdg_sw2d_rhs_curved) and held to both.
Tolerance: 1e-12 of max|RHS_i| per field, the bar of the nodal RHS.
"""
import ctypes as C
import os
import shutil
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_curved_reference_vectors as gen  # noqa: E402  (build_case / state only: nothing here reads /root/reference)

TOL = 1e-12
FIX = os.path.join(ROOT, "tests", "golden", "curved_reference_rhs.npz")


@pytest.fixture(scope="module")
def fixture():
    return np.load(FIX)


@pytest.fixture(scope="module", params=sorted(gen.CASES))
def case(request, fixture):
    name = request.param
    nx, ny, N, ncub, ng = gen.CASES[name]
    rule = fixture[name + "_rule"] if name in gen.TABLE_RULE_CASES else None  # the reference's own tabulated points
    nodes, ctx, cub, gauss, curved = gen.build_case(nx, ny, N, ncub, ng, rule=rule)
    assert cub.NumCubaturePoints == (len(rule[0]) if rule is not None else (ncub // 2 + 1) ** 2)
    d = {k: fixture["%s_%s" % (name, k)] for k in ("H", "h", "hu", "hv", "hN", "zx", "zy", "CD", "curved")}
    d["g"], d["f"] = (float(x) for x in fixture[name + "_params"][5:7])
    d["R"] = [fixture["%s_RHS%d" % (name, i + 1)] for i in range(4)]
    d["nsteps"], d["dt"] = int(fixture[name + "_step_params"][0]), float(fixture[name + "_step_params"][1])
    d["q_end"] = [fixture["%s_step_%s" % (name, k)] for k in ("h", "hu", "hv", "hN")]
    # the contexts are rebuilt here, the inputs come from the fixture: the rebuilt state must be the stored one
    H, h, hu, hv, hN, zx, zy, CD = gen.state(ctx)
    assert np.array_equal(h, d["h"]) and np.array_equal(zx, d["zx"]) and np.array_equal(curved, d["curved"])
    return name, ctx, cub, gauss, d


def relerr(a, b):
    return float(np.abs(a - b).max() / np.abs(b).max())


def test_cubature_rule_is_exact_to_its_order():
    from blitzdg_b200.curved import triangle_cubature

    for order in (3, 9, 15):
        r, s, w = triangle_cubature(order)
        rr, ss, ww = triangle_cubature(2 * order + 6)
        assert abs(w.sum() - 2.0) < 1e-14 and (w > 0).all() and (r + s < 0).all() and (r > -1).all() and (s > -1).all()
        for i in range(order + 1):
            for j in range(order + 1 - i):
                assert abs((w * r**i * s**j).sum() - (ww * rr**i * ss**j).sum()) < 2e-14, (order, i, j)


def test_contexts_hold_what_the_reference_builders_guarantee(case):
    name, ctx, cub, gauss, d = case
    K, NG3 = ctx.K, gauss.Interp.shape[0]
    gx, gy = gauss.x.flatten("F"), gauss.y.flatten("F")
    # '+' Gauss node = the same physical point seen from the neighbour (src/TriangleNodesProvisioner.cpp:335-349)
    assert np.abs(gx[gauss.mapM] - gx[gauss.mapP]).max() < 1e-13 and np.abs(gy[gauss.mapM] - gy[gauss.mapP]).max() < 1e-13
    assert np.array_equal(gauss.mapM, np.arange(NG3 * K))
    interior = gauss.mapM != gauss.mapP
    nxf, nyf = gauss.nx.flatten("F"), gauss.ny.flatten("F")
    assert np.abs(nxf**2 + nyf**2 - 1).max() < 1e-14
    assert np.abs(nxf[gauss.mapM] + nxf[gauss.mapP])[interior].max() < 1e-12  # opposite unit normals across a curved face
    assert sorted(gauss.BCmap) == [3] and np.array_equal(np.sort(gauss.BCmap[3]), np.nonzero(~interior)[0])
    assert abs(gauss.W.flatten("F")[gauss.BCmap[3]].sum() - 8.0) < 1e-12  # the deformation keeps the boundary of [-1,1]^2
    assert abs(cub.W.sum() - 4.0) < 1e-12
    # interpolation reproduces polynomials of the element's order; cubature differentiation matrices differentiate them
    r, s = ctx.r, ctx.s
    p = 1 + r - 2 * s + r * s if ctx.N >= 2 else 1 + r
    assert np.abs(cub.V @ p - (1 + cub.r - 2 * cub.s + cub.r * cub.s)).max() < 1e-13
    assert np.abs(cub.Dr @ p - (1 + cub.s)).max() < 1e-12 and np.abs(cub.Ds @ p - (-2 + cub.r)).max() < 1e-12
    for k in (0, K // 2, K - 1):
        U = cub.MMChol[:, :, k]
        assert np.abs(np.tril(U, -1)).max() == 0 and np.abs(U.T @ U - cub.MM[:, :, k]).max() < 1e-15
    # elements are genuinely curved: the nodal Jacobian varies inside an element
    assert (ctx.J.max(axis=0) - ctx.J.min(axis=0)).max() > 1e-3 * ctx.J.mean()


def test_oracle_matches_the_reference_function(case):
    from oracle.curved_numpy import rhs_curved

    name, ctx, cub, gauss, d = case
    R = rhs_curved(d["h"], d["hu"], d["hv"], d["hN"], d["zx"], d["zy"], d["g"], d["f"], d["CD"], ctx, cub, gauss, d["curved"], ctx.J,
                   gauss.mapM, gauss.mapP)
    for i in range(4):
        assert relerr(R[i], d["R"][i]) < 1e-13, (name, i)
    # both mass-matrix paths are exercised and differ on a curved element (V V^T / J is only exact on straight ones)
    Rs = rhs_curved(d["h"], d["hu"], d["hv"], d["hN"], d["zx"], d["zy"], d["g"], d["f"], d["CD"], ctx, cub, gauss, [], ctx.J, gauss.mapM, gauss.mapP)
    assert relerr(Rs[0], d["R"][0]) > 1e-6


def test_oracle_time_loop_matches_the_reference_driver_loop(case):
    """sw2d_curved.py:240-262 with the reference's own RHS function (fixture) against the same loop around the oracle."""
    from oracle.curved_numpy import rhs_curved

    name, ctx, cub, gauss, d = case
    q = [d[k].copy() for k in ("h", "hu", "hv", "hN")]

    def rhs(q):
        R = rhs_curved(q[0], q[1], q[2], q[3], d["zx"], d["zy"], d["g"], d["f"], d["CD"], ctx, cub, gauss, d["curved"], ctx.J, gauss.mapM, gauss.mapP)
        return [ctx.Filt @ r for r in R]

    for _ in range(d["nsteps"]):
        R = rhs(q)
        q1 = [q[i] + 0.5 * d["dt"] * R[i] for i in range(4)]
        R = rhs(q1)
        q = [q[i] + d["dt"] * R[i] for i in range(4)]
    for i in range(4):
        assert relerr(q[i], d["q_end"][i]) < 1e-12, (name, i)
    assert relerr(d["q_end"][0], d["h"]) > 1e-3  # the state has moved


class _CurvedOps(C.Structure):  # blitzdg_b200/csrc/sw2d_common.cuh: struct CurvedOps
    _dp, _ip, _bp = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_uint8)
    _fields_ = [("Ncub", C.c_int), ("NG3", C.c_int), ("V", _dp), ("Dr", _dp), ("Ds", _dp), ("cW", _dp), ("crx", _dp), ("csx", _dp),
                ("cry", _dp), ("csy", _dp), ("gInterp", _dp), ("gW", _dp), ("gnx", _dp), ("gny", _dp), ("mapM", _ip), ("mapP", _ip),
                ("wall", _bp), ("J", _dp), ("mminv", _dp), ("chol_idx", _ip), ("chol", _dp), ("gq", _dp), ("CD", _dp), ("out", _dp)]


def _emulation_lib():
    if not shutil.which("nvcc") and not os.path.exists("/usr/local/cuda/bin/nvcc"):
        pytest.skip("nvcc not available: the host emulation of the kernel's element functions cannot be compiled")
    out = os.path.join(ROOT, "build", "libcurved_host_emulation.so")
    src = os.path.join(ROOT, "blitzdg_b200", "csrc", "sw2d_curved.cu")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(os.path.join(os.path.dirname(src), "sw2d_common.cuh"))):
        subprocess.run(["bash", os.path.join(ROOT, "tools", "curved_host_emulation.sh")], check=True, capture_output=True)
    return C.CDLL(out)


def test_kernel_element_functions_on_the_host_match_reference_and_oracle(case):
    """The __host__ __device__ element functions of sw2d_curved.cu, one thread per element, on arrays in the device layouts."""
    name, ctx, cub, gauss, d = case
    lib = _emulation_lib()
    Np, K, Ncub, NG3 = ctx.Np, ctx.K, cub.V.shape[0], gauss.Interp.shape[0]
    keep = []

    def em(a):  # [R][K] -> element-major [K][R]
        b = np.ascontiguousarray(np.asarray(a, dtype=np.float64).T)
        keep.append(b)
        return b.ctypes.data_as(C.POINTER(C.c_double))

    def dd(a, tp=np.float64, ct=C.c_double):
        b = np.ascontiguousarray(a, dtype=tp)
        keep.append(b)
        return b.ctypes.data_as(C.POINTER(ct))

    ops = _CurvedOps()
    ops.Ncub, ops.NG3 = Ncub, NG3
    ops.V, ops.Dr, ops.Ds, ops.gInterp = dd(cub.V), dd(cub.Dr), dd(cub.Ds), dd(gauss.Interp)
    ops.cW, ops.crx, ops.csx, ops.cry, ops.csy = em(cub.W), em(cub.rx), em(cub.sx), em(cub.ry), em(cub.sy)
    ops.gW, ops.gnx, ops.gny, ops.J = em(gauss.W), em(gauss.nx), em(gauss.ny), em(ctx.J)
    ops.mapM, ops.mapP = dd(gauss.mapM, np.int32, C.c_int32), dd(gauss.mapP, np.int32, C.c_int32)
    wall = np.zeros(NG3 * K, np.uint8)
    wall[gauss.BCmap[3]] = 1
    ops.wall = dd(wall, np.uint8, C.c_uint8)
    ops.mminv = dd(ctx.V @ ctx.V.T)
    cidx = -np.ones(K, np.int32)
    cidx[d["curved"]] = np.arange(len(d["curved"]))
    ops.chol_idx = dd(cidx, np.int32, C.c_int32)
    ops.chol = dd(np.stack([cub.MMChol[:, :, k] for k in d["curved"]]))
    gq = np.zeros(4 * NG3 * K)
    ops.gq = dd(gq)
    stride = Np * K + 8
    q = np.zeros(3 * stride)
    for i, a in enumerate((d["h"], d["hu"], d["hv"])):
        q[i * stride:i * stride + Np * K] = a.T.reshape(-1)
    out = np.zeros(4 * Np * K)
    sm = np.zeros(12 * Np + 8 * Ncub + 5 * NG3)
    lib.bdg_curved_host_emulation.argtypes = [C.c_int, C.c_int, C.c_longlong, C.c_double, C.c_double, C.c_double] + [C.c_void_p] * 8
    lib.bdg_curved_host_emulation(Np, K, stride, d["g"], d["f"], 0.0, em(-d["zx"]), em(-d["zy"]), C.byref(ops), dd(q), em(d["hN"]), em(d["CD"]),
                                  dd(sm), out.ctypes.data_as(C.POINTER(C.c_double)))
    R = [out[i * Np * K:(i + 1) * Np * K].reshape(K, Np).T for i in range(4)]
    for i in range(4):
        assert relerr(R[i], d["R"][i]) < TOL, (name, i, relerr(R[i], d["R"][i]))


def test_kernel_update_function_on_the_host(case):
    """curved_update_node (the stage update of bdg_sw2d_step_curved): dst = base + c Filt R, in place as well."""
    name, ctx, cub, gauss, d = case
    lib = _emulation_lib()
    Np, K = ctx.Np, ctx.K
    nT, stride = Np * K, Np * K + 8
    rng = np.random.default_rng(11)
    R = rng.standard_normal((4, Np, K))
    base = rng.standard_normal((3, Np, K))
    baseN = rng.standard_normal((Np, K))
    em = lambda a: np.ascontiguousarray(a.T).reshape(-1)  # noqa: E731  [Np][K] -> element-major
    Rd = np.concatenate([em(R[f]) for f in range(4)])
    qd = np.zeros(3 * stride)
    for f in range(3):
        qd[f * stride:f * stride + nT] = em(base[f])
    qN = em(baseN).copy()
    Filt = np.ascontiguousarray(ctx.Filt)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    lib.bdg_curved_update_host_emulation.argtypes = [C.c_int, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p, C.c_double] + [C.c_void_p] * 4
    lib.bdg_curved_update_host_emulation(Np, K, stride, p(Filt), p(Rd), 0.37, p(qd), p(qN), p(qd), p(qN))  # in place
    for f in range(3):
        got = qd[f * stride:f * stride + nT].reshape(K, Np).T
        assert relerr(got, base[f] + 0.37 * (Filt @ R[f])) < 1e-14
    assert relerr(qN.reshape(K, Np).T, baseN + 0.37 * (Filt @ R[3])) < 1e-14


def _periodic_in_x(gauss):
    """What sw2d_curved.py:141-142 does to the Gauss maps with makeMapsPeriodic (swhelpers/maps.py): the wall Gauss nodes on
    x = -1 and x = +1 are paired by their y coordinate and become each other's '+' side; they leave the wall list."""
    import dataclasses

    gx, gy = gauss.x.flatten("F"), gauss.y.flatten("F")
    wall = np.asarray(gauss.BCmap[3], dtype=np.int64)
    left, right = wall[np.abs(gx[wall] + 1) < 1e-12], wall[np.abs(gx[wall] - 1) < 1e-12]
    assert len(left) == len(right) > 0
    mapP = gauss.mapP.copy()
    for i in left:
        j = right[np.argmin(np.abs(gy[right] - gy[i]))]
        assert abs(gy[j] - gy[i]) < 1e-12
        mapP[i], mapP[j] = j, i
    keep = np.setdiff1d(wall, np.concatenate([left, right])).astype(np.int32)
    return dataclasses.replace(gauss, mapP=mapP, BCmap={3: keep})


def test_oracle_follows_periodic_gauss_maps(case):
    from oracle.curved_numpy import rhs_curved

    name, ctx, cub, gauss, d = case
    gp = _periodic_in_x(gauss)
    args = (d["h"], d["hu"], d["hv"], d["hN"], d["zx"], d["zy"], d["g"], d["f"], d["CD"], ctx, cub)
    Rw = rhs_curved(*args, gauss, d["curved"], ctx.J, gauss.mapM, gauss.mapP)
    Rp = rhs_curved(*args, gp, d["curved"], ctx.J, gp.mapM, gp.mapP)
    touched = np.unique(np.concatenate([np.nonzero(gp.mapP != gauss.mapP)[0]]) // gauss.Interp.shape[0])
    other = np.setdiff1d(np.arange(ctx.K), touched)
    assert np.abs(Rp[0][:, touched] - Rw[0][:, touched]).max() > 1e-6 and np.array_equal(Rp[0][:, other], Rw[0][:, other])


@pytest.mark.gpu
def test_cuda_curved_rhs_with_periodic_gauss_maps(case):
    """Caller-supplied gmapM / gmapP (periodic in x) and the reduced wall list go through the element renumbering of set_curved."""
    from blitzdg_b200.sw2d import SW2dSolver, physParams
    from oracle.curved_numpy import rhs_curved

    name, ctx, cub, gauss, d = case
    gp = _periodic_in_x(gauss)
    sol = SW2dSolver(ctx, phys=physParams(g=d["g"], CD=0.0, f=d["f"]))
    sol.set_bathymetry(d["H"], -d["zx"], -d["zy"])
    sol.set_state(d["h"], d["hu"], d["hv"])
    sol.set_tracer(d["hN"])
    sol.set_curved(cub, gp, d["curved"], gmapM=gp.mapM, gmapP=gp.mapP)
    R = sol.computeRHS_curved(CD=d["CD"])
    Ro = rhs_curved(d["h"], d["hu"], d["hv"], d["hN"], d["zx"], d["zy"], d["g"], d["f"], d["CD"], ctx, cub, gp, d["curved"], ctx.J, gp.mapM, gp.mapP)
    for i in range(4):
        assert relerr(R[i], Ro[i]) < TOL, (name, i, relerr(R[i], Ro[i]))
    sol.close()


@pytest.mark.gpu
def test_cuda_curved_time_loop_matches_the_reference_driver_loop(case):
    from blitzdg_b200.sw2d import SW2dSolver, physParams

    name, ctx, cub, gauss, d = case
    sol = SW2dSolver(ctx, phys=physParams(g=d["g"], CD=0.0, f=d["f"]), use_filter=True)
    sol.set_bathymetry(d["H"], -d["zx"], -d["zy"])
    sol.set_state(d["h"], d["hu"], d["hv"])
    sol.set_tracer(d["hN"])
    sol.set_curved(cub, gauss, d["curved"])
    sol.step_curved(d["nsteps"], d["dt"], CD=d["CD"])
    h, hu, hv, t = sol.get_state()
    hN = sol.get_tracer()
    for i, a in enumerate((h, hu, hv, hN)):
        assert relerr(a, d["q_end"][i]) < 1e-10, (name, i, relerr(a, d["q_end"][i]))
    assert abs(t - d["nsteps"] * d["dt"]) < 1e-15
    sol.close()


@pytest.mark.gpu
def test_cuda_curved_rhs_matches_reference_and_oracle(case):
    from blitzdg_b200.sw2d import SW2dSolver, physParams
    from oracle.curved_numpy import rhs_curved

    name, ctx, cub, gauss, d = case
    sol = SW2dSolver(ctx, phys=physParams(g=d["g"], CD=1.5e-3, f=d["f"]))
    sol.set_bathymetry(d["H"], -d["zx"], -d["zy"])  # z = -H: Hx = -zx
    sol.set_state(d["h"], d["hu"], d["hv"])
    sol.set_tracer(d["hN"])
    sol.set_curved(cub, gauss, d["curved"])
    R = sol.computeRHS_curved(CD=d["CD"])
    for i in range(4):
        assert relerr(R[i], d["R"][i]) < TOL, (name, i, relerr(R[i], d["R"][i]))
    # a second state, scalar drag coefficient, no tracer, every element straight: against the oracle
    rng = np.random.default_rng(3)
    h2 = d["h"] * (1 + 0.05 * rng.standard_normal(d["h"].shape))
    hu2, hv2 = d["hu"] + 0.1 * rng.standard_normal(h2.shape), d["hv"] + 0.1 * rng.standard_normal(h2.shape)
    sol.set_state(h2, hu2, hv2)
    sol.set_tracer(None)
    sol.set_curved(cub, gauss, [])
    R2 = sol.computeRHS_curved()
    Ro = rhs_curved(h2, hu2, hv2, np.zeros_like(h2), d["zx"], d["zy"], d["g"], d["f"], 1.5e-3, ctx, cub, gauss, [], ctx.J, gauss.mapM, gauss.mapP)
    for i in range(3):
        assert relerr(R2[i], Ro[i]) < TOL, (name, i)
    assert np.abs(R2[3]).max() == 0.0
    # the nodal path of the same handle is untouched by the curved contexts
    sol.set_curved(None, None, None)
    with pytest.raises(RuntimeError, match="set_curved"):
        sol.computeRHS_curved()
    assert np.isfinite(sol.computeRHS()[0]).all()
    sol.close()
