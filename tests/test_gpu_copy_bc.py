"""GPU: copy rows of BC 'periodic_ISMIP-HOM' / 'infinite_SSA_icestream' and the SSA_icestream integrated test vs the CPU oracle (C ABI)."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl
from test_copy_bc import schoof_u

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    n = np.linalg.norm(b)
    return np.linalg.norm(a - b) / (n if n > 0 else 1.0)


@pytest.fixture(scope="module")
def setup():
    w = wl.ssa_icestream(32e3)
    ctx = ub.Context(device=0)
    ctx.mesh_create(w.mesh)
    ctx.calc_matrix_operators()
    o = ol.OracleMesh(w.mesh)
    o.calc_matrix_operators()
    yield w, ctx, o
    ctx.close()


def _fields(m, seed):
    rng = np.random.default_rng(seed)
    return dict(N_b=1e10 * (1 + 0.1 * rng.random(m.nTri)), dNx=1e3 * rng.standard_normal(m.nTri), dNy=1e3 * rng.standard_normal(m.nTri),
                beta=1e3 * rng.random(m.nTri), tdx=1e4 * rng.standard_normal(m.nTri), tdy=1e4 * rng.standard_normal(m.nTri),
                u=100.0 * rng.standard_normal(m.nTri), v=100.0 * rng.standard_normal(m.nTri))


def test_linearised_solve_with_copy_rows(setup):
    w, ctx, o = setup
    m = w.mesh
    f = _fields(m, 11)
    cfg = wl.scaling_config()
    cfg.visc_it_relax = 0.3; cfg.refgeo_idealised_ISMIP_HOM_L = 640e3
    for k, (bu, bv) in enumerate([(2, 2), (3, 2), (0, 0), (3, 1)]):   # north, east, south, west
        cfg.BC_u[k] = bu; cfg.BC_v[k] = bv
    u, v = f["u"].copy(), f["v"].copy()
    up = np.zeros(m.nTri); vp = np.zeros(m.nTri)
    its = C.c_int()
    ub.check(ub.lib().ufe_b200_solve_SSA_DIVA_linearised(ctx.h, C.byref(cfg), ub.dp(u), ub.dp(v), ub.dp(f["N_b"]), ub.dp(f["dNx"]), ub.dp(f["dNy"]),
                                                         ub.dp(f["beta"]), ub.dp(f["tdx"]), ub.dp(f["tdy"]), ub.dp(up), ub.dp(vp), 1e-12, 1e-30,
                                                         C.byref(its), None, None, None))
    assert np.array_equal(up, f["u"]) and np.array_equal(vp, f["v"])
    ptr, ind, val, bb = ctx.get_stiffness()
    optr, oind, oval, obb = o.assemble(cfg, f["N_b"], f["dNx"], f["dNy"], f["beta"], f["tdx"], f["tdy"], u_b_prev=f["u"], v_b_prev=f["v"])
    assert np.array_equal(ptr, optr) and np.array_equal(ind, oind)
    assert np.abs(val - oval).max() <= 1e-13 * np.abs(oval).max()
    assert np.abs(bb - obb).max() <= 1e-12 * np.abs(obb).max()
    xo, its_o, rc = ol.ksp_solve(optr, oind, oval, obb, 1e-12, 1e-30, ksp=ol.KSP_GMRES_ILU0)
    assert rc == 0
    uv = np.empty(2 * m.nTri); uv[0::2] = u; uv[1::2] = v
    assert rel_l2(uv, xo) < 1e-6, (its.value, its_o)


def _solve_ssa(ctx, cfg, w):
    m = w.mesh
    st, arrs = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    nv, na = C.c_int(), C.c_int()
    ice = w.ice_inputs()
    ub.check(ub.lib().ufe_b200_solve_SSA(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None))
    return arrs, nv.value, na.value


def test_ssa_icestream_fixed_iterations_match_oracle(setup):
    w, ctx, o = setup
    m = w.mesh
    cfg = wl.ssa_icestream(32e3).cfg
    cfg.visc_it_nit = 24; cfg.stress_balance_PETSc_rtol = 1e-12; cfg.stress_balance_PETSc_abstol = 1e-30
    g, nv, na = _solve_ssa(ctx, cfg, w)
    st, ref = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    rc, nv_o, na_o = o.solve_velocity(cfg, 0, w.ice_inputs(), st)
    assert rc == 0 and nv == nv_o
    for name in ("u_vav_b", "v_vav_b", "u_b_prev", "v_b_prev"):
        assert np.abs(g[name] - ref[name]).max() <= 1e-8 * np.abs(ref["u_vav_b"]).max(), name


def test_ssa_icestream_converged_against_schoof_2006(setup):
    w, ctx, o = setup
    m = w.mesh
    g, nv, na = _solve_ssa(ctx, w.cfg, w)
    st, ref = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    rc, nv_o, na_o = o.solve_velocity(w.cfg, 0, w.ice_inputs(), st)
    assert rc == 0 and nv < w.cfg.visc_it_nit and abs(nv - nv_o) <= 0.05 * nv_o
    assert rel_l2(g["u_vav_b"], ref["u_vav_b"]) < 1e-5
    sel = np.abs(m.TriGC[:, 0]) < 100e3
    ua = schoof_u(m.TriGC[sel, 1])
    rmse_g = np.sqrt(np.mean((g["u_vav_b"][sel] - ua) ** 2)); rmse_o = np.sqrt(np.mean((ref["u_vav_b"][sel] - ua) ** 2))
    assert rmse_g < 700.0 and abs(rmse_g - rmse_o) < 1e-3 * rmse_o


def test_unknown_bc_choice_is_an_error(setup):
    w, ctx, o = setup
    cfg = wl.ssa_icestream(32e3).cfg
    cfg.BC_v[1] = ub.BC_INFINITE_SSA_ICESTREAM   # the reference offers 'infinite_SSA_icestream' for u only
    m = w.mesh
    st, arrs = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    nv, na = C.c_int(), C.c_int()
    ice = w.ice_inputs()
    rc = ub.lib().ufe_b200_solve_SSA(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None)
    assert rc != 0 and b"unknown choice_BC_v" in ub.lib().ufe_b200_last_error()


def test_gpu_follows_the_reference_checksum_log():
    """The device path against output of the reference itself (tests/golden/ssa_icestream_small.json): mean and maximum of u_b after each of the
    first 12 relaxed viscosity iterations of integrated_test_SSA_icestream_small, and the same sequence against the oracle."""
    from test_copy_bc import check_against_reference_log, reference_sequence
    g = reference_sequence()
    w = wl.ssa_icestream(13.7e3)
    m = w.mesh
    assert abs(m.nTri - g["nTri"]) < 0.01 * g["nTri"]
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(m)
        ctx.calc_matrix_operators()
        o = ol.OracleMesh(m)
        cfg = w.cfg
        cfg.visc_it_nit = 0        # one viscosity iteration per call; the state carries u_b, v_b to the next one
        cfg.stress_balance_PETSc_rtol = 1e-11; cfg.stress_balance_PETSc_abstol = 1e-30   # tight, so device and oracle can be compared closely
        st, f = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        st_o, f_o = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        ice = w.ice_inputs()
        nv, na = C.c_int(), C.c_int()
        seq = []
        for k in range(12):
            ub.check(ub.lib().ufe_b200_solve_SSA(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None))
            rc, nv_o, na_o = o.solve_velocity(cfg, 0, ice, st_o)
            assert rc == 0 and nv.value == nv_o == 1
            assert np.abs(f["u_vav_b"] - f_o["u_vav_b"]).max() <= 1e-7 * np.abs(f_o["u_vav_b"]).max(), k
            seq.append((f["u_vav_b"].sum(), f["u_vav_b"].max()))
        check_against_reference_log(seq, m.nTri)
    finally:
        ctx.close()
