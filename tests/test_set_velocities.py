"""set_ice_velocities_to_{SIA,SSA,SIASSA,DIVA}_results (conservation_of_momentum_main.f90:311-511): the oracle against the definitions, the device
against the oracle on a hybrid SIA/SSA stress balance (solve_SIA + solve_SSA + fan-out + secondary velocities, solve_stress_balance :125-170)."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl


def rel_l2(a, b):
    n = np.linalg.norm(b)
    return np.linalg.norm(a - b) / (n if n > 0 else 1.0)


def _results(m, seed=2):
    rng = np.random.default_rng(seed)
    r3 = lambda n: np.asfortranarray(50.0 * rng.standard_normal((n, m.nz)))
    r2 = lambda n, s: s * rng.standard_normal(n)
    return dict(u_3D_b=r3(m.nTri), v_3D_b=r3(m.nTri), u_b=r2(m.nTri, 40.0), v_b=r2(m.nTri, 40.0), du_dx_a=r2(m.nV, 1e-3), du_dy_a=r2(m.nV, 1e-3),
                dv_dx_a=r2(m.nV, 1e-3), dv_dy_a=r2(m.nV, 1e-3), du_dz_3D=r3(m.nV), dv_dz_3D=r3(m.nV))


@pytest.mark.parametrize("approx,scheme", [(ub.APPROX_SIA, 0), (ub.APPROX_SSA, 0), (ub.APPROX_DIVA, 0), (ub.APPROX_SIASSA, ub.HYBRID_ADD),
                                           (ub.APPROX_SIASSA, ub.HYBRID_ADD_SIA_REDUCED)])
def test_oracle_set_ice_velocities_definitions(approx, scheme):
    w = wl.halfar(8.0e4)
    m = w.mesh
    o = ol.OracleMesh(m)
    R = _results(m)
    s, out = ub.alloc_ice_velocities(m.nV, m.nTri, m.nz, fill=np.nan)
    assert o.set_ice_velocities(approx, scheme, s, **R) == 0
    lay = lambda a: np.repeat(a[:, None], m.nz, axis=1)
    zero = np.zeros((m.nV, m.nz))
    if approx == ub.APPROX_SIA:
        want = dict(u_3D_b=R["u_3D_b"], v_3D_b=R["v_3D_b"], du_dz_3D=R["du_dz_3D"], dv_dz_3D=R["dv_dz_3D"], du_dx_3D=zero, du_dy_3D=zero, dv_dx_3D=zero,
                    dv_dy_3D=zero)
    elif approx == ub.APPROX_SSA:
        want = dict(u_3D_b=lay(R["u_b"]), v_3D_b=lay(R["v_b"]), du_dz_3D=zero, dv_dz_3D=zero)
    elif approx == ub.APPROX_DIVA:
        want = dict(u_3D_b=R["u_3D_b"], v_3D_b=R["v_3D_b"], du_dz_3D=R["du_dz_3D"], dv_dz_3D=R["dv_dz_3D"])
    else:
        wu = wv = 1.0
        if scheme == ub.HYBRID_ADD_SIA_REDUCED:
            wu = lay(1.0 - (2.0 / np.pi) * np.arctan(R["u_b"] ** 2 / 900.0)); wv = lay(1.0 - (2.0 / np.pi) * np.arctan(R["v_b"] ** 2 / 900.0))
            assert 0.0 < wu.min() < 0.5 < wu.max() <= 1.0          # both regimes of the reduction occur
        want = dict(u_3D_b=wu * R["u_3D_b"] + lay(R["u_b"]), v_3D_b=wv * R["v_3D_b"] + lay(R["v_b"]), du_dz_3D=R["du_dz_3D"], dv_dz_3D=R["dv_dz_3D"])
    if approx != ub.APPROX_SIA:
        want.update(du_dx_3D=lay(R["du_dx_a"]), du_dy_3D=lay(R["du_dy_a"]), dv_dx_3D=lay(R["dv_dx_a"]), dv_dy_3D=lay(R["dv_dy_a"]))
    want.update(dw_dx_3D=zero, dw_dy_3D=zero)
    for k, a in want.items():
        np.testing.assert_allclose(out[k], a, rtol=1e-14, atol=0.0, err_msg=k)
    assert o.set_ice_velocities(ub.APPROX_SIASSA, 7, s, **R) != 0 and o.set_ice_velocities(9, 0, s, **R) != 0


def _hybrid_workload():
    """a Halfar-type dome that slides (Weertman), so that both parts of the hybrid are non-zero (BASELINE.json configs[3])"""
    w = wl.halfar(4.0e4)
    w.cfg.choice_sliding_law = ub.SLID_WEERTMAN
    w.cfg.visc_it_nit = 6; w.cfg.stress_balance_PETSc_rtol = 1e-12; w.cfg.stress_balance_PETSc_abstol = 1e-30
    w.cfg.write_diagnostics = 1
    return w


@pytest.mark.gpu
@pytest.mark.parametrize("scheme", [ub.HYBRID_ADD, ub.HYBRID_ADD_SIA_REDUCED])
def test_gpu_hybrid_sia_ssa_stress_balance_matches_oracle(scheme):
    w = _hybrid_workload()
    m, f, cfg = w.mesh, w.f, w.cfg
    o = ol.OracleMesh(m)
    # oracle: solve_SIA, solve_SSA, set_ice_velocities_to_SIASSA_results, calc_secondary_velocities
    sia = o.solve_SIA(cfg, f["Hi"], f["Hs"], f["A_flow"])
    st_o, ssa = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    rc, nv_o, na_o = o.solve_velocity(cfg, 0, w.ice_inputs(), st_o)
    assert rc == 0 and np.abs(ssa["u_vav_b"]).max() > 1.0 and np.abs(sia["u_3D_b"]).max() > 1.0
    s_o, ref = ub.alloc_ice_velocities(m.nV, m.nTri, m.nz, fill=np.nan)
    assert o.set_ice_velocities(ub.APPROX_SIASSA, scheme, s_o, u_3D_b=sia["u_3D_b"], v_3D_b=sia["v_3D_b"], u_b=ssa["u_vav_b"], v_b=ssa["v_vav_b"],
                                du_dx_a=ssa["du_dx_a"], du_dy_a=ssa["du_dy_a"], dv_dx_a=ssa["dv_dx_a"], dv_dy_a=ssa["dv_dy_a"],
                                du_dz_3D=sia["du_dz_3D_a"], dv_dz_3D=sia["dv_dz_3D_a"]) == 0
    sec_ref = o.calc_secondary_velocities(ref["u_3D_b"], ref["v_3D_b"])
    # device: the same four calls through the C ABI, everything after the solves from the resident fields
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(m); ctx.calc_matrix_operators()
        g_sia = {k: np.zeros_like(v) for k, v in sia.items()}
        ub.check(ub.lib().ufe_b200_solve_SIA(ctx.h, C.byref(cfg), 1e5, ub.dp(f["Hi"]), ub.dp(f["Hs"]), ub.dp(f["A_flow"]),
                                             *(ub.dp(g_sia[k]) for k in ("u_3D_b", "v_3D_b", "D_3D_b", "du_dz_3D_a", "dv_dz_3D_a"))))
        st, g = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        nv, na = C.c_int(), C.c_int()
        ice = w.ice_inputs()
        ub.check(ub.lib().ufe_b200_solve_SSA(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None))
        assert nv.value == nv_o
        s, out = ub.alloc_ice_velocities(m.nV, m.nTri, m.nz, fill=np.nan)
        ub.check(ub.lib().ufe_b200_set_ice_velocities(ctx.h, ub.APPROX_SIASSA, scheme, C.byref(s)))
        sec, sec_out = ub.alloc_secondary(m.nV, m.nTri, m.nz)
        ub.check(ub.lib().ufe_b200_calc_secondary_velocities(ctx.h, None, None, C.byref(sec)))
        # a second fan-out without new solves must be refused: the SIA velocities have been consumed
        assert ub.lib().ufe_b200_set_ice_velocities(ctx.h, ub.APPROX_SIASSA, scheme, C.byref(s)) != 0
        assert ub.lib().ufe_b200_set_ice_velocities(ctx.h, ub.APPROX_DIVA, 0, C.byref(s)) != 0
    finally:
        ctx.close()
    scale_u = np.abs(ref["u_3D_b"]).max()
    for k in ub.ICE_VELOCITY_FIELDS:
        scale = scale_u if k.endswith("_b") else max(np.abs(ref[k]).max(), 1e-300)
        assert np.abs(out[k] - ref[k]).max() <= 1e-8 * scale, (k, np.abs(out[k] - ref[k]).max(), scale)
    for k in ("u_vav_b", "v_vav_b", "u_surf", "v_base", "uabs_vav"):
        assert rel_l2(sec_out[k], sec_ref[k]) < 1e-8, k


@pytest.mark.gpu
def test_gpu_set_velocities_ssa_and_diva():
    w = wl.mismip_mod(5.0e4)
    m, cfg = w.mesh, w.cfg
    cfg.visc_it_nit = 2; cfg.write_diagnostics = 1
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(m); ctx.calc_matrix_operators()
        nv, na = C.c_int(), C.c_int()
        ice = w.ice_inputs()
        lay = lambda a: np.repeat(a[:, None], m.nz, axis=1)
        st, g = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        ub.check(ub.lib().ufe_b200_solve_DIVA(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None))
        s, out = ub.alloc_ice_velocities(m.nV, m.nTri, m.nz, fill=np.nan)
        ub.check(ub.lib().ufe_b200_set_ice_velocities(ctx.h, ub.APPROX_DIVA, 0, C.byref(s)))
        assert ub.lib().ufe_b200_set_ice_velocities(ctx.h, ub.APPROX_SSA, 0, C.byref(s)) != 0       # the last solve was DIVA
        want = dict(u_3D_b=g["u_3D_b"], v_3D_b=g["v_3D_b"], du_dz_3D=g["du_dz_3D_a"], dv_dz_3D=g["dv_dz_3D_a"], du_dx_3D=lay(g["du_dx_a"]),
                    du_dy_3D=lay(g["du_dy_a"]), dv_dx_3D=lay(g["dv_dx_a"]), dv_dy_3D=lay(g["dv_dy_a"]), dw_dx_3D=np.zeros((m.nV, m.nz)),
                    dw_dy_3D=np.zeros((m.nV, m.nz)))
        for k, a in want.items():
            assert np.array_equal(out[k], a), k
        assert np.abs(out["du_dx_3D"]).max() > 0 and np.abs(out["du_dz_3D"]).max() > 0
        st2, g2 = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        ub.check(ub.lib().ufe_b200_solve_SSA(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st2), C.byref(nv), C.byref(na), None, None, None))
        s2, out2 = ub.alloc_ice_velocities(m.nV, m.nTri, m.nz, fill=np.nan)
        ub.check(ub.lib().ufe_b200_set_ice_velocities(ctx.h, ub.APPROX_SSA, 0, C.byref(s2)))
        want = dict(u_3D_b=lay(g2["u_vav_b"]), v_3D_b=lay(g2["v_vav_b"]), du_dz_3D=np.zeros((m.nV, m.nz)), dv_dz_3D=np.zeros((m.nV, m.nz)),
                    du_dx_3D=lay(g2["du_dx_a"]), dv_dy_3D=lay(g2["dv_dy_a"]))
        for k, a in want.items():
            assert np.array_equal(out2[k], a), k
        assert np.abs(out2["u_3D_b"]).max() > 0
    finally:
        ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("approx", [ub.APPROX_DIVA, ub.APPROX_SSA])
def test_gpu_set_velocities_without_the_checksum_log(approx):
    """C%do_write_checksum_log is .false. by default (model_configuration_type_and_namelist.f90:48), so the Fortran shim passes write_diagnostics = 0; the
    fan-out after the solve (set_ice_velocities_to_DIVA / SSA_results) must still find every strain rate, identical to the run with the log switched on"""
    w = wl.mismip_mod(5.0e4)
    m, cfg = w.mesh, w.cfg
    cfg.visc_it_nit = 2
    outs = []
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(m); ctx.calc_matrix_operators()
        for diag in (1, 0):
            cfg.write_diagnostics = diag
            nv, na = C.c_int(), C.c_int()
            ice = w.ice_inputs()
            st, g = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
            solve = ub.lib().ufe_b200_solve_DIVA if approx == ub.APPROX_DIVA else ub.lib().ufe_b200_solve_SSA
            ub.check(solve(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None))
            s, out = ub.alloc_ice_velocities(m.nV, m.nTri, m.nz, fill=np.nan)
            ub.check(ub.lib().ufe_b200_set_ice_velocities(ctx.h, approx, 0, C.byref(s)))
            outs.append(out)
    finally:
        ctx.close()
    for k in ub.ICE_VELOCITY_FIELDS:
        assert np.isfinite(outs[1][k]).all() and np.array_equal(outs[0][k], outs[1][k]), k
    assert np.abs(outs[1]["du_dx_3D"]).max() > 0
    if approx == ub.APPROX_DIVA:
        assert np.abs(outs[1]["du_dz_3D"]).max() > 0
