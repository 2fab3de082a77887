"""solve_SIA (SIA_main.f90:65-158), config 4 (Halfar dome).  CPU: the oracle against the closed-form column integral.  GPU: the device
kernels against the oracle, and a short Halfar run (SIA velocities -> calc_dHi_dt) against the oracle sequence and Halfar's analytic
solution (src/UFEMISM/analytical_solutions/Halfar_SIA_solution.f90; scoreboard RMSE of the reference at 5 km / 500 yr: 14.63 m)."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl

A, N_GLEN, H0, R0 = 1e-16, 3.0, 3000.0, 500e3
RHO, G = 917.0, 9.81


def rel_l2(a, b):
    n = np.linalg.norm(b)
    return np.linalg.norm(a - b) / (n if n > 0 else 1.0)


def vertical_average(zeta, f):
    """mesh_zeta.f90:269-295 (trapezoid from the surface zeta = 0 to the base zeta = 1)"""
    return np.sum(0.5 * (f[:, 1:] + f[:, :-1]) * (zeta[1:] - zeta[:-1]), axis=1)


def test_oracle_sia_matches_closed_form_column_integral():
    w = wl.halfar(4.0e4)
    m = w.mesh
    o = ol.OracleMesh(m)
    f = w.f
    r = o.solve_SIA(w.cfg, f["Hi"], f["Hs"], f["A_flow"])
    # surface velocity of an isothermal column: u_s = -2 (rho g)^n |grad h|^(n-1) A H^(n+1) / (n+1) dh/dx
    Hi_b = o.apply_operator(ub.OP["map_a_b"], f["Hi"]); dx = o.apply_operator(ub.OP["ddx_a_b"], f["Hs"]); dy = o.apply_operator(ub.OP["ddy_a_b"], f["Hs"])
    us = -2.0 * (RHO * G) ** N_GLEN * np.hypot(dx, dy) ** (N_GLEN - 1) * A * Hi_b ** (N_GLEN + 1) / (N_GLEN + 1) * dx
    inside = (np.hypot(m.TriGC[:, 0], m.TriGC[:, 1]) < 0.8 * R0) & (np.abs(us) > 1e-3)
    assert inside.sum() > 100
    assert np.abs(r["u_3D_b"][inside, 0] / us[inside] - 1.0).max() < 0.02       # trapezoid over 12 layers of zeta^3
    assert np.abs(r["u_3D_b"][:, -1]).max() == 0.0                               # no sliding: zero at the base
    assert (r["D_3D_b"] <= 0.0).all() and r["D_3D_b"].min() >= -1e5            # SIA_maximum_diffusivity
    # du/dz at the base: -2 (rho g)^n |grad h|^(n-1) A H^n dh/dx on the a-grid
    dxa = o.apply_operator(ub.OP["ddx_a_a"], f["Hs"]); dya = o.apply_operator(ub.OP["ddy_a_a"], f["Hs"])
    ref = -2.0 * (RHO * G) ** N_GLEN * np.hypot(dxa, dya) ** (N_GLEN - 1) * A * f["Hi"] ** N_GLEN * dxa
    np.testing.assert_allclose(r["du_dz_3D_a"][:, -1], ref, rtol=1e-12, atol=1e-30)


@pytest.mark.gpu
def test_gpu_sia_matches_oracle():
    w = wl.halfar(2.0e4)
    m, f = w.mesh, w.f
    rng = np.random.default_rng(3)
    A_flow = np.asfortranarray(f["A_flow"] * (0.5 + rng.random((m.nV, m.nz))))      # a non-uniform flow factor exercises map_a_b_3D
    o = ol.OracleMesh(m)
    ref = o.solve_SIA(w.cfg, f["Hi"], f["Hs"], A_flow, max_diffusivity=5e3)
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    out = {k: np.zeros_like(v) for k, v in ref.items()}
    ub.check(ub.lib().ufe_b200_solve_SIA(ctx.h, C.byref(w.cfg), 5e3, ub.dp(f["Hi"]), ub.dp(f["Hs"]), ub.dp(A_flow),
                                         *(ub.dp(out[k]) for k in ("u_3D_b", "v_3D_b", "D_3D_b", "du_dz_3D_a", "dv_dz_3D_a"))))
    ctx.close()
    assert (ref["D_3D_b"] == -5e3).any()                                            # the diffusivity limit is active somewhere
    for k in ref:
        assert rel_l2(out[k], ref[k]) < 1e-12, (k, rel_l2(out[k], ref[k]))


def _halfar_run(w, sia, thick, dt, nsteps):
    f, zeta = w.f, np.asarray(w.mesh.zeta, float)
    Hi = f["Hi"].copy()
    for _ in range(nsteps):
        Hs = wl.ice_surface_elevation(Hi, f["Hb"], f["SL"])
        u3, v3 = sia(Hi, Hs)
        u, v = np.ascontiguousarray(vertical_average(zeta, u3)), np.ascontiguousarray(vertical_average(zeta, v3))
        Hi = thick(Hi, u, v)
    return Hi


def _thickness_io(w, Hi, u, v, out):
    f = w.f
    a = dict(Hi=Hi, Hb=f["Hb"], SL=f["SL"], u_vav_b=u, v_vav_b=v, SMB=f["SMB"], BMB=f["BMB"], LMB=f["LMB"], fraction_margin=f["fraction_margin"],
             mask_noice=f["mask_noice"], dHi_dt_target=f["dHi_dt_target"])
    io = ub.ThicknessIO()
    for k, val in {**a, **out}.items():
        setattr(io, k, ub.ip(val) if val.dtype == np.int32 else ub.dp(val))
    return io, a


def _halfar_workload():
    w = wl.halfar(2.0e4)
    w.cfg.choice_ice_integration_method = ub.INTEGRATION_EXPLICIT
    w.cfg.dt_ice_min = 1e-3
    return w


def _oracle_halfar(w, o, dt, nsteps):
    m, f, cfg = w.mesh, w.f, w.cfg

    def sia(Hi, Hs):
        r = o.solve_SIA(cfg, Hi, Hs, f["A_flow"])
        return r["u_3D_b"], r["v_3D_b"]

    def thick(Hi, u, v):
        out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
        io, keep = _thickness_io(w, Hi, u, v, out)
        rc, d, its = o.calc_dHi_dt(cfg, io, dt)
        assert rc == 0 and d == dt
        return out["Hi_tplusdt"]

    return _halfar_run(w, sia, thick, dt, nsteps)


def _halfar_analytic(m, t):
    return np.array([ol.lib().orc_Halfar_H(A, N_GLEN, H0, R0, x, y, t) for x, y in m.V[:, :2]])


def test_oracle_halfar_dome_follows_the_analytic_solution():
    """config 4: SIA velocities + explicit calc_dHi_dt for 150 yr on a 20 km mesh; the reference's own scoreboard RMSE is 14.63 m (5 km, 500 yr)"""
    w = _halfar_workload()
    o = ol.OracleMesh(w.mesh)
    dt, nsteps = 1.0, 150
    Hi = _oracle_halfar(w, o, dt, nsteps)
    ana, ana0 = _halfar_analytic(w.mesh, dt * nsteps), _halfar_analytic(w.mesh, 0.0)
    rmse = np.sqrt(np.mean((Hi - ana) ** 2)); moved = np.sqrt(np.mean((ana - ana0) ** 2))
    assert moved > 50.0 and rmse < 20.0 and rmse < 0.2 * moved, (rmse, moved)


@pytest.mark.gpu
def test_gpu_halfar_dome_evolution():
    w = _halfar_workload()
    m, f, cfg = w.mesh, w.f, w.cfg
    o = ol.OracleMesh(m)
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    L = ub.lib()
    dt, nsteps = 1.0, 150

    def sia_gpu(Hi, Hs):
        u3 = np.zeros((m.nTri, m.nz), order="F"); v3 = np.zeros((m.nTri, m.nz), order="F")
        ub.check(L.ufe_b200_solve_SIA(ctx.h, C.byref(cfg), 1e5, ub.dp(Hi), ub.dp(Hs), ub.dp(f["A_flow"]), ub.dp(u3), ub.dp(v3), None, None, None))
        return u3, v3

    def thick_gpu(Hi, u, v):
        out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
        io, keep = _thickness_io(w, Hi, u, v, out)
        d = C.c_double(dt); its = C.c_int()
        ub.check(L.ufe_b200_calc_dHi_dt(ctx.h, C.byref(cfg), C.byref(io), C.cast(C.byref(d), ub.c_dp), 1, C.byref(its)))
        assert d.value == dt                                                        # the flux-limited time step does not bite
        return out["Hi_tplusdt"]

    Hi_gpu = _halfar_run(w, sia_gpu, thick_gpu, dt, nsteps)
    ctx.close()
    Hi_cpu = _oracle_halfar(w, o, dt, nsteps)
    assert rel_l2(Hi_gpu, Hi_cpu) < 1e-10, rel_l2(Hi_gpu, Hi_cpu)
    ana, ana0 = _halfar_analytic(m, dt * nsteps), _halfar_analytic(m, 0.0)
    rmse = np.sqrt(np.mean((Hi_gpu - ana) ** 2)); moved = np.sqrt(np.mean((ana - ana0) ** 2))
    assert moved > 50.0 and rmse < 20.0 and rmse < 0.2 * moved, (rmse, moved)
