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


def _random_3D_velocities(m, seed=5):
    rng = np.random.default_rng(seed)
    prof = np.linspace(1.0, 0.3, m.nz) ** 2
    u = np.asfortranarray(200.0 * rng.standard_normal((m.nTri, 1)) * prof + rng.standard_normal((m.nTri, m.nz)))
    v = np.asfortranarray(150.0 * rng.standard_normal((m.nTri, 1)) * prof + rng.standard_normal((m.nTri, m.nz)))
    return u, v


def test_oracle_secondary_velocities_definitions():
    """calc_secondary_velocities (conservation_of_momentum_main.f90:177-246) against its definitions in numpy"""
    w = wl.halfar(4.0e4)
    m = w.mesh
    o = ol.OracleMesh(m)
    u, v = _random_3D_velocities(m)
    r = o.calc_secondary_velocities(u, v)
    zeta = np.asarray(m.zeta, float)
    np.testing.assert_array_equal(r["u_surf_b"], u[:, 0]); np.testing.assert_array_equal(r["v_base_b"], v[:, -1])
    np.testing.assert_allclose(r["u_vav_b"], vertical_average(zeta, u), rtol=1e-13, atol=1e-12)
    np.testing.assert_allclose(r["uabs_vav_b"], np.hypot(r["u_vav_b"], r["v_vav_b"]), rtol=1e-14)
    np.testing.assert_allclose(r["u_3D"], o.apply_operator(ub.OP["map_b_a"], u, m.nz), rtol=1e-14, atol=1e-12)
    np.testing.assert_allclose(r["v_vav"], o.apply_operator(ub.OP["map_b_a"], r["v_vav_b"]), rtol=1e-14, atol=1e-12)
    np.testing.assert_allclose(r["R_shear"], (np.hypot(r["u_base"], r["v_base"]) + 0.1) / (np.hypot(r["u_surf"], r["v_surf"]) + 0.1), rtol=1e-14)


@pytest.mark.gpu
def test_gpu_secondary_velocities_match_oracle():
    w = wl.halfar(2.0e4)
    m = w.mesh
    o = ol.OracleMesh(m)
    u, v = _random_3D_velocities(m)
    ref = o.calc_secondary_velocities(u, v)
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    s, out = ub.alloc_secondary(m.nV, m.nTri, m.nz)
    ub.check(ub.lib().ufe_b200_calc_secondary_velocities(ctx.h, ub.dp(u), ub.dp(v), C.byref(s)))
    # second call on the resident fields (what follows a velocity solve): same result
    s2, out2 = ub.alloc_secondary(m.nV, m.nTri, m.nz)
    ub.check(ub.lib().ufe_b200_calc_secondary_velocities(ctx.h, None, None, C.byref(s2)))
    ctx.close()
    for k in ub.SECONDARY_FIELDS:
        assert rel_l2(out[k], ref[k]) < 1e-13, (k, rel_l2(out[k], ref[k]))
        assert np.array_equal(out[k], out2[k]), k
    np.testing.assert_array_equal(out["u_surf_b"], u[:, 0])


def _vertical_io(m, Hi, Hs, Hib, dHb_dt, dHi_dt, BMB, mg, mf):
    keep = [np.ascontiguousarray(a, float) for a in (Hi, Hs, Hib, dHb_dt, dHi_dt, BMB)] + [np.ascontiguousarray(a, np.int32) for a in (mg, mf)]
    w, dw = np.zeros((m.nV, m.nz), order="F"), np.zeros((m.nV, m.nz), order="F")
    io = ub.VerticalVelocityIO(*[ub.dp(a) for a in keep[:6]], ub.ip(keep[6]), ub.ip(keep[7]), ub.dp(w), ub.dp(dw))
    io._keep = keep
    return io, w, dw


def test_oracle_vertical_velocities_of_a_linear_flow():
    """calc_vertical_velocities (vertical_velocities.f90:23-196) + calc_dw_dz: for u = a x, v = b y, uniform in z, on a flat slab the Voronoi
    flux integral is exact, so dw/dz = -(a + b) and w = w_base - (a + b) Hi (1 - zeta) on every interior column"""
    w_ = wl.halfar(4.0e4)
    m = w_.mesh
    o = ol.OracleMesh(m)
    a, b, H = 2.0e-3, -0.5e-3, 1500.0
    gc = np.asarray(m.Tricc, float)                      # Voronoi vertices
    u_b = np.asfortranarray(np.repeat((a * gc[:, 0])[:, None], m.nz, 1)); v_b = np.asfortranarray(np.repeat((b * gc[:, 1])[:, None], m.nz, 1))
    u_a = np.asfortranarray(np.repeat((a * m.V[:, 0])[:, None], m.nz, 1)); v_a = np.asfortranarray(np.repeat((b * m.V[:, 1])[:, None], m.nz, 1))
    one, zero = np.ones(m.nV), np.zeros(m.nV)
    io, w, dw = _vertical_io(m, H * one, H * one, zero, zero, zero, zero, one, zero)
    o.calc_vertical_velocities(w_.cfg, io, u_b, v_b, u_a, v_a)
    zeta = np.asarray(m.zeta, float)
    interior = np.asarray(m.VBI) == 0
    for vi in np.nonzero(~interior)[0]:                  # also drop the ring next to the border (modified Voronoi cells there)
        interior[np.asarray(m.C)[vi, : m.nC[vi]] - 1] = False
    assert interior.sum() > m.nV // 2
    want = -(a + b) * H * (1.0 - zeta)
    np.testing.assert_allclose(w[interior], np.repeat(want[None, :], interior.sum(), 0), rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(dw[interior], -(a + b), rtol=1e-8)
    # ice-free columns: w = 0, dw_dz = NaN; thin ice: the basal value everywhere
    io, w, dw = _vertical_io(m, np.where(np.arange(m.nV) % 3 == 0, 5.0, H), H * one, zero, 0.3 * one, zero, -0.2 * one, np.arange(m.nV) % 2, zero)
    o.calc_vertical_velocities(w_.cfg, io, u_b, v_b, u_a, v_a)
    noice = np.arange(m.nV) % 2 == 0
    assert np.all(w[noice] == 0.0) and np.all(np.isnan(dw[noice]))
    thin = ~noice & (np.arange(m.nV) % 3 == 0)
    assert thin.any() and np.all(w[thin] == w[thin][:, -1:]) and np.allclose(w[thin][:, -1], 0.3 - 0.2)


def _vertical_case(m, seed=11):
    rng = np.random.default_rng(seed)
    x, y = m.V[:, 0], m.V[:, 1]
    L = m.xmax - m.xmin
    Hi = np.maximum(0.0, 2500.0 * np.cos(1.3 * np.pi * x / L) * np.cos(1.1 * np.pi * y / L) + 30.0 * rng.standard_normal(m.nV))
    Hi[rng.random(m.nV) < 0.02] = 4.0                      # some thin-ice columns
    Hb = -300.0 + 400.0 * np.sin(3.0 * x / L) + 20.0 * rng.standard_normal(m.nV)
    Hib = np.maximum(Hb, -0.9 * Hi)
    Hs = Hib + Hi
    mg = ((Hi > 0) & (Hib <= Hb)).astype(np.int32); mf = ((Hi > 0) & (Hib > Hb)).astype(np.int32)
    return Hi, Hs, Hib, 1e-3 * rng.standard_normal(m.nV), 0.5 * rng.standard_normal(m.nV), rng.standard_normal(m.nV), mg, mf


@pytest.mark.gpu
def test_gpu_vertical_velocities_match_oracle():
    w_ = wl.halfar(2.0e4)
    m = w_.mesh
    o = ol.OracleMesh(m)
    u, v = _random_3D_velocities(m)
    sec = o.calc_secondary_velocities(u, v)
    case = _vertical_case(m)
    assert case[6].sum() > 100 and case[7].sum() > 100 and (case[0] == 0).sum() > 100
    io, w_ref, dw_ref = _vertical_io(m, *case)
    cfg = w_.cfg
    o.calc_vertical_velocities(cfg, io, u, v, sec["u_3D"], sec["v_3D"])
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    io2, w_gpu, dw_gpu = _vertical_io(m, *case)
    assert ub.lib().ufe_b200_calc_vertical_velocities(ctx.h, C.byref(cfg), C.byref(io2)) != 0     # needs the a-grid velocities first
    s, out = ub.alloc_secondary(m.nV, m.nTri, m.nz)
    ub.check(ub.lib().ufe_b200_calc_secondary_velocities(ctx.h, ub.dp(u), ub.dp(v), C.byref(s)))
    ub.check(ub.lib().ufe_b200_calc_vertical_velocities(ctx.h, C.byref(cfg), C.byref(io2)))
    ctx.close()
    assert rel_l2(w_gpu, w_ref) < 1e-12, rel_l2(w_gpu, w_ref)
    ice = (case[6] + case[7]) > 0
    assert np.all(np.isnan(dw_gpu[~ice])) and np.all(np.isnan(dw_ref[~ice])) and np.all(w_gpu[~ice] == 0.0)
    assert rel_l2(dw_gpu[ice], dw_ref[ice]) < 1e-12, rel_l2(dw_gpu[ice], dw_ref[ice])
