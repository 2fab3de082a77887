"""GPU: the predictor-corrector time step (device phases of run_ice_dynamics_model_pc, predictor_corrector_scheme.f90:118-233).
(1) the device-resident step vs the same call sequence driven on the host with the CPU oracle;
(2) the host-buffer entry ufe_b200_pc_step (overlapped transfers) vs the resident step."""
import ctypes as C

import numpy as np
import pytest

import bench
import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl

pytestmark = pytest.mark.gpu
DT = 0.2
FIELDS = bench.STATE_FIELDS + ["u_b_prev", "v_b_prev"]


def rel_l2(a, b):
    n = np.linalg.norm(b)
    return np.linalg.norm(a - b) / (n if n > 0 else 1.0)


def _workload():
    w = wl.mismip_mod(2.5e4)
    cfg = w.cfg
    cfg.write_diagnostics = 0
    cfg.stress_balance_PETSc_rtol = 1e-11; cfg.stress_balance_PETSc_abstol = 1e-30
    cfg.dHi_PETSc_rtol = 1e-12; cfg.dHi_PETSc_abstol = 1e-30
    cfg.visc_it_norm_dUV_tol = 1e-7
    return w


def _spin_up(ctx, w):
    m, cfg = w.mesh, w.cfg
    L = ub.lib()
    ice = w.ice_inputs()
    st, A = ub.alloc_velocity_state(m.nV, m.nTri, m.nz, fields=FIELDS)
    nv, na = C.c_int(), C.c_int()
    ub.check(L.ufe_b200_velocity_upload(ctx.h, C.byref(ice), C.byref(st)))
    ub.check(L.ufe_b200_velocity_solve_resident(ctx.h, C.byref(cfg), 1, C.byref(nv), C.byref(na)))
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    return ice, st, A


@pytest.mark.parametrize("pc,ref_loop,w_lag", [(ub.PC_JACOBI, 1, 0.0), (ub.PC_MG, 1, 0.0), (ub.PC_MG, 1, 0.6), (ub.PC_MG, 0, 0.0)])
def test_resident_pc_steps_match_oracle_sequence(pc, ref_loop, w_lag):
    """the one-call step against the reference's loop body driven on the host with the oracle, for the reference's loop semantics (the default: lagged
    region%ice%Hi one whole step behind, or interpolated with the weight of the last action time) and for Eq. 30 as printed"""
    NSTEP = 3
    w = _workload()
    m, cfg = w.mesh, w.cfg
    cfg.pc_type = pc
    L = ub.lib()
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    ice, st, A = _spin_up(ctx, w)
    io, keep = bench.thickness_io(w, slice(None), w.f["Hi"].copy())
    ub.check(L.ufe_b200_thickness_upload(ctx.h, C.byref(io)))
    nv, na, nt = C.c_int(), C.c_int(), C.c_int()
    if not (ref_loop == 1 and w_lag == 0.0):            # (1, 0.0) is the default and is not set explicitly
        ub.check(L.ufe_b200_pc_reference_loop(ctx.h, ref_loop, w_lag))
    for _ in range(NSTEP):
        ub.check(L.ufe_b200_pc_step_resident(ctx.h, C.byref(cfg), 1, DT, C.byref(nv), C.byref(na), C.byref(nt)))
    Hi_g = np.zeros(m.nV); dH_g = np.zeros(m.nV)
    ub.check(L.ufe_b200_thickness_download(ctx.h, ub.dp(Hi_g), ub.dp(dH_g), None))
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    pc_g = {k: np.zeros(m.nV) for k in ("Hi_star", "d_n", "d_next", "model")}
    ub.check(L.ufe_b200_pc_download(ctx.h, ub.dp(pc_g["Hi_star"]), ub.dp(pc_g["d_n"]), ub.dp(pc_g["d_next"]), ub.dp(pc_g["model"])))
    u_g = A["u_vav_b"].copy()
    ctx.close()

    # the same two steps with the oracle behind the host call sequence (what bench.py's CPU arm times)
    o = ol.OracleMesh(m)
    w2 = _workload()
    cfg2 = w2.cfg
    st2, A2 = ub.alloc_velocity_state(m.nV, m.nTri, m.nz, fields=FIELDS)
    out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}

    def calc(Hi, u, v, dt):
        io2, _k = bench.thickness_io(w2, slice(None), Hi, u, v, out)
        rc, _dt, its = o.calc_dHi_dt(cfg2, io2, dt)
        assert rc == 0
        return out["dHi_dt"].copy(), its

    def solve(Hi, Hs):
        w2.f["Hi"], w2.f["Hs"] = Hi, Hs
        rc, a, b = o.solve_velocity(cfg2, 1, w2.ice_inputs(), st2)
        assert rc == 0
        return a, b

    rc, _a, _b = o.solve_velocity(cfg2, 1, w2.ice_inputs(), st2)
    assert rc == 0
    S = dict(Hi=w2.f["Hi"].copy(), dHdt_nm1=np.zeros(m.nV), dt_n=DT, u=A2["u_vav_b"], v=A2["v_vav_b"])
    he = lambda Hi: o.calc_effective_thickness(cfg2, Hi, w2.f["Hb"], w2.f["SL"])[0]
    for _ in range(NSTEP):
        bench.pc_step_host(calc, solve, w2, cfg2, slice(None), S, DT, hi_eff=he, reference_loop=bool(ref_loop), w_lag=w_lag)
    assert rel_l2(Hi_g, S["Hi"]) < 1e-9, rel_l2(Hi_g, S["Hi"])
    assert rel_l2(u_g, A2["u_vav_b"]) < 1e-6, rel_l2(u_g, A2["u_vav_b"])
    # what the host model keeps in type_ice_pc: pc%Hi_star_np1, the re-derived pc%dHi_dt_Hi_n_u_n, the next pc%dHi_dt_Hi_nm1_u_nm1, region%ice%Hi
    assert rel_l2(pc_g["Hi_star"], S["Hi_star"]) < 1e-9 and (not ref_loop or rel_l2(pc_g["model"], S["Hi_model"]) < 1e-9)
    scale = np.abs(S["dHdt_nm1"]).max()
    assert np.abs(pc_g["d_n"] - S["dHdt_n"]).max() < 1e-7 * scale and np.abs(pc_g["d_next"] - S["dHdt_nm1"]).max() < 1e-7 * scale
    if ref_loop and w_lag == 0.0:
        assert rel_l2(S["Hi_model"], S["Hi"]) > 1e-6          # region%ice%Hi really is one step behind


def test_host_buffer_pc_step_equals_resident():
    w = _workload()
    m, cfg = w.mesh, w.cfg
    cfg.pc_type = ub.PC_MG
    L = ub.lib()
    nv, na, nt = C.c_int(), C.c_int(), C.c_int()
    # resident
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    ice, st, A = _spin_up(ctx, w)
    A0 = {k: v.copy() for k, v in A.items()}
    io, keep = bench.thickness_io(w, slice(None), w.f["Hi"].copy())
    ub.check(L.ufe_b200_thickness_upload(ctx.h, C.byref(io)))
    for _ in range(2):
        ub.check(L.ufe_b200_pc_step_resident(ctx.h, C.byref(cfg), 1, DT, C.byref(nv), C.byref(na), C.byref(nt)))
    Hi_r = np.zeros(m.nV)
    ub.check(L.ufe_b200_thickness_download(ctx.h, ub.dp(Hi_r), None, None))
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    res = {k: v.copy() for k, v in A.items()}
    its_r = (nv.value, na.value, nt.value)
    ctx.close()
    # host buffers: everything goes up and comes back every step
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    for k in A:
        A[k][...] = A0[k]
    out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
    io2, keep2 = bench.thickness_io(w, slice(None), w.f["Hi"].copy(), out=out)
    dH_nm1 = np.zeros(m.nV)
    dt_prev = 0.0
    for _ in range(2):
        ub.check(L.ufe_b200_pc_step(ctx.h, C.byref(cfg), 1, DT, dt_prev, C.byref(ice), C.byref(st), C.byref(io2), ub.dp(dH_nm1),
                                    C.byref(nv), C.byref(na), C.byref(nt)))
        keep2["Hi"][:] = out["Hi_tplusdt"]       # the host accepts the step
        dt_prev = DT
    ctx.close()
    assert (nv.value, na.value, nt.value) == its_r
    assert np.array_equal(out["Hi_tplusdt"], Hi_r)
    for k in ("u_vav_b", "v_vav_b", "u_3D_b", "eta_3D_b", "tau_bx_b"):
        assert np.array_equal(A[k], res[k]), k


def test_host_buffer_pc_step_with_resident_solver_state():
    """ufe_b200_set_state_resident: the state is uploaded once; later calls ignore the host copies of it and only bring results back"""
    w = _workload()
    m, cfg = w.mesh, w.cfg
    cfg.pc_type = ub.PC_MG
    L = ub.lib()
    nv, na, nt = C.c_int(), C.c_int(), C.c_int()
    results = []
    for resident in (0, 1):
        ctx = ub.Context(device=0)
        ctx.mesh_create(m); ctx.calc_matrix_operators()
        ice, st, A = _spin_up(ctx, w)
        ub.check(L.ufe_b200_set_state_resident(ctx.h, resident))
        out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
        io2, keep2 = bench.thickness_io(w, slice(None), w.f["Hi"].copy(), out=out)
        dH_nm1 = np.zeros(m.nV)
        dt_prev = 0.0
        for it in range(3):
            ub.check(L.ufe_b200_pc_step(ctx.h, C.byref(cfg), 1, DT, dt_prev, C.byref(ice), C.byref(st), C.byref(io2), ub.dp(dH_nm1),
                                        C.byref(nv), C.byref(na), C.byref(nt)))
            keep2["Hi"][:] = out["Hi_tplusdt"]
            dt_prev = DT
            if resident and it == 0:
                A["eta_3D_b"][...] = np.nan; A["u_base_b"][...] = np.nan     # must not be read again
        results.append((out["Hi_tplusdt"].copy(), A["u_vav_b"].copy(), A["u_3D_b"].copy(), (nv.value, na.value, nt.value)))
        ctx.close()
    assert results[0][3] == results[1][3]
    for a, b in zip(results[0][:3], results[1][:3]):
        assert np.array_equal(a, b)


def test_resident_pc_steps_with_the_geometry_stage_on_the_device():
    """ufe_b200_geometry_resident: masks, grounded fractions and margin fractions follow the evolving thickness inside the resident step;
    the oracle runs determine_masks / calc_grounded_fractions / calc_effective_thickness between its phases as run_ice_dynamics_model_pc does"""
    w = _workload()
    m, cfg = w.mesh, w.cfg
    cfg.pc_type = ub.PC_MG
    rng = np.random.default_rng(17)
    nbins = 11
    dHb = 2.0 * np.sin(m.V[:, 0] / 3e5)
    Tri = np.asarray(m.Tri) - 1
    bins = np.linspace(-1.0, 1.0, nbins)[None, :] ** 3
    cdf = np.asfortranarray(w.f["Hb"][:, None] + (30.0 + 40.0 * rng.random(m.nV))[:, None] * bins)
    cdf_b = np.asfortranarray(w.f["Hb"][Tri].mean(axis=1)[:, None] + (30.0 + 40.0 * rng.random(m.nTri))[:, None] * bins)
    choice = ub.GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF
    NSTEP, DTL = 3, 1.0
    L = ub.lib()
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    ice, st, A = _spin_up(ctx, w)
    io, keep = bench.thickness_io(w, slice(None), w.f["Hi"].copy())
    ub.check(L.ufe_b200_thickness_upload(ctx.h, C.byref(io)))
    ub.check(L.ufe_b200_geometry_resident(ctx.h, choice, nbins, ub.dp(dHb), ub.dp(cdf), ub.dp(cdf_b)))
    nv, na, nt = C.c_int(), C.c_int(), C.c_int()
    for _ in range(NSTEP):
        ub.check(L.ufe_b200_pc_step_resident(ctx.h, C.byref(cfg), 1, DTL, C.byref(nv), C.byref(na), C.byref(nt)))
    Hi_g = np.zeros(m.nV)
    ub.check(L.ufe_b200_thickness_download(ctx.h, ub.dp(Hi_g), None, None))
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    u_g = A["u_vav_b"].copy()
    # the same steps with the geometry stage switched off again: the evolving masks must matter for this to be a test
    ctx.close()

    o = ol.OracleMesh(m)
    w2 = _workload()
    cfg2 = w2.cfg
    st2, A2 = ub.alloc_velocity_state(m.nV, m.nTri, m.nz, fields=FIELDS)
    out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
    changed = []

    def geometry(Hi, for_velocity):
        f = w2.f
        if for_velocity:
            mk = o.determine_masks(cfg2, Hi, f["Hb"], f["SL"])
            changed.append(int((mk["mask_grounded_ice"] != f["mask_grounded_ice"]).sum()))
            f["mask_grounded_ice"], f["mask_icefree_land"] = mk["mask_grounded_ice"], mk["mask_icefree_land"]
            fr, fr_b = o.calc_grounded_fractions(cfg2, choice, nbins, Hi, f["Hb"], f["SL"], dHb, mk["mask_floating_ice"], cdf, cdf_b)
            f["fraction_gr_b"] = fr_b
        else:
            he, fm = o.calc_effective_thickness(cfg2, Hi, f["Hb"], f["SL"])
            f["fraction_margin"] = fm

    def calc(Hi, u, v, dt):
        io2, _k = bench.thickness_io(w2, slice(None), Hi, u, v, out)
        rc, _dt, its = o.calc_dHi_dt(cfg2, io2, dt)
        assert rc == 0
        return out["dHi_dt"].copy(), its

    def solve(Hi, Hs):
        w2.f["Hi"], w2.f["Hs"] = Hi, Hs
        rc, a, b = o.solve_velocity(cfg2, 1, w2.ice_inputs(), st2)
        assert rc == 0
        return a, b

    rc, _a, _b = o.solve_velocity(cfg2, 1, w2.ice_inputs(), st2)
    assert rc == 0
    S = dict(Hi=w2.f["Hi"].copy(), dHdt_nm1=np.zeros(m.nV), dt_n=DTL, u=A2["u_vav_b"], v=A2["v_vav_b"])
    fgb0 = w2.f["fraction_gr_b"].copy()
    he = lambda Hi: o.calc_effective_thickness(cfg2, Hi, w2.f["Hb"], w2.f["SL"])[0]
    for _ in range(NSTEP):
        bench.pc_step_host(calc, solve, w2, cfg2, slice(None), S, DTL, geometry=geometry, hi_eff=he)
    assert rel_l2(w2.f["fraction_gr_b"], fgb0) > 1e-3          # the stage did change what the velocity solve sees
    assert rel_l2(Hi_g, S["Hi"]) < 1e-9, rel_l2(Hi_g, S["Hi"])
    assert rel_l2(u_g, A2["u_vav_b"]) < 1e-6, rel_l2(u_g, A2["u_vav_b"])


def test_host_buffer_pc_step_with_the_geometry_stage_equals_resident():
    """ufe_b200_pc_step uploads the host's masks / fractions on the copy stream; with ufe_b200_geometry_resident the device's own values
    (computed after that upload has landed) are the ones the solves use, exactly as in the resident step"""
    w = _workload()
    m, cfg = w.mesh, w.cfg
    cfg.pc_type = ub.PC_MG
    L = ub.lib()
    nv, na, nt = C.c_int(), C.c_int(), C.c_int()
    choice, DTL = ub.GF_BILIN_INTERP_TAF, 1.0
    # resident
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    ice, st, A = _spin_up(ctx, w)
    A0 = {k: v.copy() for k, v in A.items()}
    io, keep = bench.thickness_io(w, slice(None), w.f["Hi"].copy())
    ub.check(L.ufe_b200_thickness_upload(ctx.h, C.byref(io)))
    ub.check(L.ufe_b200_geometry_resident(ctx.h, choice, 0, None, None, None))
    for _ in range(2):
        ub.check(L.ufe_b200_pc_step_resident(ctx.h, C.byref(cfg), 1, DTL, C.byref(nv), C.byref(na), C.byref(nt)))
    Hi_r = np.zeros(m.nV)
    ub.check(L.ufe_b200_thickness_download(ctx.h, ub.dp(Hi_r), None, None))
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    u_r = A["u_vav_b"].copy()
    its_r = (nv.value, na.value, nt.value)
    ctx.close()
    # host buffers; the host's own (stale) masks and fractions go up every step and must be overridden on the device
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    for k in A:
        A[k][...] = A0[k]
    ub.check(L.ufe_b200_geometry_resident(ctx.h, choice, 0, None, None, None))
    out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
    io2, keep2 = bench.thickness_io(w, slice(None), w.f["Hi"].copy(), out=out)
    dH_nm1 = np.zeros(m.nV)
    dt_prev = 0.0
    for _ in range(2):
        ub.check(L.ufe_b200_pc_step(ctx.h, C.byref(cfg), 1, DTL, dt_prev, C.byref(ice), C.byref(st), C.byref(io2), ub.dp(dH_nm1),
                                    C.byref(nv), C.byref(na), C.byref(nt)))
        keep2["Hi"][:] = out["Hi_tplusdt"]
        dt_prev = DTL
    ctx.close()
    assert (nv.value, na.value, nt.value) == its_r
    assert np.array_equal(out["Hi_tplusdt"], Hi_r) and np.array_equal(A["u_vav_b"], u_r)
