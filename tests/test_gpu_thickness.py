"""GPU: ice-thickness step (explicit and semi-implicit calc_dHi_dt) vs the CPU oracle, through the C ABI."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    n = np.linalg.norm(b)
    return np.linalg.norm(a - b) / (n if n > 0 else 1.0)


@pytest.fixture(scope="module")
def setup():
    w = wl.mismip_mod(2.5e4)
    ctx = ub.Context(device=0)
    ctx.mesh_create(w.mesh)
    ctx.calc_matrix_operators()
    o = ol.OracleMesh(w.mesh)
    # a plausible velocity field: radial outflow growing with radius
    gx, gy = w.mesh.TriGC[:, 0], w.mesh.TriGC[:, 1]
    r = np.hypot(gx, gy) + 1.0
    sp = 800.0 * np.clip(r / 9e5, 0, 1) ** 2
    w.u = np.ascontiguousarray(sp * gx / r); w.v = np.ascontiguousarray(sp * gy / r)
    yield w, ctx, o
    ctx.close()


def _io(w, rng=None, partial_margin=False, noice=False, prescribed=False):
    f = w.f
    m = w.mesh
    a = dict(Hi=f["Hi"].copy(), Hb=f["Hb"].copy(), SL=f["SL"].copy(), u_vav_b=w.u, v_vav_b=w.v, SMB=f["SMB"].copy(), BMB=f["BMB"].copy(),
             LMB=f["LMB"].copy(), fraction_margin=f["fraction_margin"].copy(), mask_noice=f["mask_noice"].copy(),
             dHi_dt_target=f["dHi_dt_target"].copy(), BC_prescr_mask=None, BC_prescr_Hi=None)
    if rng is not None:
        a["BMB"] = -0.5 * rng.random(m.nV); a["LMB"] = -0.1 * rng.random(m.nV); a["dHi_dt_target"] = 0.01 * rng.standard_normal(m.nV)
    if partial_margin:
        a["fraction_margin"] = np.where(rng.random(m.nV) < 0.1, 0.5, 1.0)
    if noice:
        a["mask_noice"] = (np.hypot(m.V[:, 0], m.V[:, 1]) > 8.5e5).astype(np.int32)
    if prescribed:
        a["BC_prescr_mask"] = (rng.random(m.nV) < 0.03).astype(np.int32)
        a["BC_prescr_Hi"] = 500.0 * rng.random(m.nV)
    out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
    io = ub.ThicknessIO()
    for k, v in {**a, **out}.items():
        setattr(io, k, None if v is None else (ub.ip(v) if v.dtype == np.int32 else ub.dp(v)))
    return io, a, out


@pytest.mark.parametrize("method", [ub.INTEGRATION_EXPLICIT, ub.INTEGRATION_SEMI_IMPLICIT])
@pytest.mark.parametrize("variant", ["plain", "forcing", "margin_noice", "prescribed", "bc_infinite"])
def test_calc_dHi_dt_matches_oracle(setup, method, variant):
    w, ctx, o = setup
    cfg = wl.scaling_config()
    cfg.choice_ice_integration_method = method
    cfg.dHi_PETSc_rtol = 1e-13; cfg.dHi_PETSc_abstol = 1e-30
    rng = np.random.default_rng(4)
    kw = dict(plain={}, forcing=dict(rng=rng), margin_noice=dict(rng=rng, partial_margin=True, noice=True),
              prescribed=dict(rng=rng, prescribed=True), bc_infinite=dict(rng=rng))[variant]
    if variant == "bc_infinite":
        for k in range(4):
            cfg.BC_H[k] = ub.BC_INFINITE
    io_g, a, g = _io(w, **kw)
    dt = C.c_double(2.0)
    its = C.c_int()
    ub.check(ub.lib().ufe_b200_calc_dHi_dt(ctx.h, C.byref(cfg), C.byref(io_g), C.cast(C.byref(dt), ub.c_dp), 1, C.byref(its)))
    io_o, a2, r = _io(w, **(dict(kw, rng=np.random.default_rng(4)) if "rng" in kw else kw))
    rc, dt_o, its_o = o.calc_dHi_dt(cfg, io_o, 2.0, ksp=ol.KSP_GMRES_ILU0 if method == ub.INTEGRATION_SEMI_IMPLICIT else 0)
    assert rc == 0 and abs(dt.value - dt_o) < 1e-12
    for f in ("divQ", "Hi_tplusdt", "dHi_dt"):
        assert rel_l2(g[f], r[f]) < 1e-9, (f, rel_l2(g[f], r[f]))
    assert np.abs(g["AMB"] - r["AMB"]).max() < 1e-6 * max(1.0, np.abs(r["dHi_dt"]).max())
    assert (g["Hi_tplusdt"] >= 0).all()


def test_flux_divergence_conserves_mass(setup):
    # size-independent property: with full margins and no forcing, sum_i A_i divQ_i = net flux through the border ~ 0 here
    w, ctx, o = setup
    cfg = wl.scaling_config()
    cfg.choice_ice_integration_method = ub.INTEGRATION_EXPLICIT
    io_g, a, g = _io(w)
    a["SMB"][:] = 0.0
    dt = C.c_double(1.0); its = C.c_int()
    ub.check(ub.lib().ufe_b200_calc_dHi_dt(ctx.h, C.byref(cfg), C.byref(io_g), C.cast(C.byref(dt), ub.c_dp), 1, C.byref(its)))
    tot = (w.mesh.A * g["divQ"]).sum()
    scale = (w.mesh.A * np.abs(g["divQ"])).sum()
    assert abs(tot) < 1e-9 * scale
