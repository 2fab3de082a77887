"""calc_pc_truncation_error and the count of unstable vertices of run_ice_dynamics_model_pc (time_stepping/predictor_corrector_scheme.f90:327-361,
254-279; Robinson et al., 2020, Eq. 32): oracle against the definition, device against the oracle (bit-exact: pointwise arithmetic, max, counts)."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl


def _case(nV, seed=4):
    rng = np.random.default_rng(seed)
    Hi_np1 = 1000.0 * rng.random(nV); Hi_star = Hi_np1 + rng.standard_normal(nV) * (rng.random(nV) < 0.3)
    grounded = (rng.random(nV) < 0.7).astype(np.int32); gl_gr = (grounded & (rng.random(nV) < 0.1)).astype(np.int32)
    frac = np.where(rng.random(nV) < 0.6, 1.0, rng.random(nV))
    guilty = rng.integers(0, 5, nV).astype(np.int32)
    return Hi_np1, Hi_star, grounded, gl_gr, frac, guilty


def _oracle(nV, zeta, dt_n, eta_min, eps, Hi_np1, Hi_star, grounded, gl_gr, frac, guilty):
    tau = np.zeros(nV); g = guilty.copy(); eta = C.c_double(); nt, ng = C.c_int(), C.c_int()
    ol.lib().orc_calc_pc_truncation_error(nV, zeta, dt_n, eta_min, eps, ub.dp(Hi_np1), ub.dp(Hi_star), ub.ip(grounded), ub.ip(gl_gr), ub.dp(frac),
                                          ub.dp(tau), ub.ip(g), C.byref(eta), C.byref(nt), C.byref(ng))
    return tau, g, eta.value, nt.value, ng.value


def test_oracle_truncation_error_definition():
    nV = 5000
    case = _case(nV)
    Hi_np1, Hi_star, grounded, gl_gr, frac, guilty = case
    zeta, dt_n, eta_min, eps = 1.2, 0.5, 1e-8, 0.2
    tau, g, eta, nt, ng = _oracle(nV, zeta, dt_n, eta_min, eps, *case)
    want = zeta * np.abs(Hi_np1 - Hi_star) / ((3 * zeta + 3) * dt_n)
    np.testing.assert_allclose(tau, want, rtol=1e-15)
    sel = (grounded == 1) & (gl_gr == 0) & (frac == 1.0)
    assert eta == max(eta_min, tau[sel].max()) and tau[~sel].max() > 0.0
    full = frac >= 1.0
    assert nt == full.sum() and ng == (full & (tau > eps)).sum() and 0 < ng < nt
    assert np.array_equal(g, guilty + (full & (tau > eps)))
    # nothing changes anywhere: eta falls back to C%pc_eta_min
    assert _oracle(nV, zeta, dt_n, eta_min, eps, Hi_np1, Hi_np1, grounded, gl_gr, frac, guilty)[2:] == (eta_min, full.sum(), 0)


@pytest.mark.gpu
def test_gpu_truncation_error_matches_oracle():
    w = wl.mismip_mod(2.5e4)
    m = w.mesh
    case = _case(m.nV, seed=8)
    zeta, dt_n, eta_min, eps = 0.9, 0.2, 1e-8, 0.15
    tau_o, g_o, eta_o, nt_o, ng_o = _oracle(m.nV, zeta, dt_n, eta_min, eps, *case)
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(m); ctx.calc_matrix_operators()
        Hi_np1, Hi_star, grounded, gl_gr, frac, guilty = case
        tau = np.zeros(m.nV); g = guilty.copy(); eta = C.c_double(); nt, ng = C.c_int32(), C.c_int32()
        ub.check(ub.lib().ufe_b200_calc_pc_truncation_error(ctx.h, zeta, dt_n, eta_min, eps, ub.dp(Hi_np1), ub.dp(Hi_star), ub.ip(grounded), ub.ip(gl_gr),
                                                            ub.dp(frac), ub.dp(tau), ub.ip(g), C.byref(eta), C.byref(nt), C.byref(ng)))
        assert np.array_equal(tau, tau_o) and np.array_equal(g, g_o)
        assert (eta.value, nt.value, ng.value) == (eta_o, nt_o, ng_o) and ng_o > 0
        # optional outputs left out; no change at all -> C%pc_eta_min
        ub.check(ub.lib().ufe_b200_calc_pc_truncation_error(ctx.h, zeta, dt_n, eta_min, eps, ub.dp(Hi_np1), ub.dp(Hi_np1), ub.ip(grounded), ub.ip(gl_gr),
                                                            ub.dp(frac), None, None, C.byref(eta), C.byref(nt), C.byref(ng)))
        assert (eta.value, nt.value, ng.value) == (eta_min, nt_o, 0)
    finally:
        ctx.close()
