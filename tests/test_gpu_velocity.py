"""GPU: SSA/DIVA stiffness assembly and velocity solve vs the CPU oracle, through the C ABI (wide seam)."""
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
    o.calc_matrix_operators()
    yield w, ctx, o
    ctx.close()


def _linearised(ctx, cfg, m, N_b, dNx, dNy, beta, tdx, tdy, rtol, abstol, bc=None):
    u = np.zeros(m.nTri); v = np.zeros(m.nTri); up = np.zeros(m.nTri); vp = np.zeros(m.nTri)
    its = C.c_int()
    bm, bu, bv = bc if bc else (None, None, None)
    ub.check(ub.lib().ufe_b200_solve_SSA_DIVA_linearised(ctx.h, C.byref(cfg), ub.dp(u), ub.dp(v), ub.dp(N_b), ub.dp(dNx), ub.dp(dNy),
                                                         ub.dp(beta), ub.dp(tdx), ub.dp(tdy), ub.dp(up), ub.dp(vp), rtol, abstol,
                                                         C.byref(its), ub.ip(bm), ub.dp(bu), ub.dp(bv)))
    return u, v, its.value


@pytest.mark.parametrize("variant", ["crossterms", "sans", "bc_zero", "prescribed"])
def test_stiffness_pattern_bit_exact(setup, variant):
    w, ctx, o = setup
    m = w.mesh
    cfg = wl.scaling_config()
    rng = np.random.default_rng(2)
    N_b = 1e10 * (1 + 0.1 * rng.random(m.nTri)); dNx = 1e3 * rng.standard_normal(m.nTri); dNy = 1e3 * rng.standard_normal(m.nTri)
    beta = 1e3 * rng.random(m.nTri); tdx = 1e4 * rng.standard_normal(m.nTri); tdy = 1e4 * rng.standard_normal(m.nTri)
    bc = None
    if variant == "sans":
        cfg.do_include_SSADIVA_crossterms = 0
    if variant == "bc_zero":
        cfg.BC_u[0] = ub.BC_ZERO; cfg.BC_v[1] = ub.BC_ZERO; cfg.BC_u[3] = ub.BC_ZERO
    if variant == "prescribed":
        bm = (rng.random(m.nTri) < 0.05).astype(np.int32)
        bc = (bm, 10.0 * rng.standard_normal(m.nTri), 10.0 * rng.standard_normal(m.nTri))
    u, v, its = _linearised(ctx, cfg, m, N_b, dNx, dNy, beta, tdx, tdy, 1e-12, 1e-30, bc)
    ptr, ind, val, bb = ctx.get_stiffness()
    optr, oind, oval, obb = o.assemble(cfg, N_b, dNx, dNy, beta, tdx, tdy, *(bc or (None, None, None)))
    assert np.array_equal(ptr, optr) and np.array_equal(ind, oind)          # bit-exact pattern
    assert np.abs(val - oval).max() <= 1e-13 * np.abs(oval).max()
    assert np.abs(bb - obb).max() <= 1e-13 * np.abs(obb).max()
    xo, its_o, rc = ol.ksp_solve(optr, oind, oval, obb, 1e-12, 1e-30, ksp=ol.KSP_GMRES_ILU0)
    assert rc == 0
    uv = np.empty(2 * m.nTri); uv[0::2] = u; uv[1::2] = v
    assert rel_l2(uv, xo) < 1e-6, (variant, its, its_o)
    if variant == "prescribed":
        assert np.abs(u[bc[0] == 1] - bc[1][bc[0] == 1]).max() < 1e-9


def test_same_prescribed_mask_keeps_the_pattern_and_new_values_or_mask_take_effect(setup):
    """upload_bc (velocity.cu): the pattern depends on the mask only; new prescribed values with the same mask must reach the load vector without a pattern
    rebuild (fewer launches than the first call), a different mask must rebuild it"""
    w, ctx, o = setup
    m = w.mesh
    cfg = wl.scaling_config()
    rng = np.random.default_rng(12)
    co = (1e10 * (1 + 0.1 * rng.random(m.nTri)), 1e3 * rng.standard_normal(m.nTri), 1e3 * rng.standard_normal(m.nTri), 1e3 * rng.random(m.nTri),
          1e4 * rng.standard_normal(m.nTri), 1e4 * rng.standard_normal(m.nTri))
    mask1 = (rng.random(m.nTri) < 0.05).astype(np.int32); mask2 = (rng.random(m.nTri) < 0.08).astype(np.int32)
    launches = []
    for mask, scale in ((mask1, 10.0), (mask1, -3.0), (mask2, 5.0)):
        bc = (mask, scale * np.ones(m.nTri), 2.0 * scale * np.ones(m.nTri))
        l0 = ctx.kernel_launches()
        u, v, its = _linearised(ctx, cfg, m, *co, 1e-12, 1e-30, bc)
        launches.append((ctx.kernel_launches() - l0) / max(its, 1))
        ptr, ind, val, bb = ctx.get_stiffness()
        optr, oind, oval, obb = o.assemble(cfg, *co, *bc)
        assert np.array_equal(ptr, optr) and np.array_equal(ind, oind)
        assert np.abs(bb - obb).max() <= 1e-13 * np.abs(obb).max()
        assert np.abs(u[mask == 1] - scale).max() < 1e-9 and np.abs(v[mask == 1] - 2.0 * scale).max() < 1e-9


# field -> tolerance (relative L2) after a fixed number of viscosity iterations at tight KSP tolerance
_FIELDS = ["u_vav_b", "v_vav_b", "u_base_b", "v_base_b", "tau_bx_b", "tau_by_b", "eta_3D_b", "u_b_prev", "v_b_prev", "tau_dx_b", "tau_dy_b",
           "u_3D_b", "v_3D_b", "du_dx_a", "du_dy_a", "dv_dx_a", "dv_dy_a", "du_dz_3D_a", "dv_dz_3D_a", "eta_3D_a", "eta_vav_a", "N_a", "N_b",
           "dN_dx_b", "dN_dy_b", "F1_3D_a", "F2_3D_a", "F1_3D_b", "F2_3D_b", "basal_friction_coefficient_a", "basal_friction_coefficient_b",
           "beta_eff_a", "beta_eff_b"]


def _solve_gpu(ctx, cfg, approx, ice, m, bc=(None, None, None)):
    st, arrs = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    nv, na = C.c_int(), C.c_int()
    fn = ub.lib().ufe_b200_solve_DIVA if approx == 1 else ub.lib().ufe_b200_solve_SSA
    ub.check(fn(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), ub.ip(bc[0]), ub.dp(bc[1]), ub.dp(bc[2])))
    return arrs, nv.value, na.value


@pytest.mark.parametrize("law", [ub.SLID_ZOET_IVERSON, ub.SLID_WEERTMAN, ub.SLID_COULOMB, ub.SLID_BUDD, ub.SLID_TSAI2015, ub.SLID_SCHOOF2005])
def test_diva_every_stage_matches_oracle(setup, law):
    w, ctx, o = setup
    m = w.mesh
    cfg = wl.scaling_config()
    cfg.choice_sliding_law = law
    cfg.stress_balance_PETSc_rtol = 1e-13; cfg.stress_balance_PETSc_abstol = 1e-30
    cfg.visc_it_nit = 2          # 3 viscosity iterations: every kernel runs with non-trivial state
    cfg.visc_it_norm_dUV_tol = 1e-30
    ice = w.ice_inputs()
    g, nv, na = _solve_gpu(ctx, cfg, 1, ice, m)
    st2, r = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    rc, nv2, na2 = o.solve_velocity(cfg, 1, ice, st2)
    assert rc == 0 and nv == nv2 == 3
    for f in _FIELDS:
        assert rel_l2(g[f], r[f]) < 1e-8, (f, rel_l2(g[f], r[f]))
    assert abs(na - na2) <= max(4, na2 // 4)      # BiCGStab iteration counts at rtol 1e-13 are round-off sensitive


def test_diva_converged_solve_matches_oracle_and_petsc_default(setup):
    w, ctx, o = setup
    m = w.mesh
    cfg = wl.scaling_config()
    cfg.stress_balance_PETSc_rtol = 1e-12; cfg.stress_balance_PETSc_abstol = 1e-30
    ice = w.ice_inputs()
    g, nv, na = _solve_gpu(ctx, cfg, 1, ice, m)
    st2, r = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    rc, nv2, na2 = o.solve_velocity(cfg, 1, ice, st2, ksp=ol.KSP_GMRES_ILU0)     # the reference's default PETSc configuration
    assert rc == 0 and nv == nv2
    for f in ("u_vav_b", "v_vav_b", "u_3D_b", "tau_bx_b", "eta_3D_b"):
        assert rel_l2(g[f], r[f]) < 1e-6, (f, rel_l2(g[f], r[f]))            # north_star tolerance
    # at the config's own loose tolerance the two Krylov methods differ at the 1e-4 level; the nonlinear
    # solve still agrees to solver tolerance and needs the same number of viscosity iterations +- 1
    cfg = wl.scaling_config()
    g2, nv3, na3 = _solve_gpu(ctx, cfg, 1, ice, m)
    assert abs(nv3 - nv) <= 1 and rel_l2(g2["u_vav_b"], g["u_vav_b"]) < 5e-3


def test_blocked_and_csr_storage_agree(setup):
    # the Krylov solver streams a 2x2-blocked sliced-ELL copy of the stiffness matrix by default; the CSR matrix itself is the
    # other option.  Same algorithm, same reductions -> same iteration counts and (to round-off) the same solution.
    w, ctx, o = setup
    m = w.mesh
    ice = w.ice_inputs()
    res = []
    for fmt in (0, 1):
        cfg = wl.scaling_config()
        cfg.stiffness_format = fmt
        cfg.visc_it_nit = 3; cfg.visc_it_norm_dUV_tol = 1e-30
        cfg.stress_balance_PETSc_rtol = 1e-10; cfg.stress_balance_PETSc_abstol = 1e-30
        res.append(_solve_gpu(ctx, cfg, 1, ice, m))
    (g0, nv0, na0), (g1, nv1, na1) = res
    # the two formats sum a row in different orders (2x2 blocks / sliced-ELL with four accumulators): BiCGStab at rtol 1e-10 answers round-off with a
    # few per cent of its ~1000 iterations
    assert nv0 == nv1 and abs(na0 - na1) <= max(3, na1 // 8), (na0, na1)
    assert rel_l2(g0["u_vav_b"], g1["u_vav_b"]) < 1e-8 and rel_l2(g0["v_vav_b"], g1["v_vav_b"]) < 1e-8


def test_initial_guess_option_reaches_same_solution(setup):
    w, ctx, o = setup
    m = w.mesh
    ice = w.ice_inputs()
    res = []
    for guess in (0, 1):
        cfg = wl.scaling_config()
        cfg.ksp_initial_guess_nonzero = guess
        cfg.stress_balance_PETSc_rtol = 1e-11; cfg.stress_balance_PETSc_abstol = 1e-30
        res.append(_solve_gpu(ctx, cfg, 1, ice, m))
    (g0, nv0, na0), (g1, nv1, na1) = res
    assert nv0 == nv1 and na1 < na0                      # warm starts save Krylov iterations
    assert rel_l2(g1["u_vav_b"], g0["u_vav_b"]) < 1e-6
    st2, r = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    cfg.ksp_initial_guess_nonzero = 1
    rc, nv2, na2 = o.solve_velocity(cfg, 1, ice, st2)     # the oracle implements the same option
    assert rc == 0 and nv2 == nv1 and abs(na2 - na1) <= max(5, na2 // 20)
    assert rel_l2(g1["u_vav_b"], r["u_vav_b"]) < 1e-7


def test_ssa_matches_oracle(setup):
    w, ctx, o = setup
    m = w.mesh
    cfg = wl.scaling_config()
    cfg.choice_sliding_law = ub.SLID_SCHOOF2005
    cfg.stress_balance_PETSc_rtol = 1e-12; cfg.stress_balance_PETSc_abstol = 1e-30
    ice = w.ice_inputs()
    g, nv, na = _solve_gpu(ctx, cfg, 0, ice, m)
    st2, r = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    rc, nv2, na2 = o.solve_velocity(cfg, 0, ice, st2, ksp=ol.KSP_GMRES_ILU0)
    assert rc == 0 and nv == nv2
    for f in ("u_vav_b", "v_vav_b", "N_b", "dN_dx_b", "basal_friction_coefficient_b"):
        assert rel_l2(g[f], r[f]) < 1e-6, (f, rel_l2(g[f], r[f]))


def test_no_grounded_ice_returns_zero(setup):
    w, ctx, o = setup
    m = w.mesh
    ice = w.ice_inputs()
    zero = np.zeros(m.nV, np.int32)
    ice.mask_grounded_ice = ub.ip(zero)
    g, nv, na = _solve_gpu(ctx, wl.scaling_config(), 1, ice, m)
    assert nv == 0 and na == 0 and not g["u_vav_b"].any() and not g["u_3D_b"].any()


def test_resident_solve_equals_host_solve(setup):
    w, ctx, o = setup
    m = w.mesh
    cfg = wl.scaling_config()
    ice = w.ice_inputs()
    g, nv, na = _solve_gpu(ctx, cfg, 1, ice, m)
    st, arrs = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    L = ub.lib()
    ub.check(L.ufe_b200_velocity_upload(ctx.h, C.byref(ice), C.byref(st)))
    a, b = C.c_int(), C.c_int()
    ub.check(L.ufe_b200_velocity_solve_resident(ctx.h, C.byref(cfg), 1, C.byref(a), C.byref(b)))
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    assert (a.value, b.value) == (nv, na) and np.array_equal(arrs["u_vav_b"], g["u_vav_b"])   # deterministic reductions


@pytest.mark.parametrize("nz", [8, 10])
def test_diva_stages_with_other_layer_counts(nz):
    """the column kernels take the number of layers at run time: layer counts other than the configurations' 12"""
    w = wl.mismip_mod(3.0e4, nz=nz)
    m = w.mesh
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    o = ol.OracleMesh(m)
    cfg = wl.scaling_config()
    cfg.stress_balance_PETSc_rtol = 1e-13; cfg.stress_balance_PETSc_abstol = 1e-30
    cfg.visc_it_nit = 2; cfg.visc_it_norm_dUV_tol = 1e-30
    ice = w.ice_inputs()
    g, nv, na = _solve_gpu(ctx, cfg, 1, ice, m)
    ctx.close()
    st2, r = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    rc, nv2, na2 = o.solve_velocity(cfg, 1, ice, st2)
    assert rc == 0 and nv == nv2 == 3
    for f in _FIELDS:
        assert rel_l2(g[f], r[f]) < 1e-8, (f, rel_l2(g[f], r[f]))
