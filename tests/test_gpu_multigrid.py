"""GPU: aggregation-multigrid preconditioned BiCGStab (pc_type = UFE_PC_MG) on the SSA/DIVA stiffness matrix, through the C ABI.
The oracle solves the same linear systems with the reference's default PETSc configuration (GMRES(30) + ILU(0)); at tight
tolerance every Krylov method converges to A^-1 b, so the multigrid path is held to the same 1e-6 relative L2 as the Jacobi path."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl
from test_gpu_velocity import _linearised, _solve_gpu, rel_l2

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup():
    w = wl.mismip_mod(2.0e4)
    ctx = ub.Context(device=0)
    ctx.mesh_create(w.mesh)
    ctx.calc_matrix_operators()
    o = ol.OracleMesh(w.mesh)
    o.calc_matrix_operators()
    yield w, ctx, o
    ctx.close()


def _coefficients(m, seed=3):
    rng = np.random.default_rng(seed)
    N_b = 1e10 * (1 + 0.5 * rng.random(m.nTri)); dNx = 1e3 * rng.standard_normal(m.nTri); dNy = 1e3 * rng.standard_normal(m.nTri)
    beta = 1e3 * rng.random(m.nTri) * (rng.random(m.nTri) < 0.6)      # 40 % floating triangles (no basal friction)
    tdx = 1e4 * rng.standard_normal(m.nTri); tdy = 1e4 * rng.standard_normal(m.nTri)
    return N_b, dNx, dNy, beta, tdx, tdy


@pytest.mark.parametrize("variant", ["fp64", "fp32_storage", "bc_zero", "prescribed", "v22"])
def test_linearised_solve_with_multigrid(setup, variant):
    w, ctx, o = setup
    m = w.mesh
    co = _coefficients(m)
    cfg = wl.scaling_config()
    bc = None
    ctx.set_mg_params(nu_pre=1, nu_post=1, omega=0.8, overcorr=1.5, w_levels=3, fp32=0)
    if variant == "fp32_storage":
        ctx.set_mg_params(fp32=1)
    if variant == "v22":
        ctx.set_mg_params(nu_pre=2, nu_post=2, w_levels=0)
    if variant == "bc_zero":
        cfg.BC_u[0] = ub.BC_ZERO; cfg.BC_v[1] = ub.BC_ZERO; cfg.BC_u[3] = ub.BC_ZERO
    if variant == "prescribed":
        rng = np.random.default_rng(9)
        bm = (rng.random(m.nTri) < 0.05).astype(np.int32)
        bc = (bm, 10.0 * rng.standard_normal(m.nTri), 10.0 * rng.standard_normal(m.nTri))
    cfg.pc_type = ub.PC_JACOBI
    uj, vj, its_j = _linearised(ctx, cfg, m, *co, 1e-12, 1e-30, bc)
    cfg.pc_type = ub.PC_MG
    u, v, its = _linearised(ctx, cfg, m, *co, 1e-12, 1e-30, bc)
    info = ctx.mg_info()
    assert info["levels"] >= 3 and all(a > 2 * b for a, b in zip(info["nodes"], info["nodes"][1:])), info
    optr, oind, oval, obb = o.assemble(cfg, *co, *(bc or (None, None, None)))
    xo, its_o, rc = ol.ksp_solve(optr, oind, oval, obb, 1e-12, 1e-30, ksp=ol.KSP_GMRES_ILU0)
    assert rc == 0
    uv = np.empty(2 * m.nTri); uv[0::2] = u; uv[1::2] = v
    assert rel_l2(uv, xo) < 1e-6, (variant, its, its_o, rel_l2(uv, xo))          # north_star tolerance, against the reference's default KSP
    assert rel_l2(u, uj) < 1e-7 and rel_l2(v, vj) < 1e-7
    assert 0 < its < its_j, (its, its_j)       # friction-dominated random coefficients: Jacobi is already fast here; see the DIVA tests below
    ctx.set_mg_params(nu_pre=1, nu_post=1, omega=0.8, overcorr=1.5, w_levels=3, fp32=0)


def test_diva_solve_with_multigrid_matches_oracle(setup):
    w, ctx, o = setup
    m = w.mesh
    cfg = wl.scaling_config()
    cfg.pc_type = ub.PC_MG
    cfg.stress_balance_PETSc_rtol = 1e-12; cfg.stress_balance_PETSc_abstol = 1e-30
    ice = w.ice_inputs()
    g, nv, na = _solve_gpu(ctx, cfg, 1, ice, m)
    st2, r = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    cfg_o = wl.scaling_config()
    cfg_o.stress_balance_PETSc_rtol = 1e-12; cfg_o.stress_balance_PETSc_abstol = 1e-30
    rc, nv2, na2 = o.solve_velocity(cfg_o, 1, ice, st2, ksp=ol.KSP_GMRES_ILU0)
    assert rc == 0 and nv == nv2
    for f in ("u_vav_b", "v_vav_b", "u_3D_b", "tau_bx_b", "eta_3D_b"):
        assert rel_l2(g[f], r[f]) < 1e-6, (f, rel_l2(g[f], r[f]))


def test_multigrid_with_initial_guess_and_config_tolerance(setup):
    """the bench configuration: rtol 1e-4, warm start; the nonlinear solve lands on the same velocity field"""
    w, ctx, o = setup
    m = w.mesh
    ice = w.ice_inputs()
    res = []
    for pc in (ub.PC_JACOBI, ub.PC_MG):
        cfg = wl.scaling_config()
        cfg.pc_type = pc
        cfg.ksp_initial_guess_nonzero = 1
        res.append(_solve_gpu(ctx, cfg, 1, ice, m))
    (gj, nvj, naj), (gm, nvm, nam) = res
    assert abs(nvj - nvm) <= 1 and nam < naj / 3
    assert rel_l2(gm["u_vav_b"], gj["u_vav_b"]) < 5e-3       # two preconditioners at rtol 1e-4


def test_merged_reductions_of_the_multi_rank_recurrence(setup, monkeypatch):
    """k_ksmg_dot5 / k_ksmg_update_xr5 (two all-reduces per iteration, the default on several ranks), forced on one rank: same solution as the
    direct sums at the configuration's tolerance and at a tight one, with about the same iteration count"""
    w, ctx, o = setup
    m = w.mesh
    co = _coefficients(m, seed=6)
    cfg = wl.scaling_config()
    cfg.pc_type = ub.PC_MG
    for rtol, tol in ((1e-4, 2e-4), (1e-10, 1e-8)):
        monkeypatch.setenv("UFE_B200_KSP_MERGED", "0")
        u0, v0, its0 = _linearised(ctx, cfg, m, *co, rtol, 1e-30)
        monkeypatch.setenv("UFE_B200_KSP_MERGED", "1")
        u1, v1, its1 = _linearised(ctx, cfg, m, *co, rtol, 1e-30)
        assert rel_l2(u1, u0) < tol and rel_l2(v1, v0) < tol, (rtol, rel_l2(u1, u0), rel_l2(v1, v0))
        assert abs(its1 - its0) <= max(2, its0 // 5), (rtol, its0, its1)


@pytest.mark.parametrize("dsmem", [0, 1])
@pytest.mark.parametrize("cluster", [8, 16])
def test_coarsest_level_sweeps_as_one_cluster_kernel(setup, cluster, dsmem):
    """k_mg_coarsest_cluster (matrix slices in shared memory, iterate exchanged through L2 + cluster barrier) against the same sweeps as graph nodes"""
    w, ctx, o = setup
    m = w.mesh
    co = _coefficients(m, seed=7)
    cfg = wl.scaling_config()
    cfg.pc_type = ub.PC_MG
    try:
        ctx.set_mg_option("stop_nodes", 4000); ctx.set_mg_option("coarse_sweeps", 12)
        ctx.set_mg_option("coarsest_cluster", 0)
        u0, v0, its0 = _linearised(ctx, cfg, m, *co, 1e-10, 1e-30)
        info = ctx.mg_info()
        assert 128 < info["nodes"][-1] <= 4000, info        # the hierarchy ends in a level that is swept, not solved directly
        ctx.set_mg_option("coarsest_cluster", cluster); ctx.set_mg_option("coarsest_dsmem", dsmem)
        u1, v1, its1 = _linearised(ctx, cfg, m, *co, 1e-10, 1e-30)
        assert abs(its1 - its0) <= 1, (its0, its1)
        assert rel_l2(u1, u0) < 1e-8 and rel_l2(v1, v0) < 1e-8, (rel_l2(u1, u0), rel_l2(v1, v0))
    finally:
        ctx.set_mg_option("coarsest_cluster", 0); ctx.set_mg_option("coarsest_dsmem", 0); ctx.set_mg_option("stop_nodes", 0); ctx.set_mg_option("coarse_sweeps", 8)


def test_chebyshev_steps_on_the_coarsest_level(setup):
    """coarsest_cheb: the corrections of k_mg_coarsest_cluster as Chebyshev steps for [lmax / ratio, lmax] of Dw A (lmax from k_mg_lmax_bound); a fixed
    polynomial preconditioner, so the solve must land on the same solution, with no more iterations than twice as many stationary sweeps need"""
    w, ctx, o = setup
    m = w.mesh
    co = _coefficients(m, seed=8)
    cfg = wl.scaling_config()
    cfg.pc_type = ub.PC_MG
    try:
        ctx.set_mg_option("stop_nodes", 4000); ctx.set_mg_option("coarsest_cluster", 16)
        ctx.set_mg_option("coarse_sweeps", 24); ctx.set_mg_option("coarsest_cheb", 0)
        u0, v0, its0 = _linearised(ctx, cfg, m, *co, 1e-10, 1e-30)
        ctx.set_mg_option("coarse_sweeps", 12); ctx.set_mg_option("coarsest_cheb", 1)
        u1, v1, its1 = _linearised(ctx, cfg, m, *co, 1e-10, 1e-30)
        ub2 = _linearised(ctx, cfg, m, *co, 1e-10, 1e-30)
        assert rel_l2(u1, u0) < 1e-8 and rel_l2(v1, v0) < 1e-8, (rel_l2(u1, u0), rel_l2(v1, v0))
        assert 0 < its1 <= its0 + max(3, its0 // 4), (its0, its1)
        assert ub2[2] == its1 and np.array_equal(ub2[0], u1)          # deterministic
    finally:
        for k, v in (("coarsest_cheb", 0), ("coarsest_cluster", 0), ("stop_nodes", 0), ("coarse_sweeps", 8)):
            ctx.set_mg_option(k, v)


def test_multigrid_is_deterministic(setup):
    w, ctx, o = setup
    m = w.mesh
    co = _coefficients(m, seed=4)
    cfg = wl.scaling_config()
    cfg.pc_type = ub.PC_MG
    a = _linearised(ctx, cfg, m, *co, 1e-10, 1e-30)
    b = _linearised(ctx, cfg, m, *co, 1e-10, 1e-30)
    assert a[2] == b[2] and np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


@pytest.mark.parametrize("pc", [ub.PC_JACOBI, ub.PC_MG])
def test_config2_mismip_plus_ssa_on_graded_mesh(pc):
    """BASELINE.json configs[1]: MISMIP+ bed, SSA, Schoof2005 sliding, resolution graded towards the grounding line"""
    w = wl.mismip_plus(4.0e3)
    m = w.mesh
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    cfg = w.cfg
    cfg.pc_type = pc
    cfg.stress_balance_PETSc_rtol = 1e-12; cfg.stress_balance_PETSc_abstol = 1e-30
    ice = w.ice_inputs()
    g, nv, na = _solve_gpu(ctx, cfg, 0, ice, m)
    ctx.close()
    o = ol.OracleMesh(m)
    st2, r = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    rc, nv2, na2 = o.solve_velocity(cfg, 0, ice, st2, ksp=ol.KSP_GMRES_ILU0)
    assert rc == 0 and nv == nv2
    for f in ("u_vav_b", "v_vav_b", "N_b", "basal_friction_coefficient_b"):
        assert rel_l2(g[f], r[f]) < 1e-6, (f, rel_l2(g[f], r[f]))
    assert np.abs(g["u_vav_b"]).max() > 10.0        # a flowing ice stream, not the trivial solution


def test_single_cta_tail_equals_separate_launches():
    """the small levels run as one single-CTA kernel per visit (k_mg_tail); same arithmetic as the stand-alone kernels, so the same bits"""
    import os
    w = wl.mismip_mod(2.0e4)
    m = w.mesh
    co = _coefficients(m, seed=6)
    cfg = wl.scaling_config()
    cfg.pc_type = ub.PC_MG
    res = []
    try:
        for tail in ("0", "4096"):
            os.environ["UFE_B200_MG_TAIL"] = tail
            ctx = ub.Context(device=0)
            ctx.mesh_create(m); ctx.calc_matrix_operators()
            res.append(_linearised(ctx, cfg, m, *co, 1e-10, 1e-30))
            assert ctx.mg_info()["levels"] >= 4
            ctx.close()
    finally:
        os.environ.pop("UFE_B200_MG_TAIL", None)
    assert res[0][2] == res[1][2] and np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])


def test_assembly_writes_the_blocked_matrix_directly():
    """the assembly kernels fill the blocked sliced-ELL copy in the same pass; UFE_B200_BLK_SCATTER=1 forces the separate scatter pass -- same bits"""
    import os
    w = wl.mismip_mod(2.0e4)
    m = w.mesh
    co = _coefficients(m, seed=8)
    cfg = wl.scaling_config()
    res = []
    try:
        for scatter in (True, False):
            if scatter:
                os.environ["UFE_B200_BLK_SCATTER"] = "1"
            else:
                os.environ.pop("UFE_B200_BLK_SCATTER", None)
            ctx = ub.Context(device=0)
            ctx.mesh_create(m); ctx.calc_matrix_operators()
            for pc in (ub.PC_JACOBI, ub.PC_MG):
                cfg.pc_type = pc
                res.append(_linearised(ctx, cfg, m, *co, 1e-10, 1e-30))
                res.append(_linearised(ctx, cfg, m, *[1.1 * a for a in co], 1e-10, 1e-30))      # second assembly on the existing pattern
                ptr, ind, val, bb = ctx.get_stiffness()       # with multigrid the CSR values are re-formed from the blocked copy on demand
                res.append((val, bb, len(ptr)))
            ctx.close()
    finally:
        os.environ.pop("UFE_B200_BLK_SCATTER", None)
    n = len(res) // 2
    for a, b in zip(res[:n], res[n:]):
        assert a[2] == b[2] and np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
