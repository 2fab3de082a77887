"""CPU: the oracle and the host mesh library against the reference's own known-answer vectors (tests/golden)."""
import json
import os

import numpy as np

import oracle_lib as ol
import ufe_b200 as ub
import ufe_host as uh

G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))


def _sys(name):
    g = G[name]
    return (g["m"], g["n"], np.array(g["ptr"], np.int32), np.array(g["ind"], np.int32), np.array(g["val"], float),
            np.array(g["x"], float), np.array(g["y"], float))


def test_spmv_known_answers():
    # ut_petsc.f90:83-149 and ut_mpi_CSR_matrix_vector_multiplication.f90:186-318 (tolerance of the reference: 1e-10)
    for name in ("csr6", "csr7_1", "csr7_2"):
        m, n, ptr, ind, val, x, y = _sys(name)
        assert np.abs(ol.spmv(m, n, ptr, ind, val, x) - y).max() < 1e-10, name


def test_ksp_recovers_known_x():
    # the 7x7 systems are non-singular: the solver must recover x from y = A x.  Row 3 has no diagonal entry, so
    # ILU(0) has a zero pivot there (PETSc would stop too); only the Jacobi-preconditioned method applies.
    for name in ("csr7_1", "csr7_2"):
        m, n, ptr, ind, val, x, y = _sys(name)
        for ksp in (ol.KSP_BICGSTAB_JACOBI,):
            xs, its, rc = ol.ksp_solve(ptr, ind, val, y, 1e-13, 1e-30, ksp=ksp)
            assert rc == 0 and np.abs(xs - x).max() < 1e-8, (name, ksp, xs)


def test_partition_list_table():
    # mesh_parallelisation.f90:3-36
    g = G["partition_100_4"]
    for i in range(4):
        assert uh.partition_list(100, i, 4) == (g["vi1"][i], g["vi2"][i])
    assert uh.partition_list(7, 0, 4) == (1, 7) and uh.partition_list(7, 2, 4) == (1, 0)  # ntot <= 2n: rank 0 owns all
    for n in (1, 2, 3, 5, 8):
        tot = 0
        for i in range(n):
            a, b = uh.partition_list(1003, i, n)
            tot += b - a + 1
        assert tot == 1003


def test_dummy_mesh_connectivity():
    # mesh_dummy_meshes.f90:23-120
    m = uh.HostMesh.dummy5()
    g = G["dummy5"]
    assert (m.C[:, :4] == np.array(g["C"])).all() and (m.iTri[:, :4] == np.array(g["iTri"])).all()
    assert (m.TriC == np.array(g["TriC"])).all() and list(m.nC) == g["nC"] and list(m.niTri) == g["niTri"] and list(m.VBI) == g["VBI"]
    assert m.nE == 8 and abs(m.A.sum() - 4.0) < 1e-12


def test_synthetic_mesh_invariants():
    m = uh.HostMesh.synthetic(-1e6, 1e6, -1e6, 1e6, 5e4)
    assert m.nV - m.nE + m.nTri == 1                                  # Euler characteristic of a disc
    assert abs(m.A.sum() / 4e12 - 1) < 1e-12 and abs(m.TriA.sum() / 4e12 - 1) < 1e-12
    assert (np.diff(m.V[:, 0]) >= 0).all() and (np.diff(m.TriGC[:, 0]) >= 0).all()   # x-sorted (mesh_parallelisation.f90:14-22)
    assert ((m.VBI > 0).sum() == (m.nC != m.niTri).sum())
    # VE / C symmetry and ETri consistency
    for vi in range(0, m.nV, 97):
        for ci in range(m.nC[vi]):
            vj, ei = m.C[vi, ci] - 1, m.VE[vi, ci] - 1
            assert set(m.EV[ei, :2]) == {vi + 1, vj + 1}
    # halos are contiguous and adjacent to the ownership boundary
    h = m.halos(1, 4)
    assert h["a"]["i2_hle"] == h["a"]["i1"] - 1 and h["a"]["i1_hre"] == h["a"]["i2"] + 1


def test_operators_polynomial_exactness():
    # what ct_discretisation_mapping_derivatives checks: 1st-order operators exact on linear, M2 exact on quadratic
    m = uh.HostMesh.synthetic(-1e6, 1e6, -1e6, 1e6, 4e4)
    o = ol.OracleMesh(m)
    x, y, gx, gy = m.V[:, 0], m.V[:, 1], m.TriGC[:, 0], m.TriGC[:, 1]
    fa, fb = 3.0 + 2e-6 * x - 1e-6 * y, 3.0 + 2e-6 * gx - 1e-6 * gy
    for name in ub.OPS:
        ptr, ind, val, mm, nn = o.get_operator(ub.OP[name])
        out = o.apply_operator(ub.OP[name], np.ascontiguousarray(fa if nn == m.nV else fb))
        if name.startswith("map"):
            exact, tol = (fb if mm == m.nTri else fa), 1e-12
        elif "d2" in name:
            exact, tol = 0.0, 1e-18
        else:
            exact, tol = (2e-6 if "ddx" in name else -1e-6), 1e-15
        assert np.abs(out - exact).max() < tol, name
    qb = np.ascontiguousarray(1e-12 * (gx ** 2 + 0.5 * gx * gy - 2 * gy ** 2))
    for name, ex in (("M2_d2dx2_b_b", 2e-12), ("M2_d2dxdy_b_b", 0.5e-12), ("M2_d2dy2_b_b", -4e-12)):
        assert np.abs(o.apply_operator(ub.OP[name], qb) - ex).max() < 1e-9 * abs(ex), name
    # pattern rules: a_b rows have exactly the 3 triangle vertices in Tri order; M2 rows start with the diagonal
    ptr, ind, val, mm, nn = o.get_operator(ub.OP["map_a_b"])
    assert (np.diff(ptr) == 3).all() and (ind.reshape(-1, 3) == m.Tri).all()
    ptr, ind, val, mm, nn = o.get_operator(ub.OP["M2_ddx_b_b"])
    assert (ind[ptr[:-1] - 1] == np.arange(1, m.nTri + 1)).all() and (np.diff(ptr) - 1 >= 5).all()


def test_laplace_equation_like_reference_component_test():
    # ct_discretisation_solve_Laplace_eq.f90:103-165: d2f/dx2 + d2f/dy2 = c with f = c/4 (r^2 - r0^2) prescribed on the border
    m = uh.HostMesh.synthetic(-1e6, 1e6, -1e6, 1e6, 5e4)
    o = ol.OracleMesh(m)
    pxx, ixx, vxx, _, _ = o.get_operator(ub.OP["M2_d2dx2_b_b"])
    _, _, vyy, _, _ = o.get_operator(ub.OP["M2_d2dy2_b_b"])
    cc, r0 = 1e-9, 1e6
    gx, gy = m.TriGC[:, 0], m.TriGC[:, 1]
    f_ex = cc / 4 * (gx ** 2 + gy ** 2) - cc / 4 * r0 ** 2
    val = vxx + vyy
    b = np.full(m.nTri, cc)
    for ti in np.nonzero(m.TriBI > 0)[0]:
        k1, k2 = pxx[ti] - 1, pxx[ti + 1] - 1
        val[k1:k2] = np.where(ixx[k1:k2] == ti + 1, 1.0, 0.0)
        b[ti] = f_ex[ti]
    for ksp in (ol.KSP_BICGSTAB_JACOBI, ol.KSP_GMRES_ILU0):
        f, its, rc = ol.ksp_solve(pxx, ixx, val, b, 1e-10, 1e-12, ksp=ksp)
        assert rc == 0
        assert np.abs(f - f_ex).max() / np.abs(f_ex).max() < 1e-6, (ksp, its)


def test_analytic_solutions():
    L = ol.lib()
    # Halfar: H(0,0,0) = H0, H = 0 at r >= R0 (t = 0), volume is conserved in time (similarity solution)
    assert abs(L.orc_Halfar_H(1e-16, 3.0, 3000.0, 500e3, 0.0, 0.0, 0.0) - 3000.0) < 1e-9
    assert L.orc_Halfar_H(1e-16, 3.0, 3000.0, 500e3, 500e3, 0.0, 0.0) == 0.0
    r = np.linspace(0, 900e3, 20001)
    vol = [np.trapezoid([2 * np.pi * ri * L.orc_Halfar_H(1e-16, 3.0, 3000.0, 500e3, ri, 0.0, t) for ri in r], r) for t in (0.0, 500.0)]
    assert abs(vol[1] / vol[0] - 1) < 1e-5
    # Schoof 2006: u is even in y, maximal at y = 0, zero outside |y| > W = L (m+1)^(1/m)
    import ctypes as C
    u0, uy, um, t = C.c_double(), C.c_double(), C.c_double(), C.c_double()
    cast = lambda v: C.cast(C.byref(v), ub.c_dp)
    L.orc_Schoof2006_icestream(1e-18, 3.0, 2000.0, -0.0003, 150e3, 1.0, 0.0, cast(u0), cast(t))
    L.orc_Schoof2006_icestream(1e-18, 3.0, 2000.0, -0.0003, 150e3, 1.0, 5e4, cast(uy), cast(t))
    L.orc_Schoof2006_icestream(1e-18, 3.0, 2000.0, -0.0003, 150e3, 1.0, -5e4, cast(um), cast(t))
    assert u0.value > uy.value > 0 and abs(uy.value - um.value) < 1e-9 * abs(uy.value)
    L.orc_Schoof2006_icestream(1e-18, 3.0, 2000.0, -0.0003, 150e3, 1.0, 3.1e5, cast(uy), cast(t))
    assert uy.value == 0.0


def test_block_jacobi_ilu0_is_the_multi_rank_default():
    # PETSc's default PC on n ranks is block Jacobi with one ILU(0) per rank's row block: one block = plain ILU(0) (same iterates), more blocks = a weaker
    # preconditioner that converges to the same solution
    m = uh.HostMesh.synthetic(-1e6, 1e6, -1e6, 1e6, 5e4)
    o = ol.OracleMesh(m)
    pxx, ixx, vxx, _, _ = o.get_operator(ub.OP["M2_d2dx2_b_b"])
    _, _, vyy, _, _ = o.get_operator(ub.OP["M2_d2dy2_b_b"])
    val = vxx + vyy
    b = np.full(m.nTri, 1e-9)
    for ti in np.nonzero(m.TriBI > 0)[0]:
        k1, k2 = pxx[ti] - 1, pxx[ti + 1] - 1
        val[k1:k2] = np.where(ixx[k1:k2] == ti + 1, 1.0, 0.0)
        b[ti] = 0.0
    f1, its1, rc = ol.ksp_solve(pxx, ixx, val, b, 1e-10, 1e-30, ksp=ol.KSP_GMRES_ILU0)
    assert rc == 0
    try:
        ol.set_bjacobi_blocks(1)
        f2, its2, rc = ol.ksp_solve(pxx, ixx, val, b, 1e-10, 1e-30, ksp=ol.KSP_GMRES_BJACOBI_ILU0)
        assert rc == 0 and its2 == its1 and np.abs(f1 - f2).max() < 1e-12 * np.abs(f1).max()   # (OpenMP reductions: not bitwise)
        ol.set_bjacobi_blocks(4)
        f4, its4, rc = ol.ksp_solve(pxx, ixx, val, b, 1e-10, 1e-30, ksp=ol.KSP_GMRES_BJACOBI_ILU0)
        assert rc == 0 and its4 > its1 and np.abs(f4 - f1).max() < 1e-7 * np.abs(f1).max(), (its1, its4)
    finally:
        ol.set_bjacobi_blocks(0)
