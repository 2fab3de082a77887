"""CPU: the host mesh library and the oracle on the REFERENCE'S OWN meshes against what the reference itself computed on them.

Fixtures: tests/golden/ref_mesh_{2,3,4}.0000E+05.npz (primary + secondary mesh data the reference wrote to
automated_testing/test_meshes_and_grids/mesh_Ant_uniform_*_m_nit_Lloyd_2.nc) and tests/golden/ref_mesh_results.json (the checksums
[sum, sum|.|, min, max] of every field its component tests ct_discretisation_mapping_derivatives / ct_discretisation_solve_Laplace_eq
wrote for these meshes).  Element-wise for the mesh arrays; for the operator results the four statistics of each field, to 1e-11 of |A| |x|.
"""
import numpy as np
import pytest

import oracle_lib as ol
import ref_mesh as rm
import ufe_b200 as ub


@pytest.mark.parametrize("res", rm.RES)
def test_secondary_mesh_data_equal_the_reference(res):
    # calc_all_secondary_mesh_data (mesh_secondary.f90:31-70): index arrays bit-exact, reals to rounding
    z, m = rm.load(res)
    assert (m.nV, m.nTri, m.nE) == (rm.RESULTS[res]["nV"], rm.RESULTS[res]["nTri"], rm.RESULTS[res]["nE"])
    for name in ("TriBI", "VE", "EV", "ETri", "TriE", "EBI"):
        assert np.array_equal(getattr(m, name), z[name]), name
    for name, tol in (("E", 1e-15), ("TriGC", 1e-15), ("TriA", 1e-13), ("A", 1e-12)):
        mine, ref = getattr(m, name), z[name]
        assert np.abs(mine - ref).max() <= tol * np.abs(ref).max(), (name, np.abs(mine - ref).max() / np.abs(ref).max())


@pytest.mark.parametrize("res", rm.RES)
@pytest.mark.parametrize("kind", ("linear", "quadratic", "periodic"))
def test_operators_reproduce_the_reference_component_test(res, kind):
    # calc_all_matrix_operators_mesh + the 15 SpMVs of run_all_map_deriv_tests_on_mesh_with_function_calc_disc
    z, m = rm.load(res)
    o = ol.OracleMesh(m)
    ref = rm.RESULTS[res][kind]
    fa = rm.test_function(kind, m.V[:, 0], m.V[:, 1], m.xmin, m.xmax, m.ymin, m.ymax)
    fb = rm.test_function(kind, m.TriGC[:, 0], m.TriGC[:, 1], m.xmin, m.xmax, m.ymin, m.ymax)
    for k, name in enumerate(("d_a_ex", "ddx_a_ex", "ddy_a_ex")):
        ok, err = rm.checksum_close(rm.checksum(fa[k]), ref[name], 1e-13)
        assert ok, (name, err)
    for k, name in enumerate(("d_b_ex", "ddx_b_ex", "ddy_b_ex", "d2dx2_b_ex", "d2dxdy_b_ex", "d2dy2_b_ex")):
        ok, err = rm.checksum_close(rm.checksum(fb[k]), ref[name], 1e-13)
        assert ok, (name, err)
    worst = 0.0
    for var, (op, grid) in rm.OP_OF_RESULT.items():
        x = np.ascontiguousarray(fa[0] if grid == "a" else fb[0])
        out = o.apply_operator(ub.OP[op], x)
        # rounding is relative to |A| |x| (the quadratic function reaches 7.5e6 through its x y term while its second derivatives are 1e-13;
        # the second derivatives of the linear one are pure rounding noise, 1e-24, in the reference's file as well)
        mag = rm.spmv_magnitude(*o.get_operator(ub.OP[op])[:3], x)
        ok, err = rm.checksum_close(rm.checksum(out), ref[var], 1e-11, scale_sum=mag.sum(), scale_ext=mag.max())
        assert ok, (var, err, rm.checksum(out), ref[var])
        worst = max(worst, err.max())
    print(f"{res} {kind}: worst statistic deviation {worst:.2e} of |A||x|")


@pytest.mark.parametrize("res", rm.RES)
def test_laplace_solve_reproduces_the_reference_component_test(res):
    # ct_discretisation_solve_Laplace_eq.f90:103-165 through the KSP seam (rtol = abstol = 1e-6 in the reference: its f_disc is converged to ~1e-6)
    z, m = rm.load(res)
    o = ol.OracleMesh(m)
    pxx, ixx, vxx, _, _ = o.get_operator(ub.OP["M2_d2dx2_b_b"])
    _, _, vyy, _, _ = o.get_operator(ub.OP["M2_d2dy2_b_b"])
    ptr, ind, val, b, f_ex = rm.laplace_system(m, pxx, ixx, vxx, vyy)
    ref = rm.RESULTS[res]["Laplace"]
    ok, err = rm.checksum_close(rm.checksum(f_ex), ref["f_ex"], 1e-13)
    assert ok, err
    # The operator is exact on this quadratic f, so a converged solve returns f_ex itself; the reference's f_disc is its GMRES(30) + ILU(0) iterate
    # at rtol = abstol = 1e-6 and sits 2e-6 (4.0E+05 mesh) to 9e-5 (2.0E+05 mesh) of max|f| away from its own f_ex.  Pin: (i) with the reference's solver
    # settings the oracle lands as close to the reference's f_disc as the reference is to the exact solution; (ii) converged solves of both Krylov methods
    # reproduce the reference's f_ex statistics to 1e-10.
    _, ref_own_err = rm.checksum_close(ref["f_disc"], ref["f_ex"], 1.0)
    f, its, rc = ol.ksp_solve(ptr, ind, val, b, 1e-6, 1e-6, ksp=ol.KSP_GMRES_ILU0)
    assert rc == 0
    ok, err = rm.checksum_close(rm.checksum(f), ref["f_disc"], 1.5 * ref_own_err.max() + 1e-9)
    assert ok, (err, ref_own_err, its)
    for ksp in (ol.KSP_GMRES_ILU0, ol.KSP_BICGSTAB_JACOBI):
        f, its, rc = ol.ksp_solve(ptr, ind, val, b, 1e-12, 1e-30, ksp=ksp)
        assert rc == 0
        ok, err = rm.checksum_close(rm.checksum(f), ref["f_ex"], 1e-10)
        assert ok, (ksp, err, its)
        assert np.abs(f - f_ex).max() <= 1e-10 * np.abs(f_ex).max()


def test_hdf5_reader_refuses_what_it_does_not_understand(tmp_path):
    """tests/golden/h5lite.py reads exactly the HDF5 features the reference's files use; anything else must raise instead of returning wrong data"""
    import sys
    sys.path.insert(0, rm.GOLD)
    import h5lite
    bad = tmp_path / "not_hdf5.nc"
    bad.write_bytes(b"CDF\x01" + bytes(60))                           # a classic (non-HDF5) NetCDF header
    with pytest.raises(h5lite.H5Unsupported, match="not an HDF5 file"):
        h5lite.File(str(bad))
    v0 = tmp_path / "superblock_v0.h5"
    v0.write_bytes(b"\x89HDF\r\n\x1a\n" + bytes([0]) + bytes(87))     # HDF5 signature, superblock version 0
    with pytest.raises(h5lite.H5Unsupported, match="superblock version 0"):
        h5lite.File(str(v0))
