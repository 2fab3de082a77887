"""GPU: the device path on the REFERENCE'S OWN meshes against what the reference itself computed on them (fixtures: see test_reference_meshes.py).

The 15 operators are built on the device from the reference's mesh, applied through ufe_b200_apply_operator to the component test's three functions and
compared with the checksums of res_map_deriv_*.nc; the Laplace system of ct_discretisation_solve_Laplace_eq.f90 is assembled from the device's own
M2_d2dx2_b_b / M2_d2dy2_b_b rows and solved through ufe_b200_ksp_solve (the narrow seam replacing solve_matrix_equation_CSR_PETSc).
"""
import numpy as np
import pytest

import oracle_lib as ol
import ref_mesh as rm
import ufe_b200 as ub

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=rm.RES)
def on_device(request):
    res = request.param
    z, m = rm.load(res)
    c = ub.Context(device=0)
    c.mesh_create(m)
    c.calc_matrix_operators()
    yield res, z, m, c
    c.close()


def test_device_operators_equal_the_oracle_on_the_reference_mesh(on_device):
    res, z, m, c = on_device
    o = ol.OracleMesh(m)
    for name in ub.OPS:
        ptr, ind, val = c.get_operator(ub.OP[name])
        optr, oind, oval, mm, nn = o.get_operator(ub.OP[name])
        assert np.array_equal(ptr, optr) and np.array_equal(ind, oind), name
        assert np.abs(val - oval).max() < 1e-9 * np.abs(oval).max(), name


@pytest.mark.parametrize("kind", ("linear", "quadratic", "periodic"))
def test_device_operators_reproduce_the_reference_component_test(on_device, kind):
    res, z, m, c = on_device
    ref = rm.RESULTS[res][kind]
    fa = rm.test_function(kind, m.V[:, 0], m.V[:, 1], m.xmin, m.xmax, m.ymin, m.ymax)
    fb = rm.test_function(kind, m.TriGC[:, 0], m.TriGC[:, 1], m.xmin, m.xmax, m.ymin, m.ymax)
    for var, (op, grid) in rm.OP_OF_RESULT.items():
        x = np.ascontiguousarray(fa[0] if grid == "a" else fb[0])
        ptr, ind, val = c.get_operator(ub.OP[op])
        out = c.apply_operator(ub.OP[op], x, len(ptr) - 1)
        mag = rm.spmv_magnitude(ptr, ind, val, x)
        # the device builds the least-squares coefficients with its own 5x5 adjugate: 1e-9 of the coefficient scale (see the test above) is the bound here too
        ok, err = rm.checksum_close(rm.checksum(out), ref[var], 1e-9, scale_sum=mag.sum(), scale_ext=mag.max())
        assert ok, (var, err, rm.checksum(out), ref[var])


def test_device_laplace_solve_reproduces_the_reference_component_test(on_device):
    res, z, m, c = on_device
    pxx, ixx, vxx = c.get_operator(ub.OP["M2_d2dx2_b_b"])
    _, _, vyy = c.get_operator(ub.OP["M2_d2dy2_b_b"])
    ptr, ind, val, b, f_ex = rm.laplace_system(m, pxx, ixx, vxx, vyy)
    ref = rm.RESULTS[res]["Laplace"]
    _, ref_own_err = rm.checksum_close(ref["f_disc"], ref["f_ex"], 1.0)
    # the reference's tolerances (rtol = abstol = 1e-6): as close to its f_disc as that is to the exact solution
    f, its = c.ksp_solve(m.nTri, 1, ptr, ind, val, b, 1e-6, 1e-6)
    ok, err = rm.checksum_close(rm.checksum(f), ref["f_disc"], 1.5 * ref_own_err.max() + 1e-5)
    assert ok, (err, ref_own_err, its)
    # converged: the exact solution, whose statistics the reference stored as f_ex
    f, its = c.ksp_solve(m.nTri, 1, ptr, ind, val, b, 1e-12, 1e-30)
    ok, err = rm.checksum_close(rm.checksum(f), ref["f_ex"], 1e-8)
    assert ok, (err, its)
    assert np.abs(f - f_ex).max() <= 1e-8 * np.abs(f_ex).max()
