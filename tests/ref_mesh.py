"""The reference's own test meshes and component-test results (tests/golden/ref_mesh_*.npz, ref_mesh_results.json; written by
tests/golden/make_mesh_fixtures.py from the NetCDF files under /root/reference/automated_testing) as objects the tests share."""
import json
import os

import numpy as np

import ufe_host as uh

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RES = ("4.0000E+05", "3.0000E+05", "2.0000E+05")
RESULTS = json.load(open(os.path.join(GOLD, "ref_mesh_results.json")))["meshes"]
OP_OF_RESULT = {   # variable of res_map_deriv_*.nc -> (operator, argument grid); ct_discretisation_mapping_derivatives.f90:253-330
    "d_a_b": ("map_a_b", "a"), "d_b_a": ("map_b_a", "b"), "ddx_a_a": ("ddx_a_a", "a"), "ddx_a_b": ("ddx_a_b", "a"), "ddx_b_a": ("ddx_b_a", "b"),
    "ddx_b_b": ("ddx_b_b", "b"), "ddy_a_a": ("ddy_a_a", "a"), "ddy_a_b": ("ddy_a_b", "a"), "ddy_b_a": ("ddy_b_a", "b"), "ddy_b_b": ("ddy_b_b", "b"),
    "ddx_b_b_2nd": ("M2_ddx_b_b", "b"), "ddy_b_b_2nd": ("M2_ddy_b_b", "b"), "d2dx2_b_b_2nd": ("M2_d2dx2_b_b", "b"),
    "d2dxdy_b_b_2nd": ("M2_d2dxdy_b_b", "b"), "d2dy2_b_b_2nd": ("M2_d2dy2_b_b", "b")}


def load(res):
    """(reference arrays as a dict, HostMesh built from the reference's PRIMARY data by our own calc_secondary)"""
    z = dict(np.load(os.path.join(GOLD, f"ref_mesh_{res}.npz")))
    m = uh.HostMesh.from_primary(z["V"], z["nC"], z["C"], z["niTri"], z["iTri"], z["VBI"], z["Tri"], z["Tricc"], z["TriC"],
                                 float(z["xmin"]), float(z["xmax"]), float(z["ymin"]), float(z["ymax"]))
    return z, m


def test_function(kind, x, y, xmin, xmax, ymin, ymax):
    """test_function_{linear,quadratic,periodic}, ct_discretisation_mapping_derivatives.f90:496-560: f and its five derivatives"""
    if kind == "linear":
        c1, c2 = 2.0 / (xmax - xmin), 3.0 / (ymax - ymin)
        z = np.zeros_like(x)
        return c1 * x + c2 * y, c1 + z, c2 + z, z, z, z
    if kind == "quadratic":
        c1, c2, c3 = 2.0 / (xmax - xmin), 3.0 / (ymax - ymin), 5.0 / (ymax - ymin)
        z = np.zeros_like(x)
        return (c1 * x) ** 2 + (c2 * y) ** 2 + c3 * x * y, 2.0 * c1 ** 2 * x + c3 * y, 2.0 * c2 ** 2 * y + c3 * x, 2.0 * c1 ** 2 + z, c3 + z, 2.0 * c2 ** 2 + z
    c1, c2 = 2.0 * np.pi / (xmax - xmin), 3.0 * np.pi / (ymax - ymin)
    sx, cx, sy, cy = np.sin(c1 * (x - xmin)), np.cos(c1 * (x - xmin)), np.sin(c2 * (y - ymin)), np.cos(c2 * (y - ymin))
    return sx * sy, c1 * cx * sy, sx * c2 * cy, -(c1 ** 2) * sx * sy, c1 * cx * c2 * cy, sx * -(c2 ** 2) * sy


def checksum(a):
    """[sum, sum|.|, min, max], automated_testing/reduce_all_netcdfs_in_folder_to_checksum.py"""
    a = np.asarray(a, dtype=np.float64).ravel()
    return np.array([a.sum(), np.abs(a).sum(), a.min(), a.max()])


def checksum_close(mine, ref, rtol, scale_sum=None, scale_ext=None):
    """all four statistics agree relative to the field's magnitude: the sums relative to sum|.|, min and max relative to max|.|
    (or to the scales given: for an SpMV the magnitude that rounding is relative to is |A| |x|, not |A x|)"""
    mine, ref = np.asarray(mine), np.asarray(ref)
    scale_sum = max(abs(ref[1]), 1e-300) if scale_sum is None else scale_sum
    scale_ext = max(abs(ref[2]), abs(ref[3]), 1e-300) if scale_ext is None else scale_ext
    err = np.array([abs(mine[0] - ref[0]) / scale_sum, abs(mine[1] - ref[1]) / scale_sum, abs(mine[2] - ref[2]) / scale_ext, abs(mine[3] - ref[3]) / scale_ext])
    return bool((err <= rtol).all()), err


def spmv_magnitude(ptr, ind, val, x):
    """|A| |x| per row (1-based CSR): what the rounding error of A x is relative to"""
    row = np.repeat(np.arange(len(ptr) - 1), np.diff(ptr))
    return np.bincount(row, weights=np.abs(val) * np.abs(x[ind - 1]), minlength=len(ptr) - 1)


def laplace_system(m, pxx, ixx, vxx, vyy):
    """The system of ct_discretisation_solve_Laplace_eq.f90:103-155 from the rows of M2_d2dx2_b_b and M2_d2dy2_b_b (1-based CSR):
    identity rows with f_ex beyond r0 = 0.8 xmax, d2f/dx2 + d2f/dy2 = c inside.  The pattern of an identity row is one entry, like add_entry_CSR_dist
    produces it, so ptr/ind are rebuilt."""
    c, r0 = -1e-9, m.xmax * 0.8
    gx, gy = m.TriGC[:, 0], m.TriGC[:, 1]
    f_ex = -c / 4.0 * r0 ** 2 + c / 4.0 * (gx ** 2 + gy ** 2)
    ptr, ind, val, b = [1], [], [], np.zeros(m.nTri)
    for ti in range(m.nTri):
        if np.hypot(gx[ti], gy[ti]) >= r0:
            ind.append(ti + 1); val.append(1.0); b[ti] = f_ex[ti]
        else:
            k1, k2 = pxx[ti] - 1, pxx[ti + 1] - 1
            ind += list(ixx[k1:k2]); val += list(vxx[k1:k2] + vyy[k1:k2]); b[ti] = c
        ptr.append(len(ind) + 1)
    return np.array(ptr, np.int32), np.array(ind, np.int32), np.array(val), b, f_ex
