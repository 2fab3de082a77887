"""Extract the reference's own test meshes and the results its component tests stored for them (run in the container that has /root/reference):

    python tests/golden/make_mesh_fixtures.py

Reads, with tests/golden/h5lite.py (the files are NetCDF-4 = HDF5, superblock v2, no compression):
  /root/reference/automated_testing/test_meshes_and_grids/mesh_Ant_uniform_{2,3,4}.0000E+05_m_nit_Lloyd_2.nc
      the primary mesh (V, nC, C, niTri, iTri, VBI, Tri, Tricc, TriC) and every secondary array the reference computed and wrote
      (TriGC, TriA, TriBI, A, R, E, VE, EV, ETri, TriE, EBI, EA, the Voronoi mesh)
  /root/reference/automated_testing/UPSY/component_test_mesh_discretisation/reference/
      res_map_deriv_<mesh>_{linear,quadratic,periodic}_checksum.nc   (ct_discretisation_mapping_derivatives.f90)
      res_Laplace_<mesh>_checksum.nc                                   (ct_discretisation_solve_Laplace_eq.f90:103-165)
      per variable [sum, sum|.|, min, max] over the finite entries (automated_testing/reduce_all_netcdfs_in_folder_to_checksum.py)
Writes tests/golden/ref_mesh_<res>.npz (compressed; arrays transposed to the Fortran shapes (n, cols)) and tests/golden/ref_mesh_results.json.
"""
import json
import os

import numpy as np

import h5lite

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/automated_testing"
MESHES = ["mesh_Ant_uniform_2.0000E+05_m_nit_Lloyd_2", "mesh_Ant_uniform_3.0000E+05_m_nit_Lloyd_2", "mesh_Ant_uniform_4.0000E+05_m_nit_Lloyd_2"]
MESH_VARS = ["xmin", "xmax", "ymin", "ymax", "tol_dist", "V", "nC", "C", "niTri", "iTri", "VBI", "Tri", "Tricc", "TriC", "TriBI", "E", "VE", "EV", "ETri",
             "TriE", "EBI", "EA", "TriGC", "TriA", "R", "A", "vi2vori", "ti2vori", "ei2vori", "vori2vi", "vori2ti", "vori2ei", "Vor", "VornC", "VorC",
             "nVVor", "VVor"]
RESULT_VARS = ["d_a_ex", "ddx_a_ex", "ddy_a_ex", "d_b_ex", "ddx_b_ex", "ddy_b_ex", "d2dx2_b_ex", "d2dxdy_b_ex", "d2dy2_b_ex", "d_a_b", "d_b_a", "ddx_a_a",
               "ddx_a_b", "ddx_b_a", "ddx_b_b", "ddy_a_a", "ddy_a_b", "ddy_b_a", "ddy_b_b", "ddx_b_b_2nd", "ddy_b_b_2nd", "d2dx2_b_b_2nd", "d2dxdy_b_b_2nd",
               "d2dy2_b_b_2nd"]


def main():
    results = {"source": "/root/reference/automated_testing/UPSY/component_test_mesh_discretisation/reference/*_checksum.nc",
               "layout": "[sum, sum_abs, min, max] over finite entries", "meshes": {}}
    for name in MESHES:
        f = h5lite.File(f"{REF}/test_meshes_and_grids/{name}.nc")
        out = {}
        for k in MESH_VARS:
            a = f[k]
            out[k] = np.ascontiguousarray(a.T) if a.ndim == 2 else a   # NetCDF (cols, n) -> Fortran (n, cols)
        res = name.split("_")[3]
        np.savez_compressed(os.path.join(HERE, f"ref_mesh_{res}.npz"), **out)
        entry = {"nV": int(out["V"].shape[0]), "nTri": int(out["Tri"].shape[0]), "nE": int(out["E"].shape[0]),
                 "git_commit": str(f.root_attrs.get("git commit hash", ""))}
        for fn in ("linear", "quadratic", "periodic"):
            g = h5lite.File(f"{REF}/UPSY/component_test_mesh_discretisation/reference/res_map_deriv_{name}_{fn}_checksum.nc")
            entry[fn] = {k: [float(x) for x in g[k]] for k in RESULT_VARS}
            # the result file repeats the mesh: its checksums must be those of the mesh file we extracted
            for k in ("V", "Tri", "TriC", "C", "TriGC", "A"):
                c = g[k]
                a = out[k].astype(np.float64)
                mine = [a.sum(), np.abs(a).sum(), a.min(), a.max()]
                assert np.allclose(np.asarray(c, dtype=np.float64), mine, rtol=1e-13, atol=1e-13 * mine[1]), (name, fn, k, c, mine)   # (summation order)
        g = h5lite.File(f"{REF}/UPSY/component_test_mesh_discretisation/reference/res_Laplace_{name}_checksum.nc")
        entry["Laplace"] = {k: [float(x) for x in g[k]] for k in ("f_ex", "f_disc")}
        results["meshes"][res] = entry
        print(name, entry["nV"], entry["nTri"], entry["nE"])
    with open(os.path.join(HERE, "ref_mesh_results.json"), "w") as fh:
        json.dump(results, fh, indent=1)


if __name__ == "__main__":
    main()
