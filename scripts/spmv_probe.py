"""Times the generic CSR SpMV seam (ufe_b200_spmv on registered matrices, 1 and nz columns) and the thickness Krylov solve on a mesh of NV vertices:
kernel time per launch from the library's timing class, against the algorithmic bytes of SURVEY 8(d)
(12 nnz + 4 (rows + 1) + 8 ncols (rows + cols)) and the measured copy bandwidth.

    python scripts/spmv_probe.py [NV]                          sliced-ELL kernels (default)
    UFE_B200_CSR_KERNEL=vector python scripts/spmv_probe.py    the vector-per-row kernels
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200"))
import bench  # noqa: E402
import ufe_b200 as ub  # noqa: E402

nv = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except (OSError, KeyError):
    peak = 6650.0
w = bench.make_workload(nv, "jacobi")
m = w.mesh
ctx = ub.Context(device=0)
ctx.mesh_create(m)
ctx.calc_matrix_operators()
L = ub.lib()
print(f"kernel = {os.environ.get('UFE_B200_CSR_KERNEL', 'sliced-ELL')}, nV = {m.nV}, nTri = {m.nTri}, peak = {peak:.0f} GB/s")
rng = np.random.default_rng(1)
for name in ("map_a_b", "ddx_a_a", "ddx_b_a", "ddx_b_b", "M2_d2dx2_b_b"):
    ptr, ind, val = ctx.get_operator(ub.OP[name])
    rows = len(ptr) - 1
    cols = m.nV if name.split("_")[-2] == "a" else m.nTri
    nnz = len(ind)
    ctx.csr_register(1, rows, cols, 1, cols, 1, ptr, ind, val)
    for ncols in (1, m.nz):
        x = rng.standard_normal((cols, ncols)).copy(order="F")
        y = ctx.spmv(1, x, rows, ncols)
        # check against numpy on a sample of rows
        rs = rng.integers(0, rows, 200)
        for r in rs:
            k = slice(ptr[r] - 1, ptr[r + 1] - 1)
            want = val[k] @ x[ind[k] - 1, :]
            got = y[r]
            assert np.allclose(got, want, rtol=1e-12, atol=1e-12 * np.abs(val[k]).sum() * np.abs(x).max()), (name, r)
        B = 12 * nnz + 4 * (rows + 1) + 8 * ncols * (rows + cols)
        for variant in (("", "1"),) if ncols == 1 else (("reference-layout x (default)", "0"), ("column-interleaved x", "1")):
            os.environ["UFE_B200_MULTICOL_INTERLEAVE"] = variant[1]
            ctx.spmv(1, x, rows, ncols)
            L.ufe_b200_timing_reset(ctx.h)
            L.ufe_b200_timing_enable(ctx.h, 1)
            reps = 5
            for _ in range(reps):
                ctx.spmv(1, x, rows, ncols)
            L.ufe_b200_timing_enable(ctx.h, 0)
            ms, n = ctx.timing("operator_spmv")
            gbs = B / (ms / reps * 1e-3) / 1e9
            print(f"{name:14s} ncols={ncols:2d} rows={rows} nnz/row={nnz / rows:.1f} {variant[0]:28s}: {ms / reps:.3f} ms/product ({n // reps} launches each), "
                  f"{B / 1e9:.3f} GB algorithmic -> {gbs:.0f} GB/s = {gbs / peak:.2f} of peak")
    L.ufe_b200_csr_release(ctx.h, 1)
ctx.close()
