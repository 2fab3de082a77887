"""GPU: narrow seam (CSR SpMV, Krylov solve) and on-device operator construction vs the CPU oracle, through the C ABI."""
import json
import os

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import ufe_host as uh
import workloads as wl

pytestmark = pytest.mark.gpu
G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))


@pytest.fixture(scope="module")
def ctx():
    c = ub.Context(device=0)
    yield c
    c.close()


def _sys(name):
    g = G[name]
    return (g["m"], g["n"], np.array(g["ptr"], np.int32), np.array(g["ind"], np.int32), np.array(g["val"], float),
            np.array(g["x"], float), np.array(g["y"], float))


def test_spmv_known_answers(ctx):
    for i, name in enumerate(("csr6", "csr7_1", "csr7_2")):
        m, n, ptr, ind, val, x, y = _sys(name)
        ctx.csr_register(i, m, n, 1, n, 1, ptr, ind, val)
        assert np.abs(ctx.spmv(i, x, m) - y).max() < 1e-10, name
        X = np.asfortranarray(np.stack([x, 2 * x, -x], axis=1))
        Y = ctx.spmv(i, X, m, ncols=3)
        assert np.abs(Y - np.stack([y, 2 * y, -y], axis=1)).max() < 1e-10


def test_ksp_known_answers(ctx):
    for name in ("csr7_1", "csr7_2"):
        m, n, ptr, ind, val, x, y = _sys(name)
        xs, its = ctx.ksp_solve(m, 1, ptr, ind, val, y, 1e-13, 1e-30)
        assert np.abs(xs - x).max() < 1e-8, (name, xs, its)


def test_empty_and_ragged_rows(ctx):
    # rows of very different lengths, one empty row, duplicate column entries (legal in the reference: never merged)
    rng = np.random.default_rng(3)
    m = 1000
    lens = rng.integers(0, 40, m); lens[17] = 0; lens[500] = 300
    ptr = np.concatenate([[1], 1 + np.cumsum(lens)]).astype(np.int32)
    ind = rng.integers(1, m + 1, int(lens.sum())).astype(np.int32)
    val = rng.standard_normal(int(lens.sum()))
    x = rng.standard_normal(m)
    ctx.csr_register(7, m, m, 1, m, 1, ptr, ind, val)
    y = ctx.spmv(7, x, m)
    yo = ol.spmv(m, m, ptr, ind, val, x)
    assert np.abs(y - yo).max() < 1e-12 * max(1.0, np.abs(yo).max())


@pytest.fixture(scope="module")
def mesh_ops(ctx):
    w = wl.mismip_mod(2.5e4)
    o = ol.OracleMesh(w.mesh)
    o.calc_matrix_operators()
    ctx.mesh_create(w.mesh)
    ctx.calc_matrix_operators()
    return w, o


def test_operator_patterns_bit_exact_and_values(ctx, mesh_ops):
    w, o = mesh_ops
    for name in ub.OPS:
        ptr, ind, val = ctx.get_operator(ub.OP[name])
        optr, oind, oval, mm, nn = o.get_operator(ub.OP[name])
        assert np.array_equal(ptr, optr), name          # bit-exact CSR pattern
        assert np.array_equal(ind, oind), name
        scale = np.abs(oval).max()
        assert np.abs(val - oval).max() < 1e-9 * scale, (name, np.abs(val - oval).max() / scale)


def test_apply_operators_2D_and_3D(ctx, mesh_ops):
    w, o = mesh_ops
    m = w.mesh
    rng = np.random.default_rng(5)
    for name in ub.OPS:
        optr, oind, oval, mm, nn = o.get_operator(ub.OP[name])
        x = rng.standard_normal(nn)
        y = ctx.apply_operator(ub.OP[name], x, mm)
        yo = o.apply_operator(ub.OP[name], x)
        assert np.abs(y - yo).max() < 1e-12 * max(np.abs(yo).max(), 1e-300), name
    for name in ("map_a_b", "map_b_a", "ddx_a_b"):
        optr, oind, oval, mm, nn = o.get_operator(ub.OP[name])
        X = np.asfortranarray(rng.standard_normal((nn, m.nz)))
        Yo = o.apply_operator(ub.OP[name], X, ncols=m.nz)
        for interleave in ("0", "1"):      # k_sell_multicol on the reference layout (default) / k_sell_multicol_T on the column-interleaved copy
            os.environ["UFE_B200_MULTICOL_INTERLEAVE"] = interleave
            Y = ctx.apply_operator(ub.OP[name], X, mm, ncols=m.nz)
            assert np.abs(Y - Yo).max() < 1e-12 * np.abs(Yo).max(), (name, interleave)
        os.environ.pop("UFE_B200_MULTICOL_INTERLEAVE")


def test_ksp_on_stiffness_matrix(ctx, mesh_ops):
    # the DIVA stiffness matrix of the first viscosity iteration, assembled by the oracle; solve on the GPU
    w, o = mesh_ops
    m = w.mesh
    rng = np.random.default_rng(11)
    N_b = 1e10 * (1 + 0.1 * rng.random(m.nTri)); dNx = 1e3 * rng.standard_normal(m.nTri); dNy = 1e3 * rng.standard_normal(m.nTri)
    beta = 1e3 * rng.random(m.nTri); tdx = 1e4 * rng.standard_normal(m.nTri); tdy = 1e4 * rng.standard_normal(m.nTri)
    ptr, ind, val, bb = o.assemble(w.cfg, N_b, dNx, dNy, beta, tdx, tdy)
    x_gpu, its_gpu = ctx.ksp_solve(2 * m.nTri, 1, ptr, ind, val, bb, 1e-12, 1e-30)
    x_o, its_o, rc = ol.ksp_solve(ptr, ind, val, bb, 1e-12, 1e-30)
    x_g, its_g, rc2 = ol.ksp_solve(ptr, ind, val, bb, 1e-12, 1e-30, ksp=ol.KSP_GMRES_ILU0)
    assert rc == 0 and rc2 == 0
    n = np.linalg.norm(x_g)
    assert np.linalg.norm(x_gpu - x_g) / n < 1e-6, (its_gpu, its_o, its_g)      # vs the reference's default PETSc method
    assert np.linalg.norm(x_gpu - x_o) / n < 1e-7                                # vs the same algorithm on the CPU
    assert abs(its_gpu - its_o) <= max(3, its_o // 10), (its_gpu, its_o)
    # loose (config) tolerance: same stopping rule -> same iteration count within round-off noise
    x_gpu2, its2 = ctx.ksp_solve(2 * m.nTri, 1, ptr, ind, val, bb, 1e-4, 1e-3)
    x_o2, its_o2, _ = ol.ksp_solve(ptr, ind, val, bb, 1e-4, 1e-3)
    assert abs(its2 - its_o2) <= 2 and np.linalg.norm(x_gpu2 - x_o2) / n < 1e-3
