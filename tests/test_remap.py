"""Remapping of the velocity solver's state after a mesh update -- remap_DIVA_solver / remap_SSA_solver (DIVA_main.f90:115-222, SSA_main.f90:247-311):
map_b_a( mesh_old) -> Atlas map (apply_map_mesh_to_mesh_{2D,3D}, apply_maps.f90:531-583) -> set_border_vertices_to_interior_mean_dp
(mesh_utilities.f90:1648-1757) -> map_a_b( mesh_new) -- and the mesh destroy -> re-create -> solve sequence around it (update_mesh,
UFEMISM_main_model.f90:1211; BASELINE configs[0] has allow_mesh_updates_config = .TRUE.).

CPU: a numpy model of the chain built from the oracle's operators, its border fix checked against a loop that follows the Fortran statement by statement.
GPU: ufe_b200_remap_velocity_solver against that model, then the old handle is destroyed and the new one solves from the remapped state."""
import ctypes as C

import numpy as np
import pytest
import scipy.sparse as sp
from scipy.spatial import cKDTree

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl

FIELDS_2D = ("u_vav_b", "v_vav_b", "tau_bx_b", "tau_by_b")
FIELDS_3D = ("eta_3D_b", "u_3D_b", "v_3D_b")


def csr_of(o, name):
    ptr, ind, val, m, n = o.get_operator(ub.OP[name])
    return sp.csr_matrix((val, ind - 1, ptr - 1), shape=(m, n))


def border_mean(m, d):
    """set_border_vertices_to_interior_mean_dp_2D, vectorised: both passes read the gathered OLD values"""
    out = d.copy()
    interior = m.VBI == 0
    for vi in np.nonzero(m.VBI > 0)[0]:
        nb = m.C[vi, :m.nC[vi]] - 1
        ii = nb[interior[nb]]
        out[vi] = d[ii].mean() if len(ii) else d[nb].mean()
    return out


def border_mean_literal(m, d_partial):
    """the Fortran, statement by statement (mesh_utilities.f90:1669-1716)"""
    d_tot = d_partial.copy()
    d_partial = d_partial.copy()
    for vi in range(m.nV):
        if m.VBI[vi] > 0:
            n_int, d_sum = 0, 0.0
            for ci in range(m.nC[vi]):
                vj = m.C[vi, ci] - 1
                if m.VBI[vj] == 0:
                    n_int += 1
                    d_sum += d_tot[vj]
            if n_int > 0:
                d_partial[vi] = d_sum / n_int
    for vi in range(m.nV):
        if m.VBI[vi] > 0:
            n_int, d_sum = 0, 0.0
            for ci in range(m.nC[vi]):
                vj = m.C[vi, ci] - 1
                if m.VBI[vj] == 0:
                    n_int += 1
                d_sum += d_tot[vj]
            if n_int == 0:
                d_partial[vi] = d_sum / m.nC[vi]
    return d_partial


def nearest_vertex_map(m_old, m_new, k=3):
    """a stand-in for type_map%M with the same shape and row sums (1): inverse-distance weights of the k nearest old vertices"""
    dist, idx = cKDTree(m_old.V).query(m_new.V, k=k)
    wgt = 1.0 / np.maximum(dist, 1e-3)
    wgt /= wgt.sum(axis=1, keepdims=True)
    order = np.argsort(idx, axis=1)
    idx, wgt = np.take_along_axis(idx, order, 1), np.take_along_axis(wgt, order, 1)
    ptr = (np.arange(m_new.nV + 1) * k + 1).astype(np.int32)
    return ptr, (idx.ravel() + 1).astype(np.int32), np.ascontiguousarray(wgt.ravel())


def remap_model(m_old, m_new, M, state, names):
    o_old, o_new = ol.OracleMesh(m_old), ol.OracleMesh(m_new)
    b_a, a_b = csr_of(o_old, "map_b_a"), csr_of(o_new, "map_a_b")
    Mm = sp.csr_matrix((M[2], M[1] - 1, M[0] - 1), shape=(m_new.nV, m_old.nV))
    out = {}
    for name in names:
        f = state[name]                                        # (nTri_old,) or (nTri_old, nz)
        one = lambda col: a_b @ border_mean(m_new, Mm @ (b_a @ col))
        out[name] = one(f) if f.ndim == 1 else np.stack([one(f[:, k]) for k in range(f.shape[1])], axis=1)
    return out


def test_border_fix_model_follows_the_reference_statement_by_statement():
    m = wl.mismip_mod(1.0e5).mesh
    rng = np.random.default_rng(3)
    d = rng.standard_normal(m.nV)
    assert (m.VBI > 0).sum() > 20
    np.testing.assert_allclose(border_mean(m, d), border_mean_literal(m, d), rtol=1e-14, atol=1e-15)
    assert np.array_equal(border_mean(m, d)[m.VBI == 0], d[m.VBI == 0])


def test_remap_model_keeps_constants():
    m_old, m_new = wl.mismip_mod(1.0e5).mesh, wl.mismip_mod(8.0e4, seed=5).mesh
    M = nearest_vertex_map(m_old, m_new)
    out = remap_model(m_old, m_new, M, {"u_vav_b": np.full(m_old.nTri, 7.5)}, ("u_vav_b",))
    np.testing.assert_allclose(out["u_vav_b"], 7.5, rtol=1e-12)


def _solve(ctx, w, ice, st):
    nv, na = C.c_int(), C.c_int()
    ub.check(ub.lib().ufe_b200_velocity_solve_resident(ctx.h, C.byref(w.cfg), 1, C.byref(nv), C.byref(na)))
    ub.check(ub.lib().ufe_b200_velocity_download(ctx.h, C.byref(st)))
    return nv.value, na.value


@pytest.mark.gpu
def test_gpu_remap_velocity_solver_and_mesh_recreate():
    L = ub.lib()
    w_old, w_new = wl.mismip_mod(5.0e4), wl.mismip_mod(4.0e4, seed=11)
    for w in (w_old, w_new):
        w.cfg.stress_balance_PETSc_rtol, w.cfg.stress_balance_PETSc_abstol = 1e-10, 1e-12
    m_old, m_new = w_old.mesh, w_new.mesh
    c_old, c_new = ub.Context(device=0), ub.Context(device=0)
    try:
        c_old.mesh_create(m_old); c_old.calc_matrix_operators()
        ice_old = w_old.ice_inputs()
        st_old, A_old = ub.alloc_velocity_state(m_old.nV, m_old.nTri, m_old.nz)
        ub.check(L.ufe_b200_velocity_upload(c_old.h, C.byref(ice_old), C.byref(st_old)))
        _solve(c_old, w_old, ice_old, st_old)
        assert np.abs(A_old["u_3D_b"]).max() > 1.0 and np.abs(A_old["eta_3D_b"]).max() > 0.0
        # update_mesh: a second handle for the new mesh, the remap, then the old handle goes
        c_new.mesh_create(m_new); c_new.calc_matrix_operators()
        M = nearest_vertex_map(m_old, m_new)
        ub.check(L.ufe_b200_remap_velocity_solver(c_old.h, c_new.h, 1, m_new.nV, ub.ip(M[0]), ub.ip(M[1]), ub.dp(M[2])))
        c_old.close()
        st_new, A_new = ub.alloc_velocity_state(m_new.nV, m_new.nTri, m_new.nz)
        ub.check(L.ufe_b200_velocity_download(c_new.h, C.byref(st_new)))
        want = remap_model(m_old, m_new, M, A_old, FIELDS_2D + FIELDS_3D)
        for name in FIELDS_2D + FIELDS_3D:
            got, ref = A_new[name], want[name]
            assert got.shape == ref.shape, name
            assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max(), (name, np.abs(got - ref).max() / np.abs(ref).max())
        assert not A_new["u_b_prev"].any() and not A_new["u_base_b"].any()            # reallocate_clean / reallocate_bounds
        # the new mesh solves from the remapped state and lands where the oracle does from the same state
        remapped = {k: A_new[k].copy() for k in A_new}
        ice_new = w_new.ice_inputs()
        ub.check(L.ufe_b200_velocity_upload(c_new.h, C.byref(ice_new), None))
        nv, na = _solve(c_new, w_new, ice_new, st_new)
        st_o, A_o = ub.alloc_velocity_state(m_new.nV, m_new.nTri, m_new.nz)
        for k in FIELDS_2D + FIELDS_3D:
            A_o[k][...] = remapped[k]
        rc, nv_o, na_o = ol.OracleMesh(m_new).solve_velocity(w_new.cfg, 1, ice_new, st_o)
        assert rc == 0
        err = np.linalg.norm(A_new["u_vav_b"] - A_o["u_vav_b"]) / np.linalg.norm(A_o["u_vav_b"])
        print(f"remap: viscosity its gpu/oracle {nv}/{nv_o}, rel L2(u) {err:.2e}")
        assert err < 1e-6 and abs(nv - nv_o) <= 1
        # the same handle can also drop its mesh and take another one (ufe_b200_mesh_destroy -> ufe_b200_mesh_create): same result as a fresh handle
        ub.check(L.ufe_b200_mesh_destroy(c_new.h))
        c_new.mesh_create(m_old); c_new.calc_matrix_operators()
        st_r, A_r = ub.alloc_velocity_state(m_old.nV, m_old.nTri, m_old.nz)
        ub.check(L.ufe_b200_velocity_upload(c_new.h, C.byref(ice_old), C.byref(st_r)))
        _solve(c_new, w_old, ice_old, st_r)
        assert np.abs(A_r["u_vav_b"] - A_old["u_vav_b"]).max() <= 1e-9 * np.abs(A_old["u_vav_b"]).max()
    finally:
        c_new.close()
        c_old.close()
