"""GPU: the DIVA path against output of the reference itself (checksum log of integrated_test_MISMIP_mod_small), and against the oracle on the
same sequence of single viscosity iterations."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
from test_reference_logs import check_extremes, mismip_mod_small_workload, reference_rows

pytestmark = pytest.mark.gpu


def test_gpu_diva_follows_the_reference_checksum_log():
    nV_ref, rows = reference_rows()
    w = mismip_mod_small_workload()
    m = w.mesh
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(m)
        ctx.calc_matrix_operators()
        o = ol.OracleMesh(m)
        cfg = w.cfg
        cfg.visc_it_nit = 0
        cfg.stress_balance_PETSc_rtol = 1e-11; cfg.stress_balance_PETSc_abstol = 1e-30   # tight, so device and oracle can be compared closely
        st, f = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        st_o, f_o = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        ice = w.ice_inputs()
        nv, na = C.c_int(), C.c_int()
        seq = []
        for k in range(15):
            ub.check(ub.lib().ufe_b200_solve_DIVA(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None))
            rc, nv_o, na_o = o.solve_velocity(cfg, 1, ice, st_o)
            assert rc == 0 and nv.value == nv_o == 1
            assert np.abs(f["u_vav_b"] - f_o["u_vav_b"]).max() <= 1e-6 * np.abs(f_o["u_vav_b"]).max(), k
            seq.append((f["u_vav_b"].min(), f["u_vav_b"].max()))
        check_extremes(seq, rows)
    finally:
        ctx.close()


def test_gpu_ismip_hom_a_follows_the_reference_checksum_log():
    """ISMIP-HOM A (DIVA without sliding, periodic copy borders) on the device: all 48 viscosity iterations and every named field of the last one against
    the reference's own log; u_b of every iteration against the oracle."""
    import workloads as wl
    from test_reference_logs import check_ismip_hom_fields, check_ismip_hom_sequence, ismip_hom_reference
    g = ismip_hom_reference()
    w = wl.ismip_hom_a()
    m = w.mesh
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(m)
        ctx.calc_matrix_operators()
        o = ol.OracleMesh(m)
        cfg = w.cfg
        cfg.visc_it_nit = 0; cfg.write_diagnostics = 1
        cfg.stress_balance_PETSc_rtol = 1e-11; cfg.stress_balance_PETSc_abstol = 1e-30
        st, f = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        st_o, f_o = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        ice = w.ice_inputs()
        nv, na = C.c_int(), C.c_int()
        seq = []
        for k in range(48):
            ub.check(ub.lib().ufe_b200_solve_DIVA(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None))
            rc, nv_o, na_o = o.solve_velocity(cfg, 1, ice, st_o)
            assert rc == 0 and nv.value == nv_o == 1
            assert np.abs(f["u_vav_b"] - f_o["u_vav_b"]).max() <= 1e-6 * np.abs(f_o["u_vav_b"]).max(), k
            seq.append((f["u_vav_b"].sum(), f["u_vav_b"].min(), f["u_vav_b"].max()))
        check_ismip_hom_sequence(seq, m.nTri, g)
        check_ismip_hom_fields(f, m, g)
    finally:
        ctx.close()


def test_gpu_diva_converges_like_the_reference():
    """The whole first solve of integrated_test_MISMIP_mod_small on the device: the reference's 26 viscosity iterations and its converged maximum."""
    import json
    import os
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.json")))["mismip_mod_small"]
    w = mismip_mod_small_workload()
    m = w.mesh
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(m)
        ctx.calc_matrix_operators()
        st, f = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        ice = w.ice_inputs()
        nv, na = C.c_int(), C.c_int()
        ub.check(ub.lib().ufe_b200_solve_DIVA(ctx.h, C.byref(w.cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None))
    finally:
        ctx.close()
    assert abs(nv.value - g["n_visc_its_first_solve"]) <= 1, nv.value
    assert abs(f["u_vav_b"].max() - g["u_vav_b_converged"][3]) <= 0.02 * g["u_vav_b_converged"][3]
    assert abs(f["tau_dx_b"].max() - g["tau_dx_b"][3]) <= 0.03 * g["tau_dx_b"][3]
