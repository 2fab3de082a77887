"""Oracle against output of the reference itself: the committed checksum log of integrated_test_MISMIP_mod_small (DIVA, Zoet-Iverson sliding,
Martin2011 pore water, 1022 vertices; BASELINE.json configs[0] is the 8 km variant of the same experiment).  The reference's mesh is not
recoverable, so the pin is statistical: extremes of u_b after each relaxed viscosity iteration of the first solve (tests/golden/README.md)."""
import json
import os

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl


def mismip_mod_small_workload():
    """config of automated_testing/UFEMISM/integrated_test_MISMIP_mod_small on a uniform mesh of our generator with the reference's vertex count"""
    w = wl.mismip_mod(69e3, dome=False)         # Hi = 100 m inside r < 900 km (idealised_geometries.f90:232-236)
    f = w.f
    # calc_pore_water_pressure_Martin2011 + effective pressure (basal_hydrology_main.f90:227-266; Hb_min = 0, Hb_max = 1000 in the test's config)
    lam = np.clip(1.0 - (f["Hb"] - f["SL"] - 0.0) / (1000.0 - 0.0), 0.0, 1.0)
    f["effective_pressure"] = np.maximum(0.0, wl.ICE_DENSITY * wl.GRAV * f["Hi"] * (1.0 - 0.96 * lam))
    w.cfg.Hi_min = 0.0                          # the small test's config.cfg (the 8 km scaling file has 0.1)
    return w


def reference_rows():
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.json")))["mismip_mod_small"]
    return g["nV"], np.array(g["u_b_relaxed"])


def check_extremes(seq, rows):
    for k, (mn, mx) in enumerate(seq):
        assert abs(mx - rows[k, 3]) <= 0.03 * rows[k, 3], (k + 1, mx, rows[k, 3])
        assert abs(mn - rows[k, 2]) <= 0.20 * abs(rows[k, 2]), (k + 1, mn, rows[k, 2])     # the minimum sits on the coarser side of our mesh


def test_oracle_diva_follows_the_reference_checksum_log():
    nV_ref, rows = reference_rows()
    w = mismip_mod_small_workload()
    m = w.mesh
    assert abs(m.nV - nV_ref) <= 0.02 * nV_ref
    o = ol.OracleMesh(m)
    st, f = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    cfg = w.cfg
    cfg.visc_it_nit = 0            # one viscosity iteration per call; the state carries the DIVA fields to the next one
    seq = []
    for k in range(15):
        rc, nv, na = o.solve_velocity(cfg, 1, w.ice_inputs(), st)
        assert rc == 0 and nv == 1
        seq.append((f["u_vav_b"].min(), f["u_vav_b"].max()))
    check_extremes(seq, rows)


def test_oracle_diva_converges_like_the_reference():
    """The whole first solve of integrated_test_MISMIP_mod_small: the reference needed 26 viscosity iterations (visc_it_norm_dUV_tol = 5e-5) and ended at
    max( DIVA%u_vav_b) = 718.98 m/yr; its driving stress peaks at 488.7 Pa."""
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.json")))["mismip_mod_small"]
    assert g["n_visc_its_first_solve"] == 26 and abs(g["u_vav_b_converged"][3] - 718.979) < 1e-3
    w = mismip_mod_small_workload()
    m = w.mesh
    o = ol.OracleMesh(m)
    st, f = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    rc, n_visc, n_axb = o.solve_velocity(w.cfg, 1, w.ice_inputs(), st)
    assert rc == 0 and abs(n_visc - g["n_visc_its_first_solve"]) <= 1, n_visc
    assert abs(f["u_vav_b"].max() - g["u_vav_b_converged"][3]) <= 0.02 * g["u_vav_b_converged"][3]
    assert abs(f["tau_dx_b"].max() - g["tau_dx_b"][3]) <= 0.03 * g["tau_dx_b"][3] and abs(f["tau_dx_b"].min() - g["tau_dx_b"][2]) <= 0.03 * abs(g["tau_dx_b"][2])


def test_oracle_flux_divergence_against_the_reference_log():
    """calc_ice_flux_divergence_matrix_upwind + divQ of the first thickness step that sees velocities (conservation_of_mass_utilities.f90:23-135): the strongest
    convergence sits on the ice-free vertices just beyond the calving front and is the same on both meshes; the maximum depends on how the front is
    resolved and is not compared."""
    import bench
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.json")))["mismip_mod_small"]
    ref = g["divQ_after_first_solve"]
    w = mismip_mod_small_workload()
    m, f = w.mesh, w.f
    f["mask_noice"] = wl.mask_noice(m, "MISMIP_mod")                                   # choice_mask_noice_config = 'MISMIP_mod' (masks_mod.f90:419-428)
    o = ol.OracleMesh(m)
    st, s = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    rc, n_visc, n_axb = o.solve_velocity(w.cfg, 1, w.ice_inputs(), st)
    assert rc == 0
    out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
    io, keep = bench.thickness_io(w, slice(None), f["Hi"], s["u_vav_b"], s["v_vav_b"], out)
    rc, dt, its = o.calc_dHi_dt(w.cfg, io, 0.1)
    assert rc == 0 and dt == 0.1
    assert abs(out["divQ"].min() - ref[2]) <= 0.02 * abs(ref[2]), (out["divQ"].min(), ref[2])
    assert abs(np.abs(out["divQ"]).sum() / m.nV - ref[1] / g["nV"]) <= 0.05 * ref[1] / g["nV"]
    # no ice may appear where mask_noice holds: the flux that arrives there is taken out again
    assert not out["Hi_tplusdt"][f["mask_noice"] == 1].any() and abs(out["Hi_tplusdt"].max() - 100.03) < 1e-3          # the log: max = 0.10002999991886E+03


# ---- ISMIP-HOM experiment A, L = 160 km, DIVA without sliding, periodic borders (integrated_test_ISMIP_HOM_small) -----------------------------------
# reference name -> member of the velocity state; compared on the mean of |.| (2 %) and, where listed, on the maximum (2 %)
HOM_FIELDS = {"DIVA%du_dz_3D_a": "du_dz_3D_a", "DIVA%eta_3D_a": "eta_3D_a", "DIVA%eta_3D_b": "eta_3D_b", "DIVA%eta_vav_a": "eta_vav_a", "DIVA%N_a": "N_a",
              "DIVA%N_b": "N_b", "DIVA%F1_3D_b": "F1_3D_b", "DIVA%F2_3D_b": "F2_3D_b", "DIVA%beta_eff_a": "beta_eff_a", "DIVA%beta_eff_b": "beta_eff_b",
              "DIVA%tau_bx_b": "tau_bx_b", "DIVA%u_3D_b": "u_3D_b", "DIVA%v_3D_b": "v_3D_b", "DIVA%u_vav_b": "u_vav_b", "DIVA%v_vav_b": "v_vav_b",
              "du_dx_a": "du_dx_a", "du_dy_a": "du_dy_a"}
HOM_MAXIMA = ["DIVA%du_dz_3D_a", "DIVA%F1_3D_b", "DIVA%F2_3D_b", "DIVA%beta_eff_a", "DIVA%beta_eff_b", "DIVA%tau_bx_b", "DIVA%u_3D_b", "DIVA%u_vav_b"]


def ismip_hom_reference(L_km=160):
    return json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.json")))["ismip_hom_A_%d_DIVA" % L_km]


def check_ismip_hom_sequence(seq, nTri, g):
    """seq[k] = (sum, min, max) of u_b after relaxed viscosity iteration k + 1: all 48 iterations of the reference's solve"""
    rows = np.array(g["u_b_relaxed"])
    assert len(seq) == len(rows) >= 48
    for k, (s, mn, mx) in enumerate(seq):
        assert abs(s / nTri - rows[k, 0] / g["nTri"]) <= 0.02 * rows[k, 0] / g["nTri"], (k + 1, s / nTri, rows[k, 0] / g["nTri"])
        assert abs(mx - rows[k, 3]) <= 0.01 * rows[k, 3], (k + 1, mx, rows[k, 3])
        if k >= 4:
            assert abs(mn - rows[k, 2]) <= 0.04 * rows[k, 2], (k + 1, mn, rows[k, 2])


def check_ismip_hom_fields(f, m, g):
    for ref_name, name in HOM_FIELDS.items():
        a, r = f[name], g["last"][ref_name]
        n_ref = (g["nV"] if name.endswith("_a") else g["nTri"]) * (m.nz if a.ndim == 2 else 1)
        mean, mean_ref = np.abs(a).sum() / a.size, r[1] / n_ref
        assert abs(mean - mean_ref) <= 0.02 * mean_ref, (name, mean, mean_ref)
        if ref_name in HOM_MAXIMA:
            assert abs(a.max() - r[3]) <= 0.02 * r[3], (name, a.max(), r[3])
    assert not f["u_base_b"].any() and not f["basal_friction_coefficient_b"].any()      # no sliding: the log holds zeros too
    assert g["last"]["DIVA%u_base_b"] == [0.0] * 4 and g["last"]["DIVA%basal_friction_coefficient_b"] == [0.0] * 4


@pytest.mark.parametrize("L_km,ksp", [(160, ol.KSP_BICGSTAB_JACOBI), (160, ol.KSP_GMRES_ILU0), (80, ol.KSP_BICGSTAB_JACOBI), (40, ol.KSP_BICGSTAB_JACOBI)])
def test_oracle_ismip_hom_a_follows_the_reference_checksum_log(L_km, ksp):
    """All 48-50 viscosity iterations of the reference's ISMIP-HOM A solve, then every named DIVA field of its last iteration: means within 2 % on a mesh of
    our generator with the reference's triangle count.  The surface velocity maximum (the experiment's headline number) is 106.7 against 106.9 m/yr."""
    g = ismip_hom_reference(L_km)
    w = wl.ismip_hom_a(L=L_km * 1e3, h=6737.0 * L_km / 160.0)       # the reference scales its mesh with L (resolution L / 20)
    m = w.mesh
    assert abs(m.nTri - g["nTri"]) <= 0.02 * g["nTri"] and abs(m.nV - g["nV"]) <= 0.02 * g["nV"]
    o = ol.OracleMesh(m)
    st, f = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    cfg = w.cfg
    cfg.visc_it_nit = 0
    seq = []
    for k in range(len(g["u_b_relaxed"])):
        rc, nv, na = o.solve_velocity(cfg, 1, w.ice_inputs(), st, ksp=ksp)      # the device's Krylov method, and PETSc's default (GMRES(30) + ILU(0))
        assert rc == 0 and nv == 1
        seq.append((f["u_vav_b"].sum(), f["u_vav_b"].min(), f["u_vav_b"].max()))
    check_ismip_hom_sequence(seq, m.nTri, g)
    check_ismip_hom_fields(f, m, g)
    # the reference stopped here because the iteration had converged (visc_it_norm_dUV_tol = 5e-7); so has this one
    du = f["u_vav_b"] - f["u_b_prev"]; dv = f["v_vav_b"] - f["v_b_prev"]
    su = f["u_vav_b"] + f["u_b_prev"]; sv = f["v_vav_b"] + f["v_b_prev"]
    assert 2.0 * (du @ du + dv @ dv) / (su @ su + sv @ sv) < 2e-6
