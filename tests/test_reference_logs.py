"""Oracle against output of the reference itself: the committed checksum log of integrated_test_MISMIP_mod_small (DIVA, Zoet-Iverson sliding,
Martin2011 pore water, 1022 vertices; BASELINE.json configs[0] is the 8 km variant of the same experiment).  The reference's mesh is not
recoverable, so the pin is statistical: extremes of u_b after each relaxed viscosity iteration of the first solve (tests/golden/README.md)."""
import json
import os

import numpy as np

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
