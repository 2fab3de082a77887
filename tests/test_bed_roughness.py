"""apply_grounded_fractions_to_bed_roughness (basal_conditions/sliding_laws.f90:568-636, C%do_subgrid_friction_on_A_grid): oracle against the
definition on the masks of a marine ice sheet, device against the oracle."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl


def _case():
    w = wl.mismip_mod(4.0e4)
    m, f = w.mesh, w.f
    o = ol.OracleMesh(m)
    masks = o.determine_masks(w.cfg, f["Hi"], f["Hb"], f["SL"])
    rng = np.random.default_rng(6)
    frac = np.where(masks["mask_grounded_ice"] == 1, 1.0, 0.0)
    edge = (masks["mask_gl_gr"] | masks["mask_gl_fl"] | masks["mask_cf_gr"]) == 1
    frac[edge] = rng.random(edge.sum())
    slope = 0.02 * rng.random(m.nV) ** 3
    rough = 5.0 + 20.0 * rng.random(m.nV)
    return w, o, masks, frac, slope, rough


def _oracle(m, f, masks, frac, slope, rough):
    s = ub.Masks(*[ub.ip(masks[k]) for k in ub.MASK_FIELDS])
    out = np.full(m.nV, np.nan)
    ol.lib().orc_apply_grounded_fractions_to_bed_roughness(m.nV, ub.dp(f["Hi"]), ub.dp(slope), ub.dp(frac), C.byref(s), ub.dp(rough), ub.dp(out))
    return s, out


def test_oracle_bed_roughness_weight_definition():
    w, o, masks, frac, slope, rough = _case()
    m, f = w.mesh, w.f
    s, out = _oracle(m, f, masks, frac, slope, rough)
    expo = np.maximum(0.0, np.log10(np.maximum(1.0, f["Hi"])) - slope / 0.005)
    edge = (masks["mask_gl_gr"] | masks["mask_gl_fl"] | masks["mask_cf_gr"]) == 1
    wgt = np.ones(m.nV)                                            # ice-free land keeps its roughness
    wgt[(masks["mask_floating_ice"] == 1) | (masks["mask_icefree_ocean"] == 1)] = 0.0
    wgt[masks["mask_grounded_ice"] == 1] = 1.0
    wgt[edge] = frac[edge] ** expo[edge]
    np.testing.assert_allclose(out, rough * np.clip(wgt, 0.0, 1.0), rtol=1e-14)
    assert edge.sum() > 20 and (wgt == 0.0).sum() > 20 and ((wgt > 0.0) & (wgt < 1.0)).sum() > 10


@pytest.mark.gpu
def test_gpu_bed_roughness_weight_matches_oracle():
    w, o, masks, frac, slope, rough = _case()
    m, f = w.mesh, w.f
    s, ref = _oracle(m, f, masks, frac, slope, rough)
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(m); ctx.calc_matrix_operators()
        out = np.full(m.nV, np.nan)
        ub.check(ub.lib().ufe_b200_apply_grounded_fractions_to_bed_roughness(ctx.h, ub.dp(f["Hi"]), ub.dp(slope), ub.dp(frac), C.byref(s), ub.dp(rough),
                                                                             ub.dp(out)))
    finally:
        ctx.close()
    np.testing.assert_allclose(out, ref, rtol=1e-14, atol=0.0)
    assert np.array_equal(out == 0.0, ref == 0.0)
