"""Geometry / mask / sub-grid stage between the thickness and velocity solves: determine_masks, calc_effective_thickness,
calc_grounded_fractions (bilinear TAF + bedrock CDF), calc_critical_timestep_{adv,SIA}.  CPU: the oracle restatement against independent
numpy definitions; GPU: the CUDA path against the oracle through the C ABI (bit-exact for the masks)."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl

RHO_I, RHO_W = 917.0, 1027.0


def _case(h=4.0e4, seed=2):
    w = wl.mismip_mod(h)
    m = w.mesh
    rng = np.random.default_rng(seed)
    Hi = w.f["Hi"].copy()
    Hi[rng.random(m.nV) < 0.03] = 0.0                       # holes: margins and calving fronts inside the sheet and shelf
    thin = (rng.random(m.nV) < 0.05) & (Hi > 0)
    Hi[thin] *= 0.2
    Hb = w.f["Hb"] + 30.0 * rng.standard_normal(m.nV)
    Hi[(rng.random(m.nV) < 0.15) & (Hb > 20.0)] = 0.0        # ice-free land inside the sheet
    x, y = m.V[:, 0], m.V[:, 1]
    Hb[(np.hypot(x, y) > 0.95e6) & (np.abs(np.arctan2(y, x)) < 0.4)] = 50.0    # an island in the open ocean: coastline
    SL = np.zeros(m.nV)
    return w, m, Hi, Hb, SL


def _neigh(m):
    Cm = np.asarray(m.C).reshape(m.nV, -1, order="A") if np.asarray(m.C).ndim == 1 else np.asarray(m.C)
    return [Cm[vi, : m.nC[vi]] - 1 for vi in range(m.nV)]


def test_oracle_masks_follow_their_definitions():
    w, m, Hi, Hb, SL = _case()
    o = ol.OracleMesh(m)
    r = o.determine_masks(w.cfg, Hi, Hb, SL)
    fl = Hi < (SL - Hb) * RHO_W / RHO_I
    want = {"mask_floating_ice": fl & (Hi > 0), "mask_icefree_ocean": fl & ~(Hi > 0), "mask_grounded_ice": ~fl & (Hi > 0), "mask_icefree_land": ~fl & ~(Hi > 0)}
    for k, v in want.items():
        np.testing.assert_array_equal(r[k].astype(bool), v)
        assert v.sum() > 10, k
    nb = _neigh(m)
    ice = want["mask_floating_ice"] | want["mask_grounded_ice"]
    any_nb = lambda a: np.array([a[n].any() for n in nb])
    tr = {"mask_margin": ice & any_nb(~ice), "mask_gl_gr": want["mask_grounded_ice"] & any_nb(want["mask_floating_ice"]),
          "mask_gl_fl": want["mask_floating_ice"] & any_nb(want["mask_grounded_ice"]),
          "mask_cf_gr": want["mask_grounded_ice"] & any_nb(want["mask_icefree_ocean"]),
          "mask_cf_fl": want["mask_floating_ice"] & any_nb(want["mask_icefree_ocean"]),
          "mask_coastline": want["mask_icefree_land"] & any_nb(want["mask_icefree_ocean"])}
    code = np.select([want["mask_icefree_land"], want["mask_icefree_ocean"], want["mask_grounded_ice"], want["mask_floating_ice"]], [1, 2, 3, 4])
    for k, c in (("mask_margin", 9), ("mask_gl_gr", 5), ("mask_gl_fl", 6), ("mask_cf_gr", 7), ("mask_cf_fl", 8), ("mask_coastline", 10)):
        np.testing.assert_array_equal(r[k].astype(bool), tr[k])
        assert tr[k].sum() > 0, k
        code = np.where(tr[k], c, code)                      # later tests overwrite earlier ones (masks_mod.f90:121-203)
    np.testing.assert_array_equal(r["mask"], code)


def test_oracle_effective_thickness():
    w, m, Hi, Hb, SL = _case()
    o = ol.OracleMesh(m)
    Hi_eff, fm = o.calc_effective_thickness(w.cfg, Hi, Hb, SL)
    nb = _neigh(m)
    fl = Hi < (SL - Hb) * RHO_W / RHO_I
    margin = np.array([Hi[vi] > 0 and (Hi[nb[vi]] == 0).any() for vi in range(m.nV)])
    n_partial = 0
    for vi in range(m.nV):
        hmax = max([Hi[vj] for vj in nb[vi] if not margin[vj]], default=0.0) if (margin[vi] and fl[vi]) else 0.0
        if hmax > Hi[vi]:
            assert Hi_eff[vi] == hmax and fm[vi] == Hi[vi] / hmax
            n_partial += 1
        elif fl[vi] and Hi[vi] == 0:
            assert Hi_eff[vi] == 0 and fm[vi] == 0
        else:
            assert Hi_eff[vi] == Hi[vi] and fm[vi] == 1
    assert n_partial > 5


def _clip_area(poly, a, b, c):
    """area of polygon ∩ {a x + b y + c >= 0} (Sutherland-Hodgman), independent of the reference's corner-triangle formulas"""
    out = []
    n = len(poly)
    for i in range(n):
        p, q = poly[i], poly[(i + 1) % n]
        fp, fq = a * p[0] + b * p[1] + c, a * q[0] + b * q[1] + c
        if fp >= 0:
            out.append(p)
        if (fp >= 0) != (fq >= 0):
            t = fp / (fp - fq)
            out.append((p[0] + t * (q[0] - p[0]), p[1] + t * (q[1] - p[1])))
    if len(out) < 3:
        return 0.0
    x = np.array([p[0] for p in out]); y = np.array([p[1] for p in out])
    return 0.5 * abs(np.dot(x, np.roll(y, -1)) - np.dot(y, np.roll(x, -1)))


def test_oracle_grounded_fractions_of_a_planar_TAF_and_a_linear_CDF():
    w, m, Hi, Hb, SL = _case()
    o = ol.OracleMesh(m)
    # TAF = Hi - (SL - Hb) rho_w / rho_i planar: Hb = -500 everywhere, Hi = plane
    a, b = 3.0e-4, -1.0e-4
    Hb = np.full(m.nV, -500.0); SL = np.zeros(m.nV)
    c0 = 500.0 * RHO_W / RHO_I + 13.0
    Hi = np.maximum(0.0, a * m.V[:, 0] + b * m.V[:, 1] + c0)
    assert (Hi > 0).all()
    TAF = Hi - 500.0 * RHO_W / RHO_I                          # = a x + b y + 13
    dHb = np.zeros(m.nV)
    mf = (TAF < 0).astype(np.int32)
    fr, fr_b = o.calc_grounded_fractions(w.cfg, ub.GF_BILIN_INTERP_TAF, 11, Hi, Hb, SL, dHb, mf, None, None)
    Tri = np.asarray(m.Tri) - 1
    n_cut = 0
    for ti in range(0, m.nTri, 7):
        poly = [tuple(m.V[v]) for v in Tri[ti]]
        A = _clip_area(poly, 0, 0, 1.0)
        want = _clip_area(poly, a, b, 13.0) / A
        assert abs(fr_b[ti] - want) < 1e-9, (ti, fr_b[ti], want)
        n_cut += 0 < want < 1
    assert n_cut > 3
    # vertices: the reference takes TAF at the Voronoi vertices from map_a_b (the value at the triangle's geometric centre), so the check
    # clips every sub-triangle [vertex, cc1, cc2] with the zero line of ITS linear interpolant (general polygon clipping, no corner formulas)
    iTri = np.asarray(m.iTri) - 1
    TAF_b = o.apply_operator(ub.OP["map_a_b"], TAF)
    n_cut = 0
    for vi in np.nonzero(np.asarray(m.VBI) == 0)[0][::5]:
        ts = iTri[vi, : m.niTri[vi]]
        A = Ag = 0.0
        for k in range(len(ts)):
            t1, t2 = ts[k], ts[(k + 1) % len(ts)]
            P = np.array([m.V[vi], m.Tricc[t1], m.Tricc[t2]])
            f = np.array([TAF[vi], TAF_b[t1], TAF_b[t2]])
            pa, pb, pc = np.linalg.solve(np.c_[P, np.ones(3)], f)
            poly = [tuple(p) for p in P]
            A += _clip_area(poly, 0, 0, 1.0); Ag += _clip_area(poly, pa, pb, pc)
        want = Ag / A
        assert abs(fr[vi] - want) < 1e-9, (vi, fr[vi], want)
        n_cut += 0 < want < 1
    assert n_cut > 3
    # bedrock CDF: 11 bins linear between Hb - 100 and Hb + 100 -> grounded fraction = 1 - (Hb_float - (Hb - 100)) / 200, clipped
    nb = 11
    Hbv = -500.0 + 0.0 * m.V[:, 0]
    cdf = np.asfortranarray(Hbv[:, None] + np.linspace(-100.0, 100.0, nb)[None, :])
    Hb_b = o.apply_operator(ub.OP["map_a_b"], Hbv)
    cdf_b = np.asfortranarray(Hb_b[:, None] + np.linspace(-100.0, 100.0, nb)[None, :])
    dHb = 3.0 * np.sin(m.V[:, 0] / 2e5)
    fr, fr_b = o.calc_grounded_fractions(w.cfg, ub.GF_BEDROCK_CDF, nb, Hi, Hb, SL, dHb, mf, cdf, cdf_b)
    Hb_float = SL - Hi * RHO_I / RHO_W - dHb
    np.testing.assert_allclose(fr, np.clip(1.0 - (Hb_float - (Hbv - 100.0)) / 200.0, 0.0, 1.0), atol=1e-12)
    assert ((fr > 0) & (fr < 1)).sum() > 10
    # combined: min on the a-grid, TAF where a corner floats / CDF inland on the b-grid
    fa, fb = o.calc_grounded_fractions(w.cfg, ub.GF_BILIN_INTERP_TAF, nb, Hi, Hb, SL, dHb, mf, None, None)
    f2, f2b = o.calc_grounded_fractions(w.cfg, ub.GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF, nb, Hi, Hb, SL, dHb, mf, cdf, cdf_b)
    np.testing.assert_array_equal(f2, np.minimum(fa, fr))
    corner_floats = mf[Tri].any(axis=1)
    np.testing.assert_array_equal(f2b, np.where(corner_floats, fb, fr_b))


def test_oracle_critical_time_steps():
    w, m, Hi, Hb, SL = _case()
    o = ol.OracleMesh(m)
    rng = np.random.default_rng(4)
    u = 3.0e4 * rng.standard_normal(m.nTri); v = 2.0e4 * rng.standard_normal(m.nTri)      # fast enough to beat dt_ice_max on a 40 km mesh
    mf = ((Hi < (SL - Hb) * RHO_W / RHO_I) & (Hi > 0)).astype(np.int32)
    EV = np.asarray(m.EV)[:, :2] - 1; ETri = np.asarray(m.ETri) - 1
    for grounded_only in (0, 1):
        best = w.cfg.dt_ice_max
        for ei in range(m.nE):
            vi, vj = EV[ei]
            if Hi[vi] == 0 or Hi[vj] == 0 or (grounded_only and (mf[vi] or mf[vj])):
                continue
            ts = [t for t in ETri[ei] if t >= 0]
            uc, vc = np.mean(u[ts]), np.mean(v[ts])
            best = min(best, np.hypot(*(m.V[vi] - m.V[vj])) / max(0.1, abs(uc) + abs(vc)) * 0.9)
        got = o.calc_critical_timestep_adv(w.cfg, grounded_only, Hi, mf, u, v)
        assert abs(got - best) <= 1e-12 * best and got < w.cfg.dt_ice_max
    D = np.asfortranarray(-1e4 * rng.random((m.nTri, m.nz)))
    Tri = np.asarray(m.Tri) - 1
    d = [np.hypot(*(m.V[Tri[:, i]] - m.V[Tri[:, (i + 1) % 3]]).T) for i in range(3)]
    dmin = np.minimum(np.minimum(d[0], d[1]), d[2])
    want = np.min(dmin ** 2 / (6.0 * np.maximum(0.1, Hi[Tri].max(axis=1)) * np.maximum(1e2, np.abs(D).max(axis=1))) * 0.9)
    got = o.calc_critical_timestep_SIA(w.cfg, Hi, D)
    assert abs(got - min(want, w.cfg.dt_ice_max)) <= 1e-12 * got


def _cdfs(o, m, Hb, nb=11, seed=8):
    rng = np.random.default_rng(seed)
    spread = 40.0 + 80.0 * rng.random(m.nV)
    cdf = np.asfortranarray(Hb[:, None] + spread[:, None] * np.linspace(-1.0, 1.0, nb)[None, :] ** 3)       # uneven, increasing bins
    Hb_b = o.apply_operator(ub.OP["map_a_b"], Hb)
    spread_b = 40.0 + 80.0 * rng.random(m.nTri)
    cdf_b = np.asfortranarray(Hb_b[:, None] + spread_b[:, None] * np.linspace(-1.0, 1.0, nb)[None, :] ** 3)
    return cdf, cdf_b


@pytest.mark.gpu
def test_gpu_geometry_stage_matches_oracle():
    w, m, Hi, Hb, SL = _case(h=2.0e4)
    cfg = w.cfg
    o = ol.OracleMesh(m)
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    L = ub.lib()
    # masks: integer work, bit-exact
    ref = o.determine_masks(cfg, Hi, Hb, SL)
    s, got = ub.alloc_masks(m.nV)
    ub.check(L.ufe_b200_determine_masks(ctx.h, C.byref(cfg), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), C.byref(s)))
    for k in ub.MASK_FIELDS:
        np.testing.assert_array_equal(got[k], ref[k], err_msg=k)
    assert len(set(ref["mask"].tolist())) == 10                 # every mask type occurs in this case
    # effective thickness: comparisons, max and one division -- bit-exact
    he_ref, fm_ref = o.calc_effective_thickness(cfg, Hi, Hb, SL)
    he, fm = np.zeros(m.nV), np.zeros(m.nV)
    ub.check(L.ufe_b200_calc_effective_thickness(ctx.h, C.byref(cfg), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(he), ub.dp(fm)))
    np.testing.assert_array_equal(he, he_ref); np.testing.assert_array_equal(fm, fm_ref)
    assert ((fm > 0) & (fm < 1)).sum() > 5
    # grounded fractions, three schemes
    dHb = 5.0 * np.sin(m.V[:, 0] / 3e5)
    cdf, cdf_b = _cdfs(o, m, Hb)
    mf = ref["mask_floating_ice"]
    for choice in (ub.GF_BILIN_INTERP_TAF, ub.GF_BEDROCK_CDF, ub.GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF):
        fr_ref, frb_ref = o.calc_grounded_fractions(cfg, choice, 11, Hi, Hb, SL, dHb, mf, cdf, cdf_b)
        fr, frb = np.zeros(m.nV), np.zeros(m.nTri)
        ub.check(L.ufe_b200_calc_grounded_fractions(ctx.h, C.byref(cfg), choice, 11, ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(dHb), ub.ip(mf), ub.dp(cdf),
                                                    ub.dp(cdf_b), ub.dp(fr), ub.dp(frb)))
        assert np.max(np.abs(fr - fr_ref)) < 1e-11 and np.max(np.abs(frb - frb_ref)) < 1e-11, (choice, np.max(np.abs(fr - fr_ref)), np.max(np.abs(frb - frb_ref)))
        assert ((fr_ref > 0) & (fr_ref < 1)).sum() > 20 and ((frb_ref > 0) & (frb_ref < 1)).sum() > 20
    assert L.ufe_b200_calc_grounded_fractions(ctx.h, C.byref(cfg), 7, 11, ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(dHb), ub.ip(mf), ub.dp(cdf), ub.dp(cdf_b),
                                              ub.dp(fr), ub.dp(frb)) != 0        # unknown scheme: the reference crashes
    # critical time steps: a global minimum of products and quotients -- bit-exact
    rng = np.random.default_rng(4)
    u = 3.0e4 * rng.standard_normal(m.nTri); v = 2.0e4 * rng.standard_normal(m.nTri)
    for grounded_only in (0, 1):
        dt = C.c_double()
        ub.check(L.ufe_b200_calc_critical_timestep_adv(ctx.h, C.byref(cfg), grounded_only, ub.dp(Hi), ub.ip(mf), ub.dp(u), ub.dp(v), C.byref(dt)))
        want = o.calc_critical_timestep_adv(cfg, grounded_only, Hi, mf, u, v)
        assert abs(dt.value - want) <= 1e-14 * want and want < cfg.dt_ice_max, (dt.value, want)
    D = np.asfortranarray(-1e4 * rng.random((m.nTri, m.nz)))
    dt = C.c_double()
    ub.check(L.ufe_b200_calc_critical_timestep_SIA(ctx.h, C.byref(cfg), ub.dp(Hi), ub.dp(D), C.byref(dt)))
    want = o.calc_critical_timestep_SIA(cfg, Hi, D)
    assert abs(dt.value - want) <= 1e-14 * want and want < cfg.dt_ice_max, (dt.value, want)
    ctx.close()


def _alter_case(seed=12):
    w, m, Hi, Hb, SL = _case(h=3.0e4)
    o = ol.OracleMesh(m)
    rng = np.random.default_rng(seed)
    masks = o.determine_masks(w.cfg, Hi, Hb, SL)
    Hi_new = np.maximum(0.0, Hi + 40.0 * rng.standard_normal(m.nV) * (Hi > 0))
    ref_Hi = np.maximum(0.0, Hi + 100.0 * rng.standard_normal(m.nV) * (Hi > 0)); ref_Hi[rng.random(m.nV) < 0.1] = 0.0
    ref_Hb = Hb + 5.0 * rng.standard_normal(m.nV)
    Ti_hom = -30.0 * rng.random(m.nV)
    return w, m, o, masks, Hi, Hb, SL, Hi_new, ref_Hi, ref_Hb, Ti_hom


def _alter_variants():
    P = []
    p = ub.alter_params_defaults(); P.append(("defaults", p, 0.0))
    p = ub.alter_params_defaults(); p.Hi_min = 30.0; p.do_protect_grounded_mask = 1; p.protect_grounded_mask_t_end = 100.0; p.remove_ice_absent_at_PD = 1
    P.append(("protect+Hi_min+PD", p, 50.0))
    p = ub.alter_params_defaults(); p.do_remove_shelves = 1; P.append(("no_shelves", p, 0.0))
    p = ub.alter_params_defaults(); p.remove_shelves_larger_than_PD = 1; p.continental_shelf_calving = 1; p.continental_shelf_min_height = -300.0
    P.append(("PD_front+shelf_edge", p, 0.0))
    p = ub.alter_params_defaults(); p.fixiness_t_start, p.fixiness_t_end = 0.0, 100.0
    p.fixiness_H_gl_gr, p.fixiness_H_gl_fl, p.fixiness_H_grounded, p.fixiness_H_floating = 1.0, 0.5, 0.8, 0.2
    p.fixiness_H_freeland = 1; p.fixiness_H_freeocean = 1
    P.append(("fixiness", p, 25.0))
    p = ub.alter_params_defaults(); p.limitness_t_start, p.limitness_t_end = 0.0, 100.0; p.do_limitness_before_start = 1
    p.limitness_H_gl_gr, p.limitness_H_gl_fl, p.limitness_H_grounded, p.limitness_H_floating = 10.0, 20.0, 30.0, 40.0
    p.modiness_H_style = 1
    P.append(("limitness+Ti_hom", p, -5.0))
    p = ub.alter_params_defaults(); p.limitness_t_start, p.limitness_t_end = 0.0, 100.0; p.limitness_H_grounded = 15.0; p.limitness_H_gl_gr = 15.0
    p.limitness_H_gl_fl = p.limitness_H_floating = 25.0; p.modiness_H_style = 4
    P.append(("limitness_decaying+no_thick_inland", p, 60.0))
    return P


def test_oracle_alter_ice_thickness_properties():
    w, m, o, masks, Hi, Hb, SL, Hi_new, ref_Hi, ref_Hb, Ti_hom = _alter_case()
    V = dict((n, (p, t)) for n, p, t in _alter_variants())
    he, fm = o.calc_effective_thickness(w.cfg, Hi_new, Hb, SL)
    fl = he < (SL - Hb) * RHO_W / RHO_I
    # defaults: only the thickness-threshold calving law acts
    p, t = V["defaults"]
    out = o.alter_ice_thickness(w.cfg, p, t, masks, Hi, Hb, SL, ref_Hi, ref_Hb, Ti_hom, Hi_new)
    calved = fl & (he < 100.0)
    assert calved.sum() > 10 and (out[calved] == 0).all()
    np.testing.assert_allclose(out[~calved], Hi_new[~calved], rtol=1e-13)      # ref + (H - ref) is H up to rounding (limitness 0)
    p, t = V["no_shelves"]
    out = o.alter_ice_thickness(w.cfg, p, t, masks, Hi, Hb, SL, ref_Hi, ref_Hb, Ti_hom, Hi_new)
    assert (out[fl] == 0).all()
    np.testing.assert_allclose(out[~fl], Hi_new[~fl], rtol=1e-13)
    # full fixiness at the grounded side of the grounding line, before the interval starts it is off
    p, t = V["fixiness"]
    out = o.alter_ice_thickness(w.cfg, p, t, masks, Hi, Hb, SL, ref_Hi, ref_Hb, Ti_hom, Hi_new)
    keep = ~calved
    g = masks["mask_gl_gr"].astype(bool) & keep
    np.testing.assert_allclose(out[g], Hi[g] * 0.75 + Hi_new[g] * 0.25, rtol=1e-13)
    inl = masks["mask_grounded_ice"].astype(bool) & ~masks["mask_gl_gr"].astype(bool) & keep
    np.testing.assert_allclose(out[inl], Hi[inl] * 0.6 + Hi_new[inl] * 0.4, rtol=1e-13)
    # peak limitness: within the allowed departure from the reference thickness (modiness widens nothing: it only tightens)
    p, t = V["limitness+Ti_hom"]
    out = o.alter_ice_thickness(w.cfg, p, t, masks, Hi, Hb, SL, ref_Hi, ref_Hb, Ti_hom, Hi_new)
    lim = np.select([masks["mask_gl_gr"] > 0, masks["mask_gl_fl"] > 0, masks["mask_grounded_ice"] > 0, masks["mask_floating_ice"] > 0,
                     masks["mask_icefree_land"] > 0], [10.0, 20.0, 30.0, 40.0, 30.0], 40.0)
    assert (np.abs(out - ref_Hi) <= lim + 1e-9).all() and (np.abs(out - ref_Hi) > 0.5 * lim).sum() > 10


def test_oracle_spill_over_flux_conserves_mass():
    w, m, o, masks, Hi, Hb, SL, Hi_new, ref_Hi, ref_Hb, Ti_hom = _alter_case()
    rng = np.random.default_rng(3)
    r = np.hypot(m.V[:, 0], m.V[:, 1])
    Tri = np.asarray(m.Tri) - 1
    rc = m.V[Tri].mean(axis=1)
    u = 400.0 * rc[:, 0] / 1e6 + 20.0 * rng.standard_normal(m.nTri); v = 400.0 * rc[:, 1] / 1e6 + 20.0 * rng.standard_normal(m.nTri)   # outward flow
    cf = (masks["mask_cf_fl"] + masks["mask_cf_gr"]) > 0
    H = Hi.copy(); H[cf] *= 1.0 + rng.random(cf.sum())                  # calving-front vertices thicker than their upstream neighbours
    he, fm = o.calc_effective_thickness(w.cfg, H, Hb, SL)
    out, q = o.calc_and_apply_spill_over_flux(masks, he, u, v, 0.5, H)
    A = np.asarray(m.A)
    src, dst = q < 0, q > 0
    assert src.sum() > 10 and dst.sum() > 10 and cf[src].all() and (masks["mask_icefree_ocean"][dst] > 0).all()
    assert abs(np.sum(q * A)) < 1e-9 * np.sum(np.abs(q) * A)           # what leaves the front arrives in the ocean cells
    np.testing.assert_allclose(out, H + q * 0.5, rtol=0, atol=1e-9)
    assert (out[src] < H[src]).all() and (out[dst] > 0).all()


@pytest.mark.gpu
def test_gpu_alter_ice_thickness_and_spill_over_match_oracle():
    w, m, o, masks, Hi, Hb, SL, Hi_new, ref_Hi, ref_Hb, Ti_hom = _alter_case()
    cfg = w.cfg
    ctx = ub.Context(device=0)
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    L = ub.lib()
    ms = ub.masks_struct(masks)
    for name, p, t in _alter_variants():
        want = o.alter_ice_thickness(cfg, p, t, masks, Hi, Hb, SL, ref_Hi, ref_Hb, Ti_hom, Hi_new)
        got = Hi_new.copy()
        ub.check(L.ufe_b200_alter_ice_thickness(ctx.h, C.byref(cfg), C.byref(p), t, C.byref(ms), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(ref_Hi), ub.dp(ref_Hb),
                                                ub.dp(Ti_hom), ub.dp(got)))
        assert np.max(np.abs(got - want)) <= 1e-11 * np.max(np.abs(want)), (name, np.max(np.abs(got - want)))
        assert not np.array_equal(want, Hi_new), name
    p = ub.alter_params_defaults(); p.modiness_H_style = 9
    assert L.ufe_b200_alter_ice_thickness(ctx.h, C.byref(cfg), C.byref(p), 0.0, C.byref(ms), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(ref_Hi), ub.dp(ref_Hb),
                                          ub.dp(Ti_hom), ub.dp(Hi_new.copy())) != 0
    rng = np.random.default_rng(3)
    Tri = np.asarray(m.Tri) - 1
    rc = m.V[Tri].mean(axis=1)
    u = 400.0 * rc[:, 0] / 1e6 + 20.0 * rng.standard_normal(m.nTri); v = 400.0 * rc[:, 1] / 1e6 + 20.0 * rng.standard_normal(m.nTri)
    cf = (masks["mask_cf_fl"] + masks["mask_cf_gr"]) > 0
    H = Hi.copy(); H[cf] *= 1.0 + rng.random(cf.sum())
    he, fm = o.calc_effective_thickness(cfg, H, Hb, SL)
    want, q_want = o.calc_and_apply_spill_over_flux(masks, he, u, v, 0.5, H)
    got, q = H.copy(), np.zeros(m.nV)
    ub.check(L.ufe_b200_calc_and_apply_spill_over_flux(ctx.h, C.byref(ms), ub.dp(he), ub.dp(u), ub.dp(v), 0.5, ub.dp(got), ub.dp(q)))
    ctx.close()
    assert (q_want != 0).sum() > 20
    assert np.max(np.abs(q - q_want)) <= 1e-11 * np.max(np.abs(q_want)) and np.max(np.abs(got - want)) <= 1e-11 * np.max(want)
