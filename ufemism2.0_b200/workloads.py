"""Synthetic idealised inputs for the hot path (numpy; no external data).

Geometries follow the reference's idealised experiments
(src/UFEMISM/reference_geometries/idealised_geometries.f90:110-143 Halfar, :214-241 MISMIP_mod, :357-401 MISMIP+),
physics settings follow config-files/config_MISMIP_8km_spinup_for_scaling.cfg.
Masks follow determine_masks (src/UFEMISM/ice_dynamics/utilities/masks_mod.f90:25) at vertex level;
sub-grid fractions are simple stand-ins (they are *inputs* of the path).
"""
import numpy as np

import ufe_b200 as ub
import ufe_host as uh

ICE_DENSITY, SEAWATER_DENSITY, GRAV = 917.0, 1027.0, 9.81


def ice_surface_elevation(Hi, Hb, SL):
    return Hi + np.maximum(SL - ICE_DENSITY / SEAWATER_DENSITY * Hi, Hb)


def masks_from_geometry(Hi, Hb, SL):
    floating = Hi < (SL - Hb) * SEAWATER_DENSITY / ICE_DENSITY
    ice = Hi > 0.0
    return dict(
        mask_grounded_ice=(ice & ~floating).astype(np.int32),
        mask_floating_ice=(ice & floating).astype(np.int32),
        mask_icefree_land=(~ice & (Hb > SL)).astype(np.int32),
        mask_icefree_ocean=(~ice & (Hb <= SL)).astype(np.int32),
    )


def mask_noice(mesh, choice):
    """calc_mask_noice (src/UFEMISM/ice_dynamics/utilities/masks_mod.f90:385-450) for the idealised choices: where no ice is allowed"""
    x, y = mesh.V[:, 0], mesh.V[:, 1]
    if choice == "none":
        return np.zeros(mesh.nV, np.int32)
    if choice == "MISMIP_mod":
        return (np.hypot(x, y) > 900e3).astype(np.int32)         # kill all ice when r > 900 km
    if choice == "MISMIP+":
        return (x > 640e3).astype(np.int32)                      # kill all ice when x > 640 km
    raise ValueError(f'unknown choice_mask_noice "{choice}"')


def scaling_config(cfg=None):
    """config_MISMIP_8km_spinup_for_scaling.cfg :236-326 (DIVA, Zoet-Iverson, semi-implicit)."""
    c = cfg or ub.default_config()
    c.visc_it_norm_dUV_tol = 5e-5; c.visc_it_nit = 50; c.visc_it_relax = 0.4
    c.stress_balance_PETSc_rtol = 1e-4; c.stress_balance_PETSc_abstol = 1e-3
    c.choice_sliding_law = ub.SLID_ZOET_IVERSON; c.slid_ZI_p = 5.0; c.slid_ZI_ut = 200.0
    c.do_GL_subgrid_friction = 1; c.subgrid_friction_exponent_on_B_grid = 2.0
    c.choice_ice_integration_method = ub.INTEGRATION_SEMI_IMPLICIT; c.dHi_semiimplicit_fs = 1.5
    c.dHi_PETSc_rtol = 1e-8; c.dHi_PETSc_abstol = 1e-4
    c.dt_ice_min = 0.2; c.dt_ice_max = 10.0
    c.uniform_Glens_flow_factor = 1e-16; c.Hi_min = 0.1
    for k in range(4):
        c.BC_u[k] = ub.BC_INFINITE; c.BC_v[k] = ub.BC_INFINITE; c.BC_H[k] = ub.BC_ZERO
    return c


class Workload:
    """Mesh + fields + config for one named configuration."""

    def __init__(self, mesh, cfg, fields, approximation, name):
        self.mesh, self.cfg, self.f, self.approximation, self.name = mesh, cfg, fields, approximation, name

    def ice_inputs(self, vr=None, tr=None):
        """IceInputs struct for the vertex range vr = (vi1, vi2) and triangle range tr (1-based inclusive)."""
        f = self.f
        sv = slice(None) if vr is None else slice(vr[0] - 1, vr[1])
        st = slice(None) if tr is None else slice(tr[0] - 1, tr[1])
        self._keep = dict(
            Hi=np.ascontiguousarray(f["Hi"][sv]), Hs=np.ascontiguousarray(f["Hs"][sv]),
            A_flow=np.asfortranarray(f["A_flow"][sv, :]), fraction_gr_b=np.ascontiguousarray(f["fraction_gr_b"][st]),
            till_friction_angle=np.ascontiguousarray(f["till_friction_angle"][sv]), beta_sq=np.ascontiguousarray(f["beta_sq"][sv]),
            alpha_sq=np.ascontiguousarray(f["alpha_sq"][sv]), effective_pressure=np.ascontiguousarray(f["effective_pressure"][sv]),
            mask_grounded_ice=np.ascontiguousarray(f["mask_grounded_ice"][sv]),
            mask_icefree_land=np.ascontiguousarray(f["mask_icefree_land"][sv]))
        ice = ub.IceInputs()
        for n, a in self._keep.items():
            setattr(ice, n, ub.ip(a) if a.dtype == np.int32 else ub.dp(a))
        return ice


def _common_fields(mesh, Hi, Hb, SL, A_uniform=1e-16, phi=10.0):
    nV, nTri, nz = mesh.nV, mesh.nTri, mesh.nz
    Hs = ice_surface_elevation(Hi, Hb, SL)
    f = dict(Hi=Hi, Hb=Hb, SL=SL, Hs=Hs)
    f.update(masks_from_geometry(Hi, Hb, SL))
    gr = f["mask_grounded_ice"].astype(np.float64)
    tri = mesh.Tri - 1
    f["fraction_gr_b"] = (gr[tri[:, 0]] + gr[tri[:, 1]] + gr[tri[:, 2]]) / 3.0
    f["fraction_gr"] = gr
    f["A_flow"] = np.full((nV, nz), A_uniform, order="F")
    f["till_friction_angle"] = np.full(nV, phi)
    f["beta_sq"] = np.full(nV, 1.0e4)
    f["alpha_sq"] = np.full(nV, 0.5)
    f["effective_pressure"] = np.maximum(0.0, ICE_DENSITY * GRAV * Hi)
    f["SMB"] = np.full(nV, 0.3); f["BMB"] = np.zeros(nV); f["LMB"] = np.zeros(nV)
    f["fraction_margin"] = np.ones(nV)
    f["mask_noice"] = np.zeros(nV, np.int32)
    f["dHi_dt_target"] = np.zeros(nV)
    return f


def mismip_mod(h, half_width=1.0e6, nz=12, seed=20261104, jitter=0.2, nC_mem=32, dome=True):
    """Configs 1 and 3: MISMIP_mod cone-shaped island (Hb = 150 - 400 r / 750 km) under a dome + shelf."""
    mesh = uh.HostMesh.synthetic(-half_width, half_width, -half_width, half_width, h, jitter=jitter, seed=seed, nz=nz, nC_mem=nC_mem)
    x, y = mesh.V[:, 0], mesh.V[:, 1]
    s = half_width / 1.0e6
    r = np.sqrt(x * x + y * y) / s
    Hb = 150.0 - 400.0 * r / 750e3
    SL = np.zeros(mesh.nV)
    if dome:
        Rg = 600e3
        core = np.clip(1.0 - (r / Rg) ** (4.0 / 3.0), 0.0, None) ** (3.0 / 8.0)
        Hi = np.where(r < 900e3, np.maximum(2500.0 * core, 350.0 - 150.0 * np.clip((r - 500e3) / 400e3, 0.0, 1.0)), 0.0)
    else:
        Hi = np.where(r > 900e3, 0.0, 100.0)
    f = _common_fields(mesh, Hi, Hb, SL)
    # choice_mask_noice_config = 'MISMIP_mod' (config_MISMIP_8km_spinup_for_scaling.cfg:398; masks_mod.f90:385-450): no ice beyond r = 900 km, in the
    # coordinates of the (scaled) island like every other radius of this geometry
    f["mask_noice"] = (r > 900e3).astype(np.int32)
    return Workload(mesh, scaling_config(), f, approximation=1, name="MISMIP_mod")


def mismip_plus(h, nz=12, seed=20261104, jitter=0.2, grade_depth=0.5, nC_mem=32):
    """Config 2: MISMIP+ bed (800 x 80 km), SSA, Schoof2005 sliding, x-graded resolution around the grounding line."""
    mesh = uh.HostMesh.synthetic(0.0, 800e3, -40e3, 40e3, h, jitter=jitter, seed=seed, nz=nz, nC_mem=nC_mem,
                                 grade_x0=450e3, grade_width=120e3, grade_depth=grade_depth)
    x, y = mesh.V[:, 0], mesh.V[:, 1]
    xt = x / 300e3
    Bx = -150.0 - 728.8 * xt ** 2 + 343.91 * xt ** 4 - 50.57 * xt ** 6
    By = 500.0 / (1 + np.exp(-2.0 * (y - 24e3) / 4e3)) + 500.0 / (1 + np.exp(2.0 * (y + 24e3) / 4e3))
    Hb = np.maximum(Bx + By, -720.0)
    SL = np.zeros(mesh.nV)
    # a steady-state-like profile: thick grounded ice upstream thinning to a ~300 m shelf, open ocean beyond x = 640 km
    Hi = np.where(x > 640e3, 0.0, 300.0 + 1500.0 * np.clip(1.0 - x / 480e3, 0.0, None) ** 0.5)
    f = _common_fields(mesh, Hi, Hb, SL, A_uniform=2.0e-17)
    cfg = scaling_config()
    cfg.choice_sliding_law = ub.SLID_SCHOOF2005
    cfg.stress_balance_PETSc_rtol = 1e-7; cfg.stress_balance_PETSc_abstol = 1e-5
    return Workload(mesh, cfg, f, approximation=0, name="MISMIP+")


def halfar(h, half_width=800e3, nz=12, seed=20261104, jitter=0.2, H0=3000.0, R0=500e3, A=1e-16, n=3.0):
    """Config 4: Halfar dome at t = 0 (Hb = 0, SL = -10000)."""
    mesh = uh.HostMesh.synthetic(-half_width, half_width, -half_width, half_width, h, jitter=jitter, seed=seed, nz=nz)
    x, y = mesh.V[:, 0], mesh.V[:, 1]
    r = np.sqrt(x * x + y * y)
    G = 1.0 - np.minimum(1.0, r / R0) ** ((n + 1.0) / n)
    Hi = H0 * G ** (n / (2.0 * n + 1.0))
    f = _common_fields(mesh, Hi, np.zeros(mesh.nV), np.full(mesh.nV, -10000.0), A_uniform=A)
    f["SMB"] = np.zeros(mesh.nV)
    cfg = scaling_config()
    cfg.choice_sliding_law = ub.SLID_NO_SLIDING
    return Workload(mesh, cfg, f, approximation=2, name="Halfar")


def ssa_icestream(h, nz=12, seed=20261104, jitter=0.2):
    """The reference's integrated test idealised/SSA_icestream (automated_testing/integrated_tests/idealised/SSA_icestream/config_0*_*km.cfg):
    a 2000 m slab on a bed sloping in x (idealised_geometries.f90:186-212), till yield stress of Schoof (2006), SSA, u copied from the interior
    at the west / east borders ('infinite_SSA_icestream'), zero elsewhere.  Analytic solution: Schoof_SSA_solution.f90."""
    mesh = uh.HostMesh.synthetic(-400e3, 400e3, -400e3, 400e3, h, jitter=jitter, seed=seed, nz=nz)
    x = mesh.V[:, 0]
    Hi = np.full(mesh.nV, 2000.0)
    Hb = -0.0003 * x
    f = _common_fields(mesh, Hi, Hb, np.full(mesh.nV, -10000.0), A_uniform=1e-18)
    f["SMB"] = np.zeros(mesh.nV)
    cfg = ub.default_config()
    cfg.visc_it_norm_dUV_tol = 5e-8; cfg.visc_it_nit = 5000; cfg.visc_it_relax = 0.3; cfg.visc_eff_min = 1e4; cfg.vel_max = 5000.0
    cfg.stress_balance_PETSc_rtol = 1e-7; cfg.stress_balance_PETSc_abstol = 1e-5
    cfg.Glens_flow_law_exponent = 3.0; cfg.Glens_flow_law_epsilon_sq_0 = 1e-10; cfg.uniform_Glens_flow_factor = 1e-18
    cfg.choice_sliding_law = ub.SLID_IDEALISED_SSA_ICESTREAM
    cfg.refgeo_idealised_SSA_icestream_Hi = 2000.0; cfg.refgeo_idealised_SSA_icestream_dhdx = -0.0003
    cfg.refgeo_idealised_SSA_icestream_L = 150e3; cfg.refgeo_idealised_SSA_icestream_m = 1.0
    for k, (bu, bv) in enumerate([(ub.BC_ZERO, ub.BC_ZERO), (ub.BC_INFINITE_SSA_ICESTREAM, ub.BC_ZERO),
                                  (ub.BC_ZERO, ub.BC_ZERO), (ub.BC_INFINITE_SSA_ICESTREAM, ub.BC_ZERO)]):   # north, east, south, west
        cfg.BC_u[k] = bu; cfg.BC_v[k] = bv
    return Workload(mesh, cfg, f, approximation=0, name="SSA_icestream")


def ismip_hom_a(L=160e3, h=6737.0, approximation=1, nz=12, seed=20261104, jitter=0.2):
    """ISMIP-HOM experiment A as the reference's integrated test runs it (automated_testing/UFEMISM/integrated_test_ISMIP_HOM_small/
    all_ISMIP_HOM_config_files/config_ISMIP_HOM_A_*_DIVA.cfg): a 1000 m slab on a 0.5 degree slope over a doubly sinusoidal bed
    (idealised_geometries.f90:243-267), domain +-L, no sliding, all four borders 'periodic_ISMIP-HOM' for u and v."""
    mesh = uh.HostMesh.synthetic(-L, L, -L, L, h, jitter=jitter, seed=seed, nz=nz)
    x, y = mesh.V[:, 0], mesh.V[:, 1]
    Hs = 2000.0 - x * np.tan(0.5 * np.pi / 180.0)
    Hb = Hs - 1000.0 + 500.0 * np.sin(x * 2.0 * np.pi / L) * np.sin(y * 2.0 * np.pi / L)
    Hi = Hs - Hb
    f = _common_fields(mesh, Hi, Hb, np.full(mesh.nV, -10000.0), A_uniform=1e-16)
    f["SMB"] = np.zeros(mesh.nV)
    cfg = ub.default_config()
    cfg.visc_it_norm_dUV_tol = 5e-7; cfg.visc_it_nit = 5000; cfg.visc_it_relax = 0.4; cfg.visc_eff_min = 1e4; cfg.vel_max = 5000.0
    cfg.stress_balance_PETSc_rtol = 1e-6; cfg.stress_balance_PETSc_abstol = 1e-4
    cfg.Glens_flow_law_exponent = 3.0; cfg.Glens_flow_law_epsilon_sq_0 = 1e-12; cfg.uniform_Glens_flow_factor = 1e-16
    cfg.choice_sliding_law = ub.SLID_NO_SLIDING
    cfg.refgeo_idealised_ISMIP_HOM_L = L
    for k in range(4):
        cfg.BC_u[k] = ub.BC_PERIODIC_ISMIP_HOM; cfg.BC_v[k] = ub.BC_PERIODIC_ISMIP_HOM
    return Workload(mesh, cfg, f, approximation=approximation, name="ISMIP-HOM A")
