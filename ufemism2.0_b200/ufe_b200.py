"""ctypes binding of libufe_b200.so (include/ufe_b200.h) for tests, bench.py and smoke().

This is plumbing only: every call goes straight through the C ABI the Fortran shim binds
(ufemism2.0_b200/fortran/ufe_b200_mod.f90).  There is no fallback: if the CUDA library is
missing or no device is present the calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libufe_b200.so")

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int32)

# enumerations of include/ufe_b200.h
KSP_BICGSTAB = 0
PC_JACOBI, PC_NONE, PC_MG = 0, 1, 4
BC_INFINITE, BC_ZERO, BC_PERIODIC_ISMIP_HOM, BC_INFINITE_SSA_ICESTREAM = 0, 1, 2, 3
(SLID_NO_SLIDING, SLID_WEERTMAN, SLID_COULOMB, SLID_BUDD, SLID_TSAI2015, SLID_SCHOOF2005, SLID_ZOET_IVERSON,
 SLID_IDEALISED_SSA_ICESTREAM) = range(8)
INTEGRATION_EXPLICIT, INTEGRATION_SEMI_IMPLICIT = 0, 1
OPS = ["ddx_a_a", "ddy_a_a", "map_a_b", "ddx_a_b", "ddy_a_b", "map_b_a", "ddx_b_a", "ddy_b_a", "ddx_b_b", "ddy_b_b",
       "M2_ddx_b_b", "M2_ddy_b_b", "M2_d2dx2_b_b", "M2_d2dxdy_b_b", "M2_d2dy2_b_b"]
OP = {n: i for i, n in enumerate(OPS)}
# (row grid, column grid) of each operator
OP_GRIDS = {"ddx_a_a": "aa", "ddy_a_a": "aa", "map_a_b": "ba", "ddx_a_b": "ba", "ddy_a_b": "ba", "map_b_a": "ab",
            "ddx_b_a": "ab", "ddy_b_a": "ab"}
for _n in OPS[8:]:
    OP_GRIDS[_n] = "bb"


class Config(C.Structure):
    _fields_ = [(n, C.c_double) for n in (
        "visc_it_norm_dUV_tol", "visc_it_relax", "visc_eff_min", "vel_max", "stress_balance_PETSc_rtol",
        "stress_balance_PETSc_abstol", "Glens_flow_law_exponent", "Glens_flow_law_epsilon_sq_0", "slid_delta_v",
        "slid_Weertman_m", "slid_Budd_q_plastic", "slid_Budd_u_threshold", "slid_ZI_ut", "slid_ZI_p", "slid_beta_max",
        "subgrid_friction_exponent_on_B_grid", "Hi_min", "uniform_Glens_flow_factor",
        "refgeo_idealised_SSA_icestream_Hi", "refgeo_idealised_SSA_icestream_dhdx", "refgeo_idealised_SSA_icestream_L",
        "refgeo_idealised_SSA_icestream_m", "refgeo_idealised_ISMIP_HOM_L", "dHi_semiimplicit_fs", "dHi_PETSc_rtol", "dHi_PETSc_abstol", "dt_ice_min",
        "dt_ice_max", "ice_density", "seawater_density", "grav")] + [
        ("visc_it_nit", C.c_int32), ("do_include_SSADIVA_crossterms", C.c_int32), ("BC_u", C.c_int32 * 4),
        ("BC_v", C.c_int32 * 4), ("choice_sliding_law", C.c_int32), ("do_GL_subgrid_friction", C.c_int32),
        ("choice_ice_integration_method", C.c_int32), ("BC_H", C.c_int32 * 4), ("ksp_type", C.c_int32),
        ("pc_type", C.c_int32), ("ksp_maxit", C.c_int32), ("write_diagnostics", C.c_int32), ("ksp_initial_guess_nonzero", C.c_int32), ("stiffness_format", C.c_int32), ("reserved", C.c_int32 * 4)]


def default_config():
    """Reference namelist defaults (model_configuration_type_and_namelist.f90); mirrors ufe_b200_config_defaults."""
    c = Config()
    c.visc_it_norm_dUV_tol = 5e-5; c.visc_it_relax = 0.2; c.visc_eff_min = 1e4; c.vel_max = 5000.0
    c.stress_balance_PETSc_rtol = 1e-7; c.stress_balance_PETSc_abstol = 1e-5
    c.Glens_flow_law_exponent = 3.0; c.Glens_flow_law_epsilon_sq_0 = 1e-8
    c.slid_delta_v = 1e-3; c.slid_Weertman_m = 3.0; c.slid_Budd_q_plastic = 0.3; c.slid_Budd_u_threshold = 100.0
    c.slid_ZI_ut = 200.0; c.slid_ZI_p = 5.0; c.slid_beta_max = 1e20
    c.subgrid_friction_exponent_on_B_grid = 2.0; c.Hi_min = 0.0; c.uniform_Glens_flow_factor = 1e-16
    c.refgeo_idealised_SSA_icestream_Hi = 2000.0; c.refgeo_idealised_SSA_icestream_dhdx = -0.0003
    c.refgeo_idealised_SSA_icestream_L = 150e3; c.refgeo_idealised_SSA_icestream_m = 1.0
    c.refgeo_idealised_ISMIP_HOM_L = 0.0
    c.dHi_semiimplicit_fs = 1.5; c.dHi_PETSc_rtol = 1e-8; c.dHi_PETSc_abstol = 1e-6
    c.dt_ice_min = 0.1; c.dt_ice_max = 10.0
    c.ice_density = 917.0; c.seawater_density = 1027.0; c.grav = 9.81
    c.visc_it_nit = 50; c.do_include_SSADIVA_crossterms = 1
    for k in range(4):
        c.BC_u[k] = BC_INFINITE; c.BC_v[k] = BC_INFINITE; c.BC_H[k] = BC_ZERO
    c.choice_sliding_law = SLID_ZOET_IVERSON; c.do_GL_subgrid_friction = 1
    c.choice_ice_integration_method = INTEGRATION_SEMI_IMPLICIT
    c.ksp_type = KSP_BICGSTAB; c.pc_type = PC_JACOBI; c.ksp_maxit = 2000; c.write_diagnostics = 1
    return c


class MeshDesc(C.Structure):
    _fields_ = [("nV", C.c_int32), ("nTri", C.c_int32), ("nE", C.c_int32), ("nC_mem", C.c_int32), ("nz", C.c_int32),
                ("xmin", C.c_double), ("xmax", C.c_double), ("ymin", C.c_double), ("ymax", C.c_double),
                ("V", c_dp), ("nC", c_ip), ("C", c_ip), ("niTri", c_ip), ("iTri", c_ip), ("VBI", c_ip), ("Tri", c_ip),
                ("TriC", c_ip), ("TriBI", c_ip), ("TriGC", c_dp), ("A", c_dp), ("Cw", c_dp), ("D_x", c_dp), ("D_y", c_dp),
                ("D", c_dp), ("VE", c_ip), ("ETri", c_ip), ("zeta", c_dp), ("Tricc", c_dp)]


class IceInputs(C.Structure):
    _fields_ = [("Hi", c_dp), ("Hs", c_dp), ("A_flow", c_dp), ("fraction_gr_b", c_dp), ("till_friction_angle", c_dp),
                ("beta_sq", c_dp), ("alpha_sq", c_dp), ("effective_pressure", c_dp), ("mask_grounded_ice", c_ip),
                ("mask_icefree_land", c_ip)]


VELOCITY_FIELDS = ["u_vav_b", "v_vav_b", "u_base_b", "v_base_b", "tau_bx_b", "tau_by_b", "eta_3D_b", "u_b_prev", "v_b_prev",
                   "tau_dx_b", "tau_dy_b", "u_3D_b", "v_3D_b", "du_dx_a", "du_dy_a", "dv_dx_a", "dv_dy_a", "du_dz_3D_a",
                   "dv_dz_3D_a", "eta_3D_a", "eta_vav_a", "N_a", "N_b", "dN_dx_b", "dN_dy_b", "F1_3D_a", "F2_3D_a", "F1_3D_b",
                   "F2_3D_b", "basal_friction_coefficient_a", "basal_friction_coefficient_b", "beta_eff_a", "beta_eff_b"]
# grid ('a' vertices / 'b' triangles) and whether the field has nz layers
VELOCITY_SHAPES = {n: (("a" if n.endswith("_a") else "b"), "3D" in n) for n in VELOCITY_FIELDS}
VELOCITY_SHAPES["u_b_prev"] = ("b", False); VELOCITY_SHAPES["v_b_prev"] = ("b", False)


class VelocityState(C.Structure):
    _fields_ = [(n, c_dp) for n in VELOCITY_FIELDS]


SECONDARY_FIELDS = ["u_surf_b", "v_surf_b", "uabs_surf_b", "u_base_b", "v_base_b", "uabs_base_b", "u_vav_b", "v_vav_b", "uabs_vav_b", "u_3D", "v_3D",
                    "u_surf", "v_surf", "u_base", "v_base", "u_vav", "v_vav", "uabs_surf", "uabs_base", "uabs_vav", "R_shear"]


ICE_VELOCITY_FIELDS = ["u_3D_b", "v_3D_b", "du_dx_3D", "du_dy_3D", "du_dz_3D", "dv_dx_3D", "dv_dy_3D", "dv_dz_3D", "dw_dx_3D", "dw_dy_3D"]
APPROX_SSA, APPROX_DIVA, APPROX_SIA, APPROX_SIASSA = 0, 1, 2, 3
HYBRID_ADD, HYBRID_ADD_SIA_REDUCED = 0, 1


class IceVelocities(C.Structure):
    """outputs of set_ice_velocities_to_*_results: ice%u_3D_b, v_3D_b (ti1:ti2,nz) and the 3-D strain rates (vi1:vi2,nz)"""
    _fields_ = [(n, c_dp) for n in ICE_VELOCITY_FIELDS]


def alloc_ice_velocities(n_a, n_b, nz, fill=0.0):
    s = IceVelocities()
    arrs = {}
    for n in ICE_VELOCITY_FIELDS:
        a = np.full((n_b if n.endswith("_b") else n_a, nz), fill, order="F")
        arrs[n] = a
        setattr(s, n, dp(a))
    return s, arrs


class SecondaryVelocities(C.Structure):
    _fields_ = [(n, c_dp) for n in SECONDARY_FIELDS]


def alloc_secondary(n_a, n_b, nz):
    s = SecondaryVelocities()
    arrs = {}
    for n in SECONDARY_FIELDS:
        a = np.zeros(n_b) if n.endswith("_b") else (np.zeros((n_a, nz), order="F") if n in ("u_3D", "v_3D") else np.zeros(n_a))
        arrs[n] = a
        setattr(s, n, dp(a))
    return s, arrs


MASK_FIELDS = ["mask", "mask_icefree_land", "mask_icefree_ocean", "mask_grounded_ice", "mask_floating_ice", "mask_margin", "mask_gl_gr", "mask_gl_fl",
               "mask_cf_gr", "mask_cf_fl", "mask_coastline"]
GF_BILIN_INTERP_TAF, GF_BEDROCK_CDF, GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF = 0, 1, 2


class Masks(C.Structure):
    _fields_ = [(n, c_ip) for n in MASK_FIELDS]


def alloc_masks(n):
    arrs = {k: np.zeros(n, np.int32) for k in MASK_FIELDS}
    return Masks(*[ip(arrs[k]) for k in MASK_FIELDS]), arrs


class AlterParams(C.Structure):
    """C% scalars of alter_ice_thickness (include/ufe_b200.h: ufe_b200_alter_params)"""
    _fields_ = [("do_protect_grounded_mask", C.c_int32), ("protect_grounded_mask_t_end", C.c_double), ("Hi_min", C.c_double),
                ("calving_law_threshold_thickness", C.c_int32), ("calving_threshold_thickness_shelf", C.c_double),
                ("remove_ice_absent_at_PD", C.c_int32), ("do_remove_shelves", C.c_int32), ("remove_shelves_larger_than_PD", C.c_int32),
                ("continental_shelf_calving", C.c_int32), ("continental_shelf_min_height", C.c_double),
                ("fixiness_t_start", C.c_double), ("fixiness_t_end", C.c_double), ("do_fixiness_before_start", C.c_int32),
                ("limitness_t_start", C.c_double), ("limitness_t_end", C.c_double), ("do_limitness_before_start", C.c_int32),
                ("modiness_H_style", C.c_int32), ("modiness_T_hom_ref", C.c_double),
                ("fixiness_H_gl_gr", C.c_double), ("fixiness_H_gl_fl", C.c_double), ("fixiness_H_grounded", C.c_double), ("fixiness_H_floating", C.c_double),
                ("fixiness_H_freeland", C.c_int32), ("fixiness_H_freeocean", C.c_int32),
                ("limitness_H_gl_gr", C.c_double), ("limitness_H_gl_fl", C.c_double), ("limitness_H_grounded", C.c_double), ("limitness_H_floating", C.c_double)]


def alter_params_defaults():
    p = AlterParams()
    lib().ufe_b200_alter_params_defaults(C.byref(p))
    return p


def masks_struct(arrs):
    """Masks pointing into a dict of int32 arrays (missing keys -> NULL)"""
    return Masks(*[ip(arrs.get(k)) for k in MASK_FIELDS])


class VerticalVelocityIO(C.Structure):
    _fields_ = [(n, c_dp) for n in ("Hi", "Hs", "Hib", "dHb_dt", "dHi_dt", "BMB")] + [("mask_grounded_ice", c_ip), ("mask_floating_ice", c_ip),
                                                                                      ("w_3D", c_dp), ("dw_dz_3D", c_dp)]


class ThicknessIO(C.Structure):
    _fields_ = [("Hi", c_dp), ("Hb", c_dp), ("SL", c_dp), ("u_vav_b", c_dp), ("v_vav_b", c_dp), ("SMB", c_dp), ("BMB", c_dp),
                ("LMB", c_dp), ("fraction_margin", c_dp), ("mask_noice", c_ip), ("dHi_dt_target", c_dp),
                ("BC_prescr_mask", c_ip), ("BC_prescr_Hi", c_dp), ("AMB", c_dp), ("dHi_dt", c_dp), ("Hi_tplusdt", c_dp),
                ("divQ", c_dp)]


def dp(a):
    """double* of a contiguous float64 array (None -> NULL)."""
    if a is None:
        return None
    assert a.dtype == np.float64 and (a.flags["F_CONTIGUOUS"] or a.flags["C_CONTIGUOUS"])
    return a.ctypes.data_as(c_dp)


def ip(a):
    if a is None:
        return None
    assert a.dtype == np.int32 and (a.flags["F_CONTIGUOUS"] or a.flags["C_CONTIGUOUS"])
    return a.ctypes.data_as(c_ip)


def mesh_desc(m):
    """MeshDesc pointing into a ufe_host.HostMesh (keeps no reference: keep `m` alive)."""
    d = MeshDesc()
    d.nV, d.nTri, d.nE, d.nC_mem, d.nz = m.nV, m.nTri, m.nE, m.nC_mem, m.nz
    d.xmin, d.xmax, d.ymin, d.ymax = m.xmin, m.xmax, m.ymin, m.ymax
    for n in ("V", "TriGC", "A", "Cw", "D_x", "D_y", "D", "zeta", "Tricc"):
        setattr(d, n, dp(getattr(m, n)))
    for n in ("nC", "C", "niTri", "iTri", "VBI", "Tri", "TriC", "TriBI", "VE", "ETri"):
        setattr(d, n, ip(getattr(m, n)))
    return d


_LIB = None


def lib():
    """Load libufe_b200.so and declare prototypes. Raises if the library has not been built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
    L = C.CDLL(LIB_PATH)
    H = C.c_void_p
    L.ufe_b200_last_error.restype = C.c_char_p
    L.ufe_b200_kernel_launches.restype = C.c_longlong
    L.ufe_b200_kernel_launches.argtypes = [H]
    L.ufe_b200_config_defaults.argtypes = [C.POINTER(Config)]
    L.ufe_b200_get_unique_id.argtypes = [C.c_void_p]
    L.ufe_b200_init.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(H)]
    L.ufe_b200_finalize.argtypes = [H]
    L.ufe_b200_ksp_solve.argtypes = [H, C.c_int, C.c_int, C.c_int, c_ip, c_ip, c_dp, c_dp, c_dp, C.c_double, C.c_double,
                                     C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
    L.ufe_b200_csr_register.argtypes = [H] + [C.c_int] * 7 + [c_ip, c_ip, c_dp]
    L.ufe_b200_csr_update_values.argtypes = [H, C.c_int, c_dp]
    L.ufe_b200_spmv.argtypes = [H, C.c_int, c_dp, c_dp, C.c_int]
    L.ufe_b200_csr_release.argtypes = [H, C.c_int]
    L.ufe_b200_mesh_create.argtypes = [H, C.POINTER(MeshDesc)]
    L.ufe_b200_mesh_destroy.argtypes = [H]
    L.ufe_b200_mesh_ranges.argtypes = [H, c_ip]
    L.ufe_b200_calc_matrix_operators.argtypes = [H]
    L.ufe_b200_set_operator.argtypes = [H, C.c_int, C.c_int, c_ip, c_ip, c_dp]
    L.ufe_b200_get_operator_size.argtypes = [H, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.ufe_b200_get_operator.argtypes = [H, C.c_int, c_ip, c_ip, c_dp]
    L.ufe_b200_apply_operator.argtypes = [H, C.c_int, c_dp, c_dp, C.c_int]
    L.ufe_b200_solve_SSA_DIVA_linearised.argtypes = [H, C.POINTER(Config)] + [c_dp] * 10 + [C.c_double, C.c_double,
                                                     C.POINTER(C.c_int), c_ip, c_dp, c_dp]
    L.ufe_b200_get_stiffness_size.argtypes = [H, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.ufe_b200_get_stiffness.argtypes = [H, c_ip, c_ip, c_dp, c_dp]
    L.ufe_b200_get_stiffness_storage_bytes.argtypes = [H, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
    L.ufe_b200_set_mg_params.argtypes = [H, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int]
    L.ufe_b200_set_mg_option.argtypes = [H, C.c_char_p, C.c_double]
    L.ufe_b200_get_mg_info.argtypes = [H, C.POINTER(C.c_int), c_ip, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), c_dp]
    L.ufe_b200_mg_host_transfer.argtypes = [C.c_int, C.c_int, c_ip, c_ip, C.POINTER(C.c_int), C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
    L.ufe_b200_mg_host_transfer_get.argtypes = [c_ip] * 5
    L.ufe_b200_mg_host_transfer_parts.argtypes = [C.c_int, C.c_int] + [c_ip] * 6 + [C.POINTER(C.c_int), C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
    L.ufe_b200_mg_host_transfer_parts_get.argtypes = [c_ip] * 6
    for fn in (L.ufe_b200_solve_DIVA, L.ufe_b200_solve_SSA):
        fn.argtypes = [H, C.POINTER(Config), C.POINTER(IceInputs), C.POINTER(VelocityState), C.POINTER(C.c_int),
                       C.POINTER(C.c_int), c_ip, c_dp, c_dp]
    L.ufe_b200_velocity_upload.argtypes = [H, C.POINTER(IceInputs), C.POINTER(VelocityState)]
    L.ufe_b200_velocity_solve_resident.argtypes = [H, C.POINTER(Config), C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.ufe_b200_velocity_download.argtypes = [H, C.POINTER(VelocityState)]
    L.ufe_b200_calc_dHi_dt.argtypes = [H, C.POINTER(Config), C.POINTER(ThicknessIO), c_dp, C.c_int, C.POINTER(C.c_int)]
    L.ufe_b200_dHi_dt_finish.argtypes = [H, c_dp, c_dp, C.c_double, c_dp, c_dp]
    L.ufe_b200_thickness_upload.argtypes = [H, C.POINTER(ThicknessIO)]
    L.ufe_b200_pc_step_resident.argtypes = [H, C.POINTER(Config), C.c_int, C.c_double] + [C.POINTER(C.c_int)] * 3
    L.ufe_b200_thickness_download.argtypes = [H, c_dp, c_dp, c_dp]
    L.ufe_b200_pc_step.argtypes = [H, C.POINTER(Config), C.c_int, C.c_double, C.c_double, C.POINTER(IceInputs), C.POINTER(VelocityState),
                                   C.POINTER(ThicknessIO), c_dp] + [C.POINTER(C.c_int)] * 3
    L.ufe_b200_solve_SIA.argtypes = [H, C.POINTER(Config), C.c_double] + [c_dp] * 8
    L.ufe_b200_calc_secondary_velocities.argtypes = [H, c_dp, c_dp, C.POINTER(SecondaryVelocities)]
    L.ufe_b200_apply_grounded_fractions_to_bed_roughness.argtypes = [H, c_dp, c_dp, c_dp, C.POINTER(Masks), c_dp, c_dp]
    L.ufe_b200_calc_pc_truncation_error.argtypes = [H] + [C.c_double] * 4 + [c_dp, c_dp, c_ip, c_ip, c_dp, c_dp, c_ip, C.POINTER(C.c_double),
                                                    C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.ufe_b200_calc_pc_truncation_error_noalloc.argtypes = [H, C.c_int] + [C.c_double] * 4 + [c_dp, c_dp, c_ip, c_ip, c_dp, c_dp, c_ip, C.POINTER(C.c_double),
                                                            C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.ufe_b200_remap_velocity_solver.argtypes = [H, H, C.c_int, C.c_int, c_ip, c_ip, c_dp]
    L.ufe_b200_set_ice_velocities.argtypes = [H, C.c_int, C.c_int, C.POINTER(IceVelocities)]
    L.ufe_b200_set_state_resident.argtypes = [H, C.c_int]
    L.ufe_b200_pc_reference_loop.argtypes = [H, C.c_int, C.c_double]
    L.ufe_b200_pc_set_model_thickness.argtypes = [H, c_dp]
    L.ufe_b200_alter_resident.argtypes = [H, C.POINTER(AlterParams), C.c_double, c_dp, c_dp, c_dp]
    L.ufe_b200_pc_download.argtypes = [H, c_dp, c_dp, c_dp, c_dp]
    PC = C.POINTER(Config)
    L.ufe_b200_determine_masks.argtypes = [H, PC, c_dp, c_dp, c_dp, C.POINTER(Masks)]
    L.ufe_b200_calc_effective_thickness.argtypes = [H, PC, c_dp, c_dp, c_dp, c_dp, c_dp]
    L.ufe_b200_calc_grounded_fractions.argtypes = [H, PC, C.c_int, C.c_int, c_dp, c_dp, c_dp, c_dp, c_ip, c_dp, c_dp, c_dp, c_dp]
    L.ufe_b200_calc_critical_timestep_adv.argtypes = [H, PC, C.c_int, c_dp, c_ip, c_dp, c_dp, C.POINTER(C.c_double)]
    L.ufe_b200_calc_critical_timestep_SIA.argtypes = [H, PC, c_dp, c_dp, C.POINTER(C.c_double)]
    L.ufe_b200_geometry_resident.argtypes = [H, C.c_int, C.c_int, c_dp, c_dp, c_dp]
    L.ufe_b200_alter_params_defaults.restype = None
    L.ufe_b200_alter_params_defaults.argtypes = [C.POINTER(AlterParams)]
    L.ufe_b200_alter_ice_thickness.argtypes = [H, PC, C.POINTER(AlterParams), C.c_double, C.POINTER(Masks), c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp]
    L.ufe_b200_calc_and_apply_spill_over_flux.argtypes = [H, C.POINTER(Masks), c_dp, c_dp, c_dp, C.c_double, c_dp, c_dp]
    L.ufe_b200_calc_vertical_velocities.argtypes = [H, C.POINTER(Config), C.POINTER(VerticalVelocityIO)]
    L.ufe_b200_timing_enable.argtypes = [H, C.c_int]
    L.ufe_b200_timing_get.argtypes = [H, C.c_char_p, c_dp, C.POINTER(C.c_longlong)]
    L.ufe_b200_timing_reset.argtypes = [H]
    L.ufe_b200_timer_start.argtypes = [H]
    L.ufe_b200_l2_flush.argtypes = [H]
    L.ufe_b200_timer_stop.argtypes = [H, c_dp]
    L.ufe_b200_host_register.argtypes = [C.c_void_p, C.c_size_t]
    L.ufe_b200_host_unregister.argtypes = [C.c_void_p]
    _LIB = L
    return L


ABI_SYMBOLS = [
    "ufe_b200_config_defaults", "ufe_b200_get_unique_id", "ufe_b200_init", "ufe_b200_finalize", "ufe_b200_last_error",
    "ufe_b200_kernel_launches", "ufe_b200_ksp_solve", "ufe_b200_csr_register", "ufe_b200_csr_update_values", "ufe_b200_spmv",
    "ufe_b200_csr_release", "ufe_b200_mesh_create", "ufe_b200_mesh_destroy", "ufe_b200_mesh_ranges",
    "ufe_b200_calc_matrix_operators", "ufe_b200_set_operator", "ufe_b200_get_operator_size", "ufe_b200_get_operator",
    "ufe_b200_apply_operator", "ufe_b200_solve_SSA_DIVA_linearised", "ufe_b200_get_stiffness_size", "ufe_b200_get_stiffness", "ufe_b200_get_stiffness_storage_bytes",
    "ufe_b200_set_mg_params", "ufe_b200_get_mg_info", "ufe_b200_mg_host_transfer", "ufe_b200_mg_host_transfer_get", "ufe_b200_mg_host_transfer_parts", "ufe_b200_mg_host_transfer_parts_get",
    "ufe_b200_solve_DIVA", "ufe_b200_solve_SSA", "ufe_b200_velocity_upload", "ufe_b200_velocity_solve_resident",
    "ufe_b200_velocity_download", "ufe_b200_thickness_upload", "ufe_b200_pc_step_resident", "ufe_b200_pc_step",
    "ufe_b200_thickness_download", "ufe_b200_calc_dHi_dt", "ufe_b200_dHi_dt_finish", "ufe_b200_solve_SIA", "ufe_b200_calc_pc_truncation_error", "ufe_b200_calc_pc_truncation_error_noalloc", "ufe_b200_apply_grounded_fractions_to_bed_roughness", "ufe_b200_set_ice_velocities", "ufe_b200_remap_velocity_solver", "ufe_b200_calc_secondary_velocities", "ufe_b200_calc_vertical_velocities", "ufe_b200_set_state_resident", "ufe_b200_pc_reference_loop", "ufe_b200_set_mg_option", "ufe_b200_pc_set_model_thickness", "ufe_b200_alter_resident", "ufe_b200_pc_download", "ufe_b200_determine_masks", "ufe_b200_calc_effective_thickness",
               "ufe_b200_calc_grounded_fractions", "ufe_b200_calc_critical_timestep_adv", "ufe_b200_calc_critical_timestep_SIA", "ufe_b200_alter_params_defaults", "ufe_b200_geometry_resident", "ufe_b200_alter_ice_thickness",
               "ufe_b200_calc_and_apply_spill_over_flux",
    "ufe_b200_timing_enable", "ufe_b200_timing_get", "ufe_b200_timing_reset", "ufe_b200_timer_start", "ufe_b200_timer_stop", "ufe_b200_l2_flush",
    "ufe_b200_host_register", "ufe_b200_host_unregister"]


NOALLOC_MAX_VERTICES = 262144   # UFE_B200_NOALLOC_MAX_VERTICES of include/ufe_b200.h


class UfeError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise UfeError(f"ufe_b200 error {rc}: {lib().ufe_b200_last_error().decode(errors='replace')}")


class Context:
    """One rank <-> one GPU (ufe_b200_init / ufe_b200_finalize)."""

    def __init__(self, device=-1, rank=0, nranks=1, unique_id=None):
        L = lib()
        self.h = C.c_void_p()
        buf = None
        if unique_id is not None:
            buf = C.create_string_buffer(bytes(unique_id), 128)
        check(L.ufe_b200_init(device, rank, nranks, buf, C.byref(self.h)))
        self.rank, self.nranks = rank, nranks
        self._keep = []

    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(128)
        check(lib().ufe_b200_get_unique_id(buf))
        return buf.raw

    def close(self):
        if self.h:
            lib().ufe_b200_finalize(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- narrow seam
    def ksp_solve(self, m_glob, i1, ptr, ind, val, b, rtol, abstol, maxit=2000, ksp=KSP_BICGSTAB, pc=PC_JACOBI):
        m_loc = len(ptr) - 1
        x = np.zeros(m_loc)
        its = C.c_int()
        check(lib().ufe_b200_ksp_solve(self.h, m_glob, m_loc, i1, ip(ptr), ip(ind), dp(val), dp(b), dp(x), rtol, abstol, maxit,
                                       ksp, pc, C.byref(its)))
        return x, its.value

    def csr_register(self, id_, m_glob, n_glob, i1, n_loc, j1, ptr, ind, val):
        check(lib().ufe_b200_csr_register(self.h, id_, m_glob, n_glob, len(ptr) - 1, i1, n_loc, j1, ip(ptr), ip(ind), dp(val)))

    def spmv(self, id_, x, m_loc, ncols=1):
        y = np.zeros((m_loc, ncols), order="F") if ncols > 1 else np.zeros(m_loc)
        check(lib().ufe_b200_spmv(self.h, id_, dp(x), dp(y), ncols))
        return y

    # -- mesh
    def mesh_create(self, m):
        self._keep.append(m)
        d = mesh_desc(m)
        check(lib().ufe_b200_mesh_create(self.h, C.byref(d)))
        r = np.zeros(6, dtype=np.int32)
        check(lib().ufe_b200_mesh_ranges(self.h, ip(r)))
        self.vi1, self.vi2, self.ti1, self.ti2, self.ei1, self.ei2 = [int(v) for v in r]
        self.nz = m.nz

    def calc_matrix_operators(self):
        check(lib().ufe_b200_calc_matrix_operators(self.h))

    def get_operator(self, op):
        m, nnz = C.c_int(), C.c_int()
        check(lib().ufe_b200_get_operator_size(self.h, op, C.byref(m), C.byref(nnz)))
        ptr = np.zeros(m.value + 1, np.int32); ind = np.zeros(nnz.value, np.int32); val = np.zeros(nnz.value)
        check(lib().ufe_b200_get_operator(self.h, op, ip(ptr), ip(ind), dp(val)))
        return ptr, ind, val

    def apply_operator(self, op, x, m_loc, ncols=1):
        y = np.zeros((m_loc, ncols), order="F") if ncols > 1 else np.zeros(m_loc)
        check(lib().ufe_b200_apply_operator(self.h, op, dp(x), dp(y), ncols))
        return y

    def get_stiffness(self):
        m, nnz = C.c_int(), C.c_int()
        check(lib().ufe_b200_get_stiffness_size(self.h, C.byref(m), C.byref(nnz)))
        ptr = np.zeros(m.value + 1, np.int32); ind = np.zeros(nnz.value, np.int32); val = np.zeros(nnz.value); bb = np.zeros(m.value)
        check(lib().ufe_b200_get_stiffness(self.h, ip(ptr), ip(ind), dp(val), dp(bb)))
        return ptr, ind, val, bb

    def set_mg_params(self, nu_pre=-1, nu_post=-1, omega=-1.0, overcorr=-1.0, w_levels=-1, fp32=-1, dense_max=-1):
        check(lib().ufe_b200_set_mg_params(self.h, nu_pre, nu_post, omega, overcorr, w_levels, fp32, dense_max))

    def set_mg_option(self, name, value):
        check(lib().ufe_b200_set_mg_option(self.h, name.encode(), float(value)))

    def mg_info(self):
        nl = C.c_int(); nodes = np.zeros(16, np.int32); fb = C.c_longlong(); cb = C.c_longlong(); ss = C.c_double()
        check(lib().ufe_b200_get_mg_info(self.h, C.byref(nl), ip(nodes), C.byref(fb), C.byref(cb), C.cast(C.byref(ss), c_dp)))
        return dict(levels=nl.value, nodes=[int(x) for x in nodes[:nl.value]], fine_bytes=fb.value, coarse_bytes=cb.value, setup_s=ss.value)

    def timing(self, cls):
        ms, n = C.c_double(), C.c_longlong()
        check(lib().ufe_b200_timing_get(self.h, cls.encode(), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def kernel_launches(self):
        return int(lib().ufe_b200_kernel_launches(self.h))


def alloc_velocity_state(n_a, n_b, nz, fields=None):
    """numpy-backed VelocityState; returns (struct, dict of arrays)."""
    st = VelocityState()
    arrs = {}
    for n in VELOCITY_FIELDS:
        if fields is not None and n not in fields:
            continue
        grid, is3d = VELOCITY_SHAPES[n]
        nloc = n_a if grid == "a" else n_b
        a = np.zeros((nloc, nz), order="F") if is3d else np.zeros(nloc)
        arrs[n] = a
        setattr(st, n, dp(a))
    return st, arrs
