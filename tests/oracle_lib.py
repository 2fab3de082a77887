"""ctypes access to oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY (the CPU restatement of the reference).

Nothing under ufemism2.0_b200/ imports this module.
"""
import ctypes as C
import os

import numpy as np

import ufe_b200 as ub

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None
KSP_BICGSTAB_JACOBI, KSP_GMRES_ILU0, KSP_GMRES_BJACOBI_ILU0 = 0, 1, 2   # 2: PETSc's default on several ranks (set_bjacobi_blocks)


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
        dp, ip = ub.c_dp, ub.c_ip
        L.orc_spmv.argtypes = [C.c_int, C.c_int, ip, ip, dp, dp, dp]
        L.orc_ksp_solve.argtypes = [C.c_int, ip, ip, dp, dp, dp, C.c_double, C.c_double, C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.orc_set_bjacobi_blocks.argtypes = [C.c_int]
        L.orc_mesh_create.restype = C.c_void_p
        L.orc_mesh_create.argtypes = [C.POINTER(ub.MeshDesc)]
        L.orc_mesh_destroy.argtypes = [C.c_void_p]
        L.orc_calc_matrix_operators.argtypes = [C.c_void_p]
        L.orc_get_operator_size.argtypes = [C.c_void_p, C.c_int] + [C.POINTER(C.c_int)] * 3
        L.orc_get_operator.argtypes = [C.c_void_p, C.c_int, ip, ip, dp]
        L.orc_apply_operator.argtypes = [C.c_void_p, C.c_int, dp, dp, C.c_int]
        L.orc_assemble_SSA_DIVA.argtypes = [C.c_void_p, C.POINTER(ub.Config)] + [dp] * 6 + [ip, dp, dp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_assemble_SSA_DIVA_prev.argtypes = [C.c_void_p, C.POINTER(ub.Config)] + [dp] * 6 + [ip, dp, dp, dp, dp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_find_ti_copy.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, ip, dp]
        L.orc_get_stiffness.argtypes = [C.c_void_p, ip, ip, dp, dp]
        L.orc_solve_velocity.argtypes = [C.c_void_p, C.POINTER(ub.Config), C.c_int, C.c_int, C.POINTER(ub.IceInputs),
                                         C.POINTER(ub.VelocityState), C.POINTER(C.c_int), C.POINTER(C.c_int), ip, dp, dp, C.c_int]
        L.orc_calc_dHi_dt.argtypes = [C.c_void_p, C.POINTER(ub.Config), C.c_int, C.POINTER(ub.ThicknessIO), dp, C.POINTER(C.c_int)]
        L.orc_flux_divergence_matrix.argtypes = [C.c_void_p, dp, dp, dp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_get_flux_divergence_matrix.argtypes = [C.c_void_p, ip, ip, dp]
        L.orc_Halfar_H.restype = C.c_double
        L.orc_Halfar_H.argtypes = [C.c_double] * 7
        L.orc_Schoof2006_icestream.argtypes = [C.c_double] * 7 + [dp, dp]
        L.orc_solve_SIA.argtypes = [C.c_void_p, C.POINTER(ub.Config), C.c_double] + [dp] * 8
        L.orc_apply_grounded_fractions_to_bed_roughness.argtypes = [C.c_int, dp, dp, dp, C.POINTER(ub.Masks), dp, dp]
        L.orc_calc_pc_truncation_error.argtypes = [C.c_int] + [C.c_double] * 4 + [dp, dp, ip, ip, dp, dp, ip, C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_set_ice_velocities.argtypes = [C.c_void_p, C.c_int, C.c_int] + [dp] * 10 + [C.POINTER(ub.IceVelocities)]
        L.orc_calc_secondary_velocities.argtypes = [C.c_void_p, dp, dp, C.POINTER(ub.SecondaryVelocities)]
        L.orc_calc_vertical_velocities.argtypes = [C.c_void_p, C.POINTER(ub.Config), C.POINTER(ub.VerticalVelocityIO), dp, dp, dp, dp]
        PC = C.POINTER(ub.Config)
        L.orc_determine_masks.argtypes = [C.c_void_p, PC, dp, dp, dp, C.POINTER(ub.Masks)]
        L.orc_calc_effective_thickness.argtypes = [C.c_void_p, PC, dp, dp, dp, dp, dp]
        L.orc_calc_grounded_fractions.argtypes = [C.c_void_p, PC, C.c_int, C.c_int, dp, dp, dp, dp, ip, dp, dp, dp, dp]
        L.orc_calc_critical_timestep_adv.restype = C.c_double
        L.orc_calc_critical_timestep_adv.argtypes = [C.c_void_p, PC, C.c_int, dp, ip, dp, dp]
        L.orc_calc_critical_timestep_SIA.restype = C.c_double
        L.orc_calc_critical_timestep_SIA.argtypes = [C.c_void_p, PC, dp, dp]
        L.orc_alter_ice_thickness.argtypes = [C.c_void_p, PC, C.POINTER(ub.AlterParams), C.c_double, C.POINTER(ub.Masks), dp, dp, dp, dp, dp, dp, dp]
        L.orc_calc_and_apply_spill_over_flux.argtypes = [C.c_void_p, C.POINTER(ub.Masks), dp, dp, dp, C.c_double, dp, dp]
        _LIB = L
    return _LIB


def set_bjacobi_blocks(n):
    """number of ILU(0) blocks (= MPI ranks of the reference run) of KSP_GMRES_BJACOBI_ILU0; 0 = one per OpenMP thread"""
    lib().orc_set_bjacobi_blocks(n)


def spmv(m, n, ptr, ind, val, x):
    y = np.zeros(m)
    lib().orc_spmv(m, n, ub.ip(ptr), ub.ip(ind), ub.dp(val), ub.dp(x), ub.dp(y))
    return y


def ksp_solve(ptr, ind, val, b, rtol, abstol, maxit=2000, ksp=KSP_BICGSTAB_JACOBI):
    m = len(ptr) - 1
    x = np.zeros(m)
    its = C.c_int()
    rc = lib().orc_ksp_solve(m, ub.ip(ptr), ub.ip(ind), ub.dp(val), ub.dp(b), ub.dp(x), rtol, abstol, maxit, ksp, C.byref(its))
    return x, its.value, rc


class OracleMesh:
    def __init__(self, m):
        self.m = m
        self._d = ub.mesh_desc(m)
        self.h = lib().orc_mesh_create(C.byref(self._d))
        self.have_ops = False

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_mesh_destroy(self.h)
            self.h = None

    def calc_matrix_operators(self):
        if not self.have_ops:
            lib().orc_calc_matrix_operators(self.h)
            self.have_ops = True

    def get_operator(self, op):
        self.calc_matrix_operators()
        m, n, nnz = C.c_int(), C.c_int(), C.c_int()
        lib().orc_get_operator_size(self.h, op, C.byref(m), C.byref(n), C.byref(nnz))
        ptr = np.zeros(m.value + 1, np.int32); ind = np.zeros(nnz.value, np.int32); val = np.zeros(nnz.value)
        lib().orc_get_operator(self.h, op, ub.ip(ptr), ub.ip(ind), ub.dp(val))
        return ptr, ind, val, m.value, n.value

    def apply_operator(self, op, x, ncols=1):
        self.calc_matrix_operators()
        m, n, nnz = C.c_int(), C.c_int(), C.c_int()
        lib().orc_get_operator_size(self.h, op, C.byref(m), C.byref(n), C.byref(nnz))
        y = np.zeros((m.value, ncols), order="F") if ncols > 1 else np.zeros(m.value)
        lib().orc_apply_operator(self.h, op, ub.dp(x), ub.dp(y), ncols)
        return y

    def find_ti_copy(self, choice, L, ti):
        """find_ti_copy_ISMIP_HOM_periodic (choice 2) / find_ti_copy_SSA_icestream_infinite (choice 3) for the 1-based triangle ti"""
        tj = np.zeros(self.m.nC_mem, np.int32); w = np.zeros(self.m.nC_mem)
        lib().orc_find_ti_copy(self.h, choice, L, ti, ub.ip(tj), ub.dp(w))
        return tj, w

    def assemble(self, cfg, N_b, dN_dx_b, dN_dy_b, beta_b, tau_dx_b, tau_dy_b, BC_mask=None, BC_u=None, BC_v=None, u_b_prev=None, v_b_prev=None):
        self.calc_matrix_operators()
        m, nnz = C.c_int(), C.c_int()
        if u_b_prev is not None:
            lib().orc_assemble_SSA_DIVA_prev(self.h, C.byref(cfg), ub.dp(N_b), ub.dp(dN_dx_b), ub.dp(dN_dy_b), ub.dp(beta_b), ub.dp(tau_dx_b),
                                             ub.dp(tau_dy_b), ub.ip(BC_mask), ub.dp(BC_u), ub.dp(BC_v), ub.dp(u_b_prev), ub.dp(v_b_prev),
                                             C.byref(m), C.byref(nnz))
        else:
            lib().orc_assemble_SSA_DIVA(self.h, C.byref(cfg), ub.dp(N_b), ub.dp(dN_dx_b), ub.dp(dN_dy_b), ub.dp(beta_b), ub.dp(tau_dx_b),
                                        ub.dp(tau_dy_b), ub.ip(BC_mask), ub.dp(BC_u), ub.dp(BC_v), C.byref(m), C.byref(nnz))
        ptr = np.zeros(m.value + 1, np.int32); ind = np.zeros(nnz.value, np.int32); val = np.zeros(nnz.value); bb = np.zeros(m.value)
        lib().orc_get_stiffness(self.h, ub.ip(ptr), ub.ip(ind), ub.dp(val), ub.dp(bb))
        return ptr, ind, val, bb

    def solve_velocity(self, cfg, approximation, ice, state, ksp=KSP_BICGSTAB_JACOBI, BC_mask=None, BC_u=None, BC_v=None, max_visc_its=-1):
        self.calc_matrix_operators()
        nv, na = C.c_int(), C.c_int()
        rc = lib().orc_solve_velocity(self.h, C.byref(cfg), approximation, ksp, C.byref(ice), C.byref(state), C.byref(nv), C.byref(na),
                                      ub.ip(BC_mask), ub.dp(BC_u), ub.dp(BC_v), max_visc_its)
        return rc, nv.value, na.value

    def calc_dHi_dt(self, cfg, io, dt, ksp=KSP_BICGSTAB_JACOBI):
        d = C.c_double(dt)
        its = C.c_int()
        rc = lib().orc_calc_dHi_dt(self.h, C.byref(cfg), ksp, C.byref(io), C.cast(C.byref(d), ub.c_dp), C.byref(its))
        return rc, d.value, its.value

    def solve_SIA(self, cfg, Hi, Hs, A_flow, max_diffusivity=1e5):
        self.calc_matrix_operators()
        m = self.m
        out = dict(u_3D_b=np.zeros((m.nTri, m.nz), order="F"), v_3D_b=np.zeros((m.nTri, m.nz), order="F"), D_3D_b=np.zeros((m.nTri, m.nz), order="F"),
                   du_dz_3D_a=np.zeros((m.nV, m.nz), order="F"), dv_dz_3D_a=np.zeros((m.nV, m.nz), order="F"))
        lib().orc_solve_SIA(self.h, C.byref(cfg), max_diffusivity, ub.dp(Hi), ub.dp(Hs), ub.dp(A_flow), *(ub.dp(out[k]) for k in
                            ("u_3D_b", "v_3D_b", "D_3D_b", "du_dz_3D_a", "dv_dz_3D_a")))
        return out

    def set_ice_velocities(self, approximation, scheme, out, u_3D_b=None, v_3D_b=None, u_b=None, v_b=None, du_dx_a=None, du_dy_a=None,
                           dv_dx_a=None, dv_dy_a=None, du_dz_3D=None, dv_dz_3D=None):
        return lib().orc_set_ice_velocities(self.h, approximation, scheme, ub.dp(u_3D_b), ub.dp(v_3D_b), ub.dp(u_b), ub.dp(v_b), ub.dp(du_dx_a),
                                            ub.dp(du_dy_a), ub.dp(dv_dx_a), ub.dp(dv_dy_a), ub.dp(du_dz_3D), ub.dp(dv_dz_3D), C.byref(out))

    def calc_secondary_velocities(self, u_3D_b, v_3D_b):
        self.calc_matrix_operators()
        s, arrs = ub.alloc_secondary(self.m.nV, self.m.nTri, self.m.nz)
        lib().orc_calc_secondary_velocities(self.h, ub.dp(u_3D_b), ub.dp(v_3D_b), C.byref(s))
        return arrs

    def calc_vertical_velocities(self, cfg, io, u_3D_b, v_3D_b, u_3D, v_3D):
        self.calc_matrix_operators()
        lib().orc_calc_vertical_velocities(self.h, C.byref(cfg), C.byref(io), ub.dp(u_3D_b), ub.dp(v_3D_b), ub.dp(u_3D), ub.dp(v_3D))

    def determine_masks(self, cfg, Hi, Hb, SL):
        s, arrs = ub.alloc_masks(self.m.nV)
        lib().orc_determine_masks(self.h, C.byref(cfg), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), C.byref(s))
        return arrs

    def calc_effective_thickness(self, cfg, Hi, Hb, SL):
        Hi_eff, fm = np.zeros(self.m.nV), np.zeros(self.m.nV)
        lib().orc_calc_effective_thickness(self.h, C.byref(cfg), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(Hi_eff), ub.dp(fm))
        return Hi_eff, fm

    def calc_grounded_fractions(self, cfg, choice, nbins, Hi, Hb, SL, dHb, mask_floating_ice, cdf, cdf_b):
        self.calc_matrix_operators()
        fr, fr_b = np.zeros(self.m.nV), np.zeros(self.m.nTri)
        rc = lib().orc_calc_grounded_fractions(self.h, C.byref(cfg), choice, nbins, ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(dHb), ub.ip(mask_floating_ice),
                                               ub.dp(cdf), ub.dp(cdf_b), ub.dp(fr), ub.dp(fr_b))
        assert rc == 0
        return fr, fr_b

    def calc_critical_timestep_adv(self, cfg, grounded_only, Hi, mask_floating_ice, u_vav_b, v_vav_b):
        return lib().orc_calc_critical_timestep_adv(self.h, C.byref(cfg), grounded_only, ub.dp(Hi), ub.ip(mask_floating_ice), ub.dp(u_vav_b), ub.dp(v_vav_b))

    def calc_critical_timestep_SIA(self, cfg, Hi, D_3D_b):
        return lib().orc_calc_critical_timestep_SIA(self.h, C.byref(cfg), ub.dp(Hi), ub.dp(D_3D_b))

    def alter_ice_thickness(self, cfg, prm, time, masks, Hi_old, Hb, SL, ref_Hi, ref_Hb, Ti_hom, Hi_new):
        out = Hi_new.copy()
        ms = ub.masks_struct(masks)
        rc = lib().orc_alter_ice_thickness(self.h, C.byref(cfg), C.byref(prm), time, C.byref(ms), ub.dp(Hi_old), ub.dp(Hb), ub.dp(SL), ub.dp(ref_Hi),
                                           ub.dp(ref_Hb), ub.dp(Ti_hom), ub.dp(out))
        assert rc == 0
        return out

    def calc_and_apply_spill_over_flux(self, masks, Hi_eff, u_vav_b, v_vav_b, dt, Hi_new):
        out, q = Hi_new.copy(), np.zeros(self.m.nV)
        ms = ub.masks_struct(masks)
        lib().orc_calc_and_apply_spill_over_flux(self.h, C.byref(ms), ub.dp(Hi_eff), ub.dp(u_vav_b), ub.dp(v_vav_b), dt, ub.dp(out), ub.dp(q))
        return out, q

    def flux_divergence_matrix(self, u_b, v_b, fraction_margin):
        m, nnz = C.c_int(), C.c_int()
        lib().orc_flux_divergence_matrix(self.h, ub.dp(u_b), ub.dp(v_b), ub.dp(fraction_margin), C.byref(m), C.byref(nnz))
        ptr = np.zeros(m.value + 1, np.int32); ind = np.zeros(nnz.value, np.int32); val = np.zeros(nnz.value)
        lib().orc_get_flux_divergence_matrix(self.h, ub.ip(ptr), ub.ip(ind), ub.dp(val))
        return ptr, ind, val
