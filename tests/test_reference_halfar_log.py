"""The reference's time loop on the Halfar dome, restated around the oracle and held against the reference's own output: the checksum log of
integrated_test_Halfar_dome_small (automated_testing/UFEMISM/integrated_test_Halfar_dome_small/reference/checksum_logfile_Halfar_40km_grid.txt; 100 yr,
SIA, semi-implicit thickness steps, predictor-corrector time stepping), whose per-step statistics of `region%ice%pc%Hi_np1` and `pc%tau_np1` are committed
in tests/golden/halfar_40km.json (made by tests/golden/make_golden.py).

What is restated here is host logic the Fortran driver keeps: run_model_region's action times (UFEMISM_main_model.f90:95-160,354-435),
run_ice_dynamics_model's interpolation between Hi_prev and Hi_next (ice_dynamics_main.f90:85-151) and run_ice_dynamics_model_pc with its time-step rule
and accept / retry decision (predictor_corrector_scheme.f90:34-325).  The path's own work -- solve_SIA, calc_dHi_dt, calc_effective_thickness,
determine_masks, calc_critical_timestep_adv, calc_pc_truncation_error -- is the oracle's.

Two things the log shows about the reference's loop, both followed here:
  * the predictor evaluates calc_dHi_dt on `region%ice%Hi`, which is the thickness interpolated at the previous action time, not Hi_prev
    (the log's step-2 predictor line repeats the step-1 corrector line digit for digit);
  * `pc%dHi_dt_Hi_nm1_u_nm1` is `region%ice%dHi_dt`, i.e. the applied rate (Hi_next - Hi_prev) / dt of the previous step (ice_dynamics_main.f90:140),
    not the previous predictor rate.
The mesh is ours (uniform 40 km; the reference's is 40 km at the ice front and coarser inside, 703 vertices), so the comparison is statistical.
"""
import ctypes as C
import json
import os

import numpy as np

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl
from test_sia import _thickness_io, vertical_average

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "halfar_40km.json")
# config_Halfar_40km.cfg
PC_EPSILON, PC_K_I, PC_K_P, PC_ETA_MIN, PC_MAX_INCREASE, PC_NIT_MAX, PC_GUILTY_MAX = 0.005, 0.2, 0.2, 1e-8, 1.1, 50, 0.0
DT_MIN, DT_MAX, T_END, DT_THERMO = 0.1, 10.0, 100.0, 1.0


def halfar_40km_workload():
    w = wl.halfar(4.0e4, half_width=1000e3)
    f, c = w.f, w.cfg
    f["Hi"][f["Hi"] < 2.0] = 0.0                                        # refgeo_Hi_min_config = 2.0
    f["SL"][:] = -10.0                                                  # fixed_sealevel_config
    f["Hs"] = wl.ice_surface_elevation(f["Hi"], f["Hb"], f["SL"])
    c.choice_ice_integration_method = ub.INTEGRATION_SEMI_IMPLICIT; c.dHi_semiimplicit_fs = 1.5
    c.dHi_PETSc_rtol = 1e-8; c.dHi_PETSc_abstol = 1e-6
    c.dt_ice_min = DT_MIN; c.dt_ice_max = DT_MAX; c.Hi_min = 0.0
    for k in range(4):
        c.BC_H[k] = ub.BC_ZERO
    return w


class HalfarRun:
    """state of region%ice / region%ice%pc that the loop carries"""

    def __init__(self, w):
        self.w, self.m, self.o = w, w.mesh, ol.OracleMesh(w.mesh)
        nV, nTri = self.m.nV, self.m.nTri
        self.zeta = np.asarray(self.m.zeta, float)
        self.Hi = w.f["Hi"].copy(); self.Hi_prev = self.Hi.copy(); self.Hi_next = self.Hi.copy()
        self.t_prev = self.t_next = 0.0
        self.dHi_dt = np.zeros(nV)
        self.u, self.v = np.zeros(nTri), np.zeros(nTri)                 # no initial velocities: the log's first divQ is zero
        self.dt_n = self.dt_np1 = DT_MIN                                # initialise_pc_scheme, 'zero' (predictor_corrector_scheme.f90:421-434)
        self.eta_n = self.eta_np1 = PC_EPSILON
        self.guilty = np.zeros(nV, np.int32)
        self.frac_gr = np.ones(nV)                                      # Hb = 0, SL = -10 m: everything is grounded (the log: fraction_gr = 1 throughout)
        self.wide_seam = False                                          # True: the semantics of ufe_b200_pc_step_resident instead
        self.d_n_last = np.zeros(nV)
        self.log = []                                                   # per PC iteration: (t, dt, max Hi_np1, max tau, eta, max u_vav_b)

    # hooks for experiments with floating ice and sliding (tests/test_reference_mismip_mod_log.py); the Halfar dome needs none of them
    def begin_action(self):
        """what run_model_region does at every action time before the ice dynamics (UFEMISM_main_model.f90:118: run_basal_hydrology_model)"""

    def update_fractions(self, Hi, mk):
        """calc_grounded_fractions for the geometry the thickness step and the truncation error see"""

    def truncation_error(self, z, Hi_np1, Hi_star, mk):
        tau = np.zeros(self.m.nV); eta = C.c_double(); nt, ng = C.c_int(), C.c_int()
        ol.lib().orc_calc_pc_truncation_error(self.m.nV, z, self.dt_n, PC_ETA_MIN, PC_EPSILON, ub.dp(Hi_np1), ub.dp(Hi_star), ub.ip(mk["mask_grounded_ice"]),
                                              ub.ip(mk["mask_gl_gr"]), ub.dp(self.frac_gr), ub.dp(tau), ub.ip(self.guilty), C.byref(eta), C.byref(nt), C.byref(ng))
        return tau, eta.value, nt.value, ng.value

    def calc_dHi_dt(self, Hi, fraction_margin, dt):
        w, f = self.w, self.w.f
        f["fraction_margin"] = fraction_margin
        out = {k: np.zeros(self.m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
        io, keep = _thickness_io(w, Hi, self.u, self.v, out)
        rc, d, its = self.o.calc_dHi_dt(w.cfg, io, dt)
        assert rc == 0 and d == dt
        return out["dHi_dt"]

    def solve_stress_balance(self, Hi):
        f = self.w.f
        Hs = wl.ice_surface_elevation(Hi, f["Hb"], f["SL"])
        r = self.o.solve_SIA(self.w.cfg, Hi, Hs, f["A_flow"])           # SIA_maximum_diffusivity_config = 1e5 is the binding's default
        self.u = np.ascontiguousarray(vertical_average(self.zeta, r["u_3D_b"])); self.v = np.ascontiguousarray(vertical_average(self.zeta, r["v_3D_b"]))

    def pc(self):
        """run_ice_dynamics_model_pc (predictor_corrector_scheme.f90:34-325); alter_ice_thickness only removes negative thickness in this configuration"""
        o, cfg, f = self.o, self.w.cfg, self.w.f
        Hb, SL = f["Hb"], f["SL"]
        self.t_prev, self.Hi_prev = self.t_next, self.Hi_next.copy()
        if self.wide_seam:
            self.Hi = self.Hi_prev.copy()                                # one thickness for region%ice%Hi and Hi_prev
        self.dt_n = self.dt_np1
        dt = (PC_EPSILON / self.eta_np1) ** (PC_K_I + PC_K_P) * (PC_EPSILON / self.eta_n) ** (-PC_K_P) * self.dt_n      # Eq. 33
        dt = max(min(dt, DT_MAX, PC_MAX_INCREASE * self.dt_n), DT_MIN)
        mk = o.determine_masks(cfg, self.Hi, Hb, SL)
        dt = min(dt, o.calc_critical_timestep_adv(cfg, 0, self.Hi, mk["mask_floating_ice"], self.u, self.v))
        self.update_fractions(self.Hi, mk)
        fm = o.calc_effective_thickness(cfg, self.Hi, Hb, SL)[1]
        d_nm1 = np.where(f["mask_noice"] == 1, 0.0, self.d_n_last if self.wide_seam else self.dHi_dt)
        self.Hi_prev[f["mask_noice"] == 1] = 0.0
        self.eta_n = self.eta_np1
        for _ in range(PC_NIT_MAX):
            z = dt / self.dt_n
            d_n = np.where(f["mask_noice"] == 1, 0.0, self.calc_dHi_dt(self.Hi_prev if self.wide_seam else self.Hi, fm, dt))              # on region%ice%Hi, see the module docstring
            Hi_star = np.maximum(0.0, self.Hi_prev + dt * ((1 + z / 2) * d_n - (z / 2) * d_nm1))      # Eq. 30
            d_n = ((Hi_star - self.Hi_prev) / dt + (z / 2) * d_nm1) / (1 + z / 2)
            self.solve_stress_balance(Hi_star)
            mk = o.determine_masks(cfg, self.Hi_prev, Hb, SL)
            self.update_fractions(self.Hi_prev, mk)
            fm_c = o.calc_effective_thickness(cfg, self.Hi_prev, Hb, SL)[1]
            d_s = np.where(f["mask_noice"] == 1, 0.0, self.calc_dHi_dt(self.Hi_prev, fm_c, dt))
            Hi_np1 = np.maximum(0.0, self.Hi_prev + (dt / 2) * (d_n + d_s))                           # Eq. 31
            self.Hi = self.Hi_prev.copy()
            tau, eta, nt, ng = self.truncation_error(z, Hi_np1, Hi_star, mk)
            self.eta_np1 = eta
            self.log.append((self.t_prev, dt, Hi_np1.max(), tau.max(), eta, self.u.max()))
            if self.eta_np1 < PC_EPSILON:
                break
            if 100.0 * ng / max(nt, 1) < PC_GUILTY_MAX:
                self.eta_np1 = 0.95 * PC_EPSILON
                break
            dt *= 0.8
            if dt <= DT_MIN:
                dt = DT_MIN
                break
        self.dt_np1 = dt
        self.d_n_last = d_n
        self.t_next, self.Hi_next = self.t_prev + dt, Hi_np1

    def run(self, t_end=T_END):
        time, t_thermo = 0.0, 0.0
        while time <= t_end:
            self.begin_action()
            if time == self.t_next:
                self.pc()
            wt_prev = (self.t_next - time) / (self.t_next - self.t_prev)
            self.Hi = wt_prev * self.Hi_prev + (1.0 - wt_prev) * self.Hi_next
            self.dHi_dt = (self.Hi_next - self.Hi_prev) / (self.t_next - self.t_prev)
            if time == t_thermo:                                                                     # run_thermodynamics_model: an action every dt_thermodynamics
                t_thermo += DT_THERMO
            if time == t_end:
                break
            time = min(t_end, self.t_next, t_thermo)
        return self.Hi


def test_oracle_time_loop_follows_the_reference_halfar_log():
    g = json.load(open(GOLDEN))
    ref_np1, ref_tau = np.array(g["Hi_np1"]), np.array(g["tau_np1"])      # rows: sum, sum|.|, min, max
    assert len(ref_np1) == len(ref_tau) == 58
    run = HalfarRun(halfar_40km_workload())
    Hi = run.run()
    log = np.array(run.log)
    t, dt, hmax, taumax = log[:, 0], log[:, 1], log[:, 2], log[:, 3]
    # the start-up: two steps at dt_ice_min whose truncation error is far above pc_epsilon, then the error collapses (the log: 0.387, 0.194, 1e-4)
    assert dt[0] == dt[1] == DT_MIN
    assert abs(taumax[0] / ref_tau[0, 3] - 1.0) < 0.25 and abs(taumax[1] / ref_tau[1, 3] - 1.0) < 0.25, (taumax[:3], ref_tau[:3, 3])
    assert taumax[2] < 0.01 * taumax[1]
    # the number of predictor-corrector iterations it takes to cover the 100 years
    assert abs(len(log) - len(ref_tau)) <= 6, len(log)
    # from then on the time step grows by pc_max_time_step_increase until the error reaches the tolerance, and the error follows the log's, iteration
    # by iteration -- including the alternation that sets in once dt is limited by the error (0.0035 / 0.0042 / 0.0034 / 0.0043 ...)
    assert np.all(dt[3:20] / dt[2:19] <= PC_MAX_INCREASE + 1e-12) and dt.max() > 2.0
    ratio = taumax[2:55] / ref_tau[2:55, 3]
    assert 0.6 < ratio.min() and ratio.max() < 1.3 and abs(np.median(ratio) - 1.0) < 0.15, (ratio.min(), ratio.max(), np.median(ratio))
    late, late_ref = taumax[38:54], ref_tau[38:54, 3]
    assert (np.sign(np.diff(late)) == np.sign(np.diff(late_ref))).mean() > 0.8
    assert taumax[2:].max() < 2.0 * PC_EPSILON and ref_tau[2:, 3].max() < PC_EPSILON
    # the summit: 2994 m -> 2899 m in the log; ours starts at 2987 m (where the nearest vertex happens to sit); compare the lowering over the same iterations
    drop_ref = ref_np1[0, 3] - ref_np1[54, 3]
    drop = hmax[0] - hmax[54]
    assert abs(drop / drop_ref - 1.0) < 0.20, (drop, drop_ref)
    # and the end state against Halfar's solution at t = 100 yr
    ana = np.array([ol.lib().orc_Halfar_H(1e-16, 3.0, 3000.0, 500e3, x, y, T_END) for x, y in run.m.V[:, :2]])
    inside = ana > 500.0
    assert np.sqrt(np.mean((Hi[inside] - ana[inside]) ** 2)) < 25.0


def test_oracle_first_flux_divergence_against_the_reference_halfar_log():
    """the first thickness step that sees SIA velocities (the log's third `divQ` line: min = -4.646, max = 1.262 m/yr): thickening just inside the margin agrees
    to a few per cent; the thinning peak sits on the steepest margin triangle and follows the mesh there (ours: -3.80), like the truncation error above"""
    g = json.load(open(GOLDEN))
    ref = g["divQ"][2]
    assert g["divQ"][0][1] == 0.0 and ref[1] > 0.0                          # no velocities before the first stress balance
    w = halfar_40km_workload()
    run = HalfarRun(w)
    run.solve_stress_balance(run.Hi)
    w.f["fraction_margin"] = run.o.calc_effective_thickness(w.cfg, run.Hi, w.f["Hb"], w.f["SL"])[1]
    out = {k: np.zeros(run.m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
    io, keep = _thickness_io(w, run.Hi, run.u, run.v, out)
    rc, dt, its = run.o.calc_dHi_dt(w.cfg, io, DT_MIN)
    assert rc == 0 and dt == DT_MIN
    q, A = out["divQ"], np.asarray(run.m.A)
    assert abs(q.max() / ref[3] - 1.0) < 0.10 and abs(q.min() / ref[2] - 1.0) < 0.25, (q.min(), q.max(), ref)
    assert abs((q * A).sum()) < 1e-12 * (np.abs(q) * A).sum()               # what leaves one Voronoi cell enters its neighbour
    np.testing.assert_allclose(out["Hi_tplusdt"], np.maximum(0.0, run.Hi + DT_MIN * out["dHi_dt"]), rtol=0, atol=1e-9)


def test_wide_seam_semantics_against_the_reference_loop():
    """What the resident predictor-corrector step (ufe_b200_pc_step_resident) does differently from the reference's loop -- one thickness for `region%ice%Hi` and
    `Hi_prev`, the previous predictor rate in Eq. 30 -- and what it costs over the 100 years: the same dome to 1e-4, with a smaller
    truncation error (the reference's predictor works from a thickness one action behind), hence fewer steps"""
    ref_loop, wide = HalfarRun(halfar_40km_workload()), HalfarRun(halfar_40km_workload())
    wide.wide_seam = True
    Hi_ref, Hi_wide = ref_loop.run(), wide.run()
    assert 1e-6 < np.linalg.norm(Hi_wide - Hi_ref) / np.linalg.norm(Hi_ref) < 1e-4 and np.abs(Hi_wide - Hi_ref).max() < 0.5
    a, b = np.array(ref_loop.log), np.array(wide.log)
    assert len(b) < len(a) and np.median(b[3:30, 3]) < 0.5 * np.median(a[3:30, 3])
