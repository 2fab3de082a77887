"""The path BASELINE.json names -- DIVA velocity solve + ice-thickness step inside the predictor-corrector loop -- run for the reference's 100 model years
around the oracle and held against the reference's own output: the checksum log of integrated_test_MISMIP_mod_small (tests/golden/golden.json:
mismip_mod_small.time_loop, made by tests/golden/make_golden.py).  Per predictor-corrector iteration the log gives the number of viscosity iterations of
its stress-balance solve (26 for the first, 1 for each of the other 49), the converged DIVA%u_vav_b, the corrected thickness and the truncation error; and the
grounded fractions / mask counts at the start and at the end.

The loop itself (action times, interpolation, Eq. 33, accept / retry) is the host logic of tests/test_reference_halfar_log.py; this experiment adds floating
ice, sliding, the no-ice mask and the sub-grid grounded fractions.  The bedrock CDFs the reference builds from its 5 km geometry grid at initialisation
(bedrock_cumulative_density_functions.f90; not on the path) are replaced by percentiles of the analytic cone bed sampled inside each Voronoi cell / triangle.
"""
import json
import os

import numpy as np
import pytest

import oracle_lib as ol
import ufe_b200 as ub
import workloads as wl
from test_reference_halfar_log import HalfarRun
from test_reference_logs import mismip_mod_small_workload

NBINS = 11


def cone_bed(x, y):
    return 150.0 - 400.0 * np.hypot(x, y) / 750e3            # idealised_geometries.f90:214-241


def bedrock_cdfs(m):
    """percentiles 0, 10, ..., 100 % of the bed inside each Voronoi cell (a disc of the cell's area) and each triangle"""
    q = np.linspace(0.0, 100.0, NBINS)
    rr, th = np.meshgrid(np.sqrt((np.arange(12) + 0.5) / 12.0), 2.0 * np.pi * np.arange(24) / 24.0)
    rc = np.sqrt(np.asarray(m.A) / np.pi)
    px = m.V[:, 0][:, None] + rc[:, None] * (rr * np.cos(th)).ravel()[None, :]; py = m.V[:, 1][:, None] + rc[:, None] * (rr * np.sin(th)).ravel()[None, :]
    cdf = np.percentile(cone_bed(px, py), q, axis=1).T
    a, b = np.meshgrid(np.arange(12) + 0.5, np.arange(12) + 0.5)
    keep = (a + b) < 12
    l1, l2 = a[keep] / 12.0, b[keep] / 12.0
    T = m.V[np.asarray(m.Tri) - 1]                                     # (nTri, 3, 2)
    P = T[:, 0, None, :] * (1 - l1 - l2)[None, :, None] + T[:, 1, None, :] * l1[None, :, None] + T[:, 2, None, :] * l2[None, :, None]
    cdf_b = np.percentile(cone_bed(P[..., 0], P[..., 1]), q, axis=1).T
    return np.asfortranarray(cdf), np.asfortranarray(cdf_b)


class MismipModRun(HalfarRun):
    def __init__(self, w, ksp=ol.KSP_BICGSTAB_JACOBI):
        super().__init__(w)
        self.ksp = ksp
        m = self.m
        self.cdf, self.cdf_b = bedrock_cdfs(m)
        self.dHb = np.zeros(m.nV)
        self.st, self.state = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        self.n_visc = []
        self.first = {}
        mk = self.o.determine_masks(w.cfg, self.Hi, w.f["Hb"], w.f["SL"])
        self.update_fractions(self.Hi, mk)

    def begin_action(self):
        """calc_pore_water_pressure_Martin2011 + the effective pressure (basal_hydrology_main.f90:227-266) for region%ice%Hi"""
        f = self.w.f
        lam = np.clip(1.0 - (f["Hb"] - f["SL"]) / 1000.0, 0.0, 1.0)
        f["effective_pressure"] = np.maximum(0.0, wl.ICE_DENSITY * wl.GRAV * self.Hi * (1.0 - 0.96 * lam))

    def grounded_fractions(self, Hi, mk):
        f = self.w.f
        return self.o.calc_grounded_fractions(self.w.cfg, ub.GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF, NBINS, Hi, f["Hb"], f["SL"], self.dHb, mk["mask_floating_ice"],
                                              self.cdf, self.cdf_b)

    def update_fractions(self, Hi, mk):
        self.frac_gr, self.frac_gr_b = self.grounded_fractions(Hi, mk)
        self.masks = mk
        if not self.first:
            self.first = dict(fraction_gr=self.frac_gr.sum(), fraction_gr_b=self.frac_gr_b.sum(), grounded=int(mk["mask_grounded_ice"].sum()),
                              floating=int(mk["mask_floating_ice"].sum()))

    def solve_stress_balance(self, Hi):
        """solve_stress_balance on the predicted geometry (predictor_corrector_scheme.f90:160-186): masks and grounded fractions of Hi_star first"""
        w, f = self.w, self.w.f
        mk = self.o.determine_masks(w.cfg, Hi, f["Hb"], f["SL"])
        f["mask_grounded_ice"], f["mask_icefree_land"] = mk["mask_grounded_ice"], mk["mask_icefree_land"]
        f["fraction_gr_b"] = self.grounded_fractions(Hi, mk)[1]
        f["Hi"], f["Hs"] = Hi, wl.ice_surface_elevation(Hi, f["Hb"], f["SL"])
        rc, nv, na = self.o.solve_velocity(w.cfg, 1, w.ice_inputs(), self.st, **({} if self.ksp == ol.KSP_BICGSTAB_JACOBI else {"ksp": self.ksp}))
        assert rc == 0
        self.n_visc.append(nv)
        self.u, self.v = self.state["u_vav_b"].copy(), self.state["v_vav_b"].copy()


@pytest.mark.parametrize("ksp", [ol.KSP_BICGSTAB_JACOBI, ol.KSP_GMRES_ILU0])      # the device's Krylov method, and PETSc's default (GMRES(30) + ILU(0))
def test_oracle_diva_time_loop_follows_the_reference_mismip_mod_log(ksp):
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.json")))["mismip_mod_small"]
    T = g["time_loop"]
    ref_u, ref_H, ref_nv = np.array(T["u_vav_b"]), np.array(T["Hi_np1"]), T["n_visc_its"]
    w = mismip_mod_small_workload()
    w.f["mask_noice"] = wl.mask_noice(w.mesh, "MISMIP_mod")
    run = MismipModRun(w, ksp)
    # the initial sub-grid state.  The log's mesh is refined at the grounding line (220 grounded of 1022 vertices), ours is uniform, so counts do not compare;
    # the log's relation does (sum of fraction_gr = 218.3 for 220 grounded vertices: the grounding line cuts its cells about in half), and so does the area:
    # a 100 m slab floats where Hb < -100 rho_i / rho_w, i.e. beyond r = 448.5 km on the cone
    assert abs(T["fraction_gr"][0] / T["mask_grounded_ice"][0] - 1.0) < 0.02 and abs(run.first["fraction_gr"] / run.first["grounded"] - 1.0) < 0.02, run.first
    r_gl = (150.0 + 100.0 * wl.ICE_DENSITY / wl.SEAWATER_DENSITY) * 750e3 / 400.0
    A, Tri = np.asarray(run.m.A), np.asarray(run.m.Tri) - 1
    assert abs((run.frac_gr * A).sum() / (np.pi * r_gl ** 2) - 1.0) < 0.02
    V = run.m.V[Tri]
    A_b = 0.5 * np.abs((V[:, 1, 0] - V[:, 0, 0]) * (V[:, 2, 1] - V[:, 0, 1]) - (V[:, 2, 0] - V[:, 0, 0]) * (V[:, 1, 1] - V[:, 0, 1]))
    assert abs((run.frac_gr_b * A_b).sum() / (np.pi * r_gl ** 2) - 1.0) < 0.02
    run.run()
    log = np.array(run.log)
    dt, hmax, umax = log[:, 1], log[:, 2], log[:, 5]
    # 50 predictor-corrector iterations cover the 100 years, as in the log: the truncation error over the fully grounded interior never limits the step, so
    # dt = dt_ice_min twice and then grows by pc_max_time_step_increase every step
    assert len(log) == len(ref_nv) == 50
    np.testing.assert_allclose(dt, np.r_[0.1, 0.1 * 1.1 ** np.arange(49)], rtol=1e-12)
    assert log[:, 4].max() < 0.05 and log[1:, 4].max() < 0.005
    # viscosity iterations per stress-balance solve: 26 for the first one, then the warm start converges at once
    assert ref_nv[0] == 26 and set(ref_nv[1:]) == {1}
    assert abs(run.n_visc[0] - 26) <= 2 and max(run.n_visc[1:]) <= 4 and sum(1 for n in run.n_visc[1:] if n == 1) >= 45, run.n_visc
    # the thickest ice sits on the flat interior and follows SMB = 0.3 m/yr: identical to the log wherever the maximum is an interior vertex
    assert np.abs(hmax - ref_H[:, 3]).max() < 0.2 and (np.abs(hmax - ref_H[:, 3]) < 1e-3).sum() >= 40
    assert abs(hmax[-1] - ref_H[-1, 3]) < 1e-3                                       # 131.746 m after 105.8 yr
    # the shelf speeds up as the ice thickens (719 -> 841 m/yr in the log).  Our level is 9 % higher throughout (uniform 69 km mesh against the log's 40 km at the
    # grounding line); its growth follows the log's within 2.5 % at every one of the 50 solves
    level = umax / ref_u[:, 3]
    growth = (umax / umax[0]) / (ref_u[:, 3] / ref_u[0, 3])
    assert np.abs(level - 1.0).max() < 0.12 and np.abs(growth - 1.0).max() < 0.025, (level.min(), level.max(), growth.min(), growth.max())
    # and the grounding line advances (220 -> 244 grounded vertices in the log)
    assert T["mask_grounded_ice"][1] > T["mask_grounded_ice"][0] and T["mask_floating_ice"][1] < T["mask_floating_ice"][0]
    assert run.masks["mask_grounded_ice"].sum() > run.first["grounded"] and run.masks["mask_floating_ice"].sum() < run.first["floating"]
    assert 1.02 < run.frac_gr_b.sum() / run.first["fraction_gr_b"] < 1.15


def test_device_backend_of_the_time_loops_covers_every_call():
    """scripts/gpu_time_loop.py swaps the loops' backend for the C ABI: every method the loops call on it must exist there, with the oracle binding's arguments"""
    import ast
    import importlib.util
    import inspect
    import oracle_lib as ol
    here = os.path.dirname(os.path.abspath(__file__))
    called = set()
    for name in ("test_reference_halfar_log.py", "test_reference_mismip_mod_log.py"):
        for n in ast.walk(ast.parse(open(os.path.join(here, name)).read())):
            if isinstance(n, ast.Call) and isinstance(n.func, ast.Attribute) and isinstance(n.func.value, (ast.Name, ast.Attribute)):
                v = n.func.value
                if (isinstance(v, ast.Name) and v.id == "o") or (isinstance(v, ast.Attribute) and v.attr == "o"):
                    called.add(n.func.attr)
    assert {"determine_masks", "calc_effective_thickness", "calc_grounded_fractions", "calc_critical_timestep_adv", "calc_dHi_dt", "solve_SIA", "solve_velocity"} <= called
    spec = importlib.util.spec_from_file_location("gpu_time_loop", os.path.join(os.path.dirname(here), "scripts", "gpu_time_loop.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    for name in sorted(called):
        dev, orc = getattr(mod.DeviceBackend, name), getattr(ol.OracleMesh, name)
        a, b = list(inspect.signature(dev).parameters), list(inspect.signature(orc).parameters)
        assert a == b[:len(a)] and all(inspect.signature(orc).parameters[p].default is not inspect.Parameter.empty for p in b[len(a):]), (name, a, b)


def test_one_call_step_model_reproduces_the_loop():
    """CPU model of what timestep.cu does per call with ufe_b200_pc_reference_loop switched on (lagged thickness for the predictor, applied rate for the next
    Eq. 30, the interpolation weight of the last action time), driven the way scripts/gpu_time_loop.py::one_call_mismip_mod drives the device: one call per
    predictor-corrector step with the known time-step sequence.  It must land on the loop model's end state; without the switch it does not."""
    def workload():
        w = mismip_mod_small_workload()
        w.f["mask_noice"] = wl.mask_noice(w.mesh, "MISMIP_mod")
        return w
    loop = MismipModRun(workload())
    loop.run()
    log = np.array(loop.log)

    def one_call_run(reference_loop):
        r = MismipModRun(workload())
        f, cfg, o = r.w.f, r.w.cfg, r.o
        Hb, SL, noice = f["Hb"], f["SL"], f["mask_noice"] == 1
        Hi_n, Hi_lag, d_nm1, dt_n = r.Hi.copy(), r.Hi.copy(), np.zeros(r.m.nV), None
        for t, dt in log[:, :2]:
            last = max(t, np.ceil(t + dt) - 1.0)
            wgt = (last - t) / dt if reference_loop else 0.0
            r.Hi = Hi_lag; r.begin_action()                                    # the host: effective pressure of the thickness at the last action time
            z = dt / (dt_n if dt_n else dt)
            H_pred = Hi_lag if reference_loop else Hi_n
            d_n = np.where(noice, 0.0, r.calc_dHi_dt(H_pred, o.calc_effective_thickness(cfg, H_pred, Hb, SL)[1], dt))
            Hi_star = np.maximum(0.0, Hi_n + dt * ((1 + z / 2) * d_n - (z / 2) * d_nm1))
            r.solve_stress_balance(Hi_star)
            d_s = np.where(noice, 0.0, r.calc_dHi_dt(Hi_n, o.calc_effective_thickness(cfg, Hi_n, Hb, SL)[1], dt))
            h = np.maximum(0.0, Hi_n + (dt / 2) * (d_n + d_s))
            d_nm1 = (h - Hi_n) / dt if reference_loop else d_n
            Hi_lag = (1.0 - wgt) * Hi_n + wgt * h
            Hi_n, dt_n = h, dt
        return Hi_n, r.u.max()

    Hi_ref, u_ref = one_call_run(True)
    rel = np.linalg.norm(Hi_ref - loop.Hi_next) / np.linalg.norm(loop.Hi_next)
    assert rel < 1e-6 and abs(u_ref / loop.u.max() - 1.0) < 1e-5, (rel, u_ref, loop.u.max())
    Hi_plain, u_plain = one_call_run(False)
    rel_plain = np.linalg.norm(Hi_plain - loop.Hi_next) / np.linalg.norm(loop.Hi_next)
    assert 1e-3 < rel_plain < 2e-2, rel_plain
