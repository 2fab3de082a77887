"""The reference's 100-year time loops of integrated_test_Halfar_dome_small and integrated_test_MISMIP_mod_small (tests/test_reference_halfar_log.py,
tests/test_reference_mismip_mod_log.py) with every call of the path going through the C ABI on the GPU instead of the oracle, compared per predictor-corrector
iteration with the oracle's loop and with the reference's checksum log.

    gpurun -- 'python scripts/gpu_time_loop.py > gpurun_out/time_loop.log 2>&1'

tests/test_gpu_time_loop.py runs the three parts as `-m gpu` tests.  The host loop is the same object as in the CPU tests; only the backend differs.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200"))
import ufe_b200 as ub  # noqa: E402
import workloads as wl  # noqa: E402
import test_reference_halfar_log as H  # noqa: E402
import test_reference_mismip_mod_log as M  # noqa: E402
from test_reference_logs import mismip_mod_small_workload  # noqa: E402


class DeviceBackend:
    """the methods of tests/oracle_lib.OracleMesh the loops use, on libufe_b200.so"""

    def __init__(self, ctx, m):
        self.ctx, self.m, self.L = ctx, m, ub.lib()

    def determine_masks(self, cfg, Hi, Hb, SL):
        s, arrs = ub.alloc_masks(self.m.nV)
        ub.check(self.L.ufe_b200_determine_masks(self.ctx.h, C.byref(cfg), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), C.byref(s)))
        return arrs

    def calc_effective_thickness(self, cfg, Hi, Hb, SL):
        he, fm = np.zeros(self.m.nV), np.zeros(self.m.nV)
        ub.check(self.L.ufe_b200_calc_effective_thickness(self.ctx.h, C.byref(cfg), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(he), ub.dp(fm)))
        return he, fm

    def calc_grounded_fractions(self, cfg, choice, nbins, Hi, Hb, SL, dHb, mask_floating_ice, cdf, cdf_b):
        fr, frb = np.zeros(self.m.nV), np.zeros(self.m.nTri)
        ub.check(self.L.ufe_b200_calc_grounded_fractions(self.ctx.h, C.byref(cfg), choice, nbins, ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(dHb), ub.ip(mask_floating_ice),
                                                         ub.dp(cdf), ub.dp(cdf_b), ub.dp(fr), ub.dp(frb)))
        return fr, frb

    def calc_critical_timestep_adv(self, cfg, grounded_only, Hi, mask_floating_ice, u_vav_b, v_vav_b):
        dt = C.c_double()
        ub.check(self.L.ufe_b200_calc_critical_timestep_adv(self.ctx.h, C.byref(cfg), grounded_only, ub.dp(Hi), ub.ip(mask_floating_ice), ub.dp(u_vav_b), ub.dp(v_vav_b), C.byref(dt)))
        return dt.value

    def calc_dHi_dt(self, cfg, io, dt):
        d, its = C.c_double(dt), C.c_int()
        rc = self.L.ufe_b200_calc_dHi_dt(self.ctx.h, C.byref(cfg), C.byref(io), C.cast(C.byref(d), ub.c_dp), 1, C.byref(its))
        return rc, d.value, its.value

    def solve_SIA(self, cfg, Hi, Hs, A_flow, max_diffusivity=1e5):
        m = self.m
        r = {k: np.zeros((m.nTri if k.endswith("_b") else m.nV, m.nz), order="F") for k in ("u_3D_b", "v_3D_b", "D_3D_b", "du_dz_3D_a", "dv_dz_3D_a")}
        ub.check(self.L.ufe_b200_solve_SIA(self.ctx.h, C.byref(cfg), max_diffusivity, ub.dp(Hi), ub.dp(Hs), ub.dp(A_flow),
                                           *(ub.dp(r[k]) for k in ("u_3D_b", "v_3D_b", "D_3D_b", "du_dz_3D_a", "dv_dz_3D_a"))))
        return r

    def solve_velocity(self, cfg, approximation, ice, state):
        nv, na = C.c_int(), C.c_int()
        fn = self.L.ufe_b200_solve_DIVA if approximation == 1 else self.L.ufe_b200_solve_SSA
        rc = fn(self.ctx.h, C.byref(cfg), C.byref(ice), C.byref(state), C.byref(nv), C.byref(na), None, None, None)
        return rc, nv.value, na.value


def on_device(run, ctx):
    """swap the loop's backend: everything it calls on `run.o` and the truncation error now run on the GPU"""
    run.o = DeviceBackend(ctx, run.m)

    def truncation_error(z, Hi_np1, Hi_star, mk):
        tau = np.zeros(run.m.nV); eta = C.c_double(); nt, ng = C.c_int32(), C.c_int32()
        ub.check(ub.lib().ufe_b200_calc_pc_truncation_error(ctx.h, z, run.dt_n, H.PC_ETA_MIN, H.PC_EPSILON, ub.dp(Hi_np1), ub.dp(Hi_star), ub.ip(mk["mask_grounded_ice"]),
                                                            ub.ip(mk["mask_gl_gr"]), ub.dp(run.frac_gr), ub.dp(tau), ub.ip(run.guilty), C.byref(eta), C.byref(nt), C.byref(ng)))
        return tau, eta.value, nt.value, ng.value

    run.truncation_error = truncation_error
    return run


def compare(name, cpu, gpu, Hi_cpu, Hi_gpu):
    a, b = np.array(cpu.log), np.array(gpu.log)
    print(f"[{name}] predictor-corrector iterations: oracle {len(a)}, device {len(b)}")
    n = min(len(a), len(b))
    for k, col in ((1, "dt"), (2, "max Hi_np1"), (3, "max tau_np1"), (5, "max u_vav_b")):
        d = np.abs(a[:n, k] - b[:n, k]) / np.maximum(np.abs(a[:n, k]), 1e-30)
        print(f"[{name}]   {col:12s} largest relative difference over the run {d.max():.3e} (iteration {int(d.argmax()) + 1})")
    rel = np.linalg.norm(Hi_gpu - Hi_cpu) / np.linalg.norm(Hi_cpu)
    print(f"[{name}]   final thickness, relative L2 difference {rel:.3e}")
    return len(a) == len(b) and rel < 1e-6


def one_call_mismip_mod(cpu):
    """The same 100 years with ONE call per predictor-corrector step: ufe_b200_pc_step on host buffers with the geometry stage on the device and
    ufe_b200_pc_reference_loop switched on, driven by the time-step sequence and action times of the CPU loop model (the truncation error never limits
    the step in this experiment, so the sequence is known beforehand); compared with the CPU loop model at the end"""
    sys.path.insert(0, ROOT)
    import bench
    w = mismip_mod_workload()
    m, cfg, f, L = w.mesh, w.cfg, w.f, ub.lib()
    cdf, cdf_b = M.bedrock_cdfs(m)
    dHb = np.zeros(m.nV)
    log = np.array(cpu.log)
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(m); ctx.calc_matrix_operators()
        ub.check(L.ufe_b200_geometry_resident(ctx.h, ub.GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF, M.NBINS, ub.dp(dHb), ub.dp(cdf), ub.dp(cdf_b)))
        st, state = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
        out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
        io, keep = bench.thickness_io(w, slice(None), f["Hi"].copy(), out=out)
        d_nm1, Hi_lag, dt_prev = np.zeros(m.nV), f["Hi"].copy(), 0.0
        nv, na, nt = C.c_int(), C.c_int(), C.c_int()
        n_visc = []
        for t, dt in log[:, :2]:
            t1 = t + dt
            last = max(t, np.ceil(t1) - 1.0)                       # actions: every predictor-corrector time and every integer year (dt_thermodynamics)
            wgt = (last - t) / dt
            ub.check(L.ufe_b200_pc_reference_loop(ctx.h, 1, wgt))
            lam = np.clip(1.0 - (f["Hb"] - f["SL"]) / 1000.0, 0.0, 1.0)
            f["effective_pressure"] = np.maximum(0.0, wl.ICE_DENSITY * wl.GRAV * Hi_lag * (1.0 - 0.96 * lam))
            ice = w.ice_inputs()
            ub.check(L.ufe_b200_pc_step(ctx.h, C.byref(cfg), 1, dt, dt_prev, C.byref(ice), C.byref(st), C.byref(io), ub.dp(d_nm1),
                                        C.byref(nv), C.byref(na), C.byref(nt)))
            Hi_lag = (1.0 - wgt) * keep["Hi"] + wgt * out["Hi_tplusdt"]
            keep["Hi"][:] = out["Hi_tplusdt"]
            dt_prev = dt
            n_visc.append(nv.value)
        ub.check(L.ufe_b200_pc_reference_loop(ctx.h, 0, 0.0))
    finally:
        ctx.close()
    rel = np.linalg.norm(keep["Hi"] - cpu.Hi_next) / np.linalg.norm(cpu.Hi_next)
    relu = np.abs(state["u_vav_b"].max() / cpu.u.max() - 1.0)
    print(f"[MISMIP_mod small, one call per step] viscosity iterations {n_visc}")
    print(f"[MISMIP_mod small, one call per step] final thickness against the CPU loop model: relative L2 {rel:.3e}; max u_vav_b relative {relu:.3e}")
    return bool(rel < 1e-6 and relu < 1e-5)


def halfar_narrow_seam():
    """Halfar dome, SIA: every call of the 100-year loop through the C ABI"""
    ok = True
    cpu = H.HalfarRun(H.halfar_40km_workload())
    Hi_cpu = cpu.run()
    w = H.halfar_40km_workload()
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(w.mesh); ctx.calc_matrix_operators()
        gpu = on_device(H.HalfarRun(w), ctx)
        Hi_gpu = gpu.run()
    finally:
        ctx.close()
    ok &= compare("Halfar 40 km", cpu, gpu, Hi_cpu, Hi_gpu)
    g = json.load(open(H.GOLDEN))
    ref_tau = np.array(g["tau_np1"])[:, 3]
    ratio = np.array(gpu.log)[2:55, 3] / ref_tau[2:55]
    print(f"[Halfar 40 km]   device / reference log, max tau_np1 of iterations 3-55: {ratio.min():.3f} .. {ratio.max():.3f}, median {np.median(ratio):.3f}")
    ok &= 0.6 < ratio.min() and ratio.max() < 1.3
    return bool(ok)


def mismip_mod_workload():
    w = mismip_mod_small_workload()
    w.f["mask_noice"] = wl.mask_noice(w.mesh, "MISMIP_mod")
    return w


def mismip_mod_cpu_loop():
    cpu = M.MismipModRun(mismip_mod_workload())
    Hi_cpu = cpu.run()
    return cpu, Hi_cpu


def mismip_mod_narrow_seam(cpu, Hi_cpu):
    """MISMIP_mod, DIVA: every call of the 100-year loop through the C ABI"""
    ok = True
    w = mismip_mod_workload()
    ctx = ub.Context(device=0)
    try:
        ctx.mesh_create(w.mesh); ctx.calc_matrix_operators()
        gpu = M.MismipModRun.__new__(M.MismipModRun)
        H.HalfarRun.__init__(gpu, w)
        gpu.ksp = 0            # oracle_lib.KSP_BICGSTAB_JACOBI: the loop then calls solve_velocity without the oracle-only keyword
        on_device(gpu, ctx)
        gpu.cdf, gpu.cdf_b = M.bedrock_cdfs(w.mesh); gpu.dHb = np.zeros(w.mesh.nV)
        gpu.st, gpu.state = ub.alloc_velocity_state(w.mesh.nV, w.mesh.nTri, w.mesh.nz)
        gpu.n_visc, gpu.first = [], {}
        gpu.update_fractions(gpu.Hi, gpu.o.determine_masks(w.cfg, gpu.Hi, w.f["Hb"], w.f["SL"]))
        Hi_gpu = gpu.run()
    finally:
        ctx.close()
    ok &= compare("MISMIP_mod small", cpu, gpu, Hi_cpu, Hi_gpu)
    print(f"[MISMIP_mod small]   viscosity iterations per solve: oracle {cpu.n_visc}")
    print(f"[MISMIP_mod small]                                   device {gpu.n_visc}")
    T = json.load(open(os.path.join(ROOT, "tests", "golden", "golden.json")))["mismip_mod_small"]["time_loop"]
    ref_u, ref_H = np.array(T["u_vav_b"])[:, 3], np.array(T["Hi_np1"])[:, 3]
    log = np.array(gpu.log)
    if len(log) == 50:
        growth = (log[:, 5] / log[0, 5]) / (ref_u / ref_u[0])
        print(f"[MISMIP_mod small]   device / reference log: shelf speed level {log[0, 5] / ref_u[0]:.3f}, growth within {np.abs(growth - 1).max():.4f}; "
              f"max Hi_np1 within {np.abs(log[:, 2] - ref_H).max():.3f} m")
        ok &= np.abs(growth - 1).max() < 0.025 and np.abs(log[:, 2] - ref_H).max() < 0.2
    else:
        ok = False
    return bool(ok)


def main():
    ok = halfar_narrow_seam()
    cpu, Hi_cpu = mismip_mod_cpu_loop()
    ok &= mismip_mod_narrow_seam(cpu, Hi_cpu)
    ok &= one_call_mismip_mod(cpu)
    print("TIME LOOP PARITY", "OK" if ok else "FAILED")
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
