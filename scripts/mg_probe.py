"""Ad-hoc GPU probe: Krylov iterations and time of the first viscosity iterations of a cold DIVA solve for several preconditioner
settings (not the bench contract).  usage: mg_probe.py NV [NVISC]"""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200")); sys.path.insert(0, ROOT)
import numpy as np  # noqa
import bench  # noqa
import ufe_b200 as ub  # noqa

nv = int(sys.argv[1]); nvisc = int(sys.argv[2]) if len(sys.argv) > 2 else 2
wname = os.environ.get("PROBE_WORKLOAD", "mismip_mod")
if wname == "mismip_plus":
    import workloads as wl
    w = wl.mismip_plus(float(np.sqrt(800e3 * 80e3 / (nv * np.sqrt(3.0) / 2.0))))
    w.cfg.ksp_initial_guess_nonzero = 1
else:
    w = bench.make_workload(nv)
m, cfg = w.mesh, w.cfg
approx = w.approximation
ctx = ub.Context(device=0)
ctx.mesh_create(m); ctx.calc_matrix_operators()
cfg.visc_it_nit = nvisc - 1; cfg.visc_it_norm_dUV_tol = 1e-30; cfg.write_diagnostics = 0
cfg.ksp_initial_guess_nonzero = int(os.environ.get("PROBE_GUESS", "0"))
ice = w.ice_inputs()
L = ub.lib()
print(f"mesh nV={m.nV} nTri={m.nTri}", flush=True)

# settings: "jacobi" or "nu,w_levels,overcorr,omega,fp32" separated by ';' in argv[3]
spec = sys.argv[3] if len(sys.argv) > 3 else "jacobi;1,3,1.0,0.7,0;1,3,1.3,0.7,0"
SETTINGS = []
for tok in spec.split(";"):
    if tok == "jacobi":
        SETTINGS.append(("jacobi", None))
    else:
        f = tok.split(",")
        nu, wl_, oc, om, fp = f[:5]
        opts = dict(kv.split("=") for kv in f[5:])          # further knobs by name, e.g. adaptive=1,nu_pre=0,tail_max=12000,tail_cluster=8
        SETTINGS.append((f"mg nu={nu} w={wl_} oc={oc} om={om} fp32={fp} {opts}", dict(nu_pre=int(nu), nu_post=int(nu), w_levels=int(wl_), overcorr=float(oc), omega=float(om), fp32=int(fp), _opts=opts)))
timing = int(os.environ.get("PROBE_TIMING", "1"))

ref = None
for name, prm in SETTINGS:
    st, arrs = ub.alloc_velocity_state(m.nV, m.nTri, m.nz, fields=["u_vav_b", "v_vav_b"])
    ub.check(L.ufe_b200_velocity_upload(ctx.h, C.byref(ice), C.byref(st)))
    cfg.pc_type = ub.PC_JACOBI if prm is None else ub.PC_MG
    if prm:
        prm = dict(prm)
        opts = prm.pop("_opts", {})
        ctx.set_mg_params(**prm)
        for k0 in ("adaptive", "tail_max", "tail_cluster"):
            ctx.set_mg_option(k0, 0)
        ctx.set_mg_option("max_levels", 12); ctx.set_mg_option("coarse_sweeps", 8); ctx.set_mg_option("stop_nodes", 0); ctx.set_mg_option("coarsest_cluster", 0); ctx.set_mg_option("coarsest_dsmem", 0); ctx.set_mg_option("coarsest_threads", 256); ctx.set_mg_option("coarsest_cheb", 0); ctx.set_mg_option("cheb_ratio", 30.0)
        ctx.set_mg_option("nu_pre", prm["nu_pre"])
        for k0, v0 in opts.items():
            ctx.set_mg_option(k0, float(v0))
    L.ufe_b200_timing_reset(ctx.h); L.ufe_b200_timing_enable(ctx.h, timing)
    a, b = C.c_int(), C.c_int()
    l0 = ctx.kernel_launches()
    t0 = time.time()
    ub.check(L.ufe_b200_velocity_solve_resident(ctx.h, C.byref(cfg), approx, C.byref(a), C.byref(b)))
    dt = time.time() - t0
    L.ufe_b200_timing_enable(ctx.h, 0)
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    u = arrs["u_vav_b"].copy()
    if ref is None:
        ref = u
    cls = {c: ctx.timing(c) for c in ("spmv_stiffness", "mg_fine", "mg_vec", "mg_coarse", "mg_setup", "krylov_vec")}
    info = ctx.mg_info() if prm else {}
    print(f"{name:48s}: {a.value} visc its, {b.value:5d} Krylov its, {dt:8.3f} s, launches {ctx.kernel_launches() - l0:7d}, "
          f"rel diff vs first {np.linalg.norm(u - ref) / np.linalg.norm(ref):.2e}  "
          + " ".join(f"{k}={v[0]:.0f}ms/{v[1]}" for k, v in cls.items() if v[1])
          + (f"  levels={info['nodes']} setup_host={info['setup_s']:.2f}s" if info else ""), flush=True)
ctx.close()
