"""Ad-hoc GPU probe: kernel-class timings of a DIVA solve at a given resolution (not the bench contract)."""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200"))
import numpy as np
import ufe_b200 as ub
import workloads as wl

h = float(sys.argv[1]); nvisc = int(sys.argv[2]) if len(sys.argv) > 2 else 2
half = float(sys.argv[3]) if len(sys.argv) > 3 else 3.04e6
t0 = time.time()
w = wl.mismip_mod(h, half_width=half)
m = w.mesh
print(f"mesh nV={m.nV} nTri={m.nTri} gen {time.time()-t0:.1f}s", flush=True)
ctx = ub.Context(device=0)
t0 = time.time(); ctx.mesh_create(m); print(f"mesh_create {time.time()-t0:.2f}s")
t0 = time.time(); ctx.calc_matrix_operators(); print(f"operators {time.time()-t0:.2f}s", flush=True)
cfg = w.cfg
cfg.visc_it_nit = nvisc - 1; cfg.visc_it_norm_dUV_tol = 1e-30; cfg.write_diagnostics = 0
ice = w.ice_inputs()
st, arrs = ub.alloc_velocity_state(m.nV, m.nTri, m.nz, fields=["u_vav_b", "v_vav_b"])
L = ub.lib()
ub.check(L.ufe_b200_velocity_upload(ctx.h, C.byref(ice), C.byref(st)))
L.ufe_b200_timing_enable(ctx.h, 1)
a, b = C.c_int(), C.c_int()
t0 = time.time()
ub.check(L.ufe_b200_velocity_solve_resident(ctx.h, C.byref(cfg), 1, C.byref(a), C.byref(b)))
dt = time.time() - t0
print(f"solve: {a.value} visc its, {b.value} Krylov its, {dt:.3f}s  -> {dt/max(b.value,1)*1e3:.3f} ms/Krylov it")
R = 2 * m.nTri
nnz = ctx.get_operator(ub.OP["M2_ddx_b_b"])[1].size * 4  # approx stiffness nnz
for cls in ["spmv_stiffness", "krylov_vec", "assembly", "column", "tri", "epilogue"]:
    ms, n = ctx.timing(cls)
    extra = ""
    if cls == "spmv_stiffness" and n:
        B = 12 * nnz + 4 * (R + 1) + 8 * R + 8 * R
        extra = f"  avg {ms/n*1e3:.1f} us  algorithmic {B/(ms/n*1e-3)/1e9:.0f} GB/s (B_spmv={B/1e9:.3f} GB)"
    print(f"  {cls:16s} {ms:10.2f} ms  {n:7d} launches{extra}")
ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
print("speed max", np.hypot(arrs["u_vav_b"], arrs["v_vav_b"]).max())

# the variants of the two gather kernels of the viscosity iteration (velocity.cu), each from the same zero state: kernel time per launch and the result
if os.environ.get("PROBE_VARIANTS", "1") == "1":
    ref = None
    for name, env in (("interleaved zeta/eta copy + shared-memory columns (default)", {}), ("+ interleaved eta_3D_a copy for the assembly", {"UFE_B200_INTERLEAVE_A": "1"}),
                      ("reference-layout gathers", {"UFE_B200_INTERLEAVE": "0"}), ("local-memory columns", {"UFE_B200_VERTEX_SMEM": "0"}),
                      ("round-1 kernels", {"UFE_B200_INTERLEAVE": "0", "UFE_B200_VERTEX_SMEM": "0"})):
        for k in ("UFE_B200_INTERLEAVE", "UFE_B200_INTERLEAVE_A", "UFE_B200_VERTEX_SMEM"):
            os.environ.pop(k, None)
        os.environ.update(env)
        for a_ in arrs.values():
            a_[...] = 0.0
        ub.check(L.ufe_b200_velocity_upload(ctx.h, C.byref(ice), C.byref(st)))
        L.ufe_b200_timing_reset(ctx.h)
        ub.check(L.ufe_b200_velocity_solve_resident(ctx.h, C.byref(cfg), 1, C.byref(a), C.byref(b)))
        ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
        u = arrs["u_vav_b"].copy()
        ref = u if ref is None else ref
        col, asm = ctx.timing("column"), ctx.timing("assembly")
        print(f"{name:55s}: column {col[0] / a.value:.3f} ms / viscosity iteration ({col[1]} launches), assembly {asm[0] / a.value:.3f} ms ({asm[1]} launches), "
              f"{a.value} visc its, {b.value} Krylov its, rel diff of u vs first variant {np.linalg.norm(u - ref) / np.linalg.norm(ref):.2e}")
