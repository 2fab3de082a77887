"""Multi-GPU parity check (run under torchrun on N GPUs): every rank first solves the whole problem alone on its own GPU, then the N
ranks solve it together (row blocks from partition_list, NCCL halos + all-reduces); the rank compares its slice with the single-GPU
result.  Prints one line per case; exits non-zero on a mismatch.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/multi_gpu_parity.py
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200")); sys.path.insert(0, ROOT)
import numpy as np  # noqa
import bench  # noqa
import ufe_b200 as ub  # noqa
import workloads as wl  # noqa

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", rank))
rdv = bench.Rendezvous(rank, world)
L = ub.lib()
FIELDS = bench.STATE_FIELDS + ["u_b_prev", "v_b_prev"]
DT = 0.2


def run(ctx, w, pc, steps):
    m, cfg = w.mesh, w.cfg
    cfg.pc_type = pc
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    vr, tr = (ctx.vi1, ctx.vi2), (ctx.ti1, ctx.ti2)
    sa = slice(vr[0] - 1, vr[1])
    ice = w.ice_inputs(vr, tr)
    st, A = ub.alloc_velocity_state(vr[1] - vr[0] + 1, tr[1] - tr[0] + 1, m.nz, fields=FIELDS)
    nv, na, nt = C.c_int(), C.c_int(), C.c_int()
    ub.check(L.ufe_b200_velocity_upload(ctx.h, C.byref(ice), C.byref(st)))
    ub.check(L.ufe_b200_velocity_solve_resident(ctx.h, C.byref(cfg), 1, C.byref(nv), C.byref(na)))
    its = [(nv.value, na.value)]
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    u0 = A["u_vav_b"].copy()          # the first stress-balance solve, before any thickness step
    io, keep = bench.thickness_io(w, sa, w.f["Hi"][sa])
    ub.check(L.ufe_b200_thickness_upload(ctx.h, C.byref(io)))
    for _ in range(steps):
        ub.check(L.ufe_b200_pc_step_resident(ctx.h, C.byref(cfg), 1, DT, C.byref(nv), C.byref(na), C.byref(nt)))
        its.append((nv.value, na.value, nt.value))
    Hi = np.zeros(vr[1] - vr[0] + 1)
    ub.check(L.ufe_b200_thickness_download(ctx.h, ub.dp(Hi), None, None))
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    return dict(u=A["u_vav_b"].copy(), v=A["v_vav_b"].copy(), u3=A["u_3D_b"].copy(), Hi=Hi, vr=vr, tr=tr, its=its, u0=u0)


def rel(a, b):
    n = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / (n if n > 0 else 1.0))


ok = True
CASES = [c for c in ((ub.PC_JACOBI, "jacobi"), (ub.PC_MG, "multigrid")) if c[1] in os.environ.get("PARITY_PCS", "jacobi,multigrid").split(",")]
for pc, name in CASES:
    def workload():
        w = wl.mismip_mod(1.2e4)
        w.cfg.write_diagnostics = 0
        w.cfg.stress_balance_PETSc_rtol = 1e-11; w.cfg.stress_balance_PETSc_abstol = 1e-30
        w.cfg.dHi_PETSc_rtol = 1e-12; w.cfg.dHi_PETSc_abstol = 1e-30
        w.cfg.visc_it_norm_dUV_tol = 1e-7; w.cfg.visc_it_nit = 12
        return w
    ctx1 = ub.Context(device=local)                                  # alone
    single = run(ctx1, workload(), pc, 2)
    ctx1.close()
    uid = rdv.bcast(ub.Context.unique_id() if rank == 0 else None, 128) if world > 1 else None
    ctxn = ub.Context(device=local, rank=rank, nranks=world, unique_id=uid)   # together
    multi = run(ctxn, workload(), pc, 2)
    ctxn.close()
    sa = slice(multi["vr"][0] - 1, multi["vr"][1]); sb = slice(multi["tr"][0] - 1, multi["tr"][1])
    errs = dict(u=rel(multi["u"], single["u"][sb]), v=rel(multi["v"], single["v"][sb]), u_3D=rel(multi["u3"], single["u3"][sb, :]),
                Hi=rel(multi["Hi"], single["Hi"][sa]))
    # and against the CPU oracle (the reference's algorithm restated): rank 0 solves the first stress balance of the whole mesh with it
    if rank == 0 and os.environ.get("PARITY_ORACLE", "1") == "1":
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as ol
        wo = workload()
        st_o, A_o = ub.alloc_velocity_state(wo.mesh.nV, wo.mesh.nTri, wo.mesh.nz)
        rc, nv_o, na_o = ol.OracleMesh(wo.mesh).solve_velocity(wo.cfg, 1, wo.ice_inputs(), st_o)
        errs["u_first_solve_vs_oracle"] = rel(multi["u0"], A_o["u_vav_b"][sb]) if rc == 0 else 1.0
    worst = rdv.allreduce_max(max(errs.values()))
    good = worst < 1e-6
    ok = ok and good
    print(f"[rank {rank}/{world}] {name}: vertices {multi['vr']} triangles {multi['tr']} rel L2 vs single GPU {errs} "
          f"iterations single {single['its']} multi {multi['its']} -> {'OK' if good else 'MISMATCH'}", flush=True)
sys.exit(0 if ok else 1)
