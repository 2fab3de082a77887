"""Multi-GPU check of the copy boundary conditions ('infinite_SSA_icestream' borders of the SSA_icestream experiment, whose load vector reads the
previous solution a third of the domain away, i.e. usually on another rank): every rank first solves alone on its own GPU, then the N ranks
solve together; each rank compares its slice with the single-GPU result.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 scripts/multi_gpu_copy_bc.py
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


def run(ctx, w):
    m, cfg = w.mesh, w.cfg
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    vr, tr = (ctx.vi1, ctx.vi2), (ctx.ti1, ctx.ti2)
    ice = w.ice_inputs(vr, tr)
    st, A = ub.alloc_velocity_state(vr[1] - vr[0] + 1, tr[1] - tr[0] + 1, m.nz, fields=bench.STATE_FIELDS + ["u_b_prev", "v_b_prev"])
    nv, na = C.c_int(), C.c_int()
    ub.check(L.ufe_b200_velocity_upload(ctx.h, C.byref(ice), C.byref(st)))
    ub.check(L.ufe_b200_velocity_solve_resident(ctx.h, C.byref(cfg), 0, C.byref(nv), C.byref(na)))
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    return dict(u=A["u_vav_b"].copy(), v=A["v_vav_b"].copy(), tr=tr, its=(nv.value, na.value))


def workload():
    w = wl.ssa_icestream(16e3)
    w.cfg.write_diagnostics = 0
    w.cfg.stress_balance_PETSc_rtol = 1e-11; w.cfg.stress_balance_PETSc_abstol = 1e-30
    w.cfg.visc_it_nit = 30
    return w


ctx1 = ub.Context(device=local)
single = run(ctx1, workload())
ctx1.close()
uid = rdv.bcast(ub.Context.unique_id() if rank == 0 else None, 128) if world > 1 else None
ctxn = ub.Context(device=local, rank=rank, nranks=world, unique_id=uid)
multi = run(ctxn, workload())
ctxn.close()
sb = slice(multi["tr"][0] - 1, multi["tr"][1])
scale = float(np.abs(single["u"]).max())
errs = dict(u=float(np.abs(multi["u"] - single["u"][sb]).max() / scale), v=float(np.abs(multi["v"] - single["v"][sb]).max() / scale))
worst = rdv.allreduce_max(max(errs.values()))
good = worst < 1e-8
print(f"[rank {rank}/{world}] SSA_icestream copy BCs: triangles {multi['tr']} max |diff| / max |u| vs single GPU {errs} "
      f"iterations single {single['its']} multi {multi['its']} max u {scale:.4f} -> {'OK' if good else 'MISMATCH'}", flush=True)
sys.exit(0 if good else 1)
