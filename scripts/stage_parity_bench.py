"""The stages around the two solves on N GPUs and at full size.

  parity (default): every rank runs all stage entries alone on the whole mesh, then the N ranks run them together on their row blocks
                    (NCCL halos); each rank compares its slice.  Integer masks must be identical, fields agree to 1e-12.
      python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 scripts/stage_parity_bench.py
  timing:  STAGE_NV=4000000 python scripts/stage_parity_bench.py      (one GPU; kernel time per entry from the library's CUDA-event classes)
"""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200")); sys.path.insert(0, ROOT)
import numpy as np  # noqa
import bench  # noqa
import ufe_b200 as ub  # noqa
import workloads as wl  # noqa

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", rank))
rdv = bench.Rendezvous(rank, world)
L = ub.lib()
NV = int(os.environ.get("STAGE_NV", "0"))
KCLASSES = ("column", "tri", "thickness", "operator_spmv")


def make_data(w):
    m = w.mesh
    rng = np.random.default_rng(21)
    d = {}
    Hi = w.f["Hi"].copy()
    Hi[rng.random(m.nV) < 0.03] = 0.0
    d["Hi"] = Hi; d["Hb"] = w.f["Hb"] + 30.0 * rng.standard_normal(m.nV); d["SL"] = np.zeros(m.nV)
    d["Hs"] = wl.ice_surface_elevation(Hi, d["Hb"], d["SL"]); d["Hib"] = d["Hs"] - Hi
    d["dHb"] = 5.0 * np.sin(m.V[:, 0] / 3e5); d["dHb_dt"] = 1e-3 * rng.standard_normal(m.nV); d["dHi_dt"] = 0.5 * rng.standard_normal(m.nV)
    d["BMB"] = rng.standard_normal(m.nV); d["Ti_hom"] = -30.0 * rng.random(m.nV)
    d["Hi_new"] = np.maximum(0.0, Hi + 40.0 * rng.standard_normal(m.nV) * (Hi > 0))
    d["ref_Hi"] = np.maximum(0.0, Hi + 100.0 * rng.standard_normal(m.nV) * (Hi > 0)); d["ref_Hb"] = d["Hb"] + 5.0 * rng.standard_normal(m.nV)
    prof = np.linspace(1.0, 0.3, m.nz) ** 2
    d["u3"] = np.asfortranarray(200.0 * rng.standard_normal((m.nTri, 1)) * prof + rng.standard_normal((m.nTri, m.nz)))
    d["v3"] = np.asfortranarray(150.0 * rng.standard_normal((m.nTri, 1)) * prof + rng.standard_normal((m.nTri, m.nz)))
    d["u"] = 3.0e3 * rng.standard_normal(m.nTri); d["v"] = 2.0e3 * rng.standard_normal(m.nTri)
    d["D3"] = np.asfortranarray(-1e4 * rng.random((m.nTri, m.nz)))
    nb = 11
    d["cdf"] = np.asfortranarray(d["Hb"][:, None] + (40.0 + 80.0 * rng.random(m.nV))[:, None] * np.linspace(-1.0, 1.0, nb)[None, :] ** 3)
    Tri = np.asarray(m.Tri) - 1
    d["cdf_b"] = np.asfortranarray(d["Hb"][Tri].mean(axis=1)[:, None] + (40.0 + 80.0 * rng.random(m.nTri))[:, None] * np.linspace(-1.0, 1.0, nb)[None, :] ** 3)
    return d


def run(ctx, w, d, timing=False):
    m, cfg = w.mesh, w.cfg
    ctx.mesh_create(m); ctx.calc_matrix_operators()
    sa, sb = slice(ctx.vi1 - 1, ctx.vi2), slice(ctx.ti1 - 1, ctx.ti2)
    na, nb = ctx.vi2 - ctx.vi1 + 1, ctx.ti2 - ctx.ti1 + 1
    A = lambda k: np.ascontiguousarray(d[k][sa]); B = lambda k: np.ascontiguousarray(d[k][sb])   # noqa
    A3 = lambda k: np.asfortranarray(d[k][sa, :]); B3 = lambda k: np.asfortranarray(d[k][sb, :])   # noqa
    out, rows = {}, []

    def timed(name, bytes_alg, fn):
        if timing:
            L.ufe_b200_timing_reset(ctx.h); L.ufe_b200_timing_enable(ctx.h, 1)
        t0 = time.perf_counter()
        fn()
        wall = time.perf_counter() - t0
        if timing:
            L.ufe_b200_timing_enable(ctx.h, 0)
            k_ms = sum(ctx.timing(c)[0] for c in KCLASSES); k_n = sum(ctx.timing(c)[1] for c in KCLASSES)
            rows.append((name, k_ms, k_n, bytes_alg / 1e9 / (k_ms * 1e-3) if k_ms > 0 else 0.0, ctx.timing("h2d")[0], ctx.timing("d2h")[0], wall * 1e3))

    nz = m.nz
    u3, v3 = B3("u3"), B3("v3")
    sec_s, sec = ub.alloc_secondary(na, nb, nz)
    timed("calc_secondary_velocities", 16 * nz * nb + 72 * nb + na * (16 * nz * 6 + 16 * nz + 80),
          lambda: ub.check(L.ufe_b200_calc_secondary_velocities(ctx.h, ub.dp(u3), ub.dp(v3), C.byref(sec_s))))
    out.update({"sec_" + k: v for k, v in sec.items()})
    fa = [A(k) for k in ("Hi", "Hs", "Hib", "dHb_dt", "dHi_dt", "BMB")]
    mk_s, mk = ub.alloc_masks(na)
    Hi, Hb, SL = A("Hi"), A("Hb"), A("SL")
    timed("determine_masks", na * (24 + 4 * 7 + 44), lambda: ub.check(L.ufe_b200_determine_masks(ctx.h, C.byref(cfg), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), C.byref(mk_s))))
    out.update(mk)
    w3, dw3 = np.zeros((na, nz), order="F"), np.zeros((na, nz), order="F")
    vio = ub.VerticalVelocityIO(*[ub.dp(a) for a in fa], ub.ip(mk["mask_grounded_ice"]), ub.ip(mk["mask_floating_ice"]), ub.dp(w3), ub.dp(dw3))
    timed("calc_vertical_velocities", na * (6 * 2 * 16 * nz + 32 * nz + 200),
          lambda: ub.check(L.ufe_b200_calc_vertical_velocities(ctx.h, C.byref(cfg), C.byref(vio))))
    ice = (mk["mask_grounded_ice"] + mk["mask_floating_ice"]) > 0
    out["w_3D"] = w3; out["dw_dz_3D"] = np.where(ice[:, None], dw3, 0.0)
    he, fm = np.zeros(na), np.zeros(na)
    timed("calc_effective_thickness", na * (24 + 7 * 12 + 16),
          lambda: ub.check(L.ufe_b200_calc_effective_thickness(ctx.h, C.byref(cfg), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(he), ub.dp(fm))))
    out["Hi_eff"] = he; out["fraction_margin"] = fm
    dHb, cdf, cdf_b = A("dHb"), np.asfortranarray(d["cdf"][sa, :]), np.asfortranarray(d["cdf_b"][sb, :])
    for choice, nm in ((0, "TAF"), (1, "CDF"), (2, "TAF+CDF")):
        fr, frb = np.zeros(na), np.zeros(nb)
        timed(f"calc_grounded_fractions[{nm}]", na * (32 + 7 * 8 + 6 * 24 + 88 * (choice > 0)) + nb * (3 * 32 + 16 + 88 * (choice > 0)),
              lambda: ub.check(L.ufe_b200_calc_grounded_fractions(ctx.h, C.byref(cfg), choice, 11, ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(dHb),
                                                                  ub.ip(mk["mask_floating_ice"]), ub.dp(cdf), ub.dp(cdf_b), ub.dp(fr), ub.dp(frb))))
        out[f"fraction_gr_{nm}"] = fr; out[f"fraction_gr_b_{nm}"] = frb
    u, v, D3 = B("u"), B("v"), B3("D3")
    dts = []
    for go in (0, 1):
        dt = C.c_double()
        timed(f"calc_critical_timestep_adv[{go}]", na * (8 + 6 * (12 + 8 + 32 + 16)),
              lambda: ub.check(L.ufe_b200_calc_critical_timestep_adv(ctx.h, C.byref(cfg), go, ub.dp(Hi), ub.ip(mk["mask_floating_ice"]), ub.dp(u), ub.dp(v), C.byref(dt))))
        dts.append(dt.value)
    dt = C.c_double()
    timed("calc_critical_timestep_SIA", nb * (8 * nz + 12 + 3 * 24), lambda: ub.check(L.ufe_b200_calc_critical_timestep_SIA(ctx.h, C.byref(cfg), ub.dp(Hi), ub.dp(D3), C.byref(dt))))
    dts.append(dt.value)
    out["dts"] = np.array(dts)
    p = ub.alter_params_defaults(); p.Hi_min = 30.0; p.limitness_t_start, p.limitness_t_end = 0.0, 100.0; p.limitness_H_grounded = p.limitness_H_gl_gr = 15.0
    p.limitness_H_gl_fl = p.limitness_H_floating = 25.0; p.modiness_H_style = 1
    Hn, rHi, rHb, Th = A("Hi_new"), A("ref_Hi"), A("ref_Hb"), A("Ti_hom")
    timed("alter_ice_thickness", na * (8 * 9 + 7 * 12 + 4),
          lambda: ub.check(L.ufe_b200_alter_ice_thickness(ctx.h, C.byref(cfg), C.byref(p), 60.0, C.byref(mk_s), ub.dp(Hi), ub.dp(Hb), ub.dp(SL), ub.dp(rHi), ub.dp(rHb),
                                                         ub.dp(Th), ub.dp(Hn))))
    out["Hi_altered"] = Hn
    cf = (mk["mask_cf_fl"] + mk["mask_cf_gr"]) > 0
    Hs2 = Hi.copy(); Hs2[cf] *= 1.5
    q = np.zeros(na)
    timed("calc_and_apply_spill_over_flux", na * (8 * 4 + 4 + 7 * (12 + 48)),
          lambda: ub.check(L.ufe_b200_calc_and_apply_spill_over_flux(ctx.h, C.byref(mk_s), ub.dp(he), ub.dp(u), ub.dp(v), 0.5, ub.dp(Hs2), ub.dp(q))))
    out["Hi_spilled"] = Hs2; out["Qspill"] = q
    return out, (sa, sb), rows


if NV > 0:   # timing at full size, one GPU
    t0 = time.time()
    w = bench.make_workload(NV, "jacobi")
    print(f"mesh nV={w.mesh.nV} nTri={w.mesh.nTri} in {time.time() - t0:.0f}s", flush=True)
    d = make_data(w)
    ctx = ub.Context(device=local)
    run(ctx, w, d)                      # warm-up (allocations)
    ctx.close()
    ctx = ub.Context(device=local)
    _o, _s, rows = run(ctx, w, d, timing=True)
    ctx.close()
    print(f"{'entry':38s} {'kernels ms':>10s} {'launches':>8s} {'alg. GB/s':>10s} {'h2d ms':>8s} {'d2h ms':>8s} {'call ms':>8s}")
    for r in rows:
        print(f"{r[0]:38s} {r[1]:10.3f} {r[2]:8d} {r[3]:10.0f} {r[4]:8.2f} {r[5]:8.2f} {r[6]:8.1f}")
    sys.exit(0)

w = wl.mismip_mod(1.2e4)
d = make_data(w)
ctx1 = ub.Context(device=local)
single, _s, _r = run(ctx1, w, d)
ctx1.close()
uid = rdv.bcast(ub.Context.unique_id() if rank == 0 else None, 128) if world > 1 else None
ctxn = ub.Context(device=local, rank=rank, nranks=world, unique_id=uid)
multi, (sa, sb), _r = run(ctxn, w, d)
ctxn.close()
worst, bad = 0.0, []
for k, v in multi.items():
    ref = single[k]
    if k != "dts":
        s = sb if ref.shape[0] == w.mesh.nTri else sa
        ref = ref[s]
    if v.dtype == np.int32:
        e = float(np.max(np.abs(v - ref)))
    else:
        n = np.linalg.norm(ref)
        e = float(np.linalg.norm(v - ref) / (n if n > 0 else 1.0))
    worst = max(worst, e)
    if e > 1e-12:
        bad.append((k, e))
worst = rdv.allreduce_max(worst)
print(f"[rank {rank}/{world}] {len(multi)} stage outputs, vertices {sa.start + 1}:{sa.stop}, worst deviation from the single-GPU run {worst:.2e} "
      f"{'OK' if not bad else 'MISMATCH ' + str(bad)}", flush=True)
sys.exit(0 if worst <= 1e-12 else 1)
