#!/usr/bin/env python
"""bench.py -- DIVA velocity solves/s on the synthetic Antarctica-scale mesh (BASELINE.json configs[2]).

One "step" = one predictor-corrector time step of the hot path: calc_dHi_dt (predictor) -> DIVA velocity solve on
the predicted geometry -> calc_dHi_dt (corrector), i.e. 1 velocity solve + 2 thickness solves.
`value` is device-timed with every input resident in HBM (ufe_b200_pc_step_resident); `e2e` runs the same three
phases through the host-buffer C-ABI calls the Fortran shim binds, with the H2D/D2H copies inside the timed region.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--nv VERTICES]

For N > 1 the driver launches one process per GPU with torch.distributed.run; this script only reads
RANK/LOCAL_RANK/WORLD_SIZE/MASTER_ADDR/MASTER_PORT from the environment (PyTorch itself is not used: the path is
C-ABI + CUDA + NCCL; rendezvous of the NCCL unique id goes over a plain TCP socket).
"""
import argparse
import ctypes as C
import json
import os
import socket
import struct
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200"))

import numpy as np  # noqa: E402

HALF_WIDTH = 3.04e6          # Antarctic default domain +-3040 km
DT = 0.2                     # yr; dt_ice_min of config_MISMIP_8km_spinup_for_scaling.cfg:353


# ------------------------------------------------------------------------------------------------------------------
# tiny TCP rendezvous (rank 0 serves): broadcast bytes, all-reduce max of a float, barrier
# ------------------------------------------------------------------------------------------------------------------
class Rendezvous:
    def __init__(self, rank, world):
        self.rank, self.world = rank, world
        self.conns = []
        if world == 1:
            return
        addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
        port = int(os.environ.get("MASTER_PORT", "29500")) + 211
        if rank == 0:
            srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
            srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            srv.bind((addr, port))
            srv.listen(world)
            self.conns = [None] * world
            for _ in range(world - 1):
                c, _a = srv.accept()
                r = struct.unpack("i", self._recv(c, 4))[0]
                self.conns[r] = c
            srv.close()
        else:
            for _ in range(600):
                try:
                    s = socket.create_connection((addr, port), timeout=5)
                    break
                except OSError:
                    time.sleep(0.1)
            else:
                raise RuntimeError("rendezvous: cannot reach rank 0")
            s.settimeout(None)
            s.sendall(struct.pack("i", rank))
            self.conns = [s]

    @staticmethod
    def _recv(c, n):
        b = b""
        while len(b) < n:
            chunk = c.recv(n - len(b))
            if not chunk:
                raise RuntimeError("rendezvous: peer closed")
            b += chunk
        return b

    def bcast(self, data, n):
        if self.world == 1:
            return data
        if self.rank == 0:
            for c in self.conns[1:]:
                c.sendall(data)
            return data
        return self._recv(self.conns[0], n)

    def allreduce_max(self, x):
        if self.world == 1:
            return x
        if self.rank == 0:
            vals = [x] + [struct.unpack("d", self._recv(c, 8))[0] for c in self.conns[1:]]
            m = max(vals)
            for c in self.conns[1:]:
                c.sendall(struct.pack("d", m))
            return m
        self.conns[0].sendall(struct.pack("d", x))
        return struct.unpack("d", self._recv(self.conns[0], 8))[0]

    def barrier(self):
        self.allreduce_max(0.0)


# ------------------------------------------------------------------------------------------------------------------
# clocks sampled DURING the timed region (B200_PROFILING.md recipe)
# ------------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.lines, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i", str(self.gpu)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def bind_to_gpu_numa_node(gpu_index, log=lambda *a: None):
    """Run this process (and first-touch its page-locked buffers) on the NUMA node the GPU hangs off: host<->device copies from the far socket
    cross the inter-socket link and reach about half the PCIe rate.  No-op on single-node hosts or when sysfs / nvidia-smi say nothing."""
    try:
        bdf = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(gpu_index)], capture_output=True, text=True,
                             timeout=20).stdout.strip().lower()
        if not bdf:
            return None
        bdf = bdf[-12:] if len(bdf) > 12 else bdf          # 00000000:1B:00.0 -> 0000:1b:00.0
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            log(f"bound to NUMA node {node} of GPU {gpu_index} ({len(cpus)} cpus)")
            return node
    except (OSError, ValueError, subprocess.SubprocessError):
        pass
    return None


# ------------------------------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------------------------------
def h_for_vertices(nv):
    """target edge length giving ~nv vertices on the +-3040 km square (near-equilateral lattice: nV ~ L^2 / (h^2 sqrt(3)/2))"""
    L = 2 * HALF_WIDTH
    return float(np.sqrt(L * L / (nv * np.sqrt(3.0) / 2.0)))


def lattice_vertex_count(nv):
    """number of vertices the synthetic mesh generator produces for the target nv (ungraded lattice: columns spaced h, rows h sqrt(3)/2)"""
    h, L = h_for_vertices(nv), 2 * HALF_WIDTH
    x, ncol = -HALF_WIDTH, 1
    while x < HALF_WIDTH:
        x += h
        ncol += 1
    ny = max(2, int(np.floor(L / (h * np.sqrt(3.0) / 2.0) + 0.5)))
    return ncol * (ny + 1)


# BASELINE.json's configs as bench workloads.  The driver's line is "antarctica" (configs[2]); the others are run by hand (logs under profiles/).
CONFIGS = {
    "antarctica": dict(index=2, nv=4_000_000, approximation=1,
                       text="DIVA + ice_thickness predictor-corrector time stepping on a synthetic Antarctica-scale MISMIP_mod mesh"),
    "mismip8km": dict(index=0, nv=0, approximation=1,
                      text="MISMIP_mod island of config_MISMIP_8km_spinup_for_scaling.cfg (domain +-1000 km) at a uniform 8 km: the latency-bound regime"),
    "mismipplus_ssa": dict(index=1, nv=500_000, approximation=0,
                           text="MISMIP+ bed (800 x 80 km), SSA, Schoof2005 sliding, resolution graded towards the grounding line"),
    "sweep": dict(index=4, nv=4_000_000, approximation=1,
                  text="Krylov solve of the DIVA system of the first viscosity iteration, zero initial guess, rtol 1e-7, on the mesh generator of configs[2]"),
}


def make_workload(nv, pc="jacobi", config="antarctica"):
    import ufe_b200 as ub
    import workloads as wl
    if config == "mismip8km":
        w = wl.mismip_mod(8.0e3, half_width=1.0e6)
    elif config == "mismipplus_ssa":
        w = wl.mismip_plus(float(np.sqrt(800e3 * 80e3 / (nv * np.sqrt(3.0) / 2.0))))
    else:
        w = wl.mismip_mod(h_for_vertices(nv), half_width=HALF_WIDTH)
    cfg = w.cfg                      # config_MISMIP_8km_spinup_for_scaling.cfg physics; MISMIP+: Schoof2005 sliding, rtol 1e-7
    cfg.write_diagnostics = 0
    cfg.ksp_initial_guess_nonzero = 0  # the reference's KSP starts from zero (petsc_basic.f90:163 is commented out); B200_use_initial_guess_config is timed as an extra
    cfg.pc_type = ub.PC_MG if pc == "mg" else ub.PC_JACOBI   # the oracle ignores this key (it has its own ksp switch)
    return w


def thickness_io(w, sl_a, Hi, u=None, v=None, out=None):
    import ufe_b200 as ub
    f = w.f
    keep = dict(Hi=np.ascontiguousarray(Hi), Hb=np.ascontiguousarray(f["Hb"][sl_a]), SL=np.ascontiguousarray(f["SL"][sl_a]),
                SMB=np.ascontiguousarray(f["SMB"][sl_a]), BMB=np.ascontiguousarray(f["BMB"][sl_a]), LMB=np.ascontiguousarray(f["LMB"][sl_a]),
                fraction_margin=np.ascontiguousarray(f["fraction_margin"][sl_a]), mask_noice=np.ascontiguousarray(f["mask_noice"][sl_a]),
                dHi_dt_target=np.ascontiguousarray(f["dHi_dt_target"][sl_a]))
    io = ub.ThicknessIO()
    for k, a in keep.items():
        setattr(io, k, ub.ip(a) if a.dtype == np.int32 else ub.dp(a))
    if u is not None:
        io.u_vav_b, io.v_vav_b = ub.dp(u), ub.dp(v)
        keep["u"], keep["v"] = u, v
    if out is not None:
        for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ"):
            setattr(io, k, ub.dp(out[k]))
    return io, keep


def surf(Hi, Hb, SL, cfg):
    return Hi + np.maximum(SL - cfg.ice_density / cfg.seawater_density * Hi, Hb)


def pc_step_host(calc_dHi_dt, solve_velocity, w, cfg, sl_a, S, dt, geometry=None, hi_eff=None, alter=None, reference_loop=True, w_lag=0.0):
    """One pass of iterate_pc_timestep through host buffers, in the reference's order (predictor_corrector_scheme.f90:86-233): what ufe_b200_pc_step and
    ufe_b200_pc_step_resident do in one call.  Used by the CPU reference arm (oracle callbacks) and, as the model of the device step, by the tests.
    S: Hi (= Hi_prev), dHdt_nm1 (pc%dHi_dt_Hi_nm1_u_nm1), dt_n, u, v [, Hi_model = region%ice%Hi where it lags behind Hi_prev].
    geometry( Hi, for_velocity): optional geometry / mask / sub-grid stage (determine_masks, calc_grounded_fractions, calc_effective_thickness;
    :95-107, :172-179, :198-211); it updates the masks / fractions in w.f that the two callbacks read.
    hi_eff( Hi) -> effective thickness (calc_effective_thickness); alter( Hi_of_masks, Hi_old, Hi_new) -> Hi_new: alter_ice_thickness; default: the
    removal of ice whose effective thickness is below C%Hi_min, which is all that routine does with its optional modifications off.
    reference_loop / w_lag: see ufe_b200_pc_reference_loop."""
    f = w.f
    Hb, SL, noice = f["Hb"][sl_a], f["SL"][sl_a], f["mask_noice"][sl_a] == 1
    zeta = dt / S["dt_n"]
    if alter is None:
        def alter(_Hi_of_masks, _Hi_old, Hi_new):
            return np.where(hi_eff(Hi_new) < cfg.Hi_min, 0.0, Hi_new)
    Hi_model = S.get("Hi_model") if reference_loop else None
    if Hi_model is None:
        Hi_model = S["Hi"]
    Hi_prev = np.where(noice, 0.0, S["Hi"]); d_nm1 = np.where(noice, 0.0, S["dHdt_nm1"])                 # (:88-93)
    if geometry:
        geometry(Hi_model, False)
    d_n, its_p = calc_dHi_dt(Hi_model, S["u"], S["v"], dt)                               # predictor, on region%ice%Hi (:129)
    d_n = np.where(noice, 0.0, d_n)
    Hi_star = Hi_prev + dt * ((1 + zeta / 2) * d_n - (zeta / 2) * d_nm1)                 # Eq. 30 (:140)
    Hi_star = alter(Hi_model, Hi_prev, Hi_star)                                          # (:145)
    d_n = ((Hi_star - Hi_prev) / dt + (zeta / 2) * d_nm1) / (1 + zeta / 2)               # (:152)
    if geometry:
        geometry(Hi_star, True)
    nv, na = solve_velocity(Hi_star, surf(Hi_star, Hb, SL, cfg))                         # (:186) updates S["u"], S["v"], ...
    if geometry:
        geometry(Hi_prev, False)
    d_s, its_c = calc_dHi_dt(Hi_prev, S["u"], S["v"], dt)                                # corrector (:214)
    d_s = np.where(noice, 0.0, d_s)
    Hi_np1 = Hi_prev + (dt / 2) * (d_n + d_s)                                            # Eq. 31 (:226)
    Hi_np1 = alter(Hi_prev, Hi_prev, Hi_np1)                                             # (:233)
    S["Hi_star"], S["dHdt_n"] = Hi_star, d_n
    S["dHdt_nm1"] = (Hi_np1 - Hi_prev) / dt if reference_loop else d_n                   # ice_dynamics_main.f90:140 -> :109
    S["Hi_model"] = (1.0 - w_lag) * Hi_prev + w_lag * Hi_np1                             # ice_dynamics_main.f90:115-121
    S["Hi"] = Hi_np1
    S["dt_n"] = dt
    return nv, na, its_p + its_c


STATE_FIELDS = ["u_vav_b", "v_vav_b", "u_base_b", "v_base_b", "tau_bx_b", "tau_by_b", "eta_3D_b", "u_3D_b", "v_3D_b"]


# ------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (port of the reference's algorithm) on the host cores, on a bounded sample of the workload
# ------------------------------------------------------------------------------------------------------------------
SPINUP_VISC = 6


def host_threads():
    """threads the CPU arm may use: the cores this process is allowed on (not OMP_NUM_THREADS, which torch.distributed.run sets to 1 for its children)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def run_cpu_sample(nv_sample, nv_full, steps, warmup, log=lambda *a: None, solvers=("bicgstab_jacobi", "gmres_bjacobi_ilu0"), budget_s=150.0):
    """The oracle's predictor-corrector steps on all host cores, on a bounded sample mesh of the same workload, once per Krylov setting:
    'gmres_bjacobi_ilu0' = the reference's actual default (PETSc GMRES(30) + block Jacobi with one ILU(0) per MPI rank, one rank per core;
    petsc_basic.f90:95-105, 125-193), 'bicgstab_jacobi' = the algorithm `-ksp_type bcgs -pc_type jacobi` selects (the device's one-level method).
    Both with the reference's zero initial guess (petsc_basic.f90:163).  `value` is the faster of the two."""
    cores = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(cores)   # before liboracle.so (and libgomp with it) is loaded
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol   # CPU baseline leg: the one place bench.py may execute oracle/
    import ufe_b200 as ub
    threads_before = int(ol.lib().orc_max_threads())
    ol.lib().orc_set_num_threads(cores)
    threads_used = int(ol.lib().orc_max_threads())
    w = make_workload(nv_sample)
    m, cfg = w.mesh, w.cfg
    cfg.ksp_initial_guess_nonzero = 0            # reference semantics
    o = ol.OracleMesh(m)
    t0 = time.time()
    o.calc_matrix_operators()
    log(f"[cpu] sample mesh nV={m.nV}, operators {time.time() - t0:.1f}s, {threads_used} threads")
    sl = slice(None)
    KSP = {"bicgstab_jacobi": ol.KSP_BICGSTAB_JACOBI, "gmres_bjacobi_ilu0": ol.KSP_GMRES_BJACOBI_ILU0}
    runs = {}
    Hi0, Hs0 = w.f["Hi"].copy(), w.f["Hs"].copy()
    # spin-up (untimed, shared by both settings): SPINUP_VISC viscosity iterations from rest, Krylov solves started from the previous iterate.  A converged
    # spin-up (19 viscosity iterations) would cost more CPU time than the whole timed sample; instead every timed step runs exactly ONE viscosity iteration,
    # which is what the spun-up device run does per step (config.per_step.viscosity_its = 1.0), so the work per step is the same: one assembly, one
    # velocity Krylov solve from a zero guess, two thickness solves.
    st0, A0 = ub.alloc_velocity_state(m.nV, m.nTri, m.nz, fields=STATE_FIELDS + ["u_b_prev", "v_b_prev"])
    t0 = time.time()
    cfg.ksp_initial_guess_nonzero = 1
    rc, nv0, na0 = o.solve_velocity(cfg, 1, w.ice_inputs(), st0, ksp=ol.KSP_BICGSTAB_JACOBI, max_visc_its=SPINUP_VISC)
    cfg.ksp_initial_guess_nonzero = 0
    assert rc == 0
    log(f"[cpu] spin-up {time.time() - t0:.1f}s: {nv0} viscosity its, {na0} Krylov its")
    for name in solvers:
        ksp = KSP[name]
        st, A = ub.alloc_velocity_state(m.nV, m.nTri, m.nz, fields=STATE_FIELDS + ["u_b_prev", "v_b_prev"])
        for k in A:
            A[k][...] = A0[k]
        out = {k: np.zeros(m.nV) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
        w.f["Hi"], w.f["Hs"] = Hi0.copy(), Hs0.copy()

        def calc(Hi, u, v, dt):
            io, keep = thickness_io(w, sl, Hi, u, v, out)
            rc, dt2, its = o.calc_dHi_dt(cfg, io, dt, ksp=ksp)
            assert rc == 0
            return out["dHi_dt"].copy(), its

        def solve(Hi, Hs):
            w.f["Hi"], w.f["Hs"] = Hi, Hs
            ice = w.ice_inputs()
            rc, nv, na = o.solve_velocity(cfg, 1, ice, st, ksp=ksp, max_visc_its=1)
            assert rc == 0
            return nv, na

        S = dict(Hi=w.f["Hi"].copy(), dHdt_nm1=np.zeros(m.nV), dt_n=DT, u=A["u_vav_b"], v=A["v_vav_b"])
        he = lambda Hi: o.calc_effective_thickness(cfg, Hi, w.f["Hb"], w.f["SL"])[0]

        def geometry(Hi, for_velocity):   # the stage the b200 arm runs on the device (--geometry 1): masks, grounded fractions, effective thickness
            f = w.f
            if for_velocity:
                mk = o.determine_masks(cfg, Hi, f["Hb"], f["SL"])
                f["mask_grounded_ice"], f["mask_icefree_land"] = mk["mask_grounded_ice"], mk["mask_icefree_land"]
                f["fraction_gr_b"] = o.calc_grounded_fractions(cfg, ub.GF_BILIN_INTERP_TAF, 0, Hi, f["Hb"], f["SL"], np.zeros(m.nV), mk["mask_floating_ice"], None, None)[1]
            else:
                f["fraction_margin"] = o.calc_effective_thickness(cfg, Hi, f["Hb"], f["SL"])[1]

        for _ in range(warmup):
            pc_step_host(calc, solve, w, cfg, sl, S, DT, hi_eff=he, geometry=geometry)
        t0 = time.time()
        tot, done = [0, 0, 0], 0
        for _ in range(steps):
            r = pc_step_host(calc, solve, w, cfg, sl, S, DT, hi_eff=he, geometry=geometry)
            tot = [a + b for a, b in zip(tot, r)]
            done += 1
            if time.time() - t0 > budget_s / len(solvers):
                break
        el = time.time() - t0
        runs[name] = dict(steps_per_s_sample=done / el, ms_per_step_sample=1e3 * el / done, steps=done, viscosity_its_per_step=tot[0] / done,
                          krylov_its_velocity_per_step=tot[1] / done, krylov_its_thickness_per_step=tot[2] / done)
        log(f"[cpu] {name}: {runs[name]}")
    best = max(runs, key=lambda k: runs[k]["steps_per_s_sample"])
    factor = m.nV / nv_full
    for r in runs.values():
        r["value_extrapolated"] = r["steps_per_s_sample"] * factor
    ol.lib().orc_set_num_threads(threads_before)     # a caller that lives on (the test-suite) keeps its own thread count
    return dict(value=runs[best]["value_extrapolated"], unit="velocity solves/s", cores=threads_used, kind="port", solver=best,
                sample=(f"{runs[best]['steps']} predictor-corrector steps (after {warmup} warm-up; one viscosity iteration per step like the device run, state spun up with "
                        f"{SPINUP_VISC} viscosity iterations) of the same MISMIP_mod DIVA config on a {m.nV}-vertex mesh, "
                        f"oracle built -O3 -march=x86-64-v3 with OpenMP over rows / columns / ILU blocks on {threads_used} threads, zero initial guess; "
                        f"{runs[best]['steps_per_s_sample']:.4f} steps/s measured with {best} ({runs[best]['krylov_its_velocity_per_step']:.0f} Krylov its per step), "
                        f"scaled linearly in vertices (x {factor:.5f}) to {nv_full}; optimistic for the CPU because Krylov iterations grow with resolution"),
                same_config=False, extrapolation_factor=factor, n_vertices_sample=int(m.nV), n_vertices_full=int(nv_full),
                ms_per_step_sample=runs[best]["ms_per_step_sample"], solvers=runs,
                note=("gmres_bjacobi_ilu0 is the reference's default KSP/PC (one ILU(0) block per core, as `mpiexec -n <cores>` gives); bicgstab_jacobi is the "
                      "one-level method of the device path; the device's headline uses a multigrid preconditioner the reference does not have"))


# ------------------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------------------
def run_b200(args, rank, world, rdv, log):
    import ufe_b200 as ub
    L = ub.lib()
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    numa_node = bind_to_gpu_numa_node(local_rank, log)
    uid = None
    if world > 1:
        uid = rdv.bcast(ub.Context.unique_id() if rank == 0 else None, 128)
    ctx = ub.Context(device=local_rank, rank=rank, nranks=world, unique_id=uid)
    t0 = time.time()
    conf = CONFIGS[args.config]
    approx = conf["approximation"]
    w = make_workload(args.nv, args.pc, args.config)
    m, cfg = w.mesh, w.cfg
    dt = DT
    if args.pc == "mg":
        ctx.set_mg_params(nu_pre=args.mg_nu, nu_post=args.mg_nu, omega=args.mg_omega, overcorr=args.mg_overcorr, w_levels=args.mg_w_levels, fp32=args.mg_fp32)
        if args.mg_stop_nodes > 0:
            ctx.set_mg_option("stop_nodes", args.mg_stop_nodes); ctx.set_mg_option("coarse_sweeps", args.mg_coarse_sweeps)
            ctx.set_mg_option("coarsest_cluster", args.mg_coarsest_cluster); ctx.set_mg_option("coarsest_dsmem", args.mg_coarsest_dsmem)
            ctx.set_mg_option("coarsest_cheb", args.mg_coarsest_cheb); ctx.set_mg_option("cheb_ratio", args.mg_cheb_ratio)
    log(f"mesh nV={m.nV} nTri={m.nTri} nE={m.nE} generated in {time.time() - t0:.1f}s")
    t0 = time.time()
    ctx.mesh_create(m)
    ctx.calc_matrix_operators()
    log(f"mesh upload + on-device operator construction {time.time() - t0:.2f}s")
    vr, tr = (ctx.vi1, ctx.vi2), (ctx.ti1, ctx.ti2)
    sl_a, sl_b = slice(vr[0] - 1, vr[1]), slice(tr[0] - 1, tr[1])
    na, nb = vr[1] - vr[0] + 1, tr[1] - tr[0] + 1
    ice = w.ice_inputs(vr, tr)
    st, A = ub.alloc_velocity_state(na, nb, m.nz, fields=STATE_FIELDS + ["u_b_prev", "v_b_prev"])
    nv_, na_, nt_ = C.c_int(), C.c_int(), C.c_int()

    # ---- spin-up (untimed): cold DIVA solve to the converged velocity of the initial geometry
    t0 = time.time()
    ub.check(L.ufe_b200_velocity_upload(ctx.h, C.byref(ice), C.byref(st)))
    ub.check(L.ufe_b200_velocity_solve_resident(ctx.h, C.byref(cfg), approx, C.byref(nv_), C.byref(na_)))
    log(f"spin-up: {nv_.value} viscosity its, {na_.value} Krylov its in {time.time() - t0:.1f}s")
    if args.config == "mismipplus_ssa":
        # 0.25 of the advective limit of the finest triangles (the reference's calc_critical_timestep_adv has 0.5 built in), from the spun-up velocities
        ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
        speed = rdv.allreduce_max(float(np.hypot(A["u_vav_b"], A["v_vav_b"]).max()))
        dt = 0.25 * float(np.sqrt(m.TriA.min() * 4.0 / np.sqrt(3.0))) / max(speed, 1.0)
        log(f"MISMIP+: max speed {speed:.1f} m/yr, dt = {dt:.4f} yr")
    mg_info = ctx.mg_info() if args.pc == "mg" else None
    if mg_info:
        log(f"multigrid hierarchy: {mg_info}")
    io, keep_io = thickness_io(w, sl_a, w.f["Hi"][sl_a])
    ub.check(L.ufe_b200_thickness_upload(ctx.h, C.byref(io)))
    if args.geometry:   # determine_masks / calc_grounded_fractions / calc_effective_thickness between the solves, on the device (predictor_corrector_scheme.f90:95-211)
        ub.check(L.ufe_b200_geometry_resident(ctx.h, ub.GF_BILIN_INTERP_TAF, 0, None, None, None))

    # ---- device-resident leg (`value`)
    def step_resident():
        ub.check(L.ufe_b200_pc_step_resident(ctx.h, C.byref(cfg), approx, dt, C.byref(nv_), C.byref(na_), C.byref(nt_)))
        return nv_.value, na_.value, nt_.value

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    L.ufe_b200_timing_reset(ctx.h)
    L.ufe_b200_timing_enable(ctx.h, 1)
    launches0 = ctx.kernel_launches()
    rdv.barrier()
    tot = [0, 0, 0]
    ms = C.c_double()
    flush_l2 = m.nV < 400_000        # stiffness matrix (~1 kB / vertex) not well above the 126 MB L2: write a 512 MB buffer between the timed steps
    if flush_l2:
        ms_sum = 0.0
        for _ in range(args.steps):
            ub.check(L.ufe_b200_l2_flush(ctx.h))
            ub.check(L.ufe_b200_timer_start(ctx.h))
            tot = [a + b for a, b in zip(tot, step_resident())]
            ub.check(L.ufe_b200_timer_stop(ctx.h, C.cast(C.byref(ms), ub.c_dp)))
            ms_sum += ms.value
        ms.value = ms_sum
    else:
        ub.check(L.ufe_b200_timer_start(ctx.h))
        for _ in range(args.steps):
            tot = [a + b for a, b in zip(tot, step_resident())]
        ub.check(L.ufe_b200_timer_stop(ctx.h, C.cast(C.byref(ms), ub.c_dp)))
    rdv.barrier()
    ms_total = rdv.allreduce_max(ms.value)
    launches = ctx.kernel_launches() - launches0
    clocks = sampler.stop() if rank == 0 else None
    L.ufe_b200_timing_enable(ctx.h, 0)
    classes = {c: ctx.timing(c) for c in ("spmv_stiffness", "mg_fine", "mg_vec", "mg_coarse", "mg_setup", "krylov_vec", "assembly", "column", "tri",
                                          "epilogue", "thickness")}
    # dominant kernel family: both stream the level-0 stiffness matrix once per launch (k_spmv_blk in the Krylov recurrence,
    # k_mg_sweep = the same pass fused with the block-Jacobi update / residual inside the multigrid cycle)
    top = "mg_fine" if classes["mg_fine"][0] > classes["spmv_stiffness"][0] else "spmv_stiffness"
    spmv_ms, spmv_n = classes[top]
    log("kernel classes (ms, launches): " + ", ".join(f"{k}={v[0]:.1f}/{v[1]}" for k, v in classes.items()))

    # ---- the same steps with B200_use_initial_guess_config = .true. (each Krylov solve starts from the previous solution: the line the reference has
    # commented out, petsc_basic.f90:163).  Reported beside `value`, never as `value`.
    cfg.ksp_initial_guess_nonzero = 1
    k_warm = max(1, min(args.steps, 3))
    step_resident()
    rdv.barrier()
    ub.check(L.ufe_b200_timer_start(ctx.h))
    tot_w = [0, 0, 0]
    for _ in range(k_warm):
        tot_w = [a + b for a, b in zip(tot_w, step_resident())]
    ub.check(L.ufe_b200_timer_stop(ctx.h, C.cast(C.byref(ms), ub.c_dp)))
    ms_warm = rdv.allreduce_max(ms.value)
    cfg.ksp_initial_guess_nonzero = 0

    # ---- end-to-end leg (`e2e`): the same step through the host-buffer entry ufe_b200_pc_step -- every input the host model owns goes up
    # from page-locked host arrays and every result (thickness, rates, the whole velocity state incl. the 3-D fields) comes back, per step
    S_Hi = np.zeros(na); dH_nm1 = np.zeros(na)
    ub.check(L.ufe_b200_thickness_download(ctx.h, ub.dp(S_Hi), ub.dp(dH_nm1), None))
    ub.check(L.ufe_b200_velocity_download(ctx.h, C.byref(st)))
    out = {k: np.zeros(na) for k in ("AMB", "dHi_dt", "Hi_tplusdt", "divQ")}
    keep_io["Hi"][:] = S_Hi
    io2, _k = thickness_io(w, sl_a, keep_io["Hi"], out=out)
    for k in ("Hb", "SL", "SMB", "BMB", "LMB", "fraction_margin", "mask_noice", "dHi_dt_target"):   # the arrays that get page-locked below
        a = keep_io[k]
        setattr(io2, k, ub.ip(a) if a.dtype == np.int32 else ub.dp(a))
    io2.Hi = ub.dp(keep_io["Hi"])
    resident_state = args.e2e_state == "resident"
    st_e2e, A_e2e = st, A
    if resident_state:
        # the solver state (warm-start velocities, basal fields, eta_3D_b) stays on the device between calls, as the reference keeps it inside
        # its solver object; per step the host sends what it owns (geometry, mass balance, flow factor, friction inputs) and reads the results
        ub.check(L.ufe_b200_set_state_resident(ctx.h, 1))
        st_e2e, A_e2e = ub.alloc_velocity_state(na, nb, m.nz, fields=[k for k in STATE_FIELDS if k != "eta_3D_b"])
        for k in A_e2e:
            A_e2e[k][...] = A[k]
    pinned = list(A_e2e.values()) + list(out.values()) + [v for v in w._keep.values()] + [v for v in keep_io.values()] + [dH_nm1]
    registered = [a for a in pinned if a.nbytes >= (1 << 16) and L.ufe_b200_host_register(a.ctypes.data_as(C.c_void_p), a.nbytes) == 0]
    ice_in = [k for k in w._keep if k not in ("Hi", "Hs")]
    bytes_in = (sum(keep_io[k].nbytes for k in ("Hi", "Hb", "SL", "SMB", "BMB", "LMB", "fraction_margin", "mask_noice", "dHi_dt_target")) + dH_nm1.nbytes
                + sum(w._keep[k].nbytes for k in ice_in) + (0 if resident_state else sum(A[k].nbytes for k in STATE_FIELDS[:7])))
    bytes_out = sum(a.nbytes for a in A_e2e.values()) + sum(a.nbytes for a in out.values()) + dH_nm1.nbytes

    def step_e2e():
        ub.check(L.ufe_b200_pc_step(ctx.h, C.byref(cfg), approx, dt, dt, C.byref(ice), C.byref(st_e2e), C.byref(io2), ub.dp(dH_nm1),
                                    C.byref(nv_), C.byref(na_), C.byref(nt_)))
        keep_io["Hi"][:] = out["Hi_tplusdt"]      # the host model accepts the step: Hi_n := Hi_np1

    step_e2e()   # one untimed pass (stream creation, page faults)
    k_e2e = max(1, min(args.steps, 5))
    L.ufe_b200_timing_reset(ctx.h)
    L.ufe_b200_timing_enable(ctx.h, 2)      # times only the host<->device transfer regions
    rdv.barrier()
    t0 = time.perf_counter()
    for _ in range(k_e2e):
        step_e2e()
    e2e_s = rdv.allreduce_max(time.perf_counter() - t0)
    L.ufe_b200_timing_enable(ctx.h, 0)
    h2d_ms, d2h_ms = ctx.timing("h2d")[0] / k_e2e, ctx.timing("d2h")[0] / k_e2e
    # the same call when the host asks for the 2-D results only (NULL members of the state struct are skipped): ice%u_3D_b / v_3D_b are read by the
    # thermodynamics and the output routines, which act every dt_thermodynamics / dt_output, not every ice-dynamics step
    for k in ("u_3D_b", "v_3D_b"):
        setattr(st_e2e, k, None)
    bytes_out_2d = bytes_out - A_e2e["u_3D_b"].nbytes - A_e2e["v_3D_b"].nbytes
    step_e2e()
    rdv.barrier()
    t0 = time.perf_counter()
    for _ in range(k_e2e):
        step_e2e()
    e2e_2d_s = rdv.allreduce_max(time.perf_counter() - t0)
    for a in registered:
        L.ufe_b200_host_unregister(a.ctypes.data_as(C.c_void_p))
    bytes_in, bytes_out = [bytes_in * k_e2e], [bytes_out * k_e2e]

    res = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        csr_b, blk_b = C.c_longlong(), C.c_longlong()
        ub.check(L.ufe_b200_get_stiffness_storage_bytes(ctx.h, C.byref(csr_b), C.byref(blk_b)))
        rows_loc = 2 * nb
        # B_spmv = 12 nnz + 4 (rows + 1) + 8 rows (y) + 8 cols (x), for the rank's row block
        B_spmv = csr_b.value + 8 * rows_loc + 8 * rows_loc
        B_streamed = blk_b.value + 8 * rows_loc * 4   # blocks + y + x + Jacobi diagonal + dot-product operand
        if top == "mg_fine":
            # sweeps also read the right-hand side (8 B/row); half of them the diagonal blocks (16 B/row) and the old iterate (8 B/row)
            B_spmv += 8 * rows_loc
            B_streamed = (mg_info["fine_bytes"] if mg_info else blk_b.value) + 8 * rows_loc * 5
        kernel_name = ("k_mg_sweep level 0 (stiffness-matrix pass fused with block-Jacobi update / residual inside the multigrid cycle)" if top == "mg_fine"
                       else "k_spmv_blk (stiffness-matrix SpMV with fused Jacobi + dot products inside BiCGStab)")
        # roofline.traffic: DRAM bytes per launch of that kernel from the committed `ncu --set full` capture (profiles/r02_ncu_traffic.json), valid
        # for the mesh the capture was taken on (one rank, same triangle count); null otherwise
        traffic, traffic_src = None, None
        try:
            cap = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")))["k_mg_sweep_level0" if top == "mg_fine" else "k_spmv_blk"]
            if world == 1 and cap.get("n_triangles") == int(m.nTri) and "dram_bytes_read" in cap:
                traffic, traffic_src = int(cap["dram_bytes_read"]) + int(cap["dram_bytes_write"]), cap["report"]
        except (OSError, KeyError, ValueError):
            pass
        krylov_desc = ("BiCGStab + aggregation multigrid (%d levels %s, %d+%d block-Jacobi sweeps, omega %.2f, %s matrix storage in the preconditioner)"
                       % (mg_info["levels"], str(mg_info["nodes"]) + (f", {args.mg_coarse_sweeps} " + ("Chebyshev steps" if args.mg_coarsest_cheb else "sweeps") + " on the coarsest" + (f" in one {args.mg_coarsest_cluster}-CTA cluster kernel per visit" if args.mg_coarsest_cluster > 1 else "") if args.mg_stop_nodes > 0 else ", dense solve on the coarsest"),
                          args.mg_nu, args.mg_nu, args.mg_omega, "FP32" if args.mg_fp32 else "FP64")
                       if mg_info else "BiCGStab + Jacobi")
        avg_ms = spmv_ms / max(spmv_n, 1)
        achieved = B_spmv / (avg_ms * 1e-3) / 1e9 if spmv_n else 0.0
        res = {
            "metric": ("DIVA" if approx == 1 else "SSA") + " velocity solves/s", "value": args.steps / (ms_total * 1e-3), "unit": "velocity solves/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"configs[{conf['index']}]: {conf['text']}, "
                                   f"{m.nV} vertices / {m.nTri} triangles, nz={m.nz}, dt={dt:.4g} yr, " +
                                   ("physics of config_MISMIP_8km_spinup_for_scaling.cfg" if approx == 1 else "SSA"),
                       "name": args.config,
                       "n_vertices": int(m.nV), "partition": f"{world} contiguous row blocks (partition_list)",
                       "l2_policy": ("L2 flushed between the timed steps (512 MB written by ufe_b200_l2_flush; each step timed by its own event pair)" if flush_l2
                                     else f"inputs larger than L2 (stiffness matrix {blk_b.value / 1e9:.2f} GB >> 126 MB)"),
                       "krylov": krylov_desc + ", rtol 1e-4 / abstol 1e-3, zero initial guess (the reference's KSP semantics, petsc_basic.f90:163)",
                       "value_is": "zero initial guess; value_initial_guess_nonzero is the same step with B200_use_initial_guess_config = .true.",
                       "time_step": ("one pass of iterate_pc_timestep in the reference's order and semantics (lagged region%ice%Hi, applied rate in Eq. 30, alter_ice_thickness after Eq. 30 / 31, "
                                     "rates re-derived; ufe_b200_pc_reference_loop default) with choice_mask_noice = 'MISMIP_mod'" +
                                     (", geometry / mask / sub-grid stage on the device (bilin_interp_TAF)" if args.geometry else ", masks and sub-grid fractions as uploaded")),
                       "per_step": {"viscosity_its": tot[0] / args.steps, "krylov_its_velocity": tot[1] / args.steps, "krylov_its_thickness": tot[2] / args.steps}},
            "e2e": {"value": k_e2e / e2e_s, "unit": "velocity solves/s", "h2d_bytes_per_step": bytes_in[0] // k_e2e, "d2h_bytes_per_step": bytes_out[0] // k_e2e,
                    "steps": k_e2e, "ms_per_step": 1e3 * e2e_s / k_e2e, "h2d_ms_per_step": h2d_ms, "d2h_ms_per_step": d2h_ms,
                    "h2d_GBps": (bytes_in[0] / k_e2e) / (h2d_ms * 1e6) if h2d_ms else None, "d2h_GBps": (bytes_out[0] / k_e2e) / (d2h_ms * 1e6) if d2h_ms else None,
                    "numa_node_bound": numa_node,
                    "without_3D_results": {"value": k_e2e / e2e_2d_s, "ms_per_step": 1e3 * e2e_2d_s / k_e2e, "d2h_bytes_per_step": int(bytes_out_2d),
                                           "note": "u_3D_b / v_3D_b left on the device (NULL in the state struct); the host reads them when thermodynamics or output act"},
                    "call": "ufe_b200_pc_step (host buffers; transfers overlap the thickness solves on a second stream)",
                    "state": ("solver state resident on the device (ufe_b200_set_state_resident); per step in: geometry, mass balance, flow factor, "
                              "friction inputs; out: Hi(t+dt), rates, u/v_3D_b, depth-averaged and basal velocities, basal stresses")
                    if args.e2e_state == "resident" else "whole solver state (incl. eta_3D_b) uploaded and downloaded every step"},
            "value_initial_guess_nonzero": {"value": k_warm / (ms_warm * 1e-3), "ms_per_step": ms_warm / k_warm, "steps": k_warm,
                                            "krylov_its_velocity": tot_w[1] / k_warm, "krylov_its_thickness": tot_w[2] / k_warm},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak if peak else None, "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback",
                         "algorithmic_bytes_per_launch": B_spmv, "streamed_bytes_per_launch": B_streamed,
                         "streamed_GBps": B_streamed / (avg_ms * 1e-3) / 1e9 if spmv_n else 0.0, "avg_launch_ms": avg_ms, "launches": spmv_n, "traffic": traffic, "traffic_source": traffic_src,
                         "share_of_step": spmv_ms / ms_total if ms_total else None, "frac_of_8TBs": achieved / 8000.0},
            "kernel_classes_ms": {k: v[0] for k, v in classes.items()},
            "clocks": clocks,
        }
    ctx.close()
    return res


def run_sweep(args, rank, world, rdv, log):
    """configs[4]: the Krylov solve of the DIVA system of the FIRST viscosity iteration (velocities zero, viscosity from the
    initial strain-rate guess), zero initial guess, rtol 1e-7, on the mesh generator of configs[2] at --nv vertices, row blocks over `world` GPUs.
    One step = upload the zero state (untimed), then ufe_b200_velocity_solve_resident with visc_it_nit = 1 (timed: assembly of the first iteration,
    preconditioner set-up, the Krylov solve, epilogue)."""
    import ufe_b200 as ub
    L = ub.lib()
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    bind_to_gpu_numa_node(local_rank, log)
    uid = rdv.bcast(ub.Context.unique_id() if rank == 0 else None, 128) if world > 1 else None
    ctx = ub.Context(device=local_rank, rank=rank, nranks=world, unique_id=uid)
    w = make_workload(args.nv, args.pc, "sweep")
    m, cfg = w.mesh, w.cfg
    cfg.visc_it_nit = 1
    cfg.stress_balance_PETSc_rtol, cfg.stress_balance_PETSc_abstol = 1e-7, 1e-30
    cfg.ksp_initial_guess_nonzero = 0
    if args.pc == "mg":
        ctx.set_mg_params(nu_pre=args.mg_nu, nu_post=args.mg_nu, omega=args.mg_omega, overcorr=args.mg_overcorr, w_levels=args.mg_w_levels, fp32=args.mg_fp32)
        if args.mg_stop_nodes > 0:
            ctx.set_mg_option("stop_nodes", args.mg_stop_nodes); ctx.set_mg_option("coarse_sweeps", args.mg_coarse_sweeps)
            ctx.set_mg_option("coarsest_cluster", args.mg_coarsest_cluster); ctx.set_mg_option("coarsest_dsmem", args.mg_coarsest_dsmem)
            ctx.set_mg_option("coarsest_cheb", args.mg_coarsest_cheb); ctx.set_mg_option("cheb_ratio", args.mg_cheb_ratio)
    ctx.mesh_create(m)
    ctx.calc_matrix_operators()
    vr, tr = (ctx.vi1, ctx.vi2), (ctx.ti1, ctx.ti2)
    na, nb = vr[1] - vr[0] + 1, tr[1] - tr[0] + 1
    ice = w.ice_inputs(vr, tr)
    st, A = ub.alloc_velocity_state(na, nb, m.nz, fields=STATE_FIELDS + ["u_b_prev", "v_b_prev"])
    nv_, na_ = C.c_int(), C.c_int()
    ms = C.c_double()

    def step(timed):
        ub.check(L.ufe_b200_velocity_upload(ctx.h, C.byref(ice), C.byref(st)))          # zero velocities again: the same system every step
        if timed:
            ub.check(L.ufe_b200_timer_start(ctx.h))
        ub.check(L.ufe_b200_velocity_solve_resident(ctx.h, C.byref(cfg), 1, C.byref(nv_), C.byref(na_)))
        if timed:
            ub.check(L.ufe_b200_timer_stop(ctx.h, C.cast(C.byref(ms), ub.c_dp)))
        return ms.value, na_.value

    for _ in range(args.warmup):
        step(False)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    L.ufe_b200_timing_reset(ctx.h)
    L.ufe_b200_timing_enable(ctx.h, 1)
    launches0 = ctx.kernel_launches()
    rdv.barrier()
    ms_sum, its = 0.0, 0
    for _ in range(args.steps):
        t, k = step(True)
        ms_sum += t; its += k
    rdv.barrier()
    ms_total = rdv.allreduce_max(ms_sum)
    launches = ctx.kernel_launches() - launches0
    clocks = sampler.stop() if rank == 0 else None
    L.ufe_b200_timing_enable(ctx.h, 0)
    classes = {c: ctx.timing(c) for c in ("spmv_stiffness", "mg_fine", "mg_vec", "mg_coarse", "mg_setup", "krylov_vec", "assembly", "column", "tri", "epilogue")}
    krylov_ms = sum(classes[c][0] for c in ("spmv_stiffness", "mg_fine", "mg_vec", "mg_coarse", "krylov_vec"))
    res = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        csr_b, blk_b = C.c_longlong(), C.c_longlong()
        ub.check(L.ufe_b200_get_stiffness_storage_bytes(ctx.h, C.byref(csr_b), C.byref(blk_b)))
        top = "mg_fine" if classes["mg_fine"][0] > classes["spmv_stiffness"][0] else "spmv_stiffness"
        B_spmv = csr_b.value + 16 * 2 * nb + (8 * 2 * nb if top == "mg_fine" else 0)
        t_ms, t_n = classes[top]
        avg_ms = t_ms / max(t_n, 1)
        achieved = B_spmv / (avg_ms * 1e-3) / 1e9 if t_n else 0.0
        res = {"metric": "DIVA Krylov solves/s", "value": args.steps / (ms_total * 1e-3), "unit": "linear solves/s", "n_gpus": world, "steps": args.steps,
               "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
               "data": "synthetic",
               "config": {"workload": f"configs[4]: {CONFIGS['sweep']['text']}, {m.nV} vertices / {m.nTri} triangles, nz={m.nz}", "name": "sweep",
                          "n_vertices": int(m.nV), "partition": f"{world} contiguous row blocks (partition_list)",
                          "l2_policy": f"inputs larger than L2 (stiffness matrix {blk_b.value / 1e9:.2f} GB per rank)",
                          "krylov": ("BiCGStab + aggregation multigrid" if args.pc == "mg" else "BiCGStab + Jacobi") + ", rtol 1e-7, zero initial guess",
                          "per_step": {"krylov_its": its / args.steps, "krylov_ms": krylov_ms / args.steps,
                                       "ms_per_krylov_it": krylov_ms / max(its, 1)}},
               "e2e": None, "gpu_launches": int(launches),
               "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if peak else None,
                            "avg_launch_ms": avg_ms, "launches": t_n, "algorithmic_bytes_per_launch": B_spmv, "traffic": None},
               "kernel_classes_ms": {k: v[0] for k, v in classes.items()}, "clocks": clocks}
    ctx.close()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="antarctica", choices=sorted(CONFIGS), help="BASELINE.json config: antarctica = configs[2] (the driver's line), "
                    "mismip8km = configs[0], mismipplus_ssa = configs[1], sweep = configs[4] (one mesh size per call: --nv)")
    ap.add_argument("--nv", type=int, default=0, help="target number of vertices (0: the config's own, ~4M for configs[2])")
    ap.add_argument("--cpu-nv", type=int, default=250_000, help="vertices of the bounded CPU sample")
    ap.add_argument("--pc", default="mg", choices=["jacobi", "mg"], help="preconditioner of the stiffness-matrix solve on the device")
    ap.add_argument("--mg-nu", type=int, default=1)
    ap.add_argument("--mg-omega", type=float, default=0.8)
    ap.add_argument("--mg-overcorr", type=float, default=1.5)
    ap.add_argument("--mg-w-levels", type=int, default=3)
    ap.add_argument("--mg-fp32", type=int, default=0)
    ap.add_argument("--mg-coarsest-cheb", type=int, default=1, help="1: the corrections on the coarsest level are Chebyshev steps (k_mg_coarsest_cluster) instead of stationary sweeps")
    ap.add_argument("--mg-cheb-ratio", type=float, default=30.0)
    ap.add_argument("--mg-coarsest-dsmem", type=int, default=0, help="iterate of the coarsest-level cluster kernel in distributed shared memory (1) or through L2 (0)")
    ap.add_argument("--mg-stop-nodes", type=int, default=-1, help="stop coarsening at the first level with at most this many nodes and run --mg-coarse-sweeps "
                    "sweeps there (ufe_b200_set_mg_option); -1: 2000 for the MISMIP_mod meshes of configs[2] / [4], 0 (library default: coarsen down to the dense solve) otherwise")
    ap.add_argument("--mg-coarse-sweeps", type=int, default=-1, help="corrections on the coarsest level after the first one (-1: 8 Chebyshev steps, or 32 stationary sweeps with --mg-coarsest-cheb 0)")
    ap.add_argument("--mg-coarsest-cluster", type=int, default=-1, help="> 1: the sweeps on the coarsest level as one thread-block-cluster kernel of this many CTAs "
                    "(-1: 16 whenever --mg-stop-nodes is in effect)")
    ap.add_argument("--e2e-state", choices=["resident", "roundtrip"], default="resident",
                    help="e2e leg: solver state kept on the device between calls (default) or uploaded and downloaded every step")
    ap.add_argument("--geometry", type=int, default=1, help="1: masks / grounded fractions / effective thickness follow the thickness on the device inside the step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    args.nv = args.nv or CONFIGS[args.config]["nv"]
    if args.mg_stop_nodes < 0:
        # measured on the 4 M-vertex MISMIP_mod mesh (profiles/r02_mg_probe_stop_4M.log): 46 Krylov iterations against 63 and 17 % less solve time than
        # the full hierarchy; the graded MISMIP+ mesh and the 72 k-vertex mesh prefer the library default (r02_mg_probe_stop_{mplus,72k}.log)
        args.mg_stop_nodes = 2000 if args.config in ("antarctica", "sweep") else 0
    if args.mg_coarsest_cluster < 0:
        # 4 M vertices (profiles/r02_mg_probe_coarsest_cluster_4M.log, r02_bench_4_cluster16_cs32.log): 32 sweeps in one 16-CTA cluster kernel per visit cost
        # what 24 sweeps cost as graph nodes and save 2 more Krylov iterations per step (150.0 against 165.6 ms per step)
        args.mg_coarsest_cluster = 16 if args.mg_stop_nodes > 0 else 0
    if args.mg_coarsest_cluster <= 1:
        args.mg_coarsest_cheb = 0          # the Chebyshev steps live in the cluster kernel
    if args.mg_coarse_sweeps < 0:
        # 8 Chebyshev steps for [lmax / 30, lmax] do what 32 stationary sweeps do (42-43 Krylov iterations in the cold probe either way,
        # profiles/r02_mg_probe_cheb_4M.log) at 0.70 instead of 1.11 ms per cycle: the same step time on one GPU (150.0 ms, r02_bench_6_cheb8.log),
        # less of the replicated coarse graph on several
        args.mg_coarse_sweeps = 8 if args.mg_coarsest_cheb else 32

    def log(*a):
        if args.verbose or os.environ.get("UFE_BENCH_VERBOSE"):
            print(f"[rank {rank}]", *a, file=sys.stderr, flush=True)

    if args.impl == "reference":
        # The reference itself (Fortran + PETSc + MPI) cannot be built on this box; its CPU stand-in is the oracle port,
        # run by rank 0 alone on all host cores, on a bounded sample of the workload.
        if rank != 0:
            return
        nv_full = lattice_vertex_count(args.nv)     # the b200 arm's mesh size, without generating that mesh
        cb = run_cpu_sample(args.cpu_nv, nv_full, max(1, min(args.steps, 2)), min(args.warmup, 1), log)
        line = {"impl": "reference", "metric": "DIVA velocity solves/s", "value": cb["value"], "unit": "velocity solves/s", "n_gpus": args.gpus, "same_config": False,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / cb["value"], "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"configs[2] (see b200 arm), CPU sample: {cb['sample']}", "n_vertices": nv_full},
                "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": "velocity solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line), flush=True)
        return

    rdv = Rendezvous(rank, world)
    if args.config == "sweep":
        res = run_sweep(args, rank, world, rdv, log)
        if rank == 0:
            print(json.dumps(res), flush=True)
        return
    res = run_b200(args, rank, world, rdv, log)
    if rank == 0:
        if not args.no_cpu_baseline and world == 1 and args.config == "antarctica":
            res["cpu_baseline"] = run_cpu_sample(args.cpu_nv, res["config"]["n_vertices"], 2, 1, log)
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
