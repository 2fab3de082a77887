"""The CPU legs of bench.py (`cpu_baseline`, `--impl reference`): the bounded sample runs the oracle through the same predictor-corrector call sequence the
GPU arm times, and the reference arm prints one JSON line with the contract's keys.  (The GPU arm needs a GPU; its line is checked on the box.)"""
import json
import os
import subprocess
import sys

import bench

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpu_sample_runs_the_step_and_scales_to_the_full_workload():
    r = bench.run_cpu_sample(3000, 4005450, steps=1, warmup=0)
    assert r["kind"] == "port" and r["unit"] == "velocity solves/s" and r["cores"] >= 1
    assert 2000 < r["n_vertices_sample"] < 4500 and r["ms_per_step_sample"] > 0.0
    # value = sample rate scaled linearly in vertices to the full workload
    assert abs(r["value"] - (1e3 / r["ms_per_step_sample"]) * r["n_vertices_sample"] / 4005450) < 1e-12 * r["value"] + 1e-15
    assert "viscosity iteration" in r["sample"] and "Krylov its" in r["sample"]
    # both Krylov settings were timed: the reference's default (GMRES(30) + block Jacobi / ILU(0), one block per core) and the device's one-level method
    assert set(r["solvers"]) == {"bicgstab_jacobi", "gmres_bjacobi_ilu0"} and r["solver"] in r["solvers"] and r["same_config"] is False
    assert abs(r["extrapolation_factor"] - r["n_vertices_sample"] / 4005450) < 1e-15
    for v in r["solvers"].values():
        assert v["viscosity_its_per_step"] == 1.0 and v["krylov_its_velocity_per_step"] > 10


def test_cpu_sample_uses_all_cores_under_torchrun():
    """torch.distributed.run exports OMP_NUM_THREADS=1 to its children: the CPU arm must not inherit it (round-1 SCALE ratios were void because it did)"""
    code = ("import os, sys; sys.path.insert(0, %r); import bench; r = bench.run_cpu_sample(1500, 4005450, 1, 0, solvers=('bicgstab_jacobi',)); "
            "print('CORES', r['cores'], len(os.sched_getaffinity(0)))" % ROOT)
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert p.returncode == 0, p.stderr[-2000:]
    used, avail = [int(x) for x in p.stdout.split("CORES")[1].split()[:2]]
    assert used == avail


def test_reference_arm_prints_the_contract_line():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--cpu-nv", "3000"],
                       capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config",
              "cpu_baseline", "e2e", "gpu_launches"):
        assert k in j, k
    assert j["impl"] == "reference" and j["gpu_launches"] == 0 and j["dtype"] == "f64" and j["data"] == "synthetic" and j["vs_baseline"] is None
    assert j["e2e"] == {"value": j["value"], "unit": j["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert j["cpu_baseline"]["value"] == j["value"] and j["cpu_baseline"]["kind"] == "port" and "workload" in j["config"]


def test_reference_arm_other_ranks_exit_without_work():
    """under torchrun (N > 1) rank 0 alone runs and prints the reference arm; the other ranks exit 0 silently, without waiting for a rendezvous"""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--cpu-nv", "3000"],
                       capture_output=True, text=True, timeout=120, env=env)
    assert p.returncode == 0 and p.stdout.strip() == "", (p.returncode, p.stdout[-500:], p.stderr[-500:])


def test_every_bench_config_builds_its_workload():
    """bench.py --config: the workloads of the other BASELINE configs (mesh + fields + physics), without a GPU"""
    import bench
    w = bench.make_workload(0, "mg", "mismip8km")
    assert 60_000 < w.mesh.nV < 90_000 and w.approximation == 1 and w.f["mask_noice"].sum() > 0          # +-1000 km at 8 km, MISMIP_mod no-ice mask
    w = bench.make_workload(20_000, "mg", "mismipplus_ssa")
    assert w.approximation == 0 and abs(w.mesh.xmax - 800e3) < 1.0 and w.cfg.stress_balance_PETSc_rtol == 1e-7
    w = bench.make_workload(20_000, "mg", "sweep")
    assert w.approximation == 1 and w.cfg.ksp_initial_guess_nonzero == 0
    assert set(bench.CONFIGS) == {"antarctica", "mismip8km", "mismipplus_ssa", "sweep"} and sorted(c["index"] for c in bench.CONFIGS.values()) == [0, 1, 2, 4]
