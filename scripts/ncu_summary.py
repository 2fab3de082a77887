"""Markdown table of an `ncu --set full` report: per captured kernel the duration, DRAM bytes, achieved DRAM bandwidth against the measured copy peak,
and -- where SURVEY 8(d) / DESIGN.md 3 give a per-unit figure -- the algorithmic bytes and the traffic / algorithmic ratio.

  python scripts/ncu_summary.py gpurun_out/X.ncu-rep profiles/rNN_X_ncu.md "<command the capture came from>" [nV nTri nz]
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

rep, out, cmd = sys.argv[1:4]
nV, nTri, nz = (int(x) for x in sys.argv[4:7]) if len(sys.argv) > 6 else (0, 0, 12)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except (OSError, KeyError):
    peak = 6650.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}

# algorithmic bytes per launch (DESIGN.md 3 / SURVEY 8d), keyed by a substring of the kernel name
R = 2 * nTri
ALGO = {
    "k_vertex_diva": ("648 B / vertex (6 strain fields, A_flow, eta_3D_a, F1, F2 at nz = 12)", 648 * nV),
    "k_tri_assemble": ("836 B / triangle (5 operator rows, pattern, coefficients; 2 x 20 values out)", 836 * nTri),
    "k_th_flux": ("400 B / vertex (C, VE, Cw, D_x, D_y, D, edge velocities, 12 fields)", 400 * nV),
    "k_spmv_blk": ("12 nnz + 4 (R + 1) + 16 R, nnz ~ 20 R (the CSR figure of 8d)", 260 * R),
    "k_mg_sweep<double, 1, 2>": ("the same matrix pass + right-hand side", 268 * R),
    "k_mg_sweep<double, 0, 2>": ("the same matrix pass + right-hand side", 268 * R),
    "k_spmv_fused<CsrOp<8>": ("12 nnz + 4 (n + 1) + 16 n, thickness matrix nnz ~ 7 n", (12 * 7 + 20) * nV),
    "k_spmv_fused<SellOp": ("12 nnz + 4 (n + 1) + 16 n", None),
    "k_ks_update_xr": ("7 vector passes x 8 B x R", 56 * R),
    "k_ksmg_dot": ("2 vector passes x 8 B x R", 16 * R),
    "k_epilogue": ("13 b-grid fields x 8 B", 13 * 8 * nTri),
}


def num(r, name, scale=1.0):
    if name not in col or r[col[name]] in ("", "n/a"):
        return None
    v = float(r[col[name]].replace(",", ""))
    u = units[col[name]]
    mult = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0, "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9,
            "second": 1.0}.get(u, 1.0)
    return v * mult * scale


lines = [f"# ncu `--set full` summary of `{os.path.basename(rep)}`", "", f"Command: `{cmd}`", "",
         f"Mesh: {nV} vertices / {nTri} triangles, nz = {nz}.  Peak = measured copy bandwidth {peak:.0f} GB/s (MEASURED_PEAKS.json).  One capture per kernel "
         "(first launch inside the timed step), clocks not locked (`--clock-control none`).  `traffic` = dram__bytes_read.sum + dram__bytes_write.sum.", "",
         "| kernel | grid x block | regs | duration | DRAM read | DRAM write | DRAM GB/s | of peak | algorithmic bytes | algorithmic GB/s (of peak) | traffic / algorithmic | "
         "warps active | L1 hit | L2 hit | instructions |", "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
for r in data:
    name = re.sub(r"\(.*$", "", r[col["Kernel Name"]])
    name = re.sub(r"^void\s+|.*unnamed>::|ufe::", "", name)
    t = num(r, "gpu__time_duration.sum")
    rd, wr = num(r, "dram__bytes_read.sum"), num(r, "dram__bytes_write.sum")
    bw = (rd + wr) / t / 1e9
    algo = next(((d, b) for k, (d, b) in ALGO.items() if k in r[col["Kernel Name"]] and b), None)
    a_txt = a_bw = ratio = "-"
    if algo and nV:
        a_txt = f"{algo[1] / 1e9:.3f} GB ({algo[0]})"
        a_bw = f"{algo[1] / t / 1e9:.0f} ({algo[1] / t / 1e9 / peak:.2f})"
        ratio = f"{(rd + wr) / algo[1]:.2f}"
    g = lambda n: r[col[n]] if n in col else "-"
    lines.append(f"| `{name}` | {g('launch__grid_size')} x {g('launch__block_size')} | {g('launch__registers_per_thread')} | {t * 1e3:.3f} ms | {rd / 1e9:.3f} GB | "
                 f"{wr / 1e9:.3f} GB | {bw:.0f} | {bw / peak:.2f} | {a_txt} | {a_bw} | {ratio} | {float(g('sm__warps_active.avg.pct_of_peak_sustained_active')):.0f} % | "
                 f"{float(g('l1tex__t_sector_hit_rate.pct')):.0f} % | {float(g('lts__t_sector_hit_rate.pct')):.0f} % | {float(g('smsp__inst_executed.sum')) / 1e6:.1f} M |")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
