"""Turn `ncu --csv --metrics gpu__time_duration.sum` output into the compact launch list and per-kernel shares kept under profiles/.

  python scripts/ncu_launch_list.py gpurun_out/launches_raw.csv profiles/rNN_launches_X.csv profiles/rNN_launch_shares_X.txt "<command line>" [step_ms_without_ncu]
"""
import csv
import re
import sys
from collections import OrderedDict

raw, out_csv, out_txt, cmd = sys.argv[1:5]
step_ms = float(sys.argv[5]) if len(sys.argv) > 5 else None
rows = []
with open(raw, newline="") as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"]
    name = re.sub(r"\(.*$", "", name)                      # drop the argument list
    name = re.sub(r"^void\s+", "", name)
    name = re.sub(r"^.*unnamed>::", "", name)
    name = re.sub(r"ufe::\(anonymous namespace\)::|ufe::<unnamed>::|<unnamed>::|ufe::", "", name)
    val = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    ns = val * {"ns": 1.0, "nsecond": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "s": 1e9, "second": 1e9}.get(unit, 1.0)
    grid = int(re.sub(r"[^0-9]", " ", r["Grid Size"]).split()[0]); block = int(re.sub(r"[^0-9]", " ", r["Block Size"]).split()[0])
    rows.append((name, grid, block, ns))
with open(out_csv, "w") as f:
    f.write(f"# {cmd}\n# cold-cache, serialised launches under ncu: compare shares, not absolute times\nid,kernel,grid,block,duration_ns\n")
    for i, (n, g, b, ns) in enumerate(rows):
        f.write(f"{i},{n},{g},{b},{ns:.0f}\n")
agg = OrderedDict()
for n, g, b, ns in rows:
    key = n
    if n.startswith("k_mg_sweep"):   # U = 2 is the level-0 instantiation, U = 8 the coarse levels' (mg.cu: launch_sweep)
        key = n + (" [level 0]" if n.rstrip().endswith(", 2>") else " [levels >= 1, inside the CUDA graph]")
    elif n.startswith("k_mg_restrict") or n.startswith("k_mg_prolong") or n.startswith("k_mg_smooth0"):
        key = n + " [all levels]"
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1; a[1] += ns
tot = sum(a[1] for a in agg.values())
with open(out_txt, "w") as f:
    f.write(f"{len(rows)} launches, {tot / 1e6:.1f} ms of kernel time under ncu" + (f" (bench.py's CUDA-event time for the same region without ncu: {step_ms:.1f} ms)" if step_ms else "") + "\n")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{k:72s} n={n:6d} total={ns / 1e6:9.2f} ms avg={ns / n / 1e3:9.1f} us share={ns / tot:.4f}\n")
print(open(out_txt).read()[:1500])
