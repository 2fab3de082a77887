#!/bin/bash
# Round 2, GPU call 3 (one GPU): the gather kernels of the viscosity iteration in their variants, the multigrid hierarchy options, the GPU suite on the
# rebuilt library, PCIe / NUMA facts of the box, the default bench line and its ncu launch list.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
{ nvidia-smi -q -d PCIE | grep -A3 -i "GPU Link Info\|Generation\|Link Width" | head -40; nvidia-smi topo -m; ls /sys/devices/system/node/ | head; \
  cat /sys/bus/pci/devices/*/numa_node 2>/dev/null | sort | uniq -c; nproc; } > $O/r02_box_pcie_numa.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -x -q --durations=10 > $O/r02_pytest_gpu_3.log 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_3.log
H=$(python -c "import bench; print(bench.h_for_vertices(4000000))")
timeout 900 python scripts/perf_probe.py $H 3 > $O/r02_perf_probe_variants_4M.log 2>&1; echo "perf probe rc=$?"
timeout 900 python scripts/mg_probe.py 4000000 3 "1,3,1.5,0.8,0;1,3,1.5,0.8,0,max_levels=5,coarse_sweeps=8;1,3,1.5,0.8,0,max_levels=5,coarse_sweeps=16;1,3,1.5,0.8,0,max_levels=4,coarse_sweeps=16;1,2,1.5,0.8,0,max_levels=4,coarse_sweeps=16;1,3,1.5,0.8,0,max_levels=6" \
  > $O/r02_mg_probe_hierarchy_4M.log 2>&1; echo "mg probe rc=$?"
timeout 1200 python bench.py > $O/r02_bench_2.log 2> $O/r02_bench_2.err; echo "bench rc=$?"
ls -la $O
