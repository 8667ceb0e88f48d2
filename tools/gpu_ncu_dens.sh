#!/bin/bash
# ncu --set full of the densification-bookkeeping kernel inside one bench run.  Usage: bash tools/gpu_ncu_dens.sh <tag>
TAG=${1:-dens}
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'densification_stats_kernel' \
  --launch-skip 5 --launch-count 2 -o gpurun_out/${TAG}_dens -f \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-render > gpurun_out/${TAG}_ncu_dens.log 2>&1
echo "ncu exit $?"; ls -la gpurun_out/${TAG}_dens.ncu-rep
