#!/bin/bash
# ncu --set full of the widening-row kernels (loss, Adam, regulariser) inside one bench run.  Usage: bash tools/gpu_ncu_aux.sh <tag>
TAG=${1:-aux}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'ssim_forward_kernel|ssim_backward_kernel|adam_step_kernel|regularizer_forward_kernel|regularizer_backward_kernel' \
  --launch-skip 20 --launch-count 6 -o gpurun_out/${TAG}_aux -f \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-render > gpurun_out/${TAG}_ncu_aux.log 2>&1
ls -la gpurun_out/${TAG}_aux.ncu-rep
