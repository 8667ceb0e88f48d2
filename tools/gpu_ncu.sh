#!/bin/bash
# ncu --set full capture of selected kernels of one variant library.
# Usage: bash tools/gpu_ncu.sh <outtag> <variant|default> <kernel-regex> [launch-skip] [launch-count]
OUT=$1; V=$2; K=$3; SKIP=${4:-2}; CNT=${5:-1}
lib=gaussmart_b200/lib/libgaussmart_b200_$V.so
[ "$V" = "default" ] && lib=gaussmart_b200/lib/libgaussmart_b200.so
mkdir -p gpurun_out
GAUSSMART_B200_LIB=$PWD/$lib timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:"$K" --launch-skip $SKIP --launch-count $CNT -o gpurun_out/${OUT} -f \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-knn --no-render > gpurun_out/${OUT}_ncu.log 2>&1
tail -3 gpurun_out/${OUT}_ncu.log
ncu -i gpurun_out/${OUT}.ncu-rep --page source --csv > gpurun_out/${OUT}_source.csv 2>/dev/null
python tools/sass_profile.py gpurun_out/${OUT}_source.csv --bands
