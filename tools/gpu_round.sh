#!/bin/bash
# One GPU-box visit: parity tests, both bench arms, ncu launch list + full capture of the top kernels.
# Usage (from the repo root on the box): bash tools/gpu_round.sh <tag> [skip-tests]
TAG=${1:-r01}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
if [ "$2" != "skip-tests" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
  echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
  tail -5 gpurun_out/${TAG}_pytest.log
fi
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"
cut -c1-1500 gpurun_out/${TAG}_bench.json
cut -c1-600 gpurun_out/${TAG}_bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-knn --no-render \
  > gpurun_out/${TAG}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'render_bwd_kernel|render_fwd_kernel|preprocess_bwd_kernel|preprocess_fwd_kernel|tile_sort_kernel|scatter_kernel' \
  --launch-skip 14 --launch-count 9 -o gpurun_out/${TAG}_full -f \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-knn --no-render > gpurun_out/${TAG}_ncu_full.log 2>&1
ls -la gpurun_out/
