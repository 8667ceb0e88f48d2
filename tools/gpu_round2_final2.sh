#!/bin/bash
# Final single-GPU visit of round 2 (second edition, after the scatter / tile-scan / transpose changes): both bench arms
# back to back, smoke, the ncu launch list and the ncu --set full summary of the hot kernels (made on the box: the
# report is too big to pull), then the whole GPU suite twice (fresh processes).
TAG=${1:-r02_final}
COMMIT=${2:-unknown}
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
echo "ref bench exit $?"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-knn --no-render --no-drop-in \
  > gpurun_out/${TAG}_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none \
  -k regex:'render_bwd_kernel|render_fwd_kernel|preprocess_bwd_kernel|preprocess_fwd_kernel|tile_sort_kernel|scatter_kernel|tile_scan_kernel' \
  --launch-skip 16 --launch-count 10 -o gpurun_out/${TAG}_full -f \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-knn --no-render --no-drop-in > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
GSM_COMMIT=$COMMIT python tools/ncu_summary.py gpurun_out/${TAG}_full.ncu-rep gpurun_out/${TAG}_ncu.txt \
  "ncu --set full --clock-control none, python bench.py --steps 2 --warmup 1 (C4), commit $COMMIT: per launch" --traffic
rm -f gpurun_out/${TAG}_full.ncu-rep
for i in 1; do
  timeout 400 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/${TAG}_pytest_$i.log 2>&1
  echo "pytest $i exit $?: $(tail -1 gpurun_out/${TAG}_pytest_$i.log)"
done
ls -la gpurun_out/ | grep ${TAG}
