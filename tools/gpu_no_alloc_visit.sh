#!/bin/bash
# Last GPU visit of round 2: the allocation-free entry point's test alone in a fresh process and after other tests in
# one process, one ncu --set full capture of the hot kernels summarised on the box (the report is too big to pull),
# then the whole GPU suite with what time is left.
TAG=${1:-r02_final}
COMMIT=${2:-unknown}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/${TAG}_no_alloc_alone.log 2>&1
echo "no_alloc alone exit $?: $(tail -1 gpurun_out/${TAG}_no_alloc_alone.log)"
timeout 400 python -m pytest tests/test_loss_gpu.py tests/test_regularizer_gpu.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/${TAG}_no_alloc_after.log 2>&1
echo "no_alloc after loss tests exit $?: $(tail -1 gpurun_out/${TAG}_no_alloc_after.log)"
timeout 600 ncu --set full --clock-control none \
  -k regex:'render_bwd_kernel|render_fwd_kernel|preprocess_bwd_kernel|preprocess_fwd_kernel|tile_sort_kernel|scatter_kernel|tile_scan_kernel' \
  --launch-skip 16 --launch-count 10 -o gpurun_out/${TAG}_full -f \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-knn --no-render --no-drop-in > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
GSM_COMMIT=$COMMIT python tools/ncu_summary.py gpurun_out/${TAG}_full.ncu-rep gpurun_out/${TAG}_ncu.txt \
  "ncu --set full --clock-control none, python bench.py --steps 2 --warmup 1 (C4), commit $COMMIT: per launch" --traffic
rm -f gpurun_out/${TAG}_full.ncu-rep
timeout ${3:-600} python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?: $(tail -1 gpurun_out/${TAG}_pytest.log)"
ls -la gpurun_out/ | grep ${TAG}
