#!/bin/bash
# Final single-GPU visit of round 2: all GPU tests, both bench arms, ncu launch lists of both arms, ncu --set full of
# the hot kernels and of the auxiliary kernels.
TAG=${1:-r02_final}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
if [ "$2" != "skip-tests" ]; then
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?: $(tail -1 gpurun_out/${TAG}_pytest.log)"
fi
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
echo "ref bench exit $?"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
  --log-file gpurun_out/${TAG}_reference_launches.csv python bench.py --impl reference --steps 2 --warmup 1 --no-e2e --no-knn --no-render --no-drop-in \
  > gpurun_out/${TAG}_ncu_ref_launches.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-knn --no-render --no-drop-in \
  > gpurun_out/${TAG}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'render_bwd_kernel|render_fwd_kernel|preprocess_bwd_kernel|preprocess_fwd_kernel|tile_sort_kernel|scatter_kernel|tile_scan_kernel' \
  --launch-skip 16 --launch-count 10 -o gpurun_out/${TAG}_full -f \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-knn --no-render --no-drop-in > gpurun_out/${TAG}_ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none \
  -k regex:'knn_|sort32_|morton|minmax|tree|sh_grad_expand|densify_|tile_scan' --launch-count 40 -o gpurun_out/${TAG}_aux -f \
  python tools/aux_kernels.py > gpurun_out/${TAG}_ncu_aux.log 2>&1
# gpurun brings back at most 64 MiB: the auxiliary capture is summarised on the box and its report dropped
python tools/ncu_summary.py gpurun_out/${TAG}_aux.ncu-rep gpurun_out/${TAG}_aux_ncu.txt "ncu --set full --clock-control none, python tools/aux_kernels.py (distCUDA2 on 5 M points, SH expand 8 views, 4K tile scan, densify_and_prune 2 M): per launch" > /dev/null 2>&1
rm -f gpurun_out/${TAG}_aux.ncu-rep
ls -la gpurun_out/ | grep ${TAG}
