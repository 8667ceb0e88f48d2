#!/bin/bash
# One GPU-box visit of round 2: parity tests (all, no -x), both bench arms, ncu launch lists of BOTH arms (the
# reference's CUB stages next to ours: VERDICT r01 item 4a), one ncu --set full capture of the hot kernels.
# Usage (repo root on the box): bash tools/gpu_round2.sh <tag> [skip-tests] [skip-ncu]
TAG=${1:-r02}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
if [ "$2" != "skip-tests" ]; then
  timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/${TAG}_pytest.log 2>&1
  echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
  tail -40 gpurun_out/${TAG}_pytest.log | cut -c1-300
fi
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
echo "ref bench exit $?"; tail -3 gpurun_out/${TAG}_bench_ref.err | cut -c1-300
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"; tail -3 gpurun_out/${TAG}_bench.err | cut -c1-300
python - <<PY
import json
for f in ("gpurun_out/${TAG}_bench.json", "gpurun_out/${TAG}_bench_ref.json"):
    try:
        j = json.load(open(f))
        print(f, "value %.1f ms %.3f e2e %s" % (j["value"], j["ms_per_step"], j.get("e2e", {}) and "%.1f" % j["e2e"]["value"]),
              "drop_in", (j.get("drop_in") or {}).get("ms_per_step"), "train", (j.get("train_iteration") or {}).get("ms_per_step"),
              "api", (j.get("value_public_api") or {}).get("ms_per_step"))
        for r in j.get("stages", []):
            print("   %-16s %.3f ms  frac %.3f" % (r["kernel"], r["ms"], r["frac"]))
        print("   render", [(r["width"], round(r["fps"], 1)) for r in j.get("render", [])], "knn", (j.get("knn") or {}).get("ms"))
    except Exception as e:
        print(f, "FAILED", e)
PY
if [ "$3" != "skip-ncu" ]; then
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
fi
ls -la gpurun_out/ | tail -20
