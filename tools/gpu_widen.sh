#!/bin/bash
# Widening-rows visit: loss / Adam / regulariser parity tests, smoke, both bench arms (no ncu).  Usage: bash tools/gpu_widen.sh <tag>
TAG=${1:-widen}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_loss_gpu.py tests/test_adam_gpu.py tests/test_regularizer_gpu.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
tail -25 gpurun_out/${TAG}_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 6 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"
tail -n 3 gpurun_out/${TAG}_bench.err gpurun_out/${TAG}_bench_ref.err
python - <<PY
import json
for f in ("gpurun_out/${TAG}_bench.json", "gpurun_out/${TAG}_bench_ref.json"):
    try:
        l = [x for x in open(f) if x.startswith("{")][-1]
        d = json.loads(l)
        print(f, d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "train", d.get("train_iteration"), "loss", d.get("loss", {}).get("ms"), "adam", d.get("adam", {}).get("ms"))
    except Exception as e:
        print(f, "ERR", e)
PY
