#!/bin/bash
# Loss widening visit: loss parity tests, smoke, both bench arms (no ncu).  Usage: bash tools/gpu_loss.sh <tag>
TAG=${1:-loss}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_loss_gpu.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
tail -15 gpurun_out/${TAG}_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"
tail -3 gpurun_out/${TAG}_bench.err
python - <<PY
import json
for f in ("gpurun_out/${TAG}_bench.json", "gpurun_out/${TAG}_bench_ref.json"):
    try:
        l = [x for x in open(f) if x.startswith("{")][-1]
        d = json.loads(l)
        print(f, d["value"], d["ms_per_step"], d["e2e"], d.get("loss"))
    except Exception as e:
        print(f, "ERR", e)
PY
