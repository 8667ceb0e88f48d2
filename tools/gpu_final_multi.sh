#!/bin/bash
# Final multi-GPU lines of the round at N GPUs: (a) the default weak-scaling line the driver also takes (C4, one view per
# GPU per step, render leg included), (b) BASELINE.json configs[4]: C5 (6 M surfels), a FIXED batch of 8 views per step
# shared by the N GPUs (strong scaling); at N = 1 also the reference arm on the same C5 batch.
# Usage: bash tools/gpu_final_multi.sh <tag> <N>
TAG=$1; N=$2
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
  RUN="python"
else
  RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
fi
if [ "$N" != "1" ]; then
  timeout 600 $RUN bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
  echo "weak N=$N exit $?"
fi
timeout 600 $RUN bench.py --gpus $N --steps 20 --warmup 5 --workload C5 --views-per-step 8 --no-render --no-knn --no-cpu-baseline \
  > gpurun_out/${TAG}_c5_n$N.json 2> gpurun_out/${TAG}_c5_n$N.err
echo "C5 N=$N exit $?"
if [ "$N" = "1" ]; then
  timeout 600 python bench.py --impl reference --steps 10 --warmup 3 --workload C5 --views-per-step 8 --no-render --no-knn --no-e2e \
    > gpurun_out/${TAG}_c5_ref.json 2> gpurun_out/${TAG}_c5_ref.err
  echo "C5 reference exit $?"
fi
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/${TAG}_*n$N.json") + glob.glob("gpurun_out/${TAG}_c5_ref.json")):
    try:
        j = json.load(open(f))
        print(f, "value %.1f %s ms/step %.3f scaling %s views/step %s" % (j["value"], j["unit"], j["ms_per_step"], j["scaling"], j["config"].get("views_per_step")),
              "e2e", (j.get("e2e") or {}).get("value"), "render", [(r["width"], round(r["fps"], 1)) for r in j.get("render", [])],
              "exch", j.get("exchange_ms_rank0", {}).get("exchange_total") if j.get("exchange_ms_rank0") else None)
    except Exception as e:
        print(f, "FAILED", e)
PY
