#!/bin/bash
# A/B of run-time switches (environment variables) on one GPU box: parity subset + short device-resident bench per setting.
# Usage: bash tools/gpu_env_ab.sh <outtag> "<name>:<ENV=VAL ...>" ...      (use "default:" for no variables)
OUT=$1; shift
mkdir -p gpurun_out
for spec in "$@"; do
  name=${spec%%:*}; envs=${spec#*:}
  env $envs timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_parity_holes_gpu.py tests/test_fullsize_gpu.py tests/test_raw_gpu.py \
    -m gpu -q -p no:cacheprovider -k "not knn" > gpurun_out/${OUT}_${name}_pytest.log 2>&1
  echo "$name parity: $(tail -1 gpurun_out/${OUT}_${name}_pytest.log)"
  env $envs timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-knn --no-drop-in --no-render \
    > gpurun_out/${OUT}_$name.json 2> gpurun_out/${OUT}_$name.err
  python - <<PY
import json
try:
    j=json.load(open("gpurun_out/${OUT}_$name.json"))
    print("$name", "ms/step %.3f" % j["ms_per_step"], " ".join("%s=%.3f" % (r["kernel"], r["ms"]) for r in j["stages"]))
except Exception as e:
    print("$name", "FAILED", e, open("gpurun_out/${OUT}_$name.err").read()[-800:])
PY
done
