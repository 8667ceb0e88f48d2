#!/bin/bash
# Lean A/B of build variants on one box: short device-resident bench per variant; parity (reference-build, golden and
# full-size tests through the variant library) only for the variants named after "--parity".
# Usage: bash tools/gpu_ab2.sh <outtag> <variant>... [--parity <variant>...]
OUT=$1; shift
mkdir -p gpurun_out
par=0
for v in "$@"; do
  if [ "$v" = "--parity" ]; then par=1; continue; fi
  lib=gaussmart_b200/lib/libgaussmart_b200_$v.so
  [ "$v" = "default" ] && lib=gaussmart_b200/lib/libgaussmart_b200.so
  if [ $par = 1 ]; then
    GAUSSMART_B200_LIB=$PWD/$lib timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_parity_holes_gpu.py tests/test_fullsize_gpu.py tests/test_raw_gpu.py tests/test_reference_callers_gpu.py \
      -m gpu -q -p no:cacheprovider -k "not knn" > gpurun_out/${OUT}_${v}_pytest.log 2>&1
    echo "$v parity: $(tail -1 gpurun_out/${OUT}_${v}_pytest.log)"
    continue
  fi
  GAUSSMART_B200_LIB=$PWD/$lib timeout 600 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --no-knn --no-drop-in --no-render \
    > gpurun_out/${OUT}_$v.json 2> gpurun_out/${OUT}_$v.err
  python - <<PY
import json
try:
    j=json.load(open("gpurun_out/${OUT}_$v.json"))
    print("$v", "ms/step %.3f" % j["ms_per_step"], "api %.3f" % j["value_public_api"]["ms_per_step"], " ".join("%s=%.3f" % (r["kernel"], r["ms"]) for r in j["stages"]))
except Exception as e:
    print("$v", "FAILED", e, open("gpurun_out/${OUT}_$v.err").read()[-800:])
PY
done
