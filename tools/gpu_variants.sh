#!/bin/bash
# A/B of tuning variants built with `python -m gaussmart_b200.build --tag=<t> -D...`: short device-resident bench per variant.
# Usage: bash tools/gpu_variants.sh <outtag> <variant> [<variant> ...]
OUT=$1; shift
mkdir -p gpurun_out
for v in "$@"; do
  lib=gaussmart_b200/lib/libgaussmart_b200_$v.so
  [ "$v" = "default" ] && lib=gaussmart_b200/lib/libgaussmart_b200.so
  GAUSSMART_B200_LIB=$PWD/$lib timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-knn --no-render \
    > gpurun_out/${OUT}_$v.json 2> gpurun_out/${OUT}_$v.err
  python - <<PY
import json
try:
    j=json.load(open("gpurun_out/${OUT}_$v.json"))
    print("$v", "ms/step %.3f" % j["ms_per_step"], " ".join("%s=%.3f" % (r["kernel"], r["ms"]) for r in j["stages"]))
except Exception as e:
    print("$v", "FAILED", e, open("gpurun_out/${OUT}_$v.err").read()[-800:])
PY
done
