#!/bin/bash
# A/B of tuning variants (built with `python -m gaussmart_b200.build --tag=<t> -D...`) on one GPU box: for each variant a
# parity pass (reference-build and golden tests, through the variant library) and a short device-resident bench.
# Usage: bash tools/gpu_ab.sh <outtag> <variant> [<variant> ...]      ("default" = the production library)
OUT=$1; shift
mkdir -p gpurun_out
for v in "$@"; do
  lib=gaussmart_b200/lib/libgaussmart_b200_$v.so
  [ "$v" = "default" ] && lib=gaussmart_b200/lib/libgaussmart_b200.so
  if [ "$v" != "default" ]; then
    GAUSSMART_B200_LIB=$PWD/$lib timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_parity_holes_gpu.py tests/test_fullsize_gpu.py \
      -m gpu -q -p no:cacheprovider -k "not knn" > gpurun_out/${OUT}_${v}_pytest.log 2>&1
    echo "$v parity: $(tail -1 gpurun_out/${OUT}_${v}_pytest.log)"
  fi
  GAUSSMART_B200_LIB=$PWD/$lib timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-knn --no-drop-in \
    > gpurun_out/${OUT}_$v.json 2> gpurun_out/${OUT}_$v.err
  python - <<PY
import json
try:
    j=json.load(open("gpurun_out/${OUT}_$v.json"))
    print("$v", "ms/step %.3f" % j["ms_per_step"], " ".join("%s=%.3f" % (r["kernel"], r["ms"]) for r in j["stages"]),
          "| render", " ".join("%dx%d %.1f fps" % (r["width"], r["height"], r["fps"]) for r in j.get("render", [])))
    for r in j.get("render", []):
        print("    ", r["width"], {k: round(x, 3) for k, x in r.get("stages_ms", {}).items()})
except Exception as e:
    print("$v", "FAILED", e, open("gpurun_out/${OUT}_$v.err").read()[-800:])
PY
done
