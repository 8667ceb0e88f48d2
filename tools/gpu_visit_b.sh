#!/bin/bash
# One box visit: GPU suite, short bench, ncu --set full with source of the two compositing kernels; the SASS source
# pages are exported on the box (the report itself is too big to pull).
TAG=${1:-r02b}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?: $(tail -1 gpurun_out/${TAG}_pytest.log)"
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-knn --no-drop-in > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    j=json.load(open("gpurun_out/${TAG}_bench.json"))
    print("ms/step %.3f" % j["ms_per_step"], " ".join("%s=%.3f" % (r["kernel"], r["ms"]) for r in j["stages"]))
    for r in j.get("render", []):
        print("    ", r["width"], round(r["fps"], 1), {k: round(x, 3) for k, x in r.get("stages_ms", {}).items()})
except Exception as e:
    print("FAILED", e, open("gpurun_out/${TAG}_bench.err").read()[-800:])
PY
if [ "$2" = "sass" ]; then
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'render_bwd_kernel|render_fwd_kernel' \
  --launch-skip 4 --launch-count 2 -o gpurun_out/${TAG}_comp -f \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-knn --no-render --no-drop-in > gpurun_out/${TAG}_ncu_comp.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/${TAG}_comp.ncu-rep --page source --csv -k regex:render_bwd_kernel > gpurun_out/${TAG}_bwd_source.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_comp.ncu-rep --page source --csv -k regex:render_fwd_kernel > gpurun_out/${TAG}_fwd_source.csv 2>/dev/null
rm -f gpurun_out/${TAG}_comp.ncu-rep
fi
ls -la gpurun_out | grep ${TAG}
