#!/bin/bash
# Both bench arms back to back (reference first).  Usage: bash tools/gpu_bench_pair.sh <tag>
TAG=${1:-pair}
mkdir -p gpurun_out
timeout 400 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref exit $?"
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
tail -n 3 gpurun_out/${TAG}_bench.err; tail -n 3 gpurun_out/${TAG}_bench_ref.err; true
