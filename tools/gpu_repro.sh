#!/bin/bash
# Repro helper: the golden test alone, in fresh processes, through a chosen library (default: production).
mkdir -p gpurun_out
LIBV=${1:-default}
lib=$PWD/gaussmart_b200/lib/libgaussmart_b200_$LIBV.so
[ "$LIBV" = "default" ] && lib=$PWD/gaussmart_b200/lib/libgaussmart_b200.so
for i in 1 2 3 4 5 6; do
  GAUSSMART_B200_LIB=$lib timeout 200 python -m pytest "tests/test_parity_gpu.py::test_cuda_matches_reference_golden" -m gpu -q -p no:cacheprovider -x > gpurun_out/repro_${LIBV}_$i.log 2>&1
  echo "run $i ($LIBV): $(tail -1 gpurun_out/repro_${LIBV}_$i.log)"; grep -m2 "AssertionError" gpurun_out/repro_${LIBV}_$i.log
done
