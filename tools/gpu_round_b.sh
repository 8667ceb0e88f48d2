#!/bin/bash
# Short box visit for the densification-stats row: its parity tests, smoke, both bench arms, then the whole GPU suite.
TAG=${1:-r01b}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_densification_gpu.py tests/test_adam_gpu.py -m gpu -x -q -s > gpurun_out/${TAG}_dens_pytest.log 2>&1
echo "dens pytest exit $?"; tail -4 gpurun_out/${TAG}_dens_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cut -c1-400 gpurun_out/${TAG}_bench.json
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref exit $?"
timeout 420 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_all.log 2>&1; echo "all pytest exit $?"
tail -3 gpurun_out/${TAG}_pytest_all.log
