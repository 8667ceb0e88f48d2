#!/bin/bash
N=${1:-2}
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/overlap_probe.py 2>&1 | grep "first\|main stream" | head -6
for probe in sleep none; do
GSM_VP_PROBE=$probe GSM_VP_OVERLAP=1 GSM_VP_EXCHANGE=peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-render --no-e2e > gpurun_out/r02s_n${N}_$probe.json 2> gpurun_out/r02s_n${N}_$probe.err
python -c "
import json; j=json.load(open('gpurun_out/r02s_n${N}_$probe.json')); print('probe=$probe', j['value'], j['ms_per_step']); print(j.get('exchange_ms_rank0'))"
done
