#!/bin/bash
N=${1:-2}
for ov in 1 2 0; do
GSM_VP_OVERLAP=$ov GSM_VP_EXCHANGE=peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-render --no-e2e > gpurun_out/r02q_n${N}_ov$ov.json 2> gpurun_out/r02q_n${N}_ov$ov.err
python -c "
import json; j=json.load(open('gpurun_out/r02q_n${N}_ov$ov.json')); print('overlap=$ov', j['value'], j['ms_per_step']); print(j.get('exchange_ms_rank0'))"
done
