#!/bin/bash
# Peer-exchange probe at N GPUs: share of the SMs given to the two-shot all-reduce (GSM_PEER_SM_DIV).
N=${1:-8}
for div in 4 8 2; do
GSM_PEER_SM_DIV=$div GSM_VP_EXCHANGE=peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-render --no-e2e > gpurun_out/r02p_n${N}_d$div.json 2> gpurun_out/r02p_n${N}_d$div.err
python -c "
import json; j=json.load(open('gpurun_out/r02p_n${N}_d$div.json')); print('div=$div', j['value'], j['ms_per_step']); print(j.get('exchange_ms_rank0'))"
done
