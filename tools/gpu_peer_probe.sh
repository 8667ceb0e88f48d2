#!/bin/bash
# Peer-exchange probes at N GPUs: overlap of the SH expand with the two-shot all-reduce on / off.
N=${1:-8}
for ov in 1 0; do
GSM_VP_OVERLAP=$ov GSM_VP_EXCHANGE=peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-render --no-e2e > gpurun_out/r02l_n${N}_ov$ov.json 2> gpurun_out/r02l_n${N}_ov$ov.err
python -c "
import json; j=json.load(open('gpurun_out/r02l_n${N}_ov$ov.json')); print('overlap=$ov', j['value'], j['ms_per_step']); print(' '.join('%s=%.3f'%(r['kernel'],r['ms']) for r in j['stages'])); print(j.get('exchange_ms_rank0'))"
done
