#!/bin/bash
# Peer-exchange probe at N GPUs (timing of the parts of the exchange on rank 0).
N=${1:-8}
GSM_VP_EXCHANGE=peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-render --no-e2e > gpurun_out/r02o_n${N}.json 2> gpurun_out/r02o_n${N}.err
python -c "
import json; j=json.load(open('gpurun_out/r02o_n${N}.json')); print(j['value'], j['ms_per_step']); print(' '.join('%s=%.3f'%(r['kernel'],r['ms']) for r in j['stages'])); print(j.get('exchange_ms_rank0'))"
