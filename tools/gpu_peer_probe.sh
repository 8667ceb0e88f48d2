#!/bin/bash
# Peer-exchange probes at N GPUs: all-reduce flavour (NVLS multimem vs two-shot peer loads/stores).
N=${1:-8}
for red in multimem peer; do
GSM_VP_REDUCE=$red GSM_VP_EXCHANGE=peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-render --no-e2e > gpurun_out/r02n_n${N}_$red.json 2> gpurun_out/r02n_n${N}_$red.err
python -c "
import json; j=json.load(open('gpurun_out/r02n_n${N}_$red.json')); print('reduce=$red', j['value'], j['ms_per_step']); print(' '.join('%s=%.3f'%(r['kernel'],r['ms']) for r in j['stages'])); print(j.get('exchange_ms_rank0'))"
done
