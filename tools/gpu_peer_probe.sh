#!/bin/bash
N=${1:-2}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py --views $((2*N)) 2>&1 | grep "MULTI_GPU\|MISMATCH\|rror" | head -5
for red in push peer; do
GSM_VP_REDUCE=$red GSM_VP_EXCHANGE=peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-render --no-e2e > gpurun_out/r02t_n${N}_$red.json 2> gpurun_out/r02t_n${N}_$red.err
python -c "
import json; j=json.load(open('gpurun_out/r02t_n${N}_$red.json')); print('reduce=$red', j['value'], j['ms_per_step']); print(' '.join('%s=%.3f'%(r['kernel'],r['ms']) for r in j['stages'])); print(j.get('exchange_ms_rank0'))" || tail -5 gpurun_out/r02t_n${N}_$red.err
done
