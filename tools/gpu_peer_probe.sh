for ld in plain sys; do
GSM_PEER_LD=$ld GSM_VP_EXCHANGE=peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 --no-render --no-e2e > gpurun_out/r02i_n8_$ld.json 2> gpurun_out/r02i_n8_$ld.err
python -c "
import json; j=json.load(open('gpurun_out/r02i_n8_$ld.json')); print('$ld', j['value'], j['ms_per_step']); print(' '.join('%s=%.3f'%(r['kernel'],r['ms']) for r in j['stages'])); print(j.get('exchange_ms_rank0'))"
done
