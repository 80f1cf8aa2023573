#!/bin/bash
N=$1
for m in "" 1; do
OEVAE_DP_NCCL=$m timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 5 --no-cpu > gpurun_out/t31_n${N}_$m.json 2> gpurun_out/t31_n${N}_$m.err; echo "rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/t31_n${N}_$m.json')); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['parallelism'][:40], [(k['name'],k['launches_per_step'],round(k['us_per_launch'],1)) for k in d['kernels']], d['scoring']['value'])"
done
