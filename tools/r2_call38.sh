#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --method dopri5 --batch 256 --steps 5 --warmup 3 > gpurun_out/r2_bench38_dopri5_n2.json 2> gpurun_out/r2_bench38.err; echo "rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench38_dopri5_n2.json').read().strip().splitlines()[-1])
print(d['metric'], d['value'], d['ms_per_step'], d['config']['workload'], d.get('dopri5_stats'))
" || tail -5 gpurun_out/r2_bench38.err
