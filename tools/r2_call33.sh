#!/bin/bash
mkdir -p gpurun_out
for ov in 1 0; do
ODEX_OVERLAP_AR=$ov timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29600+ov)) bench.py --gpus 2 --steps 20 --warmup 5 --no-extras > gpurun_out/r2_bench33_ov$ov.json 2> gpurun_out/r2_bench33_ov$ov.err; echo "overlap=$ov rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench33_ov$ov.json').read().strip().splitlines()[-1])
print($ov, d['value'], d['ms_per_step'], d['e2e']['value'], d['config'].get('allreduce_overlapped'), d['config'].get('allreduce_in_graph'))
" || tail -5 gpurun_out/r2_bench33_ov$ov.err
done
