#!/bin/bash
mkdir -p gpurun_out
for n in 8 4 2; do
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 8 --warmup 3 > gpurun_out/r2_bench17_n$n.json 2> gpurun_out/r2_bench17_n$n.err; echo "n=$n rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench17_n$n.json').read().strip().splitlines()[-1])
print($n, d['value'], d['ms_per_step'], d['e2e']['value'], d['strong_scaling'])
"
done
