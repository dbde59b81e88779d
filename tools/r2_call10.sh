#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench10_n2.json 2> gpurun_out/r2_bench10_n2.err
echo "rc=$?"; tail -2 gpurun_out/r2_bench10_n2.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench10_n2.json')); print('n2', d['ms_per_step'], d['value'], d['config'].get('allreduce_in_graph'), d.get('strong_scaling'))"
ODEX_GRAPH_NCCL=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-extras > gpurun_out/r2_bench10_n2_eager.json 2> gpurun_out/r2_bench10_n2_eager.err
echo "rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2_bench10_n2_eager.json')); print('n2 eager tail', d['ms_per_step'], d['value'], d['config'].get('allreduce_in_graph'))"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 3 --warmup 3 --workload vidode --batch 32 --no-cpu-baseline > gpurun_out/r2_bench10_vid_n2.json 2> gpurun_out/r2_bench10_vid_n2.err
echo "rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2_bench10_vid_n2.json')); print('vid n2', d['ms_per_step'], d['value'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2_bench10_ref_n2.json 2>/dev/null
echo "rc=$?"; head -c 200 gpurun_out/r2_bench10_ref_n2.json
