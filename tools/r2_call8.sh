#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "bn_relu or warp_compose" 2>&1 | tail -5 > gpurun_out/r2_k8.log
tail -2 gpurun_out/r2_k8.log
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "vidode" 2>&1 | tail -5 >> gpurun_out/r2_k8.log
tail -2 gpurun_out/r2_k8.log
python bench.py --workload vidode --batch 32 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench8_vidode.json 2> gpurun_out/r2_bench8_vidode.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench8_vidode.json')); print('vidode', d['ms_per_step'], d['value'], d['hot_path']['ms_per_step'], d['gpu_launches'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-extras > gpurun_out/r2_bench8_n2.json 2> gpurun_out/r2_bench8_n2.err
tail -3 gpurun_out/r2_bench8_n2.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench8_n2.json')); print('n2', d['ms_per_step'], d['value'], d['config'].get('allreduce_in_graph'))"
ODEX_GRAPH_NCCL=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-extras > gpurun_out/r2_bench8_n2_eager.json 2> gpurun_out/r2_bench8_n2_eager.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench8_n2_eager.json')); print('n2 eager tail', d['ms_per_step'], d['value'], d['config'].get('allreduce_in_graph'))"
