#!/bin/bash
mkdir -p gpurun_out
for n in 8 4; do
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 8 --warmup 3 > gpurun_out/r2_bench17_n$n.json 2> gpurun_out/r2_bench17_n$n.err; echo "n=$n rc=$?"
tail -c 1500 gpurun_out/r2_bench17_n$n.json | head -c 700; echo
done
