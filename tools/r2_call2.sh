#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "dopri5_relu or robustness" -s 2>&1 | tail -60 > gpurun_out/r2_t2.log
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "magnitude" 2>&1 | tail -5 >> gpurun_out/r2_t2.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench2.json 2> gpurun_out/r2_bench2.err
tail -3 gpurun_out/r2_bench2.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench2_ref.json 2> /dev/null
