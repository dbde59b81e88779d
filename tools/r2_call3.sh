#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "conv16" 2>&1 | tail -40 > gpurun_out/r2_k3.log
tail -5 gpurun_out/r2_k3.log
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2_gputest3.log
tail -8 gpurun_out/r2_gputest3.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2_bench3.json 2> gpurun_out/r2_bench3.err
tail -2 gpurun_out/r2_bench3.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench3.json')); print(d['ms_per_step'], d['value'], d['gpu_launches'])"
python tools/step_profile.py > gpurun_out/r2_step_profile3.log 2>&1; tail -45 gpurun_out/r2_step_profile3.log
