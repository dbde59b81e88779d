#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "conv16" 2>&1 | tail -12 > gpurun_out/r2_k5.log
tail -4 gpurun_out/r2_k5.log
python tools/conv16_debug.py > gpurun_out/r2_conv16_timing5.log 2>&1
ODEX_CONV_DEBUG=1 python tools/conv16_debug.py 2>&1 | grep conv16 | awk 'NR%3==0' >> gpurun_out/r2_conv16_timing5.log
cat gpurun_out/r2_conv16_timing5.log
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/r2_gputest5.log
tail -6 gpurun_out/r2_gputest5.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2_bench5.json 2> gpurun_out/r2_bench5.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench5.json')); print(d['ms_per_step'], d['value'], d['gpu_launches'])"
