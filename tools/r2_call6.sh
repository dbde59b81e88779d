#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "wgrad" 2>&1 | tail -30 > gpurun_out/r2_k6.log
grep -E "passed|failed|Error|assert " gpurun_out/r2_k6.log | head -12
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/r2_gputest6.log
tail -6 gpurun_out/r2_gputest6.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2_bench6.json 2> gpurun_out/r2_bench6.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench6.json')); print(d['ms_per_step'], d['value'], d['gpu_launches'])"
ODEX_PDL=0 python tools/step_profile.py > gpurun_out/r2_step_profile6.log 2>&1; tail -32 gpurun_out/r2_step_profile6.log
