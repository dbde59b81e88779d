#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "wgrad" 2>&1 | tail -4 > gpurun_out/r2_k9.log
tail -2 gpurun_out/r2_k9.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2_bench9.json 2> gpurun_out/r2_bench9.err
ODEX_WGRAD_SPLIT=f16 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2_bench9_wf16.json 2> gpurun_out/r2_bench9_wf16.err
python -c "
import json
for n in ('r2_bench9','r2_bench9_wf16'):
    d=json.load(open('gpurun_out/%s.json'%n)); print(n, d['ms_per_step'], d['value'], d['gpu_launches'])"
ODEX_WGRAD_SPLIT=f16 ODEX_PDL=0 timeout 300 python tools/step_profile.py 2>&1 | grep -E "wgrad|absmax|device time" > gpurun_out/r2_step_profile9.log; cat gpurun_out/r2_step_profile9.log
timeout 900 bash tools/profile_pass.sh r02 > gpurun_out/r2_profile_pass.log 2>&1; tail -8 gpurun_out/r2_profile_pass.log
