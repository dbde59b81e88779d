#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "wgrad" 2>&1 | tail -12 > gpurun_out/r2_k13.log
grep -E "passed|failed|Error|assert " gpurun_out/r2_k13.log | head
for m in 0 4; do ODEX_WGRAD_PROBE=$m timeout 120 python tools/wgrad_probe.py; done > gpurun_out/r2_wgrad_probe13.log 2>&1
cat gpurun_out/r2_wgrad_probe13.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2_bench13.json 2> gpurun_out/r2_bench13.err
ODEX_WGRAD_SPLIT=f16 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2_bench13_wf16.json 2> gpurun_out/r2_bench13_wf16.err
python -c "
import json
for n in ('r2_bench13','r2_bench13_wf16'):
    try:
        d=json.load(open('gpurun_out/%s.json'%n)); print(n, d['ms_per_step'], d['value'], d['gpu_launches'])
    except Exception as e: print(n, 'failed', e)"
ODEX_WGRAD_SPLIT=f16 ODEX_PDL=0 timeout 300 python tools/step_profile.py 2>&1 | grep -E "wgrad|absmax|device time" > gpurun_out/r2_step_profile13.log; cat gpurun_out/r2_step_profile13.log
ODEX_PDL=0 timeout 300 python tools/step_profile.py 2>&1 | grep -E "wgrad|absmax|device time"
