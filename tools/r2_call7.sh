#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "bn_relu or warp_compose" 2>&1 | tail -30 > gpurun_out/r2_k7.log
grep -E "passed|failed|Error|assert " gpurun_out/r2_k7.log | head -12
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/r2_gputest7.log
tail -6 gpurun_out/r2_gputest7.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2_bench7.json 2> gpurun_out/r2_bench7.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench7.json')); print(d['ms_per_step'], d['value'], d['gpu_launches'])"
python bench.py --workload vidode --batch 32 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench7_vidode.json 2> gpurun_out/r2_bench7_vidode.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench7_vidode.json')); print('vidode', d['ms_per_step'], d['value'], d['hot_path'], d['gpu_launches'])"
ODEX_PDL=0 python tools/vidode_profile.py > gpurun_out/r2_vidode_profile7.log 2>&1; tail -30 gpurun_out/r2_vidode_profile7.log
