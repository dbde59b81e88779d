#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/r2_gputest32.log
timeout 600 python bench.py --method dopri5 --batch 512 --no-cpu-baseline --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('dopri5 train', d['value'], d['ms_per_step'])"
timeout 1500 bash tools/profile_pass.sh r02c 2>&1 | tail -6
