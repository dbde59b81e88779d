#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -6 | tee gpurun_out/r2_gputest20.log
timeout 600 python bench.py --no-extras --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/r2_bench20.json 2> gpurun_out/r2_bench20.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench20.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])
PY
ODEX_WGRAD_SCALED=0 timeout 600 python bench.py --no-extras --no-cpu-baseline --steps 20 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('scaled off:', d['value'], d['ms_per_step'])"
timeout 300 python tools/step_profile.py > gpurun_out/r2_step_profile20.log 2>&1; head -30 gpurun_out/r2_step_profile20.log
