#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -6 | tee gpurun_out/r2_gputest23.log
timeout 600 python bench.py --no-extras --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/r2_bench23.json 2> gpurun_out/r2_bench23.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench23.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])
PY
timeout 120 python tools/wgrad_probe.py 2>&1 | tee gpurun_out/r2_wgrad_probe23.log
timeout 300 python tools/step_profile.py > gpurun_out/r2_step_profile23.log 2>&1; head -16 gpurun_out/r2_step_profile23.log | cut -c1-150
