#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/r2_gputest35.log
timeout 600 python bench.py --no-extras --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/r2_bench35.json 2> gpurun_out/r2_bench35.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench35.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])
PY
