#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/r2_gputest27.log
timeout 600 python bench.py --no-extras --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/r2_bench27.json 2> gpurun_out/r2_bench27.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench27.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])
PY
timeout 600 python bench.py --workload vidode --batch 32 --no-cpu-baseline --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('vidode', d['value'], d['ms_per_step'])"
timeout 300 python tools/vidode_profile.py 2>&1 | grep -v Warn | head -24 | cut -c1-150 | tee gpurun_out/r2_vidode_profile27.log
