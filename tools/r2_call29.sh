#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "conv16 or convnet or gru" 2>&1 | tail -4
timeout 300 python tools/conv16_debug.py 2>&1 | tail -8 | cut -c1-200
timeout 600 python bench.py --no-extras --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/r2_bench29.json 2> gpurun_out/r2_bench29.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench29.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])
PY
