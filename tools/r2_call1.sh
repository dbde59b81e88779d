#!/bin/bash
# round-2 GPU call 1: kernel tests (incl. the new kind::f16 chain), full gpu suite, bench with both operand splits
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "convnet" 2>&1 | tail -30 > gpurun_out/r2_k1.log
tail -3 gpurun_out/r2_k1.log
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2_gputest1.log
tail -5 gpurun_out/r2_gputest1.log
ODEX_SPLIT=tf32 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_tf32.json 2> gpurun_out/r2_bench_tf32.err
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_f16.json 2> gpurun_out/r2_bench_f16.err
python - <<'PY'
import json
for n in ("tf32","f16"):
    try:
        d=json.load(open("gpurun_out/r2_bench_%s.json"%n)); print(n, d["ms_per_step"], d["value"], d["roofline"]["avg_launch_us"], d["roofline"]["achieved"])
    except Exception as e: print(n, "failed", e)
PY
ODEX_CONV_DEBUG=1 python tools/convnet_debug.py 2>&1 | tail -8 > gpurun_out/r2_convnet_debug.log; cat gpurun_out/r2_convnet_debug.log
python tools/convnet_debug.py 2>&1 | tail -4 >> gpurun_out/r2_convnet_debug.log; ODEX_SPLIT=tf32 python tools/convnet_debug.py 2>&1 | tail -4 >> gpurun_out/r2_convnet_debug.log; cat gpurun_out/r2_convnet_debug.log
