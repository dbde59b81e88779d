#!/bin/bash
mkdir -p gpurun_out
for t in 5 3 6 4; do ODEX_W16_TAPS3=$t timeout 120 python tools/wgrad_probe.py; done > gpurun_out/r2_wgrad_probe18.log 2>&1
cat gpurun_out/r2_wgrad_probe18.log
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "wgrad" 2>&1 | tail -4
