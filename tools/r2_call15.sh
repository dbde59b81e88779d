#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "convnet" 2>&1 | tail -8 > gpurun_out/r2_t15.log
cat gpurun_out/r2_t15.log
ODEX_CONV_DEBUG=1 timeout 120 python tools/convnet_debug.py 2>&1 | grep convnet_f16 | awk 'NR%3==0' > gpurun_out/r2_convnet_cycles15.log
timeout 120 python tools/convnet_debug.py >> gpurun_out/r2_convnet_cycles15.log 2>&1
cat gpurun_out/r2_convnet_cycles15.log
