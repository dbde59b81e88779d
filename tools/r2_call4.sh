#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "dopri5_relu" 2>&1 | grep -E "Error|assert |DopriStats|relu:" | head -20 > gpurun_out/r2_t4.log
cat gpurun_out/r2_t4.log
python tools/conv16_debug.py > gpurun_out/r2_conv16_timing.log 2>&1
ODEX_CONV_DEBUG=1 python tools/conv16_debug.py 2>&1 | grep conv16 | awk 'NR%3==0' >> gpurun_out/r2_conv16_timing.log
cat gpurun_out/r2_conv16_timing.log
