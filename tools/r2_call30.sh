#!/bin/bash
mkdir -p gpurun_out
ODEX_WGRAD_DEBUG=1 timeout 120 python tools/wgrad_debug.py 2>&1 | grep "wgrad_f16" | awk 'NR%3==0' | tee gpurun_out/r2_wgrad_f16_roles.log
