#!/bin/bash
mkdir -p gpurun_out
ODEX_WGRAD_DEBUG=1 timeout 120 python tools/wgrad_debug.py 2>&1 | grep "wgrad_tc" | awk 'NR%3==0' > gpurun_out/r2_wgrad_debug14.log
cat gpurun_out/r2_wgrad_debug14.log
for m in 15 13 6; do ODEX_WGRAD_PROBE=$m timeout 120 python tools/wgrad_probe.py; done 2>&1 | grep "R=3"
