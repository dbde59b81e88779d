#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/wgrad_scaled_debug.py > gpurun_out/r2_wgrad_scaled_debug.log 2>&1; tail -60 gpurun_out/r2_wgrad_scaled_debug.log
