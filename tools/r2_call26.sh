#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/torch_side_hunt.py 2>&1 | grep -v Warning | tail -42 | tee gpurun_out/r2_torch_side.log
