#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "gru" 2>&1 | tail -3
timeout 300 python tools/step_profile.py 2>&1 | grep -E "gru_|device time"
