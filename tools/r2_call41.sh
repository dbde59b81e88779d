#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "wgrad or weight_gradient" 2>&1 | tail -3
timeout 120 python tools/wgrad_probe.py 2>&1
