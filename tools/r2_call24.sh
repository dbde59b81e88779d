#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "any_magnitude" 2>&1 | grep -E "^E  |passed|failed" | head -20
