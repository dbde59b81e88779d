#!/bin/bash
mkdir -p gpurun_out
echo "== ODEX_WGRAD_SCALED=0"; ODEX_WGRAD_SCALED=0 timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -x 2>&1 | tail -2
echo "== ODEX_SPLIT=tf32"; ODEX_SPLIT=tf32 timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -x 2>&1 | tail -2
echo "== ODEX_CONV_IMPL=simt"; ODEX_CONV_IMPL=simt timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "not golden" 2>&1 | tail -2
