#!/bin/bash
mkdir -p gpurun_out
echo "== alone"; timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu 2>&1 | tail -4
echo "== after the others"; timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/r2_gputest19.log
