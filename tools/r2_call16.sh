#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -8 > gpurun_out/r2_gputest16.log
cat gpurun_out/r2_gputest16.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r2_bench16.json 2> gpurun_out/r2_bench16.err; echo "bench rc=$?"
tail -c 6000 gpurun_out/r2_bench16.json
tail -5 gpurun_out/r2_bench16.err
timeout 600 python bench.py --impl reference > gpurun_out/r2_bench16_ref.json 2>> gpurun_out/r2_bench16.err; echo "ref rc=$?"
cat gpurun_out/r2_bench16_ref.json
