#!/bin/bash
mkdir -p gpurun_out
for m in 0 1 8 9 2 4 6 15; do ODEX_WGRAD_PROBE=$m timeout 120 python tools/wgrad_probe.py; done > gpurun_out/r2_wgrad_probe.log 2>&1
cat gpurun_out/r2_wgrad_probe.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
