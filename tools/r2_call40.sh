#!/bin/bash
mkdir -p gpurun_out
python tools/one_step.py > gpurun_out/one_step_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r02d.csv python tools/one_step.py > gpurun_out/ncu_launches.log 2>&1
wc -l gpurun_out/launches_r02d.csv
