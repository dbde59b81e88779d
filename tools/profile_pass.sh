#!/bin/bash
# One profiling pass on the B200 (run through gpurun): plain run first, then the ncu launch list of one step, then
# --set full captures of the dominant kernels.  Outputs under gpurun_out/.
set -u
O=gpurun_out
TAG=${1:-r02}
python tools/one_step.py > $O/one_step_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/one_step_plain.log; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file $O/launches_$TAG.csv python tools/one_step.py > $O/ncu_launches.log 2>&1
for spec in "convnet_f16_kernel:10" "conv16_f16_kernel:4" "wgrad_f16_kernel:10" "lincomb_kernel:30"; do
    k=${spec%%:*}; skip=${spec##*:}
    ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:$k -s $skip -c 2 \
        -f -o $O/prof_${TAG}_$k python tools/one_step.py > $O/ncu_$k.log 2>&1
done
ls -la $O/*.ncu-rep
