#!/bin/bash
# ncu pass over one VidODE forward on the B200 (run through gpurun): plain run first, then the launch list, then --set full
# captures of the dominant kernels (ConvGRU gates convolution = conv_tc_kernel<128>, decoder convolution = conv_tc_kernel<64>,
# the bilinear upsample).  Outputs under gpurun_out/.
set -u
O=gpurun_out
TAG=${1:-r01v}
python tools/vidode_one_forward.py > $O/vidode_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/vidode_plain.log; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file $O/launches_$TAG.csv python tools/vidode_one_forward.py > $O/ncu_launches_v.log 2>&1
for spec in "conv_tc_kernel<.int.128:30" "conv_tc_kernel<.int.64:20" "upsample2x_fwd_kernel:4" "conv_tc2_kernel:20"; do
    k=${spec%%:*}; skip=${spec##*:}; name=$(echo $k | tr -d "<." | sed s/int//)
    ncu --profile-from-start off --set full --clock-control none --import-source on --kernel-name-base demangled \
        -k "regex:$k" -s $skip -c 2 \
        -f -o $O/prof_${TAG}_$name python tools/vidode_one_forward.py > $O/ncu_v_$name.log 2>&1
done
ls -la $O/*.ncu-rep
