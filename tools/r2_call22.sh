#!/bin/bash
mkdir -p gpurun_out
timeout 1500 bash tools/profile_pass.sh r02b 2>&1 | tail -8
