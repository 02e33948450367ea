#!/bin/bash
# usage: gpu_ncu.sh <kernel regex> <tag>: full ncu capture (with source) of the second launch of one kernel
K=${1:-k_encode}; TAG=${2:-enc}
mkdir -p gpurun_out
python tools/small_codec.py 48 5 > gpurun_out/small_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$K" -s 1 -c 1 -f -o gpurun_out/${TAG}_full python tools/small_codec.py 48 5 > gpurun_out/ncu_${TAG}.log 2>&1
tail -n 3 gpurun_out/small_plain.log; tail -n 3 gpurun_out/ncu_${TAG}.log
