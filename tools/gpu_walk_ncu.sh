#!/bin/bash
mkdir -p gpurun_out
python tools/small_codec.py 192 20 > gpurun_out/small_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:'k_walk|k_hdr|k_transient' -s 6 -c 6 --csv --log-file gpurun_out/ncu_walk.csv python tools/small_codec.py 192 20 > gpurun_out/ncu_walk.log 2>&1
grep -o '"k_[a-z]*[^"]*","[0-9]*","[0-9]*","([0-9, ]*)","([0-9, ]*)","0","10.0","[^"]*","[^"]*","[^"]*","[^"]*"' gpurun_out/ncu_walk.csv | sed 's/(const[^)]*)//' | cut -c1-200
