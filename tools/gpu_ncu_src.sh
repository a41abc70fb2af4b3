#!/bin/bash
# source-level ncu capture: tools/gpu_ncu_src.sh <kernel-regex> <tag> <count> <command...>
K=$1; TAG=$2; CNT=$3; shift 3
mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -k regex:$K -c $CNT -f -o /tmp/$TAG "$@" > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; ls -la /tmp/$TAG.ncu-rep
ncu -i /tmp/$TAG.ncu-rep --page source --csv --print-source sass > gpurun_out/${TAG}_sass.csv 2> gpurun_out/${TAG}_sass.err
ncu -i /tmp/$TAG.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ls -la gpurun_out/${TAG}_*
