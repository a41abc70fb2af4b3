#!/bin/bash
# source-level ncu capture of one kernel family: tools/gpu_ncu_src.sh <kernel-regex> <tag>
K=${1:-s2act_tc_kernel}; TAG=${2:-src}
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:$K -c 2 -f -o /tmp/$TAG \
  python tools/edge_bench.py --iters 1 --profile > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; ls -la /tmp/$TAG.ncu-rep
ncu -i /tmp/$TAG.ncu-rep --page source --csv --print-source sass > gpurun_out/${TAG}_sass.csv 2> gpurun_out/${TAG}_sass.err
ncu -i /tmp/$TAG.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ls -la gpurun_out/${TAG}_*
