#!/bin/bash
# torch.profiler table of one bench step (which library ops remain on the step) + host-side cProfile of a step.
set -u
mkdir -p gpurun_out
TAG=${1:-r1p}
timeout 150 python tools/profile_step.py > gpurun_out/${TAG}_torch_profile.txt 2>&1; echo "profile_step exit $?"
head -60 gpurun_out/${TAG}_torch_profile.txt | cut -c1-200
timeout 100 python tools/cpu_profile_step.py > gpurun_out/${TAG}_cpu_profile.txt 2>&1; echo "cpu_profile_step exit $?"
head -45 gpurun_out/${TAG}_cpu_profile.txt | cut -c1-160
exit 0
