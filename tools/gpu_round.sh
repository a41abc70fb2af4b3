#!/bin/bash
# One gpurun call: GPU parity tests, the bench line, a torch.profiler view, the ncu launch list and one
# ncu --set full capture of the dominant kernel (B200_PROFILING.md recipe).
set -u
mkdir -p gpurun_out
TAG=${1:-r1}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${TAG}_pytest.log
tail -5 gpurun_out/${TAG}_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json
BATCH=32 python tools/profile_step.py > gpurun_out/${TAG}_torchprof.txt 2>&1; echo "torchprof exit $?"
CMD="python bench.py --steps 1 --warmup 3 --batch 4 --no-cpu-baseline"
$CMD > gpurun_out/${TAG}_plain_b4.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 30000 --csv --log-file gpurun_out/${TAG}_launches_b4.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
$CMD > gpurun_out/${TAG}_plain_b4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -s 300 -c 4 -f -o gpurun_out/${TAG}_tc_gemm $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
