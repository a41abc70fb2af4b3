#!/bin/bash
# GPU call: parity tests, bench line, config-4 micro-bench, config-5 stress, launch list of a small bench run.
set -u
mkdir -p gpurun_out
TAG=${1:-r1f}
T0=$(date +%s)
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? ($(( $(date +%s) - T0 )) s)"
tail -8 gpurun_out/${TAG}_pytest.log
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $? ($(( $(date +%s) - T0 )) s)"
cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
timeout 200 python tools/radius_bench.py > gpurun_out/${TAG}_radius_bench.jsonl 2>&1; echo "radius_bench exit $? ($(( $(date +%s) - T0 )) s)"
cut -c1-420 gpurun_out/${TAG}_radius_bench.jsonl
timeout 300 python tools/stress_bench.py > gpurun_out/${TAG}_stress.jsonl 2> gpurun_out/${TAG}_stress.err; echo "stress exit $? ($(( $(date +%s) - T0 )) s)"
cat gpurun_out/${TAG}_stress.jsonl; tail -5 gpurun_out/${TAG}_stress.err
timeout 200 python tools/edge_bench.py > gpurun_out/${TAG}_edge_bench.jsonl 2>&1; echo "edge_bench exit $? ($(( $(date +%s) - T0 )) s)"
cut -c1-200 gpurun_out/${TAG}_edge_bench.jsonl
if [ "${NCU:-1}" = "1" ]; then
CMD2="python bench.py --steps 1 --warmup 1 --batch 4 --no-cpu-baseline"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches_b4.csv $CMD2 > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $? ($(( $(date +%s) - T0 )) s)"
fi
du -sh gpurun_out
