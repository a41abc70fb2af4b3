#!/bin/bash
# GPU call: parity tests, GEMM + edge-path micro-benchmarks, bench line, ncu --set full of one edge-layer pass.
set -u
mkdir -p gpurun_out
TAG=${1:-r1c}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/${TAG}_pytest.log
python tools/kernel_bench.py > gpurun_out/${TAG}_kernel_bench.jsonl 2>&1; echo "kernel_bench exit $?"
cat gpurun_out/${TAG}_kernel_bench.jsonl
python tools/edge_bench.py > gpurun_out/${TAG}_edge_bench.jsonl 2>&1; echo "edge_bench exit $?"
cat gpurun_out/${TAG}_edge_bench.jsonl
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json
if [ "${NCU:-1}" = "1" ]; then
CMD="python tools/edge_bench.py --profile --batch 8"
$CMD > gpurun_out/${TAG}_plain_eb.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:efl' -f -o gpurun_out/${TAG}_edge $CMD > gpurun_out/${TAG}_ncu_edge.log 2>&1
echo "ncu edge exit $?"
fi
