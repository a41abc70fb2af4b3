#!/bin/bash
# GPU call: parity tests, edge-path micro-benchmark, bench line, ncu launch list of bench, ncu --set full of one edge-layer pass.
set -u
mkdir -p gpurun_out
TAG=${1:-r1d}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/${TAG}_pytest.log
python tools/edge_bench.py > gpurun_out/${TAG}_edge_bench.jsonl 2>&1; echo "edge_bench exit $?"
cat gpurun_out/${TAG}_edge_bench.jsonl
python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json
if [ "${NCU:-1}" = "1" ]; then
CMD="python tools/edge_bench.py --profile --batch 8"
$CMD > gpurun_out/${TAG}_plain_eb.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:^(tc_gemm|s2act|rotate|segment|alpha|wigner|so3|radial|colsum|splitk|edge_degree|rmsnorm|gemm_)' -c 150 -f -o gpurun_out/${TAG}_edge $CMD > gpurun_out/${TAG}_ncu_edge.log 2>&1
echo "ncu edge exit $?"
tail -3 gpurun_out/${TAG}_ncu_edge.log
CMD2="python bench.py --steps 1 --warmup 1 --batch 4 --no-cpu-baseline"
$CMD2 > gpurun_out/${TAG}_plain_b4.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches_b4.csv $CMD2 > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
fi
