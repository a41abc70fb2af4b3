#!/bin/bash
# GPU call: parity tests, micro-benchmarks, bench line, then ncu (launch list of a small bench run; --set full of one
# edge-layer pass, raw CSV exported on the box so gpurun_out stays far below its 64 MiB limit).
set -u
mkdir -p gpurun_out
TAG=${1:-r1e}
KREGEX='regex:^(tc_gemm|s2act|rotate|segment|alpha|wigner|so3|radial|colsum|splitk|edge_degree|rmsnorm|gemm_|neighbour|rowsort|bin_|scatter_|emit_)'
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"
tail -15 gpurun_out/${TAG}_pytest.log
timeout 300 python tools/edge_bench.py > gpurun_out/${TAG}_edge_bench.jsonl 2>&1; echo "edge_bench exit $?"
cat gpurun_out/${TAG}_edge_bench.jsonl
if [ -f tools/radius_bench.py ]; then
timeout 300 python tools/radius_bench.py > gpurun_out/${TAG}_radius_bench.jsonl 2>&1; echo "radius_bench exit $?"
cat gpurun_out/${TAG}_radius_bench.jsonl
fi
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
if [ "${NCU:-1}" = "1" ]; then
CMD="python tools/edge_bench.py --profile --batch 8"
timeout 300 ncu --set full --clock-control none -k "$KREGEX" -c 80 -f -o /tmp/${TAG}_edge $CMD > gpurun_out/${TAG}_ncu_edge.log 2>&1
echo "ncu edge exit $?"; tail -2 gpurun_out/${TAG}_ncu_edge.log
if [ -f /tmp/${TAG}_edge.ncu-rep ]; then
  ncu -i /tmp/${TAG}_edge.ncu-rep --page raw --csv > gpurun_out/${TAG}_edge_raw.csv 2>/dev/null
  SZ=$(stat -c %s /tmp/${TAG}_edge.ncu-rep); echo "rep bytes $SZ"
  if [ "$SZ" -lt 30000000 ]; then cp /tmp/${TAG}_edge.ncu-rep gpurun_out/; fi
fi
CMD2="python bench.py --steps 1 --warmup 1 --batch 4 --no-cpu-baseline"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches_b4.csv $CMD2 > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
fi
du -sh gpurun_out
