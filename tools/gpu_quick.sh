#!/bin/bash
# tests + GEMM micro-bench + edge bench + bench line (no ncu)
set -u
mkdir -p gpurun_out
TAG=${1:-q}
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_pytest.log
timeout 300 python tools/kernel_bench.py > gpurun_out/${TAG}_kernel_bench.jsonl 2>&1; echo "kernel_bench exit $?"; cut -c1-220 gpurun_out/${TAG}_kernel_bench.jsonl
timeout 300 python tools/edge_bench.py > gpurun_out/${TAG}_edge_bench.jsonl 2>&1; echo "edge_bench exit $?"; cut -c1-200 gpurun_out/${TAG}_edge_bench.jsonl
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
