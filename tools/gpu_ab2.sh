#!/bin/bash
set -u
mkdir -p gpurun_out
TAG=${1:-ab2}
timeout 300 python -m pytest tests -m gpu -x -q -k "bf16 or lmax8 or reproducible" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${TAG}_pytest.log | cut -c1-600
EFL_ROTBACK_TC=1 timeout 300 python tools/edge_bench.py > gpurun_out/${TAG}_eb_tc.jsonl 2>&1; grep -E "rotate_back|summary|Error|error" gpurun_out/${TAG}_eb_tc.jsonl | cut -c1-300
EFL_ROTBACK_TC=0 timeout 300 python tools/edge_bench.py > gpurun_out/${TAG}_eb_simt.jsonl 2>&1; grep -E "rotate_back|summary" gpurun_out/${TAG}_eb_simt.jsonl | cut -c1-300
