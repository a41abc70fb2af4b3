#!/bin/bash
# quick A/B of the rotate-in backward variants + parity tests
set -u
mkdir -p gpurun_out
TAG=${1:-ab}
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_pytest.log
EFL_ROTIN_BWD=nodes timeout 300 python tools/edge_bench.py > gpurun_out/${TAG}_eb_nodes.jsonl 2>&1; grep -E "rotate_in|segment_sum|summary" gpurun_out/${TAG}_eb_nodes.jsonl
EFL_ROTIN_BWD=edges timeout 300 python tools/edge_bench.py > gpurun_out/${TAG}_eb_edges.jsonl 2>&1; grep -E "rotate_in|segment_sum|summary" gpurun_out/${TAG}_eb_edges.jsonl
