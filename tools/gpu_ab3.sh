#!/bin/bash
mkdir -p gpurun_out
for d in 0 1 2 3 4 7 8 15; do
echo "dbg $d: $(EFL_RBTC_DBG=$d EFL_ROTBACK_TC=1 timeout 300 python tools/edge_bench.py 2>&1 | grep -E 'rotate_back_aggregate_fwd' | cut -c1-120)"
done
