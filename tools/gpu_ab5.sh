#!/bin/bash
# rotate-back: one CTA per node with a warp per degree group (0) vs one single-warp CTA per (node, group) (1)
timeout 300 python -m pytest tests -m gpu -x -q -k "bf16 or lmax8 or reproducible or fp32" 2>&1 | tail -2
for w in 0 1; do
echo "split $w: $(EFL_ROTBACK_SPLIT=$w timeout 300 python tools/edge_bench.py 2>&1 | grep -E 'rotate_back' | cut -c1-70 | tr '\n' ' ')"
done
