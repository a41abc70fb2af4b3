#!/bin/bash
# s2act timing experiments: EFL_S2_STAGES caps the operand ring depth
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q -k "bf16 or node_level or lmax8 or reproducible" 2>&1 | tail -2
for d in 4 6 8; do
echo "stages $d: $(EFL_S2_STAGES=$d timeout 300 python tools/edge_bench.py 2>&1 | grep -E 's2act' | cut -c1-60 | tr '\n' ' ')"
done
