#!/bin/bash
# GPU call: parity tests, GEMM micro-bench, bench line, ncu --set full of the GEMM shapes and the HBM-bound kernels.
set -u
mkdir -p gpurun_out
TAG=${1:-r1b}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/${TAG}_pytest.log
python tools/kernel_bench.py > gpurun_out/${TAG}_kernel_bench.jsonl 2>&1; echo "kernel_bench exit $?"
cat gpurun_out/${TAG}_kernel_bench.jsonl
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json
CMD="python tools/kernel_bench.py --profile --ec 131072"
$CMD > gpurun_out/${TAG}_plain_kb.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -f -o gpurun_out/${TAG}_gemm $CMD > gpurun_out/${TAG}_ncu_gemm.log 2>&1
echo "ncu gemm exit $?"
CMD2="python bench.py --steps 1 --warmup 3 --batch 8 --no-cpu-baseline"
$CMD2 > gpurun_out/${TAG}_plain_b8.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:rotate_back|alpha_logits|rotate_in|segment_sum|so3_|gemm_nt_kernel|gemm_tn_kernel' -s 60 -c 24 -f -o gpurun_out/${TAG}_hbm $CMD2 > gpurun_out/${TAG}_ncu_hbm.log 2>&1
echo "ncu hbm exit $?"
