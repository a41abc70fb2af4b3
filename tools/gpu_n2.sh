#!/bin/bash
# 2-GPU call: bench line and config-5 stress, one rank per GPU over NCCL.
set -u
mkdir -p gpurun_out
TAG=${1:-r1n2}
N=${2:-2}
T0=$(date +%s)
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $? ($(( $(date +%s) - T0 )) s)"
cat gpurun_out/${TAG}_bench.json | cut -c1-900; tail -3 gpurun_out/${TAG}_bench.err
[ "${STRESS:-1}" = "1" ] && timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/stress_bench.py > gpurun_out/${TAG}_stress.jsonl 2> gpurun_out/${TAG}_stress.err; echo "stress exit $? ($(( $(date +%s) - T0 )) s)"
cat gpurun_out/${TAG}_stress.jsonl; tail -3 gpurun_out/${TAG}_stress.err
exit 0
