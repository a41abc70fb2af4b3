#!/bin/bash
# GPU call: parity tests, bench line (dominant-only profile in the timed region), lmax-8 stress with the GEMM route.
set -u
mkdir -p gpurun_out
TAG=${1:-r1h}
T0=$(date +%s)
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? ($(( $(date +%s) - T0 )) s)"
tail -6 gpurun_out/${TAG}_pytest.log
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $? ($(( $(date +%s) - T0 )) s)"
cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
timeout 300 python tools/stress_bench.py --profile > gpurun_out/${TAG}_stress.jsonl 2> gpurun_out/${TAG}_stress.err; echo "stress exit $? ($(( $(date +%s) - T0 )) s)"
cat gpurun_out/${TAG}_stress.jsonl; tail -5 gpurun_out/${TAG}_stress.err
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "reference arm exit $? ($(( $(date +%s) - T0 )) s)"
cut -c1-400 gpurun_out/${TAG}_bench_ref.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
du -sh gpurun_out
