#!/bin/bash
# One parameterised GPU call (replaces the round-1 gpu_*.sh scratch scripts).  Usage on the box, through gpurun:
#   gpurun --timeout 1500 -- 'bash tools/gpu_call.sh TAG [steps...]'
# steps (default "test smoke bench"): test testall smoke bench config3 edge kernel radius stress ref probe ncu_edge ncu_launch ncu_src prof
# Every step writes gpurun_out/${TAG}_<step>.*; a number printed under ncu is never a bench value.
set -u
mkdir -p gpurun_out
TAG=${1:-r2a}; shift || true
STEPS=${*:-test smoke bench}
T0=$(date +%s)
el() { echo "[$(( $(date +%s) - T0 )) s] $*"; }
KREGEX='regex:^(tc_gemm|s2act|rotate|segment|alpha|wigner|so3|radial|colsum|splitk|edge_degree|rmsnorm|gemm_|neighbour|rowsort|bin_|scatter_|emit_|knn|mha_|grouped)'
for S in $STEPS; do
case $S in
test)
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; el "pytest exit $?"
  tail -8 gpurun_out/${TAG}_pytest.log ;;
testall)
  timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; el "pytest (no -x) exit $?"
  tail -25 gpurun_out/${TAG}_pytest.log ;;
smoke)
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; el "smoke exit $?"
  tail -3 gpurun_out/${TAG}_smoke.log ;;
bench)
  timeout 600 python bench.py --steps ${BENCH_STEPS:-10} --warmup 3 ${BENCH_ARGS:-} > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; el "bench exit $?"
  cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err ;;
bench_x3)
  timeout 600 python bench.py --steps 3 --warmup 3 --precision bf16x3 > gpurun_out/${TAG}_bench_bf16x3.json 2> gpurun_out/${TAG}_bench_bf16x3.err; el "bench bf16x3 exit $?"
  cat gpurun_out/${TAG}_bench_bf16x3.json | cut -c1-1500; tail -3 gpurun_out/${TAG}_bench_bf16x3.err ;;
bench_fp32)
  timeout 900 python bench.py --steps 2 --warmup 3 --precision fp32 --no-cpu-baseline > gpurun_out/${TAG}_bench_fp32.json 2> gpurun_out/${TAG}_bench_fp32.err; el "bench fp32 exit $?"
  cat gpurun_out/${TAG}_bench_fp32.json | cut -c1-1500; tail -3 gpurun_out/${TAG}_bench_fp32.err ;;
bench2)
  timeout 600 python bench.py --steps ${BENCH_STEPS:-10} --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench_run2.json 2> gpurun_out/${TAG}_bench_run2.err; el "bench (second run) exit $?"
  python -c "import json;b=json.load(open('gpurun_out/${TAG}_bench_run2.json'));print('run2 ms/step',b['ms_per_step'],'e2e',b['e2e']['ms_per_step'],b['allocator'],b['clocks'])" ;;
config5)
  timeout 600 python bench.py --config 5 --steps 3 --warmup 2 > gpurun_out/${TAG}_bench_config5.json 2> gpurun_out/${TAG}_bench_config5.err; el "bench config 5 exit $?"
  cut -c1-1800 gpurun_out/${TAG}_bench_config5.json; tail -3 gpurun_out/${TAG}_bench_config5.err ;;
ref)
  timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; el "reference arm exit $?"
  cut -c1-500 gpurun_out/${TAG}_bench_ref.json ;;
edge)
  timeout 300 python tools/edge_bench.py > gpurun_out/${TAG}_edge_bench.jsonl 2>&1; el "edge_bench exit $?"
  cat gpurun_out/${TAG}_edge_bench.jsonl ;;
kernel)
  timeout 300 python tools/kernel_bench.py > gpurun_out/${TAG}_kernel_bench.jsonl 2>&1; el "kernel_bench exit $?"
  cat gpurun_out/${TAG}_kernel_bench.jsonl ;;
radius)
  timeout 300 python tools/radius_bench.py > gpurun_out/${TAG}_radius_bench.jsonl 2>&1; el "radius_bench exit $?"
  cat gpurun_out/${TAG}_radius_bench.jsonl ;;
stress)
  timeout 400 python tools/stress_bench.py --profile > gpurun_out/${TAG}_stress.jsonl 2> gpurun_out/${TAG}_stress.err; el "stress exit $?"
  cat gpurun_out/${TAG}_stress.jsonl; tail -3 gpurun_out/${TAG}_stress.err ;;
ncu_edge)   # ncu --set full of ONE edge-layer pass (fwd+bwd) at the bench's chunk size, raw CSV exported on the box
  timeout 600 ncu --set full --clock-control none -k "$KREGEX" -c ${NCU_COUNT:-120} -f -o /tmp/${TAG}_edge python tools/edge_bench.py --profile --batch ${NCU_BATCH:-9} > gpurun_out/${TAG}_ncu_edge.log 2>&1
  el "ncu edge exit $?"; tail -2 gpurun_out/${TAG}_ncu_edge.log
  if [ -f /tmp/${TAG}_edge.ncu-rep ]; then
    ncu -i /tmp/${TAG}_edge.ncu-rep --page raw --csv > gpurun_out/${TAG}_edge_raw.csv 2>/dev/null
    SZ=$(stat -c %s /tmp/${TAG}_edge.ncu-rep); echo "rep bytes $SZ"
    if [ "$SZ" -lt 40000000 ]; then cp /tmp/${TAG}_edge.ncu-rep gpurun_out/; fi
  fi ;;
ncu_launch)  # launch list (gpu__time_duration only) of a small whole-step run
  timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches_b4.csv python bench.py --steps 1 --warmup 3 --batch 4 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launches.log 2>&1
  el "ncu launches exit $?" ;;
ncu_src)    # source-level capture of one kernel: NCU_KERNEL=regex NCU_CMD="python tools/..." 
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:${NCU_KERNEL} -c ${NCU_COUNT:-3} -f -o /tmp/${TAG}_src ${NCU_CMD} > gpurun_out/${TAG}_ncu_src.log 2>&1
  el "ncu src exit $?"
  ncu -i /tmp/${TAG}_src.ncu-rep --page source --csv --print-source sass > gpurun_out/${TAG}_src_sass.csv 2>/dev/null
  ncu -i /tmp/${TAG}_src.ncu-rep --page raw --csv > gpurun_out/${TAG}_src_raw.csv 2>/dev/null ;;
overlap)
  timeout 300 python tools/overlap_probe.py > gpurun_out/${TAG}_overlap_probe.jsonl 2>&1; el "overlap_probe exit $?"
  cat gpurun_out/${TAG}_overlap_probe.jsonl | tail -5 ;;
probe)
  timeout 300 python tools/e2e_probe.py > gpurun_out/${TAG}_e2e_probe.txt 2>&1; el "e2e_probe exit $?"
  cat gpurun_out/${TAG}_e2e_probe.txt | tail -6 ;;
config3)
  timeout 600 python bench.py --config 3 --steps ${BENCH_STEPS:-5} --warmup 3 > gpurun_out/${TAG}_bench_config3.json 2> gpurun_out/${TAG}_bench_config3.err; el "bench config 3 exit $?"
  cat gpurun_out/${TAG}_bench_config3.json; tail -5 gpurun_out/${TAG}_bench_config3.err ;;
prof)
  timeout 200 python tools/profile_step.py > gpurun_out/${TAG}_torch_profile.txt 2>&1; el "profile_step exit $?"
  timeout 120 python tools/cpu_profile_step.py > gpurun_out/${TAG}_cpu_profile.txt 2>&1; el "cpu_profile_step exit $?"
  head -40 gpurun_out/${TAG}_cpu_profile.txt | cut -c1-160 ;;
*) echo "unknown step $S" ;;
esac
done
du -sh gpurun_out; el done
