#!/bin/bash
# GPU call: parity tests, Wigner store A/B (staged vs direct), lmax-8 stress with per-entry-point times, bench line,
# launch list of a small bench run.
set -u
mkdir -p gpurun_out
TAG=${1:-r1g}
T0=$(date +%s)
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $? ($(( $(date +%s) - T0 )) s)"
tail -4 gpurun_out/${TAG}_pytest.log
timeout 200 python tools/radius_bench.py > gpurun_out/${TAG}_radius_bench.jsonl 2>&1; echo "radius_bench exit $? ($(( $(date +%s) - T0 )) s)"
cut -c1-420 gpurun_out/${TAG}_radius_bench.jsonl
EFL_WIGNER_DIRECT=1 timeout 200 python tools/radius_bench.py --radii 8 > gpurun_out/${TAG}_radius_bench_direct.jsonl 2>&1; echo "radius_bench (direct stores) exit $?"
grep wigner gpurun_out/${TAG}_radius_bench_direct.jsonl | cut -c1-300
timeout 300 python tools/stress_bench.py --profile > gpurun_out/${TAG}_stress.jsonl 2> gpurun_out/${TAG}_stress.err; echo "stress exit $? ($(( $(date +%s) - T0 )) s)"
cat gpurun_out/${TAG}_stress.jsonl; tail -5 gpurun_out/${TAG}_stress.err
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $? ($(( $(date +%s) - T0 )) s)"
cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
if [ "${NCU:-1}" = "1" ]; then
timeout 200 ncu --set full --clock-control none -k regex:'wigner|neighbour|rowsort' -c 6 -f -o /tmp/${TAG}_wig python tools/radius_bench.py --radii 8 --profile > gpurun_out/${TAG}_ncu_wig.log 2>&1
echo "ncu wigner/radius exit $? ($(( $(date +%s) - T0 )) s)"
if [ -f /tmp/${TAG}_wig.ncu-rep ]; then ncu -i /tmp/${TAG}_wig.ncu-rep --page raw --csv > gpurun_out/${TAG}_wig_raw.csv 2>/dev/null; fi
CMD2="python bench.py --steps 1 --warmup 1 --batch 4 --no-cpu-baseline"
timeout 480 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches_b4.csv $CMD2 > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $? ($(( $(date +%s) - T0 )) s)"
fi
du -sh gpurun_out
