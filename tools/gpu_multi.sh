#!/bin/bash
# Multi-GPU measurements on one box: gpurun --gpus N -- 'bash tools/gpu_multi.sh N TAG [steps...]'
# steps (default "weak strong config3"): weak strong config3 config5 test
set -u
N=${1:-2}; TAG=${2:-r2n$N}; shift 2 || true
STEPS=${*:-weak strong config3}
mkdir -p gpurun_out
T0=$(date +%s)
el() { echo "[$(( $(date +%s) - T0 )) s] $*"; }
run() {  # run NAME ARGS...
  local name=$1; shift
  if [ "$N" = "1" ]; then
    timeout 420 python bench.py --gpus 1 "$@" > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err
  else
    timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N "$@" > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err
  fi
  el "$name exit $?"
  python - <<PY
import json
try:
    b=json.load(open("gpurun_out/${TAG}_${name}.json"))
    print("${name}", "n_gpus", b["n_gpus"], "ms/step", round(b["ms_per_step"],2), "edges/s", round(b["value"]), "e2e", round(b["e2e"]["value"]), "complexes/s", b["config"].get("complexes_per_s"), b.get("clocks",{}).get("sm_mhz") if b.get("clocks") else None)
except Exception as e:
    print("${name}: no JSON line", e); print(open("gpurun_out/${TAG}_${name}.err").read()[-1500:])
PY
}
for S in $STEPS; do
case $S in
weak)    run weak --steps 10 --warmup 5 --no-cpu-baseline ;;
strong)  run strong --steps 10 --warmup 5 --scaling strong --global-batch 64 --no-cpu-baseline ;;
config3) run config3 --config 3 --steps 5 --warmup 3 ;;
config5) run config5 --config 5 --steps 3 --warmup 2 ;;
test)    timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; el "pytest exit $?"; tail -3 gpurun_out/${TAG}_pytest.log ;;
*) echo "unknown step $S" ;;
esac
done
el done
