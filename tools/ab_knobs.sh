set -u
for st in 4 6 8; do
  echo "== EFL_S2_STAGES=$st"; EFL_S2_STAGES=$st python tools/edge_bench.py 2>/dev/null | grep -E "s2act|summary" | cut -c1-200
done
echo "== EFL_ROTBACK_SPLIT=0"; EFL_ROTBACK_SPLIT=0 python tools/edge_bench.py 2>/dev/null | grep -E "rotate_back|summary" | cut -c1-200
echo "== default again"; python tools/edge_bench.py 2>/dev/null | grep -E "rotate_back|s2act|summary" | cut -c1-200
for ch in 524288 1048576; do
  echo "== EFL_MAX_CHUNK_EDGES=$ch"; EFL_MAX_CHUNK_EDGES=$ch python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; b=json.loads(sys.stdin.read()); print(b['ms_per_step'], b['e2e']['ms_per_step'], b['gpu_launches'], b['clocks']['sm_mhz'])"
done
echo "== default chunk"; python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; b=json.loads(sys.stdin.read()); print(b['ms_per_step'], b['e2e']['ms_per_step'], b['gpu_launches'], b['clocks']['sm_mhz'])"
