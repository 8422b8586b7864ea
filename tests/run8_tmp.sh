python -m pytest tests/test_gpu_lag.py tests/test_gpu_configs.py -m gpu -q --tb=short -s 2>&1 | grep -v "^$" | tail -30
cd gaussian-splatting-mpm_b200
for v in "MPMK_LAG=0" "MPMK_LAG=1" "MPMK_LAG_SERIAL=1" "MPMK_LAG_P2G_CTAS=3"; do
  echo "== $v"; env MPMK_TRACE=1 $v python -m harness.stagebench --warm 20 --reps 20 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['graph_ms_per_substep'], d['trace_us'])"
done
