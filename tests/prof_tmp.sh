cd gaussian-splatting-mpm_b200
for lib in "" lib/variants/libmpmk_g2pbulk.so lib/variants/libmpmk_p2gbulk.so; do
  echo "== lib=$lib"; MPMK_LIB=$lib timeout 120 python -m harness.stagebench --warm 30 --reps 30 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['graph_ms_per_substep'], d['stage_ms'], d['finite'])"
done
cd ..
MPMK_LIB=$PWD/gaussian-splatting-mpm_b200/lib/variants/libmpmk_g2pbulk.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "stagewise or trajectory or lagged" 2>&1 | tail -3
MPMK_LIB=$PWD/gaussian-splatting-mpm_b200/lib/variants/libmpmk_p2gbulk.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "stagewise or trajectory or lagged" 2>&1 | tail -3
