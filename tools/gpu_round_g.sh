# 1 GPU: parity suite + C4 bench + one traced e2e pass
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_g.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_g.log
timeout 300 python bench.py --steps 20 > gpurun_out/g_C4_n1.json 2> gpurun_out/g_C4_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/g_C4_n1.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e'],'stage',d['config']['stage_ms'])
PY
AMRVR_TRACE=1 timeout 300 python bench.py --steps 3 --no-cpu-baseline 2>&1 >/dev/null | grep 'amrvr trace' | tail -8 | cut -c1-260
