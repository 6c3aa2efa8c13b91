# 1 GPU: final parity suite, default bench line, and the 4096-box ordering time (C5-16 at one rank)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_k.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_k.log
timeout 300 python bench.py > gpurun_out/k_C4_n1.json 2> gpurun_out/k_C4_n1.err; echo "bench rc=$?"
timeout 200 python bench.py --workload C5-16 --steps 5 --no-cpu-baseline > gpurun_out/k_C5-16_n1.json 2> gpurun_out/k_C5-16_n1.err; echo "c5 rc=$?"
python - <<'PY'
import json
for f in ("k_C4_n1","k_C5-16_n1"):
    d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f,'value',round(d['value'],1),'ms',round(d['ms_per_step'],4),'e2e',round(d['e2e']['ms_per_step'],2),'launches',d.get('gpu_launches'),{k:round(v,4) for k,v in d['config']['stage_ms'].items()})
PY
