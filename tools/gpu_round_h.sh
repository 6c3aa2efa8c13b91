# 8 GPUs: multi-rank parity tests, C4 at 8 and 2, C5-16 at 8, one traced C4 run at 8
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu_h.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_h.log
timeout 200 $TR --nproc-per-node 8 --master-port 29801 bench.py --gpus 8 --steps 20 > gpurun_out/h_C4_n8.json 2> gpurun_out/h_C4_n8.err; echo "C4 n8 rc=$?"
timeout 200 $TR --nproc-per-node 2 --master-port 29802 bench.py --gpus 2 --steps 20 > gpurun_out/h_C4_n2.json 2> gpurun_out/h_C4_n2.err; echo "C4 n2 rc=$?"
timeout 200 $TR --nproc-per-node 4 --master-port 29805 bench.py --gpus 4 --steps 20 > gpurun_out/h_C4_n4.json 2> gpurun_out/h_C4_n4.err; echo "C4 n4 rc=$?"
timeout 200 $TR --nproc-per-node 8 --master-port 29803 bench.py --gpus 8 --workload C5-16 --steps 10 > gpurun_out/h_C5-16_n8.json 2> gpurun_out/h_C5-16_n8.err; echo "C5-16 n8 rc=$?"
AMRVR_TRACE=1 timeout 200 $TR --nproc-per-node 8 --master-port 29804 bench.py --gpus 8 --steps 5 > gpurun_out/h_trace_n8.json 2> gpurun_out/h_trace_n8.err; echo "trace rc=$?"
python - <<'PY'
import json
for f in ("h_C4_n8","h_C4_n4","h_C4_n2","h_C5-16_n8"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e_ms", round(d["e2e"]["ms_per_step"],2), d["config"]["stage_ms"])
    except Exception as e: print(f, "ERR", e)
PY
grep 'trace' gpurun_out/h_trace_n8.err | tail -6 | cut -c1-400
