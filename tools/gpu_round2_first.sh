# First GPU calls of the next round (cheap by construction; see the lesson in
# profiles/r1_scaling_summary.md: an 8-GPU call is charged 8x its box time, so it
# runs bench lines only -- the multi-rank parity tests go in the 2-GPU call).
#
#   gpurun --timeout 300 -- 'bash tools/gpu_round2_first.sh one'          # 1 GPU
#   gpurun --gpus 2 --timeout 300 -- 'bash tools/gpu_round2_first.sh two' # 2 GPUs
#   gpurun --gpus 8 --timeout 240 -- 'bash tools/gpu_round2_first.sh eight'
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
line() {  # one-line digest of a bench JSON
python - "$@" <<'PY'
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 4), "e2e_ms",
              round(d["e2e"]["ms_per_step"], 2), "launches", d.get("gpu_launches"),
              {k: round(v, 4) for k, v in d["config"]["stage_ms"].items()},
              "nvlink", (d.get("composite") or {}).get("nvlink_frac"))
    except Exception as e:
        print(f, "ERR", e)
PY
}
case "$1" in
one)
  timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_1.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu_1.log
  timeout 300 python bench.py > gpurun_out/r2_C4_n1.json 2> gpurun_out/r2_C4_n1.err; echo "bench rc=$?"
  line r2_C4_n1
  ;;
two)
  timeout 200 python -m pytest tests/test_multi_gpu.py tests/test_dropin.py -m gpu -x -q > gpurun_out/r2_pytest_gpu_2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu_2.log
  timeout 120 $TR --nproc-per-node 2 --master-port 29812 bench.py --gpus 2 --steps 20 > gpurun_out/r2_C4_n2.json 2> gpurun_out/r2_C4_n2.err; echo "C4 n2 rc=$?"
  line r2_C4_n2
  ;;
eight)
  timeout 100 $TR --nproc-per-node 8 --master-port 29818 bench.py --gpus 8 --steps 20 --no-cpu-baseline > gpurun_out/r2_C4_n8.json 2> gpurun_out/r2_C4_n8.err; echo "C4 n8 rc=$?"
  timeout 100 $TR --nproc-per-node 4 --master-port 29814 bench.py --gpus 4 --steps 20 --no-cpu-baseline > gpurun_out/r2_C4_n4.json 2> gpurun_out/r2_C4_n4.err; echo "C4 n4 rc=$?"
  timeout 100 $TR --nproc-per-node 8 --master-port 29819 bench.py --gpus 8 --workload C5-16 --steps 10 --no-cpu-baseline > gpurun_out/r2_C5-16_n8.json 2> gpurun_out/r2_C5-16_n8.err; echo "C5-16 n8 rc=$?"
  line r2_C4_n8 r2_C4_n4 r2_C5-16_n8
  ;;
*) echo "usage: $0 one|two|eight"; exit 2 ;;
esac
