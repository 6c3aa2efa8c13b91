#!/bin/bash
# The GPU calls of a round, one script (an N-GPU call is charged N x its box time:
# the 8-GPU call runs bench lines only, multi-rank parity has calls of its own).
# Every inner `timeout` is bounded so that their SUM stays below the gpurun --timeout
# written here (BT = seconds per bench line, PT = per pytest run, NT = per ncu capture):
# a slow step is cut by its own timeout, its rc is printed, and the steps after it still run.
#
#   gpurun --timeout 720  -- 'bash tools/gpu.sh one'              # PT + 3 BT: tests + C4/C3/C2 lines
#   gpurun --timeout 600  -- 'bash tools/gpu.sh launches'         # BT + NT: launch list
#   gpurun --timeout 1000 -- 'bash tools/gpu.sh profile'          # BT + 2 NT: ncu --set full of march and fold
#   gpurun --timeout 500  -- 'bash tools/gpu.sh c5one'            # 3 BT: C5 sweep at one GPU
#   gpurun --gpus 2 --timeout 300 -- 'bash tools/gpu.sh twotests' # PT: multi-rank parity + drop-in
#   gpurun --gpus 2 --timeout 660 -- 'bash tools/gpu.sh two'      # 4 BT: C4/C5 lines
#   gpurun --gpus 4 --timeout 660 -- 'bash tools/gpu.sh four'     # 4 BT
#   gpurun --gpus 8 --timeout 660 -- 'bash tools/gpu.sh eight'    # 4 BT (eighttests: PT, the 8-rank parity cases)
# TAG (env) prefixes the output files under gpurun_out/ (default r3).
BT=${BT:-150}
PT=${PT:-240}
NT=${NT:-400}
mkdir -p gpurun_out
TAG=${TAG:-r3}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
line() {  # one-line digest of bench JSON lines
python - "$TAG" "$@" <<'PY'
import json, sys
tag = sys.argv[1]
for f in sys.argv[2:]:
    try:
        d = json.loads(open(f"gpurun_out/{tag}_{f}.json").read().strip().splitlines()[-1])
        c = d.get("composite") or {}
        print(f, "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 4), "e2e_ms",
              round(d["e2e"]["ms_per_step"], 2), "launches", d.get("gpu_launches"),
              {k: round(v, 4) for k, v in d["config"]["stage_ms"].items()},
              "parity", d.get("parity"), "composite", {k: c.get(k) for k in
              ("composite_ms_max_rank", "fold_ms_max_rank", "pushed_bytes_per_rank", "stolen_tiles_per_rank")})
    except Exception as e:
        print(f, "ERR", e)
PY
}
bench() {  # bench N WORKLOAD NAME [extra args]
  local n=$1 w=$2 name=$3; shift 3
  if [ "$n" = 1 ]; then
    timeout $BT python bench.py --workload "$w" "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  else
    timeout $BT $TR --nproc-per-node "$n" --master-port $((29800 + RANDOM % 100)) bench.py --gpus "$n" \
      --workload "$w" "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  fi
  echo "$name rc=$?"; tail -2 gpurun_out/${TAG}_$name.err
}
case "$1" in
one)
  timeout $PT python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu_1.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${TAG}_pytest_gpu_1.log
  bench 1 C4 C4_n1
  bench 1 C3 C3_n1 --no-cpu-baseline --steps 10
  bench 1 C2 C2_n1 --no-cpu-baseline --steps 10
  line C4_n1 C3_n1 C2_n1
  ;;
launches)
  bench 1 C4 C4_n1 --steps 2 --warmup 3
  timeout $NT ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
    > gpurun_out/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
  ;;
profile)
  # one capture of the march (timed instantiation <0>) and one of the fold, C4, after the bench exits 0
  bench 1 C4 C4_n1 --steps 3 --no-cpu-baseline
  timeout $NT ncu --set full --clock-control none --import-source on -k regex:raymarch_bricks_kernel \
    -s 4 -c 1 -o gpurun_out/${TAG}_march python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
    > gpurun_out/${TAG}_ncu_march.log 2>&1; echo "ncu march rc=$?"
  timeout $NT ncu --set full --clock-control none --import-source on -k regex:fold_layers_kernel \
    -s 4 -c 1 -o gpurun_out/${TAG}_fold python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
    > gpurun_out/${TAG}_ncu_fold.log 2>&1; echo "ncu fold rc=$?"
  ;;
c5one)
  for w in C5-4 C5-8 C5-16; do bench 1 $w ${w}_n1 --steps 10 --no-cpu-baseline; done
  line C5-4_n1 C5-8_n1 C5-16_n1
  ;;
twotests)
  timeout $PT python -m pytest tests/test_multi_gpu.py tests/test_dropin.py -m gpu -q > gpurun_out/${TAG}_pytest_gpu_2.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${TAG}_pytest_gpu_2.log
  ;;
two)
  bench 2 C4 C4_n2
  for w in C5-4 C5-8 C5-16; do bench 2 $w ${w}_n2 --steps 10 --no-cpu-baseline; done
  line C4_n2 C5-4_n2 C5-8_n2 C5-16_n2
  ;;
four)
  bench 4 C4 C4_n4
  for w in C5-4 C5-8 C5-16; do bench 4 $w ${w}_n4 --steps 10 --no-cpu-baseline; done
  line C4_n4 C5-4_n4 C5-8_n4 C5-16_n4
  ;;
twoc4)
  bench 2 C4 C4_n2 --no-cpu-baseline
  line C4_n2
  ;;
fourc4)
  bench 4 C4 C4_n4 --no-cpu-baseline
  line C4_n4
  ;;
eightc4)
  bench 8 C4 C4_n8 --no-cpu-baseline
  line C4_n8
  ;;
eighttests)
  timeout $PT python -m pytest tests/test_multi_gpu.py -m gpu -q -k "eight_ranks" > gpurun_out/${TAG}_pytest_gpu_8.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${TAG}_pytest_gpu_8.log
  ;;
eight)
  bench 8 C4 C4_n8
  for w in C5-4 C5-8 C5-16; do bench 8 $w ${w}_n8 --steps 10 --no-cpu-baseline; done
  line C4_n8 C5-4_n8 C5-8_n8 C5-16_n8
  ;;
*) echo "usage: $0 one|launches|profile|c5one|twotests|two|four|eight|eighttests|twoc4|fourc4|eightc4"; exit 2 ;;
esac
