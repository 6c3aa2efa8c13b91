mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu2.log
for wl in C5-4 C5-8 C5-16; do
  timeout 300 python bench.py --workload $wl --steps 10 --no-cpu-baseline > gpurun_out/bench_${wl}_n1.json 2> gpurun_out/bench_${wl}_n1.err; echo "$wl n1 rc=$?"
  timeout 300 $TR --nproc-per-node 2 --master-port 29511 bench.py --gpus 2 --workload $wl --steps 10 > gpurun_out/bench_${wl}_n2.json 2> gpurun_out/bench_${wl}_n2.err; echo "$wl n2 rc=$?"
done
timeout 300 $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 > gpurun_out/bench_C4_n2.json 2> gpurun_out/bench_C4_n2.err; echo "C4 n2 rc=$?"
tail -3 gpurun_out/*.err | tail -40
