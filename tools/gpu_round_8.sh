# 8 GPUs: 4/8-rank parity tests, then the scaling lines (C4 at 2/4/8, C5 sweep at 8, C5-8 at 4)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -k "four or eight" > gpurun_out/pytest_gpu8.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu8.log
port=29600
for n in 2 4 8; do
  port=$((port+1))
  timeout 300 $TR --nproc-per-node $n --master-port $port bench.py --gpus $n --steps 20 > gpurun_out/s_C4_n$n.json 2> gpurun_out/s_C4_n$n.err; echo "C4 n$n rc=$?"
done
for wl in C5-4 C5-8 C5-16; do
  port=$((port+1))
  timeout 300 $TR --nproc-per-node 8 --master-port $port bench.py --gpus 8 --workload $wl --steps 10 > gpurun_out/s_${wl}_n8.json 2> gpurun_out/s_${wl}_n8.err; echo "$wl n8 rc=$?"
done
port=$((port+1))
timeout 300 $TR --nproc-per-node 4 --master-port $port bench.py --gpus 4 --workload C5-8 --steps 10 > gpurun_out/s_C5-8_n4.json 2> gpurun_out/s_C5-8_n4.err; echo "C5-8 n4 rc=$?"
port=$((port+1))
AMRVR_TRACE=1 timeout 300 $TR --nproc-per-node 8 --master-port $port bench.py --gpus 8 --steps 5 > gpurun_out/s_trace_n8.json 2> gpurun_out/s_trace_n8.err; echo "trace rc=$?"
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
