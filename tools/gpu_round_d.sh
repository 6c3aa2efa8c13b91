# 2 GPUs: all parity tests (1- and 2-rank), C4 at N=1 and N=2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_d.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_d.log
timeout 300 python bench.py --steps 20 > gpurun_out/d_C4_n1.json 2> gpurun_out/d_C4_n1.err; echo "C4 n1 rc=$?"
timeout 300 $TR --nproc-per-node 2 --master-port 29711 bench.py --gpus 2 --steps 20 > gpurun_out/d_C4_n2.json 2> gpurun_out/d_C4_n2.err; echo "C4 n2 rc=$?"
timeout 300 $TR --nproc-per-node 2 --master-port 29712 bench.py --gpus 2 --workload C5-8 --steps 10 > gpurun_out/d_C5-8_n2.json 2> gpurun_out/d_C5-8_n2.err; echo "C5-8 n2 rc=$?"
