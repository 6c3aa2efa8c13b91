# 1 GPU: parity tests, then C4 / C5 bench lines (fold kernel work)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu3.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu3.log
for wl in C4 C5-4 C5-8 C5-16; do
  timeout 300 python bench.py --workload $wl --steps 10 --no-cpu-baseline > gpurun_out/b_${wl}_n1.json 2> gpurun_out/b_${wl}_n1.err; echo "$wl n1 rc=$?"
done
