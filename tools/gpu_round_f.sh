# 1 GPU: full parity suite (incl. the plotfile/render API tests), C4 bench, then the
# ncu launch list and one --set full capture of the two dominant kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_f.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_f.log
timeout 300 python bench.py --steps 20 > gpurun_out/f_C4_n1.json 2> gpurun_out/f_C4_n1.err; echo "bench rc=$?"; cat gpurun_out/f_C4_n1.json | cut -c1-600
if [ "$1" = "ncu" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_f.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'raymarch_bricks_kernel|fold_layers' -c 4 -o gpurun_out/prof_f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
