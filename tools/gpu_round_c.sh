mkdir -p gpurun_out
for wl in C4 C5-4 C5-16; do
  timeout 300 python bench.py --workload $wl --steps 10 --no-cpu-baseline > gpurun_out/c_${wl}_n1.json 2> gpurun_out/c_${wl}_n1.err; echo "$wl n1 rc=$?"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fold_layers_kernel -s 3 -c 1 -o gpurun_out/prof_fold_v3 python bench.py --workload C5-4 --steps 2 --no-cpu-baseline > gpurun_out/ncu_fold.log 2>&1; echo "ncu rc=$?"
