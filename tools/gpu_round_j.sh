# 1 GPU: one --set full capture of the march and the fold (run after bench.py exited 0 without ncu)
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'raymarch_bricks_kernel|fold_layers' -c 4 -o gpurun_out/prof_j2 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full_j2.log 2>&1; echo "ncu full rc=$?"
