CMD="python bench.py --no-configs --no-cpu --e2e-steps 0 --steps 1 --warmup 1"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"ask_walk_kernel" -s 2 -c 2 -o gpurun_out/r02_ask_walk2 $CMD > gpurun_out/ncu_aw.log 2>&1; echo rc=$?; tail -1 gpurun_out/ncu_aw.log
