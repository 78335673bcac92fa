CMD="python bench.py --no-configs --no-cpu --e2e-steps 0 --steps 1 --warmup 1"
ncu --set full --clock-control none --import-source on -k regex:"ask_" -s 16 -c 4 -o gpurun_out/r02_ask $CMD > gpurun_out/ncu_ask.log 2>&1; echo rc=$?; tail -2 gpurun_out/ncu_ask.log
