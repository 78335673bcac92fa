timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
CMD="python bench.py --no-cpu --e2e-steps 0 --no-ask --no-configs --steps 1 --warmup 1"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"scan_tc_kernel" -s 1 -c 1 -o gpurun_out/r2_prof_b $CMD > gpurun_out/ncu.log 2>&1; tail -2 gpurun_out/ncu.log
