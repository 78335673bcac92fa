ncu --set full --import-source on --clock-control none -k regex:ask_tc_kernel -c 1 -o gpurun_out/asktc54 python bench.py --no-configs --no-links --e2e-steps 0 --no-cpu --steps 1 --warmup 1 > /dev/null 2>&1
ls -la gpurun_out/asktc54.ncu-rep
