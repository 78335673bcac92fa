# round 2, call 5: GPU tests, the default bench line, the launch list and --set full captures of the three hot kernels
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r02_gputests.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo bench rc=$?; tail -3 gpurun_out/r02_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_ref.json 2>/dev/null; echo ref rc=$?
CMD="python bench.py --no-cpu --e2e-steps 0 --no-ask --no-configs --steps 2 --warmup 1"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_l.log 2>&1; echo launches rc=$?
ncu --set full --clock-control none --import-source on -k regex:"scan_tc_kernel|cluster_kernel|walk_kernel" -s 3 -c 3 -o gpurun_out/r02_full $CMD > gpurun_out/ncu_f.log 2>&1; echo full rc=$?; tail -2 gpurun_out/ncu_f.log
