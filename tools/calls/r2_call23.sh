# round 2, call 16: whole GPU suite, default bench line, reference arm, launch list
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | grep -v Warning | tail -6 | tee gpurun_out/r02_gputests_e.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r02_bench_e.json 2> gpurun_out/r02_bench_e.err; echo bench rc=$?; tail -3 gpurun_out/r02_bench_e.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_ref_e.json 2>/dev/null; echo ref rc=$?
CMD="python bench.py --no-cpu --e2e-steps 0 --no-configs --steps 2 --warmup 1"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_e.csv $CMD > gpurun_out/ncu_l.log 2>&1; echo launches rc=$?
ncu --set full --clock-control none --import-source on -k regex:"scan_tc_kernel|cluster_kernel|walk_kernel|ask_tc_kernel|ask_power_kernel|ask_xmax_kernel|ask_walk_kernel|mcm_demod" -s 4 -c 12 -o gpurun_out/r02_final_full $CMD > gpurun_out/ncu_ff.log 2>&1; echo full rc=$?; tail -1 gpurun_out/ncu_ff.log
