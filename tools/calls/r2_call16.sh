# round 2, call 16: whole GPU suite, default bench line, reference arm, launch list
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | grep -v Warning | tail -6 | tee gpurun_out/r02_gputests_c.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r02_bench_c.json 2> gpurun_out/r02_bench_c.err; echo bench rc=$?; tail -3 gpurun_out/r02_bench_c.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_ref_c.json 2>/dev/null; echo ref rc=$?
CMD="python bench.py --no-cpu --e2e-steps 0 --no-configs --steps 2 --warmup 1"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_c.csv $CMD > gpurun_out/ncu_l.log 2>&1; echo launches rc=$?
