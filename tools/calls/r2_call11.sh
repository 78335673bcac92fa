# round 2, call 11: whole GPU suite + default bench line + reference arm
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | grep -v Warning | tail -6 | tee gpurun_out/r02_gputests_b.txt
python bench.py > gpurun_out/r02_bench_b.json 2> gpurun_out/r02_bench_b.err; echo bench rc=$?; tail -3 gpurun_out/r02_bench_b.err
