# round 2, call 6: the no-alloc entry point's test (alone in a fresh process, then the whole suite) and a source-level profile of the walk on C4
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
CMD="python bench.py --configs c4 --no-cpu --e2e-steps 0 --no-ask --steps 1 --warmup 1"
$CMD > gpurun_out/c4_plain.json 2> gpurun_out/c4_plain.err && ncu --set full --clock-control none --import-source on -k regex:"walk_kernel" -s 4 -c 1 -o gpurun_out/r02_c4_walk $CMD > gpurun_out/ncu_c4.log 2>&1; echo c4 rc=$?; tail -2 gpurun_out/ncu_c4.log
