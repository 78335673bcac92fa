python bench.py > gpurun_out/r02_bench_f.json 2> gpurun_out/r02_bench_f.err; echo bench rc=$?; tail -2 gpurun_out/r02_bench_f.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_ref_f.json 2>/dev/null; echo ref rc=$?
