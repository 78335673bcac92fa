# final measurements of the round: GPU suite, default bench line, reference arm, launch list, --set full of the three headline kernels
( time timeout 1500 python -m pytest tests -m gpu -q 2>&1 | grep -v Warning | tail -4 ) > gpurun_out/r02_gputests_final.txt 2>&1
( time python bench.py > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err ) 2> gpurun_out/r02_bench_final.time
( time python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_ref_final.json 2> gpurun_out/r02_bench_ref_final.err ) 2> gpurun_out/r02_bench_ref_final.time
CMD="python bench.py --no-cpu --e2e-steps 0 --no-pcm16 --no-configs --no-links --steps 2 --warmup 1"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_final.csv $CMD > /dev/null 2>&1
CMD2="python bench.py --no-cpu --e2e-steps 0 --no-pcm16 --no-configs --no-links --no-ask --steps 1 --warmup 1"
ncu --set full --import-source on --clock-control none -k regex:"scan_tc_kernel|cluster_kernel|walk_kernel" -c 6 -o gpurun_out/r02_full_final $CMD2 > /dev/null 2>&1
cat gpurun_out/r02_gputests_final.txt gpurun_out/r02_bench_final.time gpurun_out/r02_bench_ref_final.time
head -c 600 gpurun_out/r02_bench_final.json; echo; head -c 400 gpurun_out/r02_bench_ref_final.json
