# round 2, call 13: the tensor-core chirp filter: parity tests, A/B against the FFT filter, ncu of the new kernel
timeout 900 python -m pytest tests/test_gpu_ask.py tests/test_gpu_mcm.py tests/test_gpu_configs.py tests/test_gpu_ingest.py -m gpu -x -q 2>&1 | grep -v Warning | tail -12
for v in 1 0; do
  if [ $v = 1 ]; then export TMK_ASK_FFT=1; else unset TMK_ASK_FFT; fi
  timeout 600 python bench.py --no-configs --no-cpu --e2e-steps 0 --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('FFT=$v', json.dumps(d['chirp_bpsk']))"
done
CMD="python bench.py --no-configs --no-cpu --e2e-steps 0 --steps 1 --warmup 1"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"ask_tc_kernel|ask_power_kernel|ask_walk_kernel" -s 6 -c 3 -o gpurun_out/r02_ask_tc $CMD > gpurun_out/ncu_ask_tc.log 2>&1; echo rc=$?; tail -2 gpurun_out/ncu_ask_tc.log
