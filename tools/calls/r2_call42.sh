for v in "" _pwG _pwH; do
  export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so
  timeout 600 python -m pytest tests/test_gpu_ask.py -m gpu -x -q 2>&1 | grep -v Warning | tail -2
  python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 3 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); c=d['chirp_bpsk']; print('VAR', '$v', round(c['value']), round(c['filter_ms'],2), round(c['walk_ms'],2), c['frames'])"
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ask_power_kernel -c 3 --csv --log-file gpurun_out/pw42$v.csv python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 1 --warmup 1 > /dev/null 2>&1
  grep ask_power_kernel gpurun_out/pw42$v.csv | awk -F'","' '{print "   ncu ns", $NF}' | tr '\n' ' '; echo
done
