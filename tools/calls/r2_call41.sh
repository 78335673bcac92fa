for v in "" _pwB _pwC _pwD _pwE _pwF; do
  export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so
  python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 3 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); c=d['chirp_bpsk']; print('VAR', '$v', round(c['value']), round(c['filter_ms'],2), round(c['walk_ms'],2), c['frames'])"
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ask_power_kernel -c 3 --csv --log-file gpurun_out/pw41$v.csv python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 1 --warmup 1 > /dev/null 2>&1
  grep ask_power_kernel gpurun_out/pw41$v.csv | awk -F'","' '{print "   ncu ns", $NF}' | tr '\n' ' '; echo
done
unset TMKPHY_LIB
ncu --set full --import-source on --clock-control none -k regex:ask_power_kernel -c 1 -o gpurun_out/pw41_full python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 1 --warmup 1 > /dev/null 2>&1
ls -la gpurun_out/pw41_full.ncu-rep
