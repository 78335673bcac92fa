for v in "" _xmma _xema _xen _xall; do
  export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ask_tc_kernel -c 2 --csv --log-file gpurun_out/x51$v.csv python bench.py --no-configs --no-links --e2e-steps 0 --no-cpu --steps 1 --warmup 1 > /dev/null 2>&1
  echo "VAR $v" $(grep ask_tc_kernel gpurun_out/x51$v.csv | awk -F'","' '{print $NF}')
done
