timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | grep -v Warning | tail -4
for n in 1 256 4096; do python tools/bench_links.py $n 100; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/links45.csv python tools/bench_links.py 256 12 > /dev/null 2>&1
grep -v "^==" gpurun_out/links45.csv | awk -F'","' 'NR>1{print substr($5,1,40), $NF}' | tail -14
