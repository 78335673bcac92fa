python tools/bench_links.py 256 60 2>&1 | tail -1
python tools/bench_links.py 1 60 2>&1 | tail -1
ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 60 --csv --log-file gpurun_out/links_launches.csv python tools/bench_links.py 256 40 > /dev/null 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/links_launches.csv')) if len(r)>5]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value')
for r in rows[1:31]:
    print(r[ki].split('(')[0][-40:], r[vi])
PY
