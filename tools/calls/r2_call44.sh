ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size --clock-control none -c 200 --csv --log-file gpurun_out/links44.csv python tools/bench_links.py 256 12 > /dev/null 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/links44.csv')))
hi=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[hi]; 
from collections import defaultdict
d=defaultdict(dict)
for r in rows[hi+1:]:
    if len(r)<len(h): continue
    rec=dict(zip(h,r)); d[int(rec['ID'])][rec['Metric Name']]=rec['Metric Value']; d[int(rec['ID'])]['k']=rec['Kernel Name'][:60]
ids=sorted(d)
for i in ids[-40:]:
    print(i, d[i]['k'], d[i].get('gpu__time_duration.sum'), d[i].get('launch__grid_size'), d[i].get('launch__block_size'))
PY
