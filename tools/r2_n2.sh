python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02_bench_n2_final.json 2> gpurun_out/r02_bench_n2_final.err; echo "rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r02_bench_n2_final.json').readline()); c=d['configs']['C5_mixed']
print(d['n_gpus'], round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'C5', round(c['msamples_per_s']), round(c['ms_per_step'],2), c.get('per_rank_ms'), c['parity'], list(d['configs'].keys()))"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | head -c 300
