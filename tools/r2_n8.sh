for n in 2 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520+n)) bench.py --gpus $n --steps 5 --warmup 3 --configs c5 --e2e-steps 2 --no-links > gpurun_out/r02_bench_n$n.json 2> gpurun_out/r02_bench_n$n.err; echo "n=$n rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r02_bench_n$n.json').readline()); c=d['configs']['C5_mixed']
print($n, round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), round(d.get('e2e_pcm16',{}).get('value',0)), 'C5', round(c['msamples_per_s']), round(c['ms_per_step'],2), c.get('per_rank_ms'), c['parity'])"
done
