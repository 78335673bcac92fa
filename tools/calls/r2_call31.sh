python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/comm_check.py 2>&1 | grep -E "comm ok|Error|assert" | head -5
timeout 300 python -m pytest tests/test_gpu_ingest.py -m gpu -x -q 2>&1 | grep -v Warning | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 --configs c5 --e2e-steps 1 > gpurun_out/r02_bench_n2_b.json 2> gpurun_out/r02_bench_n2_b.err; echo rc=$?
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_n2_b.json')); c=d['configs']['C5_mixed']; print(round(d['value']), round(c['msamples_per_s']), c['ms_per_step'], c['per_rank_ms'], c['per_rank_records'], c['parity'])"
