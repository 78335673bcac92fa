# round 2, call 12 (8 GPUs): bare concurrent H2D copies at 1/2/4/8 ranks, then the bench at 8 / 4 / 2 ranks
nvidia-smi topo -m 2>/dev/null | head -14 > gpurun_out/r02_topo.txt; nproc >> gpurun_out/r02_topo.txt; lscpu | grep -E "Socket|NUMA node\(s\)|Model name" >> gpurun_out/r02_topo.txt
python tools/h2d_sweep.py 2>/dev/null | tee gpurun_out/r02_h2d.txt
for n in 2 4 8; do python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) tools/h2d_sweep.py 2>/dev/null | grep H2D | tee -a gpurun_out/r02_h2d.txt; done
for n in 8 4 2; do python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r02_bench_n$n.json 2> gpurun_out/r02_bench_n$n.err; echo "n=$n rc=$?"; tail -2 gpurun_out/r02_bench_n$n.err; done
