# round 2, call 1: toolchain probe on the GPU box + the configurations that were never timed in round 1
{ echo "== toolchain probe"; for t in rustc cargo go javac node; do echo -n "$t: "; (command -v $t && $t --version 2>&1 | head -1) || echo absent; done; ls ~/.cargo/bin 2>&1 | head; nproc; lscpu | grep -E 'Model name|Socket|NUMA node\(s\)'; nvidia-smi -L | head -2; } > gpurun_out/r2_probe.txt 2>&1
cat gpurun_out/r2_probe.txt
run() { tag=$1; shift; python bench.py --no-cpu --e2e-steps 0 --no-ask --steps 5 --warmup 3 "$@" 2>gpurun_out/r2_$tag.err | python -c "
import json,sys
l=sys.stdin.readline()
open('gpurun_out/r2_$tag.json','w').write(l)
d=json.loads(l)
print('$tag', round(d['value']), round(d['ms_per_step'],3), {k: round(v,3) for k,v in d['stage_ms_per_step'].items()}, d['decode']['walk_last_step'])"; }
run c2_4b5b
run c2_man --line-code manchester
run c2_4b5b_chunk1200 --chunk 1200
run c2_man_chunk1200 --line-code manchester --chunk 1200
for s in -3 0 3 6 10 30; do run snr_$s --snrs=$s --streams 512; done
for s in 0 3 10; do run man_snr_$s --snrs=$s --streams 512 --line-code manchester; done
