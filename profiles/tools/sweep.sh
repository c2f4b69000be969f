B="python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-c3"
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" $B > gpurun_out/s_$name.json 2> gpurun_out/s_$name.err; python -c "import json; d=json.loads(open('gpurun_out/s_$name.json').read().strip().splitlines()[-1]); print('$name', round(d['value'],1), round(d['stage_ms']['trace'],1), d['gpu_launches'], d['rays_per_frame'])" || tail -3 gpurun_out/s_$name.err; }
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for f in 0 48 24 12 6; do run f$f RTB200_FINISH_DIVISOR=$f; done
run f24g4 RTB200_FINISH_DIVISOR=24 RTB200_GROUPS=4
run f24g16 RTB200_FINISH_DIVISOR=24 RTB200_GROUPS=16
