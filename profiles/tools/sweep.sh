B="python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-c3"
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" $B > gpurun_out/s_$name.json 2> gpurun_out/s_$name.err; python -c "import json; d=json.load(open('gpurun_out/s_$name.json')); print('$name', round(d['value'],1), round(d['stage_ms']['trace'],1), d['gpu_launches'])" || tail -3 gpurun_out/s_$name.err; }
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run base X=1
run nm16 RTB200_NODE_MIN=16
run nm24 RTB200_NODE_MIN=24
