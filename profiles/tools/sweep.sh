B="python bench.py --steps 2 --warmup 2 --no-cpu-baseline"
P=real-time-gpu-ray-tracer_b200
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" $B > gpurun_out/s_$name.json 2> gpurun_out/s_$name.err; python -c "import json; d=json.load(open('gpurun_out/s_$name.json')); print('$name', round(d['value'],1), round(d['stage_ms']['trace'],1), d['roofline']['per_ray']['node_fetches'], d['roofline']['per_ray']['prim_fetches'])" || tail -3 gpurun_out/s_$name.err; }
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run base X=1
for lm in 1 2; do run lm$lm RTB200_LEAF_MAX=$lm; done
run lm1_nm20 RTB200_LEAF_MAX=1 RTB200_NODE_MIN=20
run lm1_nm24 RTB200_LEAF_MAX=1 RTB200_NODE_MIN=24
run lm1_b6 RTB200_LEAF_MAX=1 RTB200_TRACE_BLOCKS_PER_SM=6
