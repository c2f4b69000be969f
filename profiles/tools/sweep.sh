B="python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-c3"
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" $B > gpurun_out/s_$name.json 2> gpurun_out/s_$name.err; python -c "import json; d=json.load(open('gpurun_out/s_$name.json')); print('$name', round(d['value'],1), round(d['stage_ms']['trace'],1), d['lbvh_build_ms'], d['roofline']['per_ray'])" || tail -3 gpurun_out/s_$name.err; }
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run base X=1
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-ref-cuda --no-c3"
ncu --metrics gpu__time_duration.sum --clock-control none -c 20 --csv --log-file gpurun_out/launches_build.csv $CMD > gpurun_out/ncu1.log 2>&1
grep -E "collapse_kernel|refit|karras" gpurun_out/launches_build.csv | head -4 | cut -c1-60,200-400
