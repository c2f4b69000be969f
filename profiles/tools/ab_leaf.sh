# leaf-phase variants of the path kernel (prefetch of the first leaf record at the end of a node step; all pending
# primitives in one LEAF round) against the default, on C2 and C5; then the config-3 frame-sharding test and bench
B="timeout 200 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-c3"
p() { python -c "import json,sys; d=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],2))" || tail -3 $1.err; }
for v in "" pfl drain pfdrain; do
  for c in c2 c5; do
    o=gpurun_out/leaf_${v:-default}_$c.json
    if [ -z "$v" ]; then $B --config $c > $o 2> $o.err; else RTB200_LIB=$PWD/real-time-gpu-ray-tracer_b200/librtb200_$v.so $B --config $c > $o 2> $o.err; fi
    p $o
  done
done
for v in pfl drain pfdrain; do
  RTB200_LIB=$PWD/real-time-gpu-ray-tracer_b200/librtb200_$v.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "frame_matches_oracle or comparison_modes" 2>&1 | tail -1
done
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "vtk" 2>&1 | tail -2
timeout 300 python bench.py --config c3 --steps 3 --warmup 3 > gpurun_out/c3_n1.json 2> gpurun_out/c3_n1.err; tail -c 1500 gpurun_out/c3_n1.json; tail -3 gpurun_out/c3_n1.err
