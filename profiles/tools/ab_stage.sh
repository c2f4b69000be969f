# shared-memory staging of the top BVH levels (variant build) vs default (nodes through L1/L2)
B="timeout 200 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-c3"
p() { python -c "import json,sys; d=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],2))" || tail -3 $1.err; }
$B > gpurun_out/st_off.json 2> gpurun_out/st_off.json.err; p gpurun_out/st_off.json
RTB200_LIB=$PWD/real-time-gpu-ray-tracer_b200/librtb200_stage.so $B > gpurun_out/st_on.json 2> gpurun_out/st_on.json.err; p gpurun_out/st_on.json
RTB200_LIB=$PWD/real-time-gpu-ray-tracer_b200/librtb200_stage.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "frame_matches_oracle" 2>&1 | tail -2
