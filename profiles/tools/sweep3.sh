# path kernel tuning after the cubic-grid change: voting thresholds, 7 CTAs x 72 registers
B="timeout 120 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-c3"
p() { python -c "import json,sys; d=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],2))" || tail -3 $1.err; }
$B > gpurun_out/s_base.json 2> gpurun_out/s_base.json.err; p gpurun_out/s_base.json
RTB200_LIB=$PWD/real-time-gpu-ray-tracer_b200/librtb200_pb7.so $B > gpurun_out/s_pb7.json 2> gpurun_out/s_pb7.json.err; p gpurun_out/s_pb7.json
for nm in 14 17 23; do RTB200_NODE_MIN=$nm $B > gpurun_out/s_n$nm.json 2> gpurun_out/s_n$nm.json.err; p gpurun_out/s_n$nm.json; done
for sm in 6 10; do RTB200_SHADE_MIN=$sm $B > gpurun_out/s_s$sm.json 2> gpurun_out/s_s$sm.json.err; p gpurun_out/s_s$sm.json; done
