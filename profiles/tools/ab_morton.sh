# cubic Morton grid (default) vs per-axis scaling (variant lib): nodes per ray, Mrays/s, C3 frame
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; tail -3 gpurun_out/t.log
B="timeout 200 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-ref-cuda"
p() { python -c "
import json,sys; d=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],2), d['roofline']['per_ray'], d['lbvh_build_ms'], d.get('c3_particle_rebuild'))" || tail -3 $1.err; }
$B > gpurun_out/m_cubic.json 2> gpurun_out/m_cubic.json.err; p gpurun_out/m_cubic.json
RTB200_LIB=$PWD/real-time-gpu-ray-tracer_b200/librtb200_mpa.so $B > gpurun_out/m_peraxis.json 2> gpurun_out/m_peraxis.json.err; p gpurun_out/m_peraxis.json
