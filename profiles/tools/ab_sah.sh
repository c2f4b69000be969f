# SAH tree rotations in the refit vs plain Karras LBVH
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="timeout 200 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-ref-cuda"
p() { python -c "
import json,sys; d=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],2), d['roofline']['per_ray'], d['lbvh_build_ms'], {k:d['c3_particle_rebuild'][k] for k in ('lbvh_rebuild_ms','tlas_ms','render_ms_1080p_1spp')} if 'c3_particle_rebuild' in d else '')" || tail -3 $1.err; }
$B > gpurun_out/sah_on.json 2> gpurun_out/sah_on.json.err; p gpurun_out/sah_on.json
$B --no-sah > gpurun_out/sah_off.json 2> gpurun_out/sah_off.json.err; p gpurun_out/sah_off.json
$B --config c5 > gpurun_out/sah_on_c5.json 2> gpurun_out/sah_on_c5.json.err; p gpurun_out/sah_on_c5.json
$B --config c5 --no-sah > gpurun_out/sah_off_c5.json 2> gpurun_out/sah_off_c5.json.err; p gpurun_out/sah_off_c5.json
