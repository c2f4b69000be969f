mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for gr in 1 0; do for h in 1080 270 135; do
RTB200_GRAPHS=$gr python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-c3 --height $h > gpurun_out/s_h$h.json 2> gpurun_out/s_h$h.err
python -c "import json; d=json.loads(open('gpurun_out/s_h$h.json').read().strip().splitlines()[-1]); print('graphs$gr h$h', round(d['value'],1), round(d['stage_ms']['trace'],2), d['rays_per_frame'], round(d['wall_ms_per_step'],2))" || tail -3 gpurun_out/s_h$h.err
done; done
