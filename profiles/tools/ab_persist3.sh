# persistent path kernel: tile hand-out order (dear-first from the previous frame's cost vs scan order) on one GPU rendering
# the full frame and rank 0's tiles of an 8/4/2-way split; wave loop beside it
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; tail -3 gpurun_out/t.log
B="timeout 120 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-c3"
p() { python -c "import json,sys; d=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],2), d['stage_ms'], d['gpu_launches'])" || tail -3 $1.err; }
for o in 1 0; do
RTB200_TILE_ORDER=$o $B > gpurun_out/p_full_o$o.json 2> gpurun_out/p_full_o$o.json.err; p gpurun_out/p_full_o$o.json
for w in 8 4 2; do
RTB200_TILE_ORDER=$o $B --emulate-shard $w > gpurun_out/p_s${w}_o$o.json 2> gpurun_out/p_s${w}_o$o.json.err; p gpurun_out/p_s${w}_o$o.json
done; done
$B --emulate-shard 8 --wavefront > gpurun_out/w_s8.json 2> gpurun_out/w_s8.json.err; p gpurun_out/w_s8.json
for b in 3 4 5; do
RTB200_TRACE_BLOCKS_PER_SM=$b $B --emulate-shard 8 > gpurun_out/p_s8_b$b.json 2> gpurun_out/p_s8_b$b.json.err; p gpurun_out/p_s8_b$b.json
done
