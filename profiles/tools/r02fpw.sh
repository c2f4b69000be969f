# 1/8 shard on one GPU, frames per window 1..8: fit ms_per_frame = w + C / frames_per_window (fixed cost per window)
O=gpurun_out; L=$O/r02_fpw_shard8.txt; : > $L
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 32 --warmup 16 --emulate-shard 8"
for f in 1 2 4 8; do
  timeout 300 python bench.py $S --frames-per-window $f > $O/r02_fpw$f.json 2>/dev/null
  python -c "
import json; d=json.loads(open('$O/r02_fpw$f.json').read().strip().splitlines()[-1]); print('shard8 frames_per_window $f ms_per_frame', round(d['ms_per_step'],3), 'sync', round(d['frame_loop']['synchronous_frame_ms'],3), d['stage_ms'])" >> $L
done
cat $L
