# CTA size of the path kernel (slots are released per CTA when a window drains): 128 (default) / 64 / 32 threads, full frame and 1/8 shard
O=gpurun_out; L=$O/r02_pt_ab.txt; : > $L
P=real-time-gpu-ray-tracer_b200
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 32 --warmup 16"
for v in "" _pt64 _pt32 ""; do
  RTB200_LIB=$PWD/$P/librtb200$v.so timeout 300 python bench.py $S --emulate-shard 8 > $O/r02_pt_shard8$v.json 2>/dev/null
  RTB200_LIB=$PWD/$P/librtb200$v.so timeout 300 python bench.py $S --steps 16 --warmup 16 > $O/r02_pt_full$v.json 2>/dev/null
  python -c "
import json
a=json.loads(open('$O/r02_pt_shard8$v.json').read().strip().splitlines()[-1]); b=json.loads(open('$O/r02_pt_full$v.json').read().strip().splitlines()[-1])
print('variant [$v] shard8 ms/frame', round(a['ms_per_step'],3), 'sync', round(a['frame_loop']['synchronous_frame_ms'],2), '| full', round(b['ms_per_step'],2), round(b['value'],1), 'sync', round(b['frame_loop']['synchronous_frame_ms'],2))" >> $L
done
cat $L
