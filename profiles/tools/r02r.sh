O=gpurun_out; S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 32 --warmup 3"; P=$PWD/real-time-gpu-ray-tracer_b200
for sh in 8 1; do
timeout 300 python bench.py --emulate-shard $sh $S > $O/r02r_w8_s$sh.json 2>/dev/null; python -c "
import json; d=json.loads(open('$O/r02r_w8_s$sh.json').read().strip().splitlines()[-1]); print('window 8 shard $sh', round(d['ms_per_step'],3), round(d['value'],1))"
RTB200_LIB=$P/librtb200_win16.so RTB200_MAX_WINDOW=16 timeout 300 python bench.py --emulate-shard $sh --frames-per-window 16 $S > $O/r02r_w16_s$sh.json 2>$O/r02r_w16.err; python -c "
import json; d=json.loads(open('$O/r02r_w16_s$sh.json').read().strip().splitlines()[-1]); print('window 16 shard $sh', round(d['ms_per_step'],3), round(d['value'],1), d['frame_loop']['frames_per_window'])" || tail -3 $O/r02r_w16.err
done
