# A/B of LEAF-step variants of the path kernel (csrc/Makefile VARIANT=lloop / lpf1 / lpf2 / lboth) on the headline workload, C5 and C4
TAG=${1:-r02z}
O=gpurun_out
mkdir -p $O
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 16 --warmup 3"
P=real-time-gpu-ray-tracer_b200
for v in "" _lloop _lpf1 _lpf2 _lboth ""; do
  RTB200_LIB=$PWD/$P/librtb200$v.so timeout 300 python bench.py $S > $O/${TAG}_ab$v.json 2> $O/${TAG}_ab$v.err
  python -c "
import json; d=json.loads(open('$O/${TAG}_ab$v.json').read().strip().splitlines()[-1]); print('variant [$v]', round(d['value'],1), round(d['ms_per_step'],2), 'sync', round(d['frame_loop']['synchronous_frame_ms'],2))" || tail -3 $O/${TAG}_ab$v.err
done
for v in "" _lloop _lpf1 _lboth; do
  RTB200_LIB=$PWD/$P/librtb200$v.so timeout 300 python bench.py --config c5 --steps 8 --warmup 3 > $O/${TAG}_ab_c5$v.json 2> $O/${TAG}_ab_c5$v.err
  python -c "
import json; d=json.loads(open('$O/${TAG}_ab_c5$v.json').read().strip().splitlines()[-1]); print('c5 variant [$v]', round(d['value'],1), round(d['ms_per_step'],2))" || tail -3 $O/${TAG}_ab_c5$v.err
done
