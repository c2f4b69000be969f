# A/B of the path-kernel variants (csrc/Makefile VARIANT=...) on the headline workload + the GPU tests that changed
TAG=${1:-r02c}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -k "headline or quoted_size or fourth_type or ploc or frames or ring or window" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 16 --warmup 3"
P=real-time-gpu-ray-tracer_b200
for v in "" _ffma2 _pop1 _pop2 _pop3 ""; do
  RTB200_LIB=$PWD/$P/librtb200$v.so timeout 300 python bench.py $S > $O/${TAG}_ab$v.json 2> $O/${TAG}_ab$v.err
  python -c "
import json; d=json.loads(open('$O/${TAG}_ab$v.json').read().strip().splitlines()[-1]); print('variant [$v]', round(d['value'],1), round(d['ms_per_step'],2), 'sync', round(d['frame_loop']['synchronous_frame_ms'],2))" || tail -3 $O/${TAG}_ab$v.err
done
for v in "" _ffma2 _pop1; do
  RTB200_LIB=$PWD/$P/librtb200$v.so timeout 300 python bench.py --config c5 --steps 8 --warmup 3 > $O/${TAG}_ab_c5$v.json 2> $O/${TAG}_ab_c5$v.err
  python -c "
import json; d=json.loads(open('$O/${TAG}_ab_c5$v.json').read().strip().splitlines()[-1]); print('c5 variant [$v]', round(d['value'],1), round(d['ms_per_step'],2))" || tail -3 $O/${TAG}_ab_c5$v.err
done
