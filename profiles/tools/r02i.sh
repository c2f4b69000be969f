TAG=${1:-r02i}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -k "twin or sah or ploc or pressure or frame_matches or fourth or c4 or c5 or vtk or full_size" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
( RTB200_COLLAPSE_LEVELS=0 timeout 200 python profiles/tools/build_ab.py | sed 's/^/ticket-only /'
for l in 6 8 12; do for b in 2 4; do RTB200_COLLAPSE_LEVELS=$l RTB200_COLLAPSE_BLOCKS_PER_SM=$b timeout 200 python profiles/tools/build_ab.py | sed "s/^/levels=$l /"; done; done ) 2>&1 | tee $O/${TAG}_build_ab.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $O/${TAG}_build_launches.csv python profiles/tools/build_ab.py > /dev/null 2>&1
timeout 300 python bench.py --config c3 --frames 64 --steps 3 --warmup 3 > $O/${TAG}_bench_c3_n1.json 2> $O/${TAG}_bench_c3_n1.err; python -c "
import json; d=json.loads(open('$O/${TAG}_bench_c3_n1.json').read().strip().splitlines()[-1]); print('c3', d['value'], d['frames_per_s'], d['e2e']['frames_per_s'], d['slowest_rebuild_ms'])"
