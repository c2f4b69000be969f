# the split kernel (dedicated shade warp): parity under RTB200_SPLIT=1, then A/B on C2 / shard / C5
TAG=${1:-r02l}; O=gpurun_out; mkdir -p $O
RTB200_SPLIT=1 timeout 900 python -m pytest tests -m gpu -q -x -k "frame_matches or bitwise or window or two_windows or ring or headline or dielectric or quoted or c5_inst or beyond" > $O/${TAG}_pytest_split.log 2>&1; echo "pytest rc=$?"; tail -5 $O/${TAG}_pytest_split.log
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 16 --warmup 3"
for v in 0 1; do
  RTB200_SPLIT=$v timeout 300 python bench.py $S > $O/${TAG}_c2_split$v.json 2> $O/${TAG}_c2_split$v.err
  python -c "
import json; d=json.loads(open('$O/${TAG}_c2_split$v.json').read().strip().splitlines()[-1]); print('c2 split=$v', round(d['value'],1), round(d['ms_per_step'],2), 'sync', round(d['frame_loop']['synchronous_frame_ms'],2))" || tail -3 $O/${TAG}_c2_split$v.err
done
for nm in 12 16 24 28; do RTB200_NODE_MIN=$nm RTB200_SPLIT=1 timeout 300 python bench.py $S > $O/${TAG}_c2_split1_nm$nm.json 2>/dev/null; python -c "
import json; d=json.loads(open('$O/${TAG}_c2_split1_nm$nm.json').read().strip().splitlines()[-1]); print('c2 split=1 node_min=$nm', round(d['value'],1))"; done
for v in 0 1; do
  RTB200_SPLIT=$v timeout 300 python bench.py --config c5 --steps 8 --warmup 3 > $O/${TAG}_c5_split$v.json 2> $O/${TAG}_c5_split$v.err
  python -c "
import json; d=json.loads(open('$O/${TAG}_c5_split$v.json').read().strip().splitlines()[-1]); print('c5 split=$v', round(d['value'],1), round(d['ms_per_step'],2))" || tail -3 $O/${TAG}_c5_split$v.err
  RTB200_SPLIT=$v timeout 300 python bench.py --emulate-shard 8 $S > $O/${TAG}_shard8_split$v.json 2> $O/${TAG}_shard8_split$v.err
  python -c "
import json; d=json.loads(open('$O/${TAG}_shard8_split$v.json').read().strip().splitlines()[-1]); print('shard8 split=$v', round(d['ms_per_step'],2))" || tail -3 $O/${TAG}_shard8_split$v.err
done
