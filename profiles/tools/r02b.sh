# Round-2 second record on one GPU: full GPU test suite, bench line with both windows warmed, 1/8 and 1/4 shard lines.
TAG=${1:-r02b}
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/${TAG}_pytest.log
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4"
timeout 300 python bench.py --steps 16 --warmup 3 $S > $O/${TAG}_bench_n1_short.json 2> $O/${TAG}_bench_n1_short.err; tail -c 600 $O/${TAG}_bench_n1_short.json; tail -3 $O/${TAG}_bench_n1_short.err
for n in 8 4 2; do
timeout 300 python bench.py --emulate-shard $n --steps 16 --warmup 3 $S > $O/${TAG}_bench_shard$n.json 2> $O/${TAG}_bench_shard$n.err; python -c "
import json; d=json.loads(open('$O/${TAG}_bench_shard$n.json').read().strip().splitlines()[-1]); print('shard $n', d['value'], d['ms_per_step'], d['e2e']['value'], d['frame_loop'])" || tail -3 $O/${TAG}_bench_shard$n.err
done
RTB200_PLOC=1 timeout 300 python bench.py --steps 8 --warmup 3 $S > $O/${TAG}_bench_n1_ploc.json 2> $O/${TAG}_bench_n1_ploc.err; tail -c 300 $O/${TAG}_bench_n1_ploc.json
timeout 300 python bench.py --config c5 --steps 8 --warmup 3 > $O/${TAG}_bench_c5_n1.json 2> $O/${TAG}_bench_c5_n1.err; tail -c 300 $O/${TAG}_bench_c5_n1.json
timeout 300 python bench.py --config c3 --frames 64 --steps 3 --warmup 3 > $O/${TAG}_bench_c3_n1.json 2> $O/${TAG}_bench_c3_n1.err; tail -c 400 $O/${TAG}_bench_c3_n1.json
