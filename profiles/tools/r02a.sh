# Round-2 first record on one GPU: the GPU test suite, the bench line, the 1/8-shard line, launch list, and ncu --set full
# captures of the path kernel on (a) the C2 bench workload, (b) rank 0's 1/8 shard of it, (c) C4 at 16 spp.
TAG=${1:-r02a}
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/${TAG}_pytest.log
timeout 600 python bench.py > $O/${TAG}_bench_n1.json 2> $O/${TAG}_bench_n1.err; tail -c 400 $O/${TAG}_bench_n1.json; tail -3 $O/${TAG}_bench_n1.err
S="--no-cpu-baseline --no-ref-cuda --no-c3"
timeout 300 python bench.py --emulate-shard 8 $S > $O/${TAG}_bench_shard8.json 2> $O/${TAG}_bench_shard8.err; tail -c 300 $O/${TAG}_bench_shard8.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:persist8 -s 2 -c 1 -o $O/${TAG}_persist8_c2 python bench.py --steps 1 --warmup 2 $S > $O/ncu_c2.log 2>&1; tail -1 $O/ncu_c2.log | cut -c1-150
timeout 600 ncu --set full --clock-control none --import-source on -k regex:persist8 -s 2 -c 1 -o $O/${TAG}_persist8_c2_shard8 python bench.py --emulate-shard 8 --steps 1 --warmup 2 $S > $O/ncu_shard.log 2>&1; tail -1 $O/ncu_shard.log | cut -c1-150
timeout 900 ncu --set full --clock-control none --import-source on -k regex:persist8 -s 1 -c 1 -o $O/${TAG}_persist8_c4_spp16 python bench.py --config c4 --spp 16 --steps 1 --warmup 1 > $O/ncu_c4.log 2>&1; tail -1 $O/ncu_c4.log | cut -c1-150
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv python bench.py --steps 2 --warmup 1 $S > $O/ncu_l.log 2>&1; tail -1 $O/ncu_l.log | cut -c1-150
timeout 300 python bench.py --config c4 --steps 2 --warmup 1 > $O/${TAG}_bench_c4_n1.json 2> $O/${TAG}_bench_c4_n1.err; tail -c 300 $O/${TAG}_bench_c4_n1.json
ls -la $O | head -40
