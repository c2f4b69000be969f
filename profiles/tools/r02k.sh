# Final one-GPU record of round 2 with the final binary: full GPU suite, the default bench line (as the driver runs it), ncu --set full
# captures of the path kernel (C2 frame, window of frames in flight, 1/8 shard, C4 at 16 spp) with the DRAM traffic tied to the kernel
# source hash, launch list.
TAG=${1:-r02k}; O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:persist8 -s 2 -c 1 -o $O/${TAG}_persist8_c2 python bench.py --sync --steps 1 --warmup 2 $S > $O/ncu_c2.log 2>&1; tail -1 $O/ncu_c2.log | cut -c1-120
timeout 600 ncu --set full --clock-control none --import-source on -k regex:persist8 -s 2 -c 1 -o $O/${TAG}_persist8_c2_shard8 python bench.py --sync --emulate-shard 8 --steps 1 --warmup 2 $S > $O/ncu_shard.log 2>&1; tail -1 $O/ncu_shard.log | cut -c1-120
timeout 900 ncu --set full --clock-control none --import-source on -k regex:persist8 -s 1 -c 1 -o $O/${TAG}_persist8_c4_spp16 python bench.py --config c4 --spp 16 --sync --steps 1 --warmup 1 > $O/ncu_c4.log 2>&1; tail -1 $O/ncu_c4.log | cut -c1-120
timeout 600 ncu --set full --clock-control none --import-source on -k regex:persist8 -s 4 -c 1 -o $O/${TAG}_persist8_c2_window4 python bench.py --steps 4 --warmup 2 --frames-per-window 4 $S > $O/ncu_win.log 2>&1; tail -1 $O/ncu_win.log | cut -c1-120
python profiles/tools/ncu_summary.py $O/${TAG}_persist8_c2.ncu-rep persist8 $O/${TAG}_persist8_c2_ncu_summary.json profiles/traffic_latest.json c2 > /dev/null
python profiles/tools/ncu_summary.py $O/${TAG}_persist8_c4_spp16.ncu-rep persist8 $O/${TAG}_persist8_c4_spp16_ncu_summary.json profiles/traffic_latest.json c4 16 > /dev/null
python profiles/tools/ncu_summary.py $O/${TAG}_persist8_c2_shard8.ncu-rep persist8 $O/${TAG}_persist8_c2_shard8_ncu_summary.json > /dev/null
python profiles/tools/ncu_summary.py $O/${TAG}_persist8_c2_window4.ncu-rep persist8 $O/${TAG}_persist8_c2_window4_ncu_summary.json > /dev/null
cp profiles/traffic_latest.json $O/${TAG}_traffic_latest.json; cat profiles/traffic_latest.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv python bench.py --steps 8 --warmup 2 $S > $O/ncu_l.log 2>&1
( time timeout 900 python bench.py --steps 20 --warmup 5 > $O/${TAG}_bench_n1.json 2> $O/${TAG}_bench_n1.err ) 2>&1 | grep real; tail -c 300 $O/${TAG}_bench_n1.json
( time timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err ) 2>&1 | grep real; tail -c 300 $O/${TAG}_bench_reference.json
