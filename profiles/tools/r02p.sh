# final-binary refresh: full GPU suite, ncu --set full of the path kernel on C2 and C4 (DRAM traffic -> traffic_latest.json with the
# kernel source hash), launch list, short bench line
TAG=${1:-r02p}; O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:persist8_kernel -s 2 -c 1 -o $O/${TAG}_persist8_c2 python bench.py --sync --steps 1 --warmup 2 $S > $O/ncu_c2.log 2>&1; tail -1 $O/ncu_c2.log | cut -c1-120
timeout 900 ncu --set full --clock-control none --import-source on -k regex:persist8_kernel -s 1 -c 1 -o $O/${TAG}_persist8_c4_spp16 python bench.py --config c4 --spp 16 --sync --steps 1 --warmup 1 > $O/ncu_c4.log 2>&1; tail -1 $O/ncu_c4.log | cut -c1-120
echo '{}' > $O/${TAG}_traffic_latest.json
python profiles/tools/ncu_summary.py $O/${TAG}_persist8_c2.ncu-rep persist8_kernel $O/${TAG}_persist8_c2_ncu_summary.json $O/${TAG}_traffic_latest.json c2 > /dev/null
python profiles/tools/ncu_summary.py $O/${TAG}_persist8_c4_spp16.ncu-rep persist8_kernel $O/${TAG}_persist8_c4_spp16_ncu_summary.json $O/${TAG}_traffic_latest.json c4 16 > /dev/null
cat $O/${TAG}_traffic_latest.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv python bench.py --steps 8 --warmup 2 $S > $O/ncu_l.log 2>&1
timeout 300 python bench.py --steps 16 --warmup 3 $S > $O/${TAG}_bench_n1_short.json 2> $O/${TAG}_bench_n1_short.err; tail -c 200 $O/${TAG}_bench_n1_short.json
