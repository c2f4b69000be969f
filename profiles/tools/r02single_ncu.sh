# ncu --set full of the SINGLE instantiation of the path kernel on the C2 frame (DRAM traffic for the new source hash); the same
# command exited 0 without ncu in r02single.sh
O=gpurun_out; mkdir -p $O; TAG=r02s1
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4"
timeout 160 ncu --set full --clock-control none --import-source on -k regex:persist8_kernel -s 2 -c 1 -o $O/${TAG}_persist8_c2 python bench.py --sync --steps 1 --warmup 2 $S > $O/ncu_c2.log 2>&1; echo "ncu rc=$?"; tail -1 $O/ncu_c2.log | cut -c1-160
echo '{}' > $O/${TAG}_traffic_latest.json
python profiles/tools/ncu_summary.py $O/${TAG}_persist8_c2.ncu-rep persist8_kernel $O/${TAG}_persist8_c2_ncu_summary.json $O/${TAG}_traffic_latest.json c2 > /dev/null; echo "sum rc=$?"
cat $O/${TAG}_traffic_latest.json
