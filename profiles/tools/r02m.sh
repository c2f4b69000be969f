TAG=${1:-r02m}; O=gpurun_out; mkdir -p $O
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4"
RTB200_SPLIT=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:persist8_split -s 2 -c 1 -o $O/${TAG}_split_c2 python bench.py --sync --steps 1 --warmup 2 $S > $O/ncu_split.log 2>&1; tail -1 $O/ncu_split.log | cut -c1-120
