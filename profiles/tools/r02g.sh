TAG=${1:-r02g}; O=gpurun_out; mkdir -p $O
RTB200_COLLAPSE=thread timeout 300 ncu --set full --clock-control none --import-source on -k regex:collapse_kernel -c 1 -o $O/${TAG}_collapse_thread python profiles/tools/build_ab.py > $O/ncu_collapse_t.log 2>&1; tail -2 $O/ncu_collapse_t.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:collapse_octet -c 1 -o $O/${TAG}_collapse_octet python profiles/tools/build_ab.py > $O/ncu_collapse_o.log 2>&1; tail -2 $O/ncu_collapse_o.log
RTB200_COLLAPSE=thread timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/${TAG}_build_launches.csv python profiles/tools/build_ab.py > /dev/null 2>&1
