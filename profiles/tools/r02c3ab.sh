# config-3 rebuild under the collapse variants (octets, level launches, grid sizes)
O=gpurun_out; mkdir -p $O; L=$O/r02_c3_build_ab.txt; : > $L
run() { env "$@" timeout 200 python profiles/tools/c3_build_ab.py 2>/dev/null | tail -1 >> $L; }
run A=1
for b in 2 4 6 8; do run RTB200_COLLAPSE_BLOCKS_PER_SM=$b; done
for b in 4 8 12; do run RTB200_COLLAPSE=octet RTB200_COLLAPSE_BLOCKS_PER_SM=$b; done
for n in 2 3 4; do run RTB200_COLLAPSE_LEVELS=$n; done
run A=1
cat $L
