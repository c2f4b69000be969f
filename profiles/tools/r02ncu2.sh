# ncu --set full of the two build-side kernels added late in round 2: the four-pixel RNG init (C2 frame, 1080p) and the one-block TLAS
# kernel (a config-3 frame of 505 instances); each only after the same command exited 0 without ncu (r02final2, r02rng2)
O=gpurun_out; mkdir -p $O
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:rng_init_tables4_kernel -s 2 -c 1 -o $O/r02_rng4 python bench.py --sync --steps 1 --warmup 2 $S > $O/ncu_rng4.log 2>&1; echo "rng4 rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:tlas_small_kernel -s 6 -c 1 -o $O/r02_tlas_small python bench.py --config c3 --frames 4 --steps 1 --warmup 1 --renderers-per-gpu 1 > $O/ncu_tlas.log 2>&1; echo "tlas rc=$?"
python profiles/tools/ncu_summary.py $O/r02_rng4.ncu-rep rng_init_tables4_kernel $O/r02_rng4_ncu_summary.json > /dev/null; echo "sum rc=$?"
python profiles/tools/ncu_summary.py $O/r02_tlas_small.ncu-rep tlas_small_kernel $O/r02_tlas_small_ncu_summary.json > /dev/null; echo "sum rc=$?"
cat $O/r02_rng4_ncu_summary.json $O/r02_tlas_small_ncu_summary.json | head -60
