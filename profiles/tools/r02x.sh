# fused small-TLAS kernel: full GPU suite, config 3 with the fused kernel and with the general kernels (RTB200_TLAS_FUSED=0),
# C2 short line (window prep), build A/B
TAG=${1:-r02x}; O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
timeout 400 python bench.py --config c3 --steps 2 --warmup 1 > $O/${TAG}_bench_c3_fused.json 2> $O/${TAG}_c3_fused.err; tail -c 1500 $O/${TAG}_bench_c3_fused.json
RTB200_TLAS_FUSED=0 timeout 400 python bench.py --config c3 --steps 2 --warmup 1 > $O/${TAG}_bench_c3_general.json 2> $O/${TAG}_c3_general.err; tail -c 1500 $O/${TAG}_bench_c3_general.json
S="--no-cpu-baseline --no-ref-cuda --no-c4"
timeout 300 python bench.py --steps 16 --warmup 3 $S > $O/${TAG}_bench_n1_short.json 2> $O/${TAG}_bench_n1_short.err; tail -c 1200 $O/${TAG}_bench_n1_short.json
