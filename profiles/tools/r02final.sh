# what the driver runs at round end, on the final HEAD: smoke, the default bench line, the reference arm
TAG=${1:-r02final}; O=gpurun_out; mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/${TAG}_smoke.log
SECONDS=0; python bench.py --steps 20 --warmup 5 > $O/${TAG}_bench_n1.json 2> $O/${TAG}_bench_n1.err; echo "bench rc=$?"; echo "elapsed ${SECONDS}s"
SECONDS=0; python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "ref rc=$?"; echo "elapsed ${SECONDS}s"; cut -c1-400 $O/${TAG}_bench_reference.json
