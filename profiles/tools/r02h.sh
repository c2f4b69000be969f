TAG=${1:-r02h}; O=gpurun_out; mkdir -p $O; P=$PWD/real-time-gpu-ray-tracer_b200
( for b in 2 3 4 6; do RTB200_LIB=$P/librtb200_nopf.so RTB200_COLLAPSE_BLOCKS_PER_SM=$b timeout 200 python profiles/tools/build_ab.py | sed 's/^/noprefetch /'; done
for b in 2 3 4 6; do RTB200_COLLAPSE_BLOCKS_PER_SM=$b timeout 200 python profiles/tools/build_ab.py | sed 's/^/prefetch /'; done ) 2>&1 | tee $O/${TAG}_build_ab.txt
timeout 600 python -m pytest tests -m gpu -q -k "twin or sah or ploc or frame_matches" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
