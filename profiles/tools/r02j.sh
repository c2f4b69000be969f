TAG=${1:-r02j}; O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q -k "twin or sah or ploc or pressure or frame_matches" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
( for b in 2 3 4; do RTB200_COLLAPSE_BLOCKS_PER_SM=$b timeout 200 python profiles/tools/build_ab.py | sed 's/^/thread-nofence /'; done
for b in 4 8 12; do RTB200_COLLAPSE=octet RTB200_COLLAPSE_BLOCKS_PER_SM=$b timeout 200 python profiles/tools/build_ab.py | sed 's/^/octet-nofence /'; done ) 2>&1 | tee $O/${TAG}_build_ab.txt
