# octet collapse: twin-equality tests, then A/B against the one-thread-per-item collapse
TAG=${1:-r02f}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -k "twin or sah or ploc or pressure or frame_matches or fourth or c4 or c5 or vtk" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
( RTB200_COLLAPSE=thread RTB200_COLLAPSE_BLOCKS_PER_SM=3 timeout 200 python profiles/tools/build_ab.py | sed 's/^/thread /'
for b in 2 4 6 8 12 16; do RTB200_COLLAPSE_BLOCKS_PER_SM=$b timeout 200 python profiles/tools/build_ab.py | sed 's/^/octet /'; done ) 2>&1 | tee $O/${TAG}_build_ab.txt
