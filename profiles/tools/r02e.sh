# full GPU suite + A/B of the collapse kernel's grid size
TAG=${1:-r02e}; O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
for b in 1 2 3 4 6 8; do RTB200_COLLAPSE_BLOCKS_PER_SM=$b timeout 200 python profiles/tools/build_ab.py; done 2>&1 | tee $O/${TAG}_build_ab.txt
