# closing record of round 2 on one GPU: smoke, the driver's bench command, the reference arm
TAG=${1:-r02u}; O=gpurun_out; mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${TAG}_pytest.log
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/${TAG}_bench_n1.json 2> $O/${TAG}_bench_n1.err ) 2>&1 | grep real
python -c "
import json; d=json.loads(open('$O/${TAG}_bench_n1.json').read().strip().splitlines()[-1]); print(round(d['value'],1), round(d['e2e']['value'],1), d['roofline']['bound'], round(d['roofline']['frac'],4), d['roofline']['traffic'], d['lbvh_build_ms']['blas_rebuild'], d['c2_dielectric']['value'], d['c4']['value'], d['reference_cuda'].get('median'), d['vs_reference_cuda'])"
