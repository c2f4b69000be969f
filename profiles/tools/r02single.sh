# SINGLE instantiation of the path kernel (one-instance scenes skip the top level): GPU suite with it on (default), then A/B on C2
O=gpurun_out; mkdir -p $O; L=$O/r02_single_ab.txt; : > $L
timeout 1400 python -m pytest tests -m gpu -x -q > $O/r02_single_pytest.log 2>&1; echo "pytest rc=$?" >> $L; tail -1 $O/r02_single_pytest.log >> $L
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 16 --warmup 3"
for v in 0 1 0 1; do
  RTB200_SINGLE_INSTANCE=$v timeout 300 python bench.py $S > $O/r02_single_c2_$v.json 2> $O/r02_single_c2_$v.err
  python -c "
import json; d=json.loads(open('$O/r02_single_c2_$v.json').read().strip().splitlines()[-1]); print('c2 single=$v', round(d['value'],1), round(d['ms_per_step'],2), 'sync', round(d['frame_loop']['synchronous_frame_ms'],2), 'nodes/ray', round(d['roofline']['per_ray']['node_fetches'],2), 'e2e', round(d['e2e']['value'],1))" >> $L || tail -3 $O/r02_single_c2_$v.err >> $L
done
cat $L
