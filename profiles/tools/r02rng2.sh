O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "parity or fullsize or frames" > $O/r02rng2_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02rng2_pytest.log
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 8 --warmup 3"
timeout 300 python bench.py $S > $O/r02rng2_c2.json 2> $O/r02rng2_c2.err
python -c "
import json; d=json.loads(open('$O/r02rng2_c2.json').read().strip().splitlines()[-1]); print('c2', round(d['value'],1), d['stage_ms'])"
timeout 400 python bench.py --config c3 --steps 2 --warmup 1 > $O/r02rng2_c3.json 2> $O/r02rng2_c3.err
python -c "
import json; d=json.loads(open('$O/r02rng2_c3.json').read().strip().splitlines()[-1]); print('c3', round(d['value'],1), round(d['frames_per_s'],1), 'e2e', round(d['e2e']['frames_per_s'],1))"
