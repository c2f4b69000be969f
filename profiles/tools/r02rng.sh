# RNG init with four pixels per thread: full GPU suite, stage times on C2 (1080p) with it and with the one-pixel kernel, config 3
TAG=${1:-r02rng}; O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 8 --warmup 3"
for m in k4 single; do
  RTB200_RNG_INIT=$m timeout 300 python bench.py $S > $O/${TAG}_c2_$m.json 2> $O/${TAG}_c2_$m.err
  python -c "
import json; d=json.loads(open('$O/${TAG}_c2_$m.json').read().strip().splitlines()[-1]); print('c2 rng=$m', round(d['value'],1), d['stage_ms'])" || tail -3 $O/${TAG}_c2_$m.err
  RTB200_RNG_INIT=$m timeout 400 python bench.py --config c3 --steps 2 --warmup 1 > $O/${TAG}_c3_$m.json 2> $O/${TAG}_c3_$m.err
  python -c "
import json; d=json.loads(open('$O/${TAG}_c3_$m.json').read().strip().splitlines()[-1]); print('c3 rng=$m', round(d['value'],1), round(d['frames_per_s'],1), 'e2e', round(d['e2e']['frames_per_s'],1))" || tail -3 $O/${TAG}_c3_$m.err
done
RTB200_RNG_INIT=k4 timeout 300 python bench.py --emulate-shard 8 $S > $O/${TAG}_shard8.json 2>/dev/null; python -c "
import json; d=json.loads(open('$O/${TAG}_shard8.json').read().strip().splitlines()[-1]); print('shard8', round(d['ms_per_step'],2), d['stage_ms'])"
