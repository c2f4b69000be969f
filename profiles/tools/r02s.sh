O=gpurun_out; S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 16 --warmup 3"
run() { env "$@" timeout 300 python bench.py $S > $O/r02s_tmp.json 2>/dev/null; python -c "
import json; d=json.loads(open('$O/r02s_tmp.json').read().strip().splitlines()[-1]); print('$*', round(d['value'],1))"; }
run RTB200_NODE_MIN=20
for nm in 16 18 22 24; do run RTB200_NODE_MIN=$nm; done
for sm in 4 6 10 12 16; do run RTB200_SHADE_MIN=$sm; done
run RTB200_NODE_MIN=18 RTB200_SHADE_MIN=10
run RTB200_NODE_MIN=22 RTB200_SHADE_MIN=6
