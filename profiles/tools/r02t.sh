O=gpurun_out; S="--no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 --steps 16 --warmup 3"
run() { cfg=$1; shift; env "$@" timeout 300 python bench.py $cfg $S > $O/r02t_tmp.json 2>/dev/null; python -c "
import json; d=json.loads(open('$O/r02t_tmp.json').read().strip().splitlines()[-1]); print('[$cfg] $*', round(d['value'],1), round(d['ms_per_step'],3))"; }
for c in "18 9" "18 10" "18 11" "19 10" "19 11" "20 10" "20 11" "17 10"; do set -- $c; run "" RTB200_NODE_MIN=$1 RTB200_SHADE_MIN=$2; done
for c in "20 8" "18 10" "19 10" "20 10"; do set -- $c; run "--config c5" RTB200_NODE_MIN=$1 RTB200_SHADE_MIN=$2; run "--emulate-shard 8" RTB200_NODE_MIN=$1 RTB200_SHADE_MIN=$2; done
