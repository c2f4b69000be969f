# Multi-GPU record (usage: gpurun --gpus N -- 'bash profiles/tools/r02d.sh r02d N'): IPC frame-ring check, then the bench at N ranks
TAG=${1:-r02d}; N=${2:-2}
O=gpurun_out; mkdir -p $O
T() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+RANDOM%200)) "$@"; }
T profiles/tools/ring_check.py > $O/${TAG}_ring_check_n$N.log 2>&1; echo "ring_check rc=$?"; grep -a "RING_CHECK\|Error\|error" $O/${TAG}_ring_check_n$N.log | head -5
timeout 300 python bench.py --steps 16 --warmup 3 --no-cpu-baseline --no-ref-cuda --no-c3 --no-c4 > $O/${TAG}_bench_n1.json 2> $O/${TAG}_bench_n1.err
T bench.py --gpus $N --steps 16 --warmup 3 > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err
python - <<PY
import json
a=json.loads(open('$O/${TAG}_bench_n1.json').read().strip().splitlines()[-1])
try:
    b=json.loads(open('$O/${TAG}_bench_n$N.json').read().strip().splitlines()[-1])
    print('n1',round(a['value'],1),'n$N',round(b['value'],1),'eff',round(b['value']/a['value']/$N,3),'e2e',round(b['e2e']['value'],1), b['parallelism'][:120])
    print('c4', {k:b.get('c4',{}).get(k) for k in ('value','ms_per_step','unavailable')})
except Exception as e:
    print('N>1 line missing', e); print(open('$O/${TAG}_bench_n$N.err').read()[-1500:])
PY
