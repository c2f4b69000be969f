mkdir -p gpurun_out
for g in 8; do
RTB200_GROUPS=$g python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/b2_g$g.json 2> gpurun_out/b2_g$g.err
python -c "import json; d=json.loads(open('gpurun_out/b2_g$g.json').read().strip().splitlines()[-1]); print('g$g', round(d['value'],1), round(d['ms_per_step'],1), d['stage_ms'], round(d['e2e']['value'],1))" || tail -3 gpurun_out/b2_g$g.err
done
