# Measurements left open at the end of round 1, in one call:
#   gpurun --gpus 8 --timeout 900 -- 'bash profiles/tools/next_round.sh r02a'
# Writes gpurun_out/<tag>_*.json; copy what should be judged into profiles/.
TAG=${1:-r02a}
B="python bench.py"
T() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29700+$1)) bench.py --gpus $1 "${@:2}"; }
p() { python -c "import json; d=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1))" || tail -3 $1.err; }
# config 3 (animation) with the faster RNG init and the centre-spanned TLAS grid: 1 and 8 GPUs (r01k: 1159 frames/s device, 1191 e2e at 8)
timeout 200 $B --config c3 --frames 128 --steps 3 --warmup 3 > gpurun_out/${TAG}_bench_c3_n1.json 2> gpurun_out/${TAG}_bench_c3_n1.json.err; p gpurun_out/${TAG}_bench_c3_n1.json
T 8 --config c3 --frames 128 --steps 3 --warmup 3 > gpurun_out/${TAG}_bench_c3_n8.json 2> gpurun_out/${TAG}_bench_c3_n8.json.err; p gpurun_out/${TAG}_bench_c3_n8.json
# config 4 (10 M triangles, 4K, 256 spp) at 1 and 8 GPUs (r01i: 1793 / 12 416 Mrays/s)
timeout 300 $B --config c4 --steps 2 --warmup 2 > gpurun_out/${TAG}_bench_c4_n1.json 2> gpurun_out/${TAG}_bench_c4_n1.json.err; p gpurun_out/${TAG}_bench_c4_n1.json
T 8 --config c4 --steps 3 --warmup 2 > gpurun_out/${TAG}_bench_c4_n8.json 2> gpurun_out/${TAG}_bench_c4_n8.json.err; p gpurun_out/${TAG}_bench_c4_n8.json
# config 2 scaling line by line (the driver's own run does the same)
for n in 2 4 8; do T $n > gpurun_out/${TAG}_bench_n$n.json 2> gpurun_out/${TAG}_bench_n$n.json.err; p gpurun_out/${TAG}_bench_n$n.json; done
