# config 3 (animation, frames sharded i mod N) on one box; 128 frames so that every GPU has 16 frames at N = 8
for N in ${C3_NS:-8 4 2}; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --config c3 --frames 128 --steps 3 --warmup 3 > gpurun_out/c3_n$N.json 2> gpurun_out/c3_n$N.err
  python -c "import json; d=json.loads(open('gpurun_out/c3_n$N.json').read().strip().splitlines()[-1]); print('N=$N value', round(d['value'],1), 'fps(dev)', round(d['frames_per_s'],1), 'e2e fps', round(d['e2e']['frames_per_s'],1), 'e2e Mrays/s', round(d['e2e']['value'],1), d['e2e']['renderers_per_gpu'], d['e2e']['host_threads_per_renderer'], d['e2e']['host_cores'], d['e2e']['rank0_host_ms_per_frame'])" || tail -5 gpurun_out/c3_n$N.err
done
