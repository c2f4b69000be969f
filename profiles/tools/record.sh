# round record on one box: N=1 bench line, ncu full capture of the path kernel on the bench workload, launch list,
# C5 / C4 lines; with more GPUs visible also the N=2/4/8 lines (usage: record.sh [max_gpus])
MAXN=${1:-1}
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 300 gpurun_out/bench_n1.json; tail -3 gpurun_out/bench_n1.err
ncu --set full --clock-control none --import-source on -k regex:persist8 -s 2 -c 1 -o gpurun_out/r01i_persist8 python bench.py --steps 1 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-c3 > gpurun_out/ncu_full.log 2>&1; tail -1 gpurun_out/ncu_full.log | cut -c1-150
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r01i_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-ref-cuda --no-c3 > gpurun_out/ncu2.log 2>&1; tail -1 gpurun_out/ncu2.log | cut -c1-150
python bench.py --config c5 > gpurun_out/bench_c5_n1.json 2> gpurun_out/bench_c5_n1.err; tail -c 200 gpurun_out/bench_c5_n1.json
python bench.py --config c4 --steps 2 --warmup 2 > gpurun_out/bench_c4_n1.json 2> gpurun_out/bench_c4_n1.err; tail -c 200 gpurun_out/bench_c4_n1.json
for n in 2 4 8; do
  if [ $n -le $MAXN ]; then
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; tail -c 300 gpurun_out/bench_n$n.json
  fi
done
if [ 8 -le $MAXN ]; then
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --config c4 --steps 3 --warmup 2 > gpurun_out/bench_c4_n8.json 2> gpurun_out/bench_c4_n8.err; tail -c 300 gpurun_out/bench_c4_n8.json
fi
