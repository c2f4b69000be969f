# extra bench lines: config 5 (10k instances) and config 4 (10M triangles, 4K, 256 spp) on one GPU
python bench.py --config c5 --steps 5 --warmup 3 > gpurun_out/bench_c5_n1.json 2> gpurun_out/bench_c5_n1.err; tail -c 1200 gpurun_out/bench_c5_n1.json; tail -2 gpurun_out/bench_c5_n1.err
python bench.py --config c4 --steps 2 --warmup 2 > gpurun_out/bench_c4_n1.json 2> gpurun_out/bench_c4_n1.err; tail -c 1200 gpurun_out/bench_c4_n1.json; tail -2 gpurun_out/bench_c4_n1.err
