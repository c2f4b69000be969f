O=gpurun_out
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/r02x_launches_c3.csv python bench.py --config c3 --frames 4 --steps 1 --warmup 1 --renderers-per-gpu 1 > $O/ncu_c3.log 2>&1
tail -2 $O/ncu_c3.log | cut -c1-200
