# config 3 (animation) on one GPU: the library's frame loop with R renderer contexts per GPU; strips expanded on the device (default) or on the host
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "vtk or animation" 2>&1 | tail -12
RTB200_VTK_EXPAND=host timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "vtk or animation" 2>&1 | tail -3
for R in 1 3; do
  for M in device host; do
  RTB200_VTK_EXPAND=$M timeout 200 python bench.py --config c3 --steps 3 --warmup 3 --renderers-per-gpu $R > gpurun_out/c3_n1_r${R}_$M.json 2> gpurun_out/c3_n1_r${R}_$M.err
  python -c "import json; d=json.loads(open('gpurun_out/c3_n1_r${R}_$M.json').read().strip().splitlines()[-1]); print('R=$R $M value', round(d['value'],1), 'fps(dev)', round(d['frames_per_s'],1), 'fps(wall,1 renderer)', round(d['wall_frames_per_s_one_renderer'],1), 'e2e fps', round(d['e2e']['frames_per_s'],1), 'slowest rebuild', round(d['slowest_rebuild_ms'],3), d['e2e']['rank0_host_ms_per_frame'])" || tail -5 gpurun_out/c3_n1_r${R}_$M.err
  done
done
