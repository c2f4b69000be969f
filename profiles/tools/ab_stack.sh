# does the shared-memory stack size (L1 carve-out) matter?  variants with 16 / 12 entries instead of 24
B="timeout 200 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-ref-cuda --no-c3"
p() { python -c "import json,sys; d=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],2))" || tail -3 $1.err; }
$B > gpurun_out/sk24.json 2> gpurun_out/sk24.json.err; p gpurun_out/sk24.json
for v in st16 st12; do
RTB200_LIB=$PWD/real-time-gpu-ray-tracer_b200/librtb200_$v.so $B > gpurun_out/$v.json 2> gpurun_out/$v.json.err; p gpurun_out/$v.json
RTB200_LIB=$PWD/real-time-gpu-ray-tracer_b200/librtb200_$v.so python - <<'PY'
import importlib, sys
sys.path.insert(0, '.')
pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
for name, sc in (("c2", pkg.scenes.c2(spp=4)), ("c5", pkg.scenes.c5(spp=1))):
    r = pkg.Renderer(device=0); r.commitScene(sc); out = r.renderFrame(0, 0x5EED); print(name, "stack_overflow", out.stack_overflow); r.cleanup()
PY
done
