"""Multi-process check of the frame ring (run under torchrun with N >= 2 ranks on N GPUs):
every rank renders its tile shard of a few frames in flight and stores its texels into rank 0's ring through the CUDA IPC
mapping; rank 0 compares each assembled frame, bit for bit, with the frame of an unsharded context of its own.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 profiles/tools/ring_check.py"""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
SEED = 0x5EED
S = pkg.scenes
sc = S.c2(width=640, height=360, spp=8, depth=10, nx=300, nz=150)
frames = 6
r = pkg.Renderer(device=local)
r.commitScene(sc)
r.setTileShard(rank, world, 8)
h = torch.zeros(64, dtype=torch.uint8, device=f"cuda:{local}")
if rank == 0:
    h = torch.frombuffer(bytearray(r.frameRingCreate()), dtype=torch.uint8).to(f"cuda:{local}")
dist.broadcast(h, src=0)
if rank != 0:
    r.frameRingOpen(bytes(h.cpu().numpy().tobytes()))
dist.barrier()
ok = True
for rounds in range(3):
    w = r.submitFrames(0, frames, SEED)
    got = []
    for k in range(frames):
        r.waitFrame(w, k)
        if rank == 0:
            got.append(r.readFrame(w, k, pkg.abi.BUF_RGBA8).copy())
    r.releaseWindow(w)
    dist.barrier()
    if rank == 0:
        ref = pkg.Renderer(device=local)
        ref.commitScene(sc)
        for k in range(frames):
            ref.renderFrame(k, SEED)
            want = ref.readFramebuffer(pkg.abi.BUF_RGBA8)
            same = np.array_equal(got[k], want)
            ok = ok and same
            if not same:
                print(f"round {rounds} frame {k}: {int((got[k] != want).sum())} bytes differ")
        ref.cleanup()
# the library's loop with a SLOW consumer on rank 0 (its sink sleeps): the other ranks run ahead and must wait for rank 0 to release
# a window slot before they store into it again (the ring's back-pressure), or rank 0 would read overwritten frames
import time
n_loop = 24
delivered = {}


def sink(frame, img, out):
    if rank == 0:
        time.sleep(0.03)
        delivered[frame] = img.copy()
    return 0


r.runFrames(0, n_loop, SEED, 4, sink)
dist.barrier()
if rank == 0:
    ref = pkg.Renderer(device=local)
    ref.commitScene(sc)
    for k in range(n_loop):
        ref.renderFrame(k, SEED)
        same = k in delivered and np.array_equal(delivered[k].reshape(-1), ref.readFramebuffer(pkg.abi.BUF_RGBA8))
        ok = ok and same
        if not same:
            print(f"loop frame {k}: differs or missing")
    ref.cleanup()
# the synchronous frame through the ring as well
out = r.renderFrame(0, SEED)
torch.cuda.synchronize()
dist.barrier()
if rank == 0:
    ref = pkg.Renderer(device=local)
    ref.commitScene(sc)
    ref.renderFrame(0, SEED)
    ok = ok and np.array_equal(r.readFramebuffer(pkg.abi.BUF_RGBA8), ref.readFramebuffer(pkg.abi.BUF_RGBA8))
    ref.cleanup()
    print("RING_CHECK", "OK" if ok else "FAILED", f"world={world}")
dist.barrier()
r.frameRingClose()
r.cleanup()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
