"""`not gpu` test of the N > 1 path: world_size-2 gloo ranks each render their image-tile shard with the CPU twin of
the device cores (tests/emul), pack it like librtb200 does, gather to rank 0 and de-interleave with the same helper
bench.py uses; the result must equal the single-rank frame bit for bit (pixel RNG is keyed by the global pixel)."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SEED = 0x5EED
TILE_ROWS = 8


def _rank_main(rank, world, port, height, out_path):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    import emul_ctypes as ec
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    sh = pkg.sharding
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    sc = pkg.scenes.c1a(frame=3, width=64, height=height, spp=1, depth=4)
    e = ec.Emul(sc)
    W = sc.width
    plane = np.zeros((sh.local_rows(height, 0, world, TILE_ROWS), W * 4), np.uint8)     # padded to the largest shard
    full = e.render(SEED)["rgba8"].reshape(height, W * 4)           # the twin renders whole frames; a rank keeps its rows
    for ly in range(sh.local_rows(height, rank, world, TILE_ROWS)):
        gy = sh.global_row(ly, rank, world, TILE_ROWS)
        if gy < height:
            plane[ly] = full[gy]
    assert sorted(sum((sh.owned_rows(height, r, world, TILE_ROWS) for r in range(world)), [])) == list(range(height))
    t = torch.from_numpy(plane.reshape(-1))
    gathered = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
    dist.gather(t, gathered, dst=0)
    if rank == 0:
        frame = sh.assemble(gathered, height, W * 4, world, TILE_ROWS).numpy()
        np.save(out_path, np.stack([frame, full]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("height", [36, 44])          # 44 rows = 5.5 tiles: ragged last tile, unequal shards
def test_tile_shards_gather_to_the_single_rank_frame(tmp_path, height):
    import torch.multiprocessing as mp
    out = str(tmp_path / "frame.npy")
    port = 29500 + (os.getpid() + height) % 2000
    mp.spawn(_rank_main, args=(2, port, height, out), nprocs=2, join=True)
    frame, full = np.load(out)
    assert np.array_equal(frame, full)


def test_shard_arithmetic_matches_the_library_rule():
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    sh = pkg.sharding
    for H, world, tr in [(1080, 1, 8), (1080, 2, 8), (1080, 8, 8), (2160, 4, 16), (37, 8, 8)]:
        rows = sum((sh.owned_rows(H, r, world, tr) for r in range(world)), [])
        assert sorted(rows) == list(range(H))
        assert sh.padded_plane_bytes(H, 4, world, tr) >= max(sh.local_rows(H, r, world, tr) for r in range(world)) * 4


def _frame_rank_main(rank, world, port, n_frames, out_path):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    import emul_ctypes as ec
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    sh = pkg.sharding
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    W, H = 48, 27

    def render(f):          # frame f of the animation: the demo instances move with the frame number (Main.cu:6-42)
        return ec.Emul(pkg.scenes.c1a(frame=f, width=W, height=H, spp=1, depth=4)).render(SEED)["rgba8"].reshape(-1)

    mine = sh.frames_of_rank(n_frames, rank, world)
    slots = len(sh.frames_of_rank(n_frames, 0, world))                 # rank 0 owns the most frames: pad to that
    buf = np.zeros((slots, W * H * 4), np.uint8)
    for j, f in enumerate(mine):
        buf[j] = render(f)
    t = torch.from_numpy(buf.reshape(-1))
    gathered = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
    dist.gather(t, gathered, dst=0)                                     # only the finished images travel, once
    if rank == 0:
        per_rank = [g.numpy().reshape(slots, -1) for g in gathered]
        got = np.stack(sh.interleave_frames(per_rank, n_frames))
        want = np.stack([render(f) for f in range(n_frames)])
        np.save(out_path, np.stack([got, want]))
    dist.barrier()
    dist.destroy_process_group()


def test_frame_shards_cover_the_animation(tmp_path):
    """Config 3's partitioning: frame i -> rank i % world, no per-frame exchange; the frames the ranks rendered,
    put back in order, are the frames of the single-rank loop."""
    import torch.multiprocessing as mp
    out = str(tmp_path / "frames.npy")
    port = 31500 + os.getpid() % 2000
    mp.spawn(_frame_rank_main, args=(2, port, 5, out), nprocs=2, join=True)
    got, want = np.load(out)
    assert np.array_equal(got, want)
    assert not np.array_equal(want[0], want[1])        # the frames do differ


def test_frame_shard_arithmetic_and_rank_local_series(tmp_path):
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    sh, S = pkg.sharding, pkg.scenes
    for n, world in [(401, 8), (5, 2), (3, 8), (64, 4)]:
        owned = [sh.frames_of_rank(n, q, world) for q in range(world)]
        assert sorted(sum(owned, [])) == list(range(n))
        assert sh.interleave_frames(owned, n) == list(range(n))
    counts = S.c3_particle_counts(401)
    assert counts[0] == 25 and counts[-1] == 500 and all(b >= a for a, b in zip(counts, counts[1:]))
    # a rank's own series holds byte-identical copies of the frames it owns
    fixture = os.path.join(ROOT, "tests", "golden", "particle_frame0.vtk")
    cnt = [26, 27, 30, 28]
    S.synth_vtk_series(fixture, str(tmp_path / "all"), cnt, frame_ids=list(range(4)))
    mine = sh.frames_of_rank(4, 1, 2)
    series = S.synth_vtk_series(fixture, str(tmp_path / "r1"), [cnt[f] for f in mine], frame_ids=mine)
    assert open(series).read().count('"name"') == 2
    for f in mine:
        name = f"particle_mesh/particle_{f * 100:015d}.vtk"
        assert (tmp_path / "all" / name).read_bytes() == (tmp_path / "r1" / name).read_bytes()


def test_sharding_properties_hypothesis():
    """Property tests of the two partitions (image tiles by row, animation frames): every unit is owned exactly once,
    the packed planes de-interleave back to the frame for ragged heights / more ranks than tiles, and the frame
    interleave inverts frames_of_rank."""
    hyp = pytest.importorskip("hypothesis")
    st = pytest.importorskip("hypothesis.strategies")
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    sh = pkg.sharding

    @hyp.settings(max_examples=150, deadline=None, derandomize=True)
    @hyp.given(st.integers(1, 300), st.integers(1, 9), st.sampled_from([1, 2, 4, 8, 16, 32]), st.integers(1, 5))
    def tiles(height, world, tile_rows, row_bytes):
        owned = [sh.owned_rows(height, r, world, tile_rows) for r in range(world)]
        assert sorted(sum(owned, [])) == list(range(height))
        frame = np.arange(height * row_bytes, dtype=np.int64).reshape(height, row_bytes) % 251
        pad = sh.padded_plane_bytes(height, row_bytes, world, tile_rows)
        planes = []
        for r in range(world):
            plane = np.zeros(pad, np.uint8)
            rows = owned[r]
            plane[:len(rows) * row_bytes] = frame[rows].astype(np.uint8).reshape(-1)
            assert sh.local_rows(height, r, world, tile_rows) * row_bytes <= pad
            planes.append(plane)
        assert np.array_equal(sh.assemble(planes, height, row_bytes, world, tile_rows), frame.astype(np.uint8))

    @hyp.settings(max_examples=150, deadline=None, derandomize=True)
    @hyp.given(st.integers(0, 500), st.integers(1, 16))
    def frames(n, world):
        per_rank = [sh.frames_of_rank(n, r, world) for r in range(world)]
        assert sorted(sum(per_rank, [])) == list(range(n))
        assert max(len(p) for p in per_rank) - min(len(p) for p in per_rank) <= 1
        assert sh.interleave_frames(per_rank, n) == list(range(n))

    tiles()
    frames()
