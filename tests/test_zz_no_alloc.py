"""rt_build_tlas_no_alloc (include/rtb200.h): the one-block TLAS build with its scratch in a file-scope static device
array, for callers that cannot allocate.  No counterpart in the reference.

The GPU test calls it through the C-ABI (ctypes, plain device pointers) on several scenes IN A ROW — 100 instances, then 5,
then 100 others, so a call finds the previous call's larger tree in the shared scratch — and compares each tree with the one
the existing entry point builds for the same leaf boxes (rt_render_frame's TLAS rebuild, read back with
rt_read_wide_bvh(ctx, 1)): node count, node contents (layout-independent walk, bit for bit: this is integer / byte work) and
leaf order.  This file is named to be collected last; it passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

SEED = 0x5EED
f32 = np.float32


def bind(lib):
    lib.rt_build_tlas_no_alloc.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
    lib.rt_build_tlas_no_alloc.restype = C.c_int
    return lib.rt_build_tlas_no_alloc


def test_no_alloc_rejects_sizes_beyond_the_static_scratch(pkg):
    """Argument checks come before any CUDA call: sizes beyond the stated maximum (and 0) are refused with a clear message."""
    fn = bind(pkg.abi.load_library())
    err = C.create_string_buffer(256)
    dummy = C.c_void_p(256)                           # never dereferenced: the size check comes first
    assert fn(dummy, 1025, dummy, dummy, None, None, err, 256) == pkg.abi.RT_ERR_ARGUMENT
    assert b"1025" in err.value and b"1024" in err.value, err.value
    assert fn(dummy, 0, dummy, dummy, None, None, err, 256) == pkg.abi.RT_ERR_ARGUMENT
    assert b"at least 1" in err.value, err.value
    assert fn(None, 8, dummy, dummy, None, None, err, 256) == pkg.abi.RT_ERR_ARGUMENT
    assert fn(dummy, 8, C.c_void_p(264), dummy, None, None, err, 256) == pkg.abi.RT_ERR_ARGUMENT and b"16-byte" in err.value
    assert fn(dummy, 1025, dummy, dummy, None, None, None, 0) == pkg.abi.RT_ERR_ARGUMENT      # no error buffer is fine too


def leaf_boxes(sc):
    """The world boxes the frame path gives the TLAS builder: Instance::updateTransformArguments' transformedBoundingBox
    (CPU twin, bit-identical to the device's instance update: tests/test_gpu_parity.py) padded by aabb_pad
    (csrc/rt_build_core.h) in float32 — the library is built with -fmad=false, so numpy's float32 ops round the same.
    The scenes used here have no instance whose box the frame path tightens from its vertices (rotated triangle meshes)."""
    import emul_ctypes as ec
    e = ec.Emul(sc)
    try:
        w = e.instances()[:, 24:30].astype(f32)
    finally:
        e.close()
    lo, hi = w[:, :3], w[:, 3:]
    lo = lo - (np.abs(lo) * f32(4e-7) + f32(1e-6))
    hi = hi + (np.abs(hi) * f32(4e-7) + f32(1e-6))
    return np.ascontiguousarray(np.concatenate([lo, hi], axis=1), dtype=f32)


def scenes_in_a_row(S):
    a = S.c1b(width=64, height=36, spp=1, depth=2)                      # 100 instances
    b = S.c1a(frame=7, width=64, height=36, spp=1, depth=2)             # 5 instances: a small tree after a large one
    c = S.c1b(width=64, height=36, spp=1, depth=2, seed=5)              # 100 other instances, moved, turned and scaled
    rs = np.random.RandomState(11)
    c.xforms["shift"] = rs.uniform(-3, 3, (100, 3)).astype(f32)
    c.xforms["rotate_deg"] = rs.uniform(-180, 180, (100, 3)).astype(f32)
    c.xforms["scale"] = rs.uniform(0.5, 2.0, (100, 3)).astype(f32)
    c.xforms[0] = ((0, 0, 0), (0, 0, 0), (1, 1, 1))                     # the ground stays where it is
    c.name = "C1b-moved"
    return [a, b, c, a]


@pytest.mark.gpu
def test_no_alloc_tlas_equals_the_frame_paths(pkg):
    import torch
    assert torch.cuda.is_available()
    dev = torch.device("cuda", 0)
    fn = bind(pkg.abi.load_library())
    canon = pkg.abi.canonical_wide_bvh
    err = C.create_string_buffer(256)
    # the caller's buffers, allocated once, before the calls (the entry point itself allocates nothing)
    cap = 1024
    boxes_d = torch.empty(cap * 6, dtype=torch.float32, device=dev)
    nodes_d = torch.empty((cap + 1) * 20, dtype=torch.int32, device=dev)
    leaf_d = torch.empty(cap, dtype=torch.int32, device=dev)
    count_d = torch.zeros(1, dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    for sc in scenes_in_a_row(pkg.scenes):
        # existing entry point: the TLAS the frame path builds
        r = pkg.Renderer(device=0)
        try:
            r.commitScene(sc)
            r.renderFrame(0, SEED)
            ref_nodes, ref_leaf = r.readWideBVH(1)
        finally:
            r.cleanup()
        n = len(sc.instances)
        boxes = leaf_boxes(sc)
        assert boxes.shape == (n, 6)
        boxes_d[: n * 6].copy_(torch.from_numpy(boxes.ravel()))
        nodes_d.fill_(-1); leaf_d.fill_(-1); count_d.fill_(-1)
        torch.cuda.synchronize(dev)
        with torch.cuda.device(dev):
            rc = fn(boxes_d.data_ptr(), n, nodes_d.data_ptr(), leaf_d.data_ptr(), count_d.data_ptr(), C.c_void_p(stream.cuda_stream), err, 256)
        assert rc == pkg.abi.RT_OK, err.value
        stream.synchronize()
        n_nodes = int(count_d.cpu().numpy().view(np.uint32)[0])
        nodes = nodes_d.cpu().numpy().view(np.uint32).reshape(-1, 20)[:n_nodes]
        leaf = leaf_d.cpu().numpy().view(np.uint32)[:n]
        assert n_nodes == len(ref_nodes), (sc.name, n_nodes, len(ref_nodes))
        assert sorted(leaf.tolist()) == list(range(n)), sc.name
        assert canon(nodes, leaf, [0]) == canon(ref_nodes, ref_leaf, [0]), f"{sc.name}: no-alloc TLAS differs from the frame path's"
    # the maximum size runs too (1024 boxes on a jittered grid): every instance lands in exactly one leaf
    rs = np.random.RandomState(3)
    c = rs.uniform(-50, 50, (1024, 3)).astype(f32)
    h = rs.uniform(0.1, 2.0, (1024, 3)).astype(f32)
    big = np.ascontiguousarray(np.concatenate([c - h, c + h], axis=1), dtype=f32)
    boxes_d.copy_(torch.from_numpy(big.ravel()))
    torch.cuda.synchronize(dev)
    with torch.cuda.device(dev):
        rc = fn(boxes_d.data_ptr(), 1024, nodes_d.data_ptr(), leaf_d.data_ptr(), count_d.data_ptr(), C.c_void_p(stream.cuda_stream), err, 256)
    assert rc == pkg.abi.RT_OK, err.value
    stream.synchronize()
    assert sorted(leaf_d.cpu().numpy().view(np.uint32).tolist()) == list(range(1024))
    assert 1 <= int(count_d.cpu().numpy().view(np.uint32)[0]) <= 1024
