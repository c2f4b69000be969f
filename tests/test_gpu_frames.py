"""GPU tests of the frame loop behind the ABI: frames in flight (rt_submit_frames .. rt_run_frames:
a window of frames served by ONE launch of the path kernel, reference src/Global/Renderer.cu:189-207,278-317 keeps two
frames in flight with double-buffered instances / TLAS) and the multi-GPU frame ring (ranks store finished texels into one
full frame on rank 0).  Every pipelined frame must be, bit for bit, the frame rt_render_frame produces, which the parity
tests pin to the oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SEED = 0x5EED


def moving_scene(pkg, **kw):
    """the reference's demo scene with its per-frame instance update (src/Global/Main.cu:6-42)"""
    S = pkg.scenes
    sc = S.c1a(frame=0, **kw)

    def update(xf, frame):
        xf[:] = S.main_cu_xforms(frame)
    return sc, update


def one_by_one(pkg, sc, update, frames, seeds, flags=0, shard=None):
    r = pkg.Renderer(device=0, flags=flags | pkg.abi.FLAG_RECORD_PRIMARY_AOV)
    res = []
    try:
        r.commitScene(sc, update)
        if shard:
            r.setTileShard(*shard)
        for f, s in zip(frames, seeds):
            out = r.renderFrame(f, s)
            assert out.stack_overflow == 0
            res.append(dict(rays=out.rays, rgba8=r.readFramebuffer(pkg.abi.BUF_RGBA8).copy(), radiance=r.readFramebuffer(pkg.abi.BUF_RADIANCE).copy(),
                            t=r.readFramebuffer(pkg.abi.BUF_PRIMARY_T).copy(), inst=r.readFramebuffer(pkg.abi.BUF_PRIMARY_INSTANCE).copy()))
    finally:
        r.cleanup()
    return res


def test_window_frames_equal_single_frames(pkg, port):
    """Frames 3..9 of the moving demo scene (new instance transforms, new TLAS, new seed every frame) as one window of 7
    frames in flight: each equals rt_render_frame's frame bit for bit; frame 5 is also checked against the oracle."""
    sc, update = moving_scene(pkg, width=160, height=90, spp=4, depth=6)
    frames = list(range(3, 10))
    seeds = [SEED + 17 * f for f in frames]
    want = one_by_one(pkg, sc, update, frames, seeds)
    assert not np.array_equal(want[0]["radiance"], want[1]["radiance"])
    r = pkg.Renderer(device=0, flags=pkg.abi.FLAG_RECORD_PRIMARY_AOV)
    try:
        r.commitScene(sc, update)
        w = r.submitFrames(frames[0], len(frames), seeds=seeds)
        for k in reversed(range(len(frames))):          # any order: a finished frame stays readable until the window is released
            out = r.waitFrame(w, k)
            assert out.rays == want[k]["rays"] and out.stack_overflow == 0 and out.render_ms > 0
            assert np.array_equal(r.readFrame(w, k, pkg.abi.BUF_RGBA8), want[k]["rgba8"]), f"frame {frames[k]}"
            assert np.array_equal(r.readFrame(w, k, pkg.abi.BUF_RADIANCE), want[k]["radiance"])
            assert np.array_equal(r.readFrame(w, k, pkg.abi.BUF_PRIMARY_T), want[k]["t"])
            assert np.array_equal(r.readFrame(w, k, pkg.abi.BUF_PRIMARY_INSTANCE), want[k]["inst"])
        st = r.releaseWindow(w)
        assert st.frames == len(frames) and st.rays == sum(x["rays"] for x in want) and st.device_ms > 0 and st.stack_overflow == 0
        # the context is still good for a synchronous frame afterwards
        out = r.renderFrame(frames[2], seeds[2])
        assert out.rays == want[2]["rays"] and np.array_equal(r.readFramebuffer(pkg.abi.BUF_RGBA8), want[2]["rgba8"])
    finally:
        r.cleanup()
    # ... and the chain to the reference: frame 5 of the window against the oracle
    S = pkg.scenes
    sc5 = S.c1a(frame=5, width=160, height=90, spp=4, depth=6)
    img, rad, rays = port.scene(sc5, 3).render(seeds[2])
    assert rays == want[2]["rays"] and np.array_equal(img.reshape(-1), want[2]["rgba8"])


def test_two_windows_in_flight_and_the_loop(pkg):
    """Two windows in flight on their own streams, then rt_run_frames: 19 frames in windows of 4, delivered in order."""
    sc, update = moving_scene(pkg, width=128, height=72, spp=2, depth=5)
    frames = list(range(19))
    want = one_by_one(pkg, sc, update, frames, [SEED] * len(frames))
    r = pkg.Renderer(device=0)
    try:
        r.commitScene(sc, update)
        a = r.submitFrames(0, 4)
        b = r.submitFrames(4, 3)
        assert {a, b} == {0, 1}
        with pytest.raises(pkg.RtError) as e:
            r.submitFrames(7, 2)                    # both windows are in flight
        assert e.value.code == -3
        for call in (lambda: r.buildAccelerationStructure(), lambda: r.configureCamera(sc.window(), sc.camera_input()),
                     lambda: r.commitGeometryData(sc.spheres, sc.parallelograms, sc.triangles)):
            with pytest.raises(pkg.RtError) as e:   # the scene a window's kernel reads cannot be replaced under it
                call()
            assert e.value.code == -3
        for w, first, n in ((b, 4, 3), (a, 0, 4)):
            for k in range(n):
                assert r.waitFrame(w, k).rays == want[first + k]["rays"]
                assert np.array_equal(r.readFrame(w, k), want[first + k]["rgba8"])
            r.releaseWindow(w)
        with pytest.raises(pkg.RtError):
            r.waitFrame(a, 0)                       # released
        got = []
        st = r.runFrames(0, len(frames), SEED, 4, lambda f, img, out: got.append((f, img.copy(), out.rays)) or 0)
        assert [g[0] for g in got] == frames and st.frames == len(frames) and st.rays == sum(x["rays"] for x in want)
        assert st.wall_ms > 0 and st.device_ms > 0 and st.max_latency_ms > 0
        for f, img, rays in got:
            assert rays == want[f]["rays"] and np.array_equal(img.reshape(-1), want[f]["rgba8"]), f"frame {f}"
        seen = []
        with pytest.raises(pkg.RtError) as e:
            r.runFrames(0, 12, SEED, 3, lambda f, img, out: seen.append(f) or (1 if f == 4 else 0))
        assert e.value.code == -3 and seen == [0, 1, 2, 3, 4]
        r.runFrames(0, 3, SEED, 8, None)            # usable after a stop
        with pytest.raises(pkg.RtError):
            r.submitFrames(0, 9)                    # more than RT_MAX_FRAMES_IN_FLIGHT
    finally:
        r.cleanup()


def test_window_on_the_headline_geometry(pkg):
    """Config 2's geometry at 480x270, 16 spp: a window of 3 frames equals three synchronous frames; the dear-tiles-first
    order (taken from the previous launch's costs) must not change a bit either — second window, same frames."""
    sc = pkg.scenes.c2(width=480, height=270, spp=16, depth=10)
    seeds = [SEED, SEED + 1, SEED + 2]
    want = one_by_one(pkg, sc, None, [0, 1, 2], seeds)
    r = pkg.Renderer(device=0)
    try:
        r.commitScene(sc)
        for rounds in range(2):
            w = r.submitFrames(0, 3, seeds=seeds)
            for k in range(3):
                assert r.waitFrame(w, k).rays == want[k]["rays"]
                assert np.array_equal(r.readFrame(w, k, pkg.abi.BUF_RADIANCE), want[k]["radiance"])
            r.releaseWindow(w)
    finally:
        r.cleanup()


@pytest.mark.parametrize("world,tile_rows", [(2, 8), (3, 16)])
def test_frame_ring_assembles_the_single_gpu_frame(pkg, world, tile_rows):
    """The fused gather: `world` tile-shard contexts (here all on device 0, driven by one process) store their finished
    texels straight into rank 0's ring of full frames and raise their flags; what rank 0 reads back is, bit for bit, the
    frame of one unsharded context — for the synchronous rt_render_frame and for windows of frames in flight."""
    sc, update = moving_scene(pkg, width=160, height=100, spp=2, depth=6)      # 100 rows: the last tile row is ragged
    frames = [2, 3, 4, 5, 6]
    want = one_by_one(pkg, sc, update, frames, [SEED] * len(frames))
    ranks = []
    try:
        for q in range(world):
            r = pkg.Renderer(device=0)
            r.commitScene(sc, update)
            r.setTileShard(q, world, tile_rows)
            ranks.append(r)
        handle = ranks[0].frameRingCreate()
        assert len(handle) == pkg.abi.IPC_HANDLE_BYTES
        for r in ranks[1:]:
            r.frameRingAttach(ranks[0])
        with pytest.raises(pkg.RtError) as e:
            ranks[1].setTileShard(0, world, tile_rows)      # the ring belongs to the shard
        assert e.value.code == -3
        # synchronous frames: every rank renders, then rank 0 reads slot 0 of window 0
        rays = sum(r.renderFrame(frames[0], SEED).rays for r in ranks)
        assert rays == want[0]["rays"]
        full = ranks[0].readFramebuffer(pkg.abi.BUF_RGBA8)
        assert full.size == sc.width * sc.height * 4 and np.array_equal(full, want[0]["rgba8"])
        # frames in flight: every rank submits the same window, rank 0 waits for all of them per frame
        for rounds in range(2):
            ws = [r.submitFrames(frames[0], len(frames)) for r in ranks]
            assert len(set(ws)) == 1
            for k in range(len(frames)):
                outs = [r.waitFrame(ws[0], k) for r in reversed(ranks)]        # rank 0 last: it also waits for the others' flags
                assert sum(o.rays for o in outs) == want[k]["rays"]
                img = ranks[0].readFrame(ws[0], k, pkg.abi.BUF_RGBA8)
                assert img.size == sc.width * sc.height * 4
                assert np.array_equal(img, want[k]["rgba8"]), f"world {world}: frame {frames[k]} differs from the unsharded frame"
            for r in ranks:
                r.releaseWindow(ws[0])
        for r in ranks[1:]:
            r.frameRingClose()
        ranks[0].frameRingClose()
        out = ranks[0].renderFrame(frames[0], SEED)          # back to the packed local plane
        assert out.rays > 0
    finally:
        for r in ranks:
            r.cleanup()


def test_frames_in_flight_need_the_default_kernel(pkg):
    sc, _ = moving_scene(pkg, width=64, height=36, spp=1, depth=2)
    for flags in (pkg.abi.FLAG_WAVEFRONT, pkg.abi.FLAG_MEGAKERNEL, pkg.abi.FLAG_COUNT_TRAVERSAL):
        r = pkg.Renderer(device=0, flags=flags)
        try:
            r.commitScene(sc)
            with pytest.raises(pkg.RtError) as e:
                r.submitFrames(0, 2)
            assert e.value.code == -3
        finally:
            r.cleanup()


def test_window_with_a_camera_per_frame(pkg):
    """rt_set_window_cameras: a scripted camera walk (rt_camera_step, the reference's controller) inside ONE window of frames in
    flight — the reference re-reads its controller every frame (src/Global/Renderer.cu:247-255).  Every frame equals the
    synchronous frame rendered after rt_configure_camera with that frame's camera."""
    sc, update = moving_scene(pkg, width=160, height=90, spp=2, depth=6)
    cams = [sc.camera_input()]
    for dx, dy, keys in [(40, -10, pkg.abi.KEY_W), (0, 0, pkg.abi.KEY_D | pkg.abi.KEY_SPACE), (-90, 25, 0), (15, 5, pkg.abi.KEY_S), (0, -30, pkg.abi.KEY_A)]:
        nxt, moved = pkg.cameraStep(cams[-1], dx, dy, keys, move_speed=0.4)
        assert moved
        cams.append(nxt)
    frames = list(range(len(cams)))
    want = []
    r = pkg.Renderer(device=0)
    try:
        r.commitScene(sc, update)
        for f, cam in zip(frames, cams):
            r.configureCamera(sc.window(), cam)
            out = r.renderFrame(f, SEED + f)
            want.append((out.rays, r.readFramebuffer(pkg.abi.BUF_RGBA8).copy()))
        assert not np.array_equal(want[0][1], want[3][1])
        r.configureCamera(sc.window(), cams[0])
        r.setWindowCameras(cams)
        w = r.submitFrames(0, len(cams), seeds=[SEED + f for f in frames])
        for k in frames:
            assert r.waitFrame(w, k).rays == want[k][0]
            assert np.array_equal(r.readFrame(w, k), want[k][1]), f"frame {k}"
        r.releaseWindow(w)
        # the cameras were this window's: the next one uses the configured camera again
        w = r.submitFrames(0, 2, seeds=[SEED, SEED + 1])
        assert np.array_equal(r.readFrame(w, 0), want[0][1]) and not np.array_equal(r.readFrame(w, 1), want[1][1])
        r.releaseWindow(w)
        r.setWindowCameras(cams[:3])
        with pytest.raises(pkg.RtError) as e:
            r.submitFrames(0, 4)                   # three cameras for four frames
        assert e.value.code == -2
        bad = sc.camera_input()
        bad.sample_count = 16
        with pytest.raises(pkg.RtError):
            r.setWindowCameras([bad])
    finally:
        r.cleanup()
