"""GPU parity tests proper: the CUDA path, called through the C-ABI, against the CPU oracle on the same seeded
inputs.  Tolerances are the north-star's: primary-hit ids >= 99.99 % of pixels, t within 1e-5 relative,
image RMSE <= 1e-3 (fp32) — in practice the path is bit-exact and the tests report that too."""
import numpy as np
import pytest

from conftest import N_PARITY, parity_scenes

pytestmark = pytest.mark.gpu

SEED = 0x5EED


def gpu_frame(pkg, sc, flags=0, seed=SEED):
    r = pkg.Renderer(device=0, flags=flags | pkg.abi.FLAG_RECORD_PRIMARY_AOV)
    try:
        st = r.commitScene(sc)
        out = r.renderFrame(0, seed)
        H, W = sc.height, sc.width
        res = dict(out=out, stats=st,
                   rgba8=r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(H, W, 4).copy(),
                   radiance=r.readFramebuffer(pkg.abi.BUF_RADIANCE).reshape(H, W, 3).copy(),
                   t=r.readFramebuffer(pkg.abi.BUF_PRIMARY_T).reshape(H, W).copy(),
                   inst=r.readFramebuffer(pkg.abi.BUF_PRIMARY_INSTANCE).reshape(H, W).copy(),
                   prim=r.readFramebuffer(pkg.abi.BUF_PRIMARY_PRIM).reshape(H, W).copy(),
                   lbvh=r.readLBVH(0), tlas=r.readLBVH(1), wide=r.readWideBVH(0), wide_tlas=r.readWideBVH(1))
    finally:
        r.cleanup()
    return res


def check_against_oracle(g, h, sc):
    img, rad, rays = h.render(SEED)
    t, inst, prim = h.primary(SEED)
    assert g["out"].stack_overflow == 0
    same = (g["inst"] == inst) & (g["prim"] == prim)
    agree = same.mean()
    assert agree >= 0.9999, f"{sc.name}: primary-hit id agreement {agree:.6f} < 99.99 %"
    hit = same & (inst >= 0)
    if hit.any():
        rel = np.abs(g["t"][hit] - t[hit]) / np.abs(t[hit])
        assert rel.max() <= 1e-5, f"{sc.name}: hit t relative error {rel.max():.3e} > 1e-5"
    rmse = float(np.sqrt(np.mean((g["radiance"] - rad) ** 2)))
    assert rmse <= 1e-3, f"{sc.name}: radiance RMSE {rmse:.3e} > 1e-3"
    diff8 = np.abs(g["rgba8"].astype(int) - img.astype(int)).max()
    # The bar for byte work is bit-exact: castToUchar4's output (reference include/Basic/Color3.cuh:99-114) compared directly.
    # Where two primitives tie within epsilon (Q1) the hit the REFERENCE keeps follows its traversal order, and its host
    # builder draws the split axes at random (src/AS/BLAS.cu:84-89): on those pixels its own frames differ from build to
    # build.  A pixel therefore has to equal the oracle's under one of a few build seeds; all the others must be equal
    # under every seed.
    bad = (g["rgba8"] != img).any(axis=-1)
    if bad.any():
        assert bad.mean() <= 1e-4, f"{sc.name}: {int(bad.sum())} pixels differ from the oracle (max byte difference {diff8})"
        for build_seed in (101, 202, 303, 404):
            alt = h.o.scene(sc, build_seed).render(SEED)[0]
            bad &= (g["rgba8"] != alt).any(axis=-1)
        # One more order dependence is the reference's BUG, not its epsilon: a parallelogram is reached only through its
        # wrong bounding box (Q10, src/Geometry/Parallelogram.cu:48-50), so whether it or a farther surface is kept depends on
        # what the reference's TLAS walk found first.  rt_intersect.h quad_hit_t reproduces the near-first case; on a path
        # where the reference's walk goes the other way the kept surface differs (C1a at 640x360: 2 of 230 400 pixels, same
        # primary hit, same ray count).  Those pixels are tolerated on scenes that contain bug-compatible parallelograms,
        # within the north-star's own 0.01 % bar, and reported.
        if bad.any() and not len(sc.parallelograms):
            from tie_check import is_epsilon_tie
            for y, x in np.argwhere(bad).tolist():       # ties the reference resolves the same way under every build seed
                assert is_epsilon_tie(sc, h, int(x), int(y), SEED), f"{sc.name} pixel ({x},{y}) differs from the oracle and not through an epsilon-tie"
        elif bad.any():
            print(f"{sc.name}: {int(bad.sum())} pixels keep another surface than the reference behind a bug-compatible parallelogram (Q10): {np.argwhere(bad)[:4].tolist()}")
            assert bad.mean() <= 1e-4
    bitexact = np.array_equal(g["radiance"], rad) and np.array_equal(g["t"], t) and g["out"].rays == rays
    print(f"{sc.name}: id-agreement={agree:.6f} rmse={rmse:.3e} max|rgba8 diff|={diff8} rays gpu={g['out'].rays} cpu={rays} bit-exact={bitexact}")
    return bitexact


@pytest.mark.parametrize("idx", range(N_PARITY))
def test_frame_matches_oracle(pkg, port, idx):
    sc = parity_scenes(pkg.scenes)[idx]
    g = gpu_frame(pkg, sc)
    h = port.scene(sc, 4321)
    check_against_oracle(g, h, sc)


@pytest.mark.parametrize("variant", ["c1a", "c1b"])
def test_config1_at_its_quoted_size(pkg, port, variant):
    """BASELINE.json configs[0] whole: 640x360, 1 spp, depth 4, seed 0x5EED (SURVEY 8d C1a = the reference's Main.cu
    fixture, C1b = ~100 primitives), every pixel against the oracle (reference src/Global/Kernel.cu:105-147)."""
    sc = getattr(pkg.scenes, variant)(width=640, height=360, spp=1, depth=4)
    g = gpu_frame(pkg, sc)
    h = port.scene(sc, 2024)
    check_against_oracle(g, h, sc)
    h.close()


@pytest.mark.parametrize("idx", [0, 3, 4])
def test_device_lbvh_equals_cpu_twin(pkg, idx):
    """BVH topology is bit-exact for the same Morton ordering: device radix sort + Karras + refit vs the serial twin."""
    import emul_ctypes as ec
    sc = parity_scenes(pkg.scenes)[idx]
    g = gpu_frame(pkg, sc)
    e = ec.Emul(sc)
    for which, key in ((0, "lbvh"), (1, "tlas")):
        twin = e.lbvh(which)
        for name in ("keys", "prim_ids", "left", "right", "parent", "bounds"):
            assert np.array_equal(g[key][name], twin[name]), f"{sc.name}: {key}.{name} differs from the CPU twin"


@pytest.mark.parametrize("idx", range(N_PARITY))
def test_device_wide_bvh_equals_cpu_twin(pkg, idx):
    """The compressed 8-wide BVH built on the device equals the serial twin's node for node (memory placement aside)."""
    import emul_ctypes as ec
    sc = parity_scenes(pkg.scenes)[idx]
    g = gpu_frame(pkg, sc)
    e = ec.Emul(sc)
    canon = pkg.abi.canonical_wide_bvh
    assert canon(*g["wide"], range(e.n_blas())) == canon(*e.wide(0), range(e.n_blas())), sc.name
    assert canon(*g["wide_tlas"], [0]) == canon(*e.wide(1), [0]), sc.name
    assert sorted(g["wide"][1].tolist()) == sorted(e.wide(0)[1].tolist())


@pytest.mark.parametrize("idx", range(N_PARITY))
@pytest.mark.parametrize("mode", ["bvh2", "megakernel", "megakernel+bvh2", "wavefront", "wavefront+bvh2"])
def test_comparison_modes(pkg, port, idx, mode):
    """The comparison modes (binary LBVH traversal, one-thread-per-pixel kernel, wave loop) give the same frame as the
    default path (persistent path kernel over the wide BVH)."""
    flags = (pkg.abi.FLAG_TRAVERSE_BVH2 if "bvh2" in mode else 0) | (pkg.abi.FLAG_MEGAKERNEL if "megakernel" in mode else 0) \
        | (pkg.abi.FLAG_WAVEFRONT if "wavefront" in mode else 0)
    sc = parity_scenes(pkg.scenes)[idx]
    check_against_oracle(gpu_frame(pkg, sc, flags=flags), port.scene(sc, 99), sc)


def test_device_equals_cpu_emulation_bitwise(pkg):
    """Same source, two compilers: the CUDA kernels and the g++ build of the same cores must agree bit for bit."""
    import emul_ctypes as ec
    for sc in parity_scenes(pkg.scenes):
        g = gpu_frame(pkg, sc)
        r = ec.Emul(sc).render(SEED)
        assert np.array_equal(g["radiance"], r["radiance"]), sc.name
        assert np.array_equal(g["rgba8"], r["rgba8"]), sc.name
        assert np.array_equal(g["t"], r["t"]) and np.array_equal(g["inst"], r["inst"]) and np.array_equal(g["prim"], r["prim"]), sc.name
        assert g["out"].rays == r["rays"], sc.name


def test_parallelogram_fix_flag(pkg, port):
    """RT_FLAG_FIX_PARALLELOGRAM_BBOX renders the whole quad; default is bug-compatible (Q10)."""
    sc = pkg.scenes.c1b(width=128, height=72, spp=1, depth=2)
    compat = gpu_frame(pkg, sc)
    fixed = gpu_frame(pkg, sc, flags=pkg.abi.FLAG_FIX_PARALLELOGRAM_BBOX)
    assert (fixed["inst"] == 0).sum() > (compat["inst"] == 0).sum()


@pytest.mark.parametrize("which", ["mixed", "c2"])
def test_dielectric_matches_port(pkg, port, which):
    """EXTENSION — PARITY UNPINNED (the reference has ROUGH and METAL only, include/Basic/BasicTypes.cuh:14-16): the CUDA
    path against the oracle port's independent restatement of the same Snell / Schlick scatter, all three materials on
    all three primitive types, and the extended config 2 (glass instead of metal) at test size."""
    sc = pkg.scenes.c1_mixed() if which == "mixed" else pkg.scenes.c2_dielectric(width=160, height=90, spp=4, depth=10, nx=100, nz=50)
    for flags in (0, pkg.abi.FLAG_WAVEFRONT, pkg.abi.FLAG_MEGAKERNEL):
        assert check_against_oracle(gpu_frame(pkg, sc, flags=flags), port.scene(sc, 77), sc)


def test_argument_limits(pkg):
    """what the kernels cannot represent is refused with RT_ERR_ARGUMENT instead of hanging or reading out of bounds."""
    r = pkg.Renderer(device=0)
    try:
        sc = pkg.scenes.c1b(width=32, height=18, spp=1, depth=2)
        r.commitScene(sc)
        for spp, depth in ((256 * 256, 4), (4, 65536)):
            bad = pkg.scenes.c1b(width=32, height=18, spp=spp, depth=depth)
            with pytest.raises(pkg.RtError) as e:
                r.configureCamera(bad.window(), bad.camera_input())
            assert e.value.code == -2
        # a material index beyond the committed materials
        sc.spheres[5]["material_type"] = pkg.abi.METAL
        sc.spheres[5]["material_index"] = 4096
        r.commitScene(sc)
        with pytest.raises(pkg.RtError) as e:
            r.renderFrame(0, SEED)
        assert e.value.code == -2 and "metal" in str(e.value)
        sc.spheres[5]["material_type"] = 9
        with pytest.raises(pkg.RtError) as e:
            r.commitGeometryData(sc.spheres, sc.parallelograms, sc.triangles)
        assert e.value.code == -2
    finally:
        r.cleanup()


def test_stack_overflow_is_an_error(pkg):
    """a traversal stack overflow drops subtrees, so the frame is wrong — rt_render_frame / rt_release_window return
    RT_ERR_OVERFLOW instead of only setting a flag.  Forced with the test-only build whose stacks hold 2 entries."""
    import os
    lib = os.path.join(os.path.dirname(os.path.abspath(pkg.__file__)), "librtb200_tinystack.so")
    sc = pkg.scenes.c2(width=160, height=90, spp=2, depth=6, nx=100, nz=50)
    r = pkg.Renderer(device=0, lib_path=lib)
    try:
        r.commitScene(sc)
        with pytest.raises(pkg.RtError) as e:
            r.renderFrame(0, SEED)
        assert e.value.code == -4
        w = r.submitFrames(0, 2)
        assert r.waitFrame(w, 1).rays > 0
        with pytest.raises(pkg.RtError) as e:
            r.releaseWindow(w)
        assert e.value.code == -4
        r.submitFrames(2, 1)                       # the handle is free again
    finally:
        r.cleanup()
    ok = pkg.Renderer(device=0)                    # the product build has room: same scene, no error
    try:
        ok.commitScene(sc)
        assert ok.renderFrame(0, SEED).stack_overflow == 0
    finally:
        ok.cleanup()


def test_depth_zero_is_white(pkg, port):
    """ray_trace_depth = 0: rayColor's loop never runs (reference src/Global/Kernel.cu:64), every sample is (1,1,1)."""
    sc = pkg.scenes.c1a(width=64, height=36, spp=4, depth=0)
    g = gpu_frame(pkg, sc)
    img, rad, rays = port.scene(sc, 1).render(SEED)
    assert g["out"].rays == rays == 0
    assert np.array_equal(g["radiance"], rad) and np.array_equal(g["rgba8"], img) and (rad == 1.0).all()


def test_error_codes(pkg):
    r = pkg.Renderer(device=0)
    with pytest.raises(pkg.RtError) as e:
        r.renderFrame()
    assert e.value.code == -3
    with pytest.raises(pkg.RtError) as e:
        r.configureInstances(np.zeros(1, pkg.abi.instance_desc_dt))
    assert e.value.code == -3
    r.cleanup()


@pytest.mark.parametrize("expand", ["device", "host"])
def test_vtk_series_frames(pkg, port, tmp_path, monkeypatch, expand):
    """Config 3: rt_load_vtk_series + rt_set_vtk_frame (C++ reader, per-frame rebuild of all particle BLASes + TLAS)
    gives the same frame as committing the numpy-parsed scene directly, and matches the oracle — with the strips
    expanded into triangle records on the device (default) and on the host (RTB200_VTK_EXPAND=host)."""
    import os
    monkeypatch.setenv("RTB200_VTK_EXPAND", expand)
    S = pkg.scenes
    fixture = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "particle_frame0.vtk")
    series = S.synth_vtk_series(fixture, str(tmp_path), [25, 40])
    base = S.c1a(frame=12, width=160, height=90, spp=1, depth=10)
    r = pkg.Renderer(device=0, flags=pkg.abi.FLAG_RECORD_PRIMARY_AOV)
    try:
        r.commitScene(base)
        assert r.configureVTKFiles(series) == 2
        for k, name in enumerate(["particle_mesh/particle_000000000000000.vtk", "particle_mesh/particle_000000000000100.vtk"]):
            if k == 1:
                r.prefetchVTKFrame(1)       # read ahead on a host thread: setVTKFrame(1) must take that result
            st = r.setVTKFrame(k)
            sc = S.c3(os.path.join(str(tmp_path), name), frame=12, width=160, height=90, spp=1, depth=10)
            assert st.primitive_count == len(sc.triangles) - 1 + 4           # particle triangles + the four distinct single-primitive BLASes of the demo scene
            out = r.renderFrame(0, SEED)
            g = dict(out=out, radiance=r.readFramebuffer(pkg.abi.BUF_RADIANCE).reshape(90, 160, 3).copy(),
                     rgba8=r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(90, 160, 4).copy(),
                     t=r.readFramebuffer(pkg.abi.BUF_PRIMARY_T).reshape(90, 160).copy(),
                     inst=r.readFramebuffer(pkg.abi.BUF_PRIMARY_INSTANCE).reshape(90, 160).copy(),
                     prim=r.readFramebuffer(pkg.abi.BUF_PRIMARY_PRIM).reshape(90, 160).copy())
            assert (g["inst"] >= 5).any(), "no particle visible"
            check_against_oracle(g, port.scene(sc, 31 + k), sc)
    finally:
        r.cleanup()
    with pytest.raises(pkg.RtError) as e:
        rr = pkg.Renderer(device=0)
        rr.configureVTKFiles(str(tmp_path / "missing.series"))
    assert e.value.code == -1


def test_vtk_frame_shards_equal_the_single_loop(pkg, tmp_path):
    """Config 3's multi-GPU partitioning (frame i -> rank i % world): each rank's renderer reads, rebuilds and renders
    only its own frames from its own copy of the series; put back in frame order the images are, bit for bit, those
    of one renderer looping over the whole animation (a frame does not depend on the frames rendered before it)."""
    import os
    S, sh = pkg.scenes, pkg.sharding
    fixture = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "particle_frame0.vtk")
    counts = [25, 31, 28, 40, 33]
    F = len(counts)

    def loop(series, frames):
        r = pkg.Renderer(device=0)
        imgs, tris = [], []
        try:
            r.commitScene(S.c1a(frame=0, width=160, height=90, spp=2, depth=10))
            assert r.configureVTKFiles(series) == len(frames)
            r.prefetchVTKFrame(0)
            for j, f in enumerate(frames):
                st = r.setVTKFrame(j)
                if j + 1 < len(frames):
                    r.prefetchVTKFrame(j + 1)
                r.renderFrame(f, SEED)
                imgs.append(r.readFramebuffer(pkg.abi.BUF_RADIANCE).copy())
                tris.append(int(st.primitive_count))
        finally:
            r.cleanup()
        return imgs, tris

    whole, tris = loop(S.synth_vtk_series(fixture, str(tmp_path / "all"), counts, frame_ids=list(range(F))), list(range(F)))
    assert len({t for t in tris}) == len(set(counts))            # every frame rebuilt its own particle set
    for world in (2, 3):
        per_rank = []
        for rank in range(world):
            mine = sh.frames_of_rank(F, rank, world)
            series = S.synth_vtk_series(fixture, str(tmp_path / f"w{world}r{rank}"), [counts[f] for f in mine], frame_ids=mine)
            per_rank.append(loop(series, mine)[0])
        got = sh.interleave_frames(per_rank, F)
        for f in range(F):
            assert np.array_equal(got[f], whole[f]), f"world {world}: frame {f} differs"
    assert not np.array_equal(whole[0], whole[1])


def test_animation_loop_in_the_library(pkg, tmp_path):
    """rt_run_vtk_animation (the headless frame loop behind the ABI, several renderer contexts per GPU): every frame
    it hands to the sink is the frame rt_set_vtk_frame + rt_render_frame give on one context, bit for bit; frame
    sharding is first_frame = rank, frame_stride = world."""
    import os
    S = pkg.scenes
    fixture = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "particle_frame0.vtk")
    counts = [25, 31, 28, 40, 33, 26]
    F = len(counts)
    series = S.synth_vtk_series(fixture, str(tmp_path), counts, frame_ids=list(range(F)))
    base = S.c1a(frame=0, width=160, height=90, spp=2, depth=10)
    want, rays = [], []
    r = pkg.Renderer(device=0)
    try:
        r.commitScene(base)
        assert r.configureVTKFiles(series) == F
        for f in range(F):
            r.setVTKFrame(f)
            rays.append(r.renderFrame(f, SEED).rays)
            want.append(r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(90, 160, 4).copy())
    finally:
        r.cleanup()
    for lanes, first, stride, limit in [(1, 0, 1, 0), (2, 0, 1, 0), (3, 0, 1, 0), (2, 1, 2, 0), (2, 0, 1, 3), (8, 4, 1, 0)]:
        frames = list(range(first, F, stride))[: (limit or F)]
        got = {}
        r = pkg.Renderer(device=0)
        try:
            r.commitScene(base)
            r.configureVTKFiles(series)

            def sink(frame, img, out):
                assert frame not in got
                got[frame] = (img.copy(), out.rays)

            st = r.runVTKAnimation(first, stride, limit, SEED, lanes, sink)
            assert sorted(got) == frames and st.frames == len(frames)
            assert st.rays == sum(rays[f] for f in frames) and st.wall_ms > 0 and st.device_ms > 0 and 0 < st.slowest_rebuild_ms < 50
            for f in frames:
                assert got[f][1] == rays[f]
                assert np.array_equal(got[f][0], want[f]), f"lanes={lanes}: frame {f} differs"
            # the context stays usable after the loop (lane 0 is the caller's context)
            r.setVTKFrame(0)
            r.renderFrame(0, SEED)
            assert np.array_equal(r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(90, 160, 4), want[0])
            # a sink that returns non-zero stops the loop
            seen = []
            with pytest.raises(pkg.RtError) as e:
                r.runVTKAnimation(0, 1, 0, SEED, lanes, lambda frame, img, out: seen.append(frame) or 1)
            assert e.value.code == -3 and 1 <= len(seen) <= lanes
        finally:
            r.cleanup()
    with pytest.raises(pkg.RtError):
        rr = pkg.Renderer(device=0)
        rr.commitScene(base)
        rr.runVTKAnimation()                      # no series loaded


def test_scripted_camera_path(pkg):
    """Frame loop with scripted camera input (rt_camera_step -> rt_configure_camera between frames, what
    Renderer.cu:247-255 does with the WASD / mouse state): every frame of the walk equals the frame of a fresh renderer
    given that camera, and the walk does change the image."""
    sc = parity_scenes(pkg.scenes)[0]
    cam = sc.camera_input()
    r = pkg.Renderer(device=0)
    try:
        r.commitScene(sc)
        r.renderFrame(0, SEED)
        prev = r.readFramebuffer(pkg.abi.BUF_RGBA8).copy()
        for step, (dx, dy, keys) in enumerate([(60, -15, pkg.abi.KEY_W), (0, 0, pkg.abi.KEY_D | pkg.abi.KEY_SPACE), (-120, 30, 0)]):
            cam, moved = pkg.cameraStep(cam, dx, dy, keys, move_speed=0.4)
            assert moved
            r.configureCamera(sc.window(), cam)
            r.renderFrame(step + 1, SEED)
            img = r.readFramebuffer(pkg.abi.BUF_RGBA8).copy()
            assert not np.array_equal(img, prev)
            fresh = pkg.Renderer(device=0)
            try:
                fresh.commitScene(sc)
                fresh.configureCamera(sc.window(), cam)
                fresh.renderFrame(step + 1, SEED)
                assert np.array_equal(fresh.readFramebuffer(pkg.abi.BUF_RGBA8), img)
            finally:
                fresh.cleanup()
            prev = img
    finally:
        r.cleanup()


def test_output_writers(pkg, tmp_path):
    """rt_write_image: PPM / PFM / AOV dumps hold exactly the planes rt_read_framebuffer returns."""
    sc = parity_scenes(pkg.scenes)[0]
    r = pkg.Renderer(device=0, flags=pkg.abi.FLAG_RECORD_PRIMARY_AOV)
    try:
        r.commitScene(sc)
        r.renderFrame(0, SEED)
        H, W = sc.height, sc.width
        for kind, name in [(pkg.abi.BUF_RGBA8, "f.ppm"), (pkg.abi.BUF_RADIANCE, "f.pfm"), (pkg.abi.BUF_PRIMARY_T, "t.pfm"), (pkg.abi.BUF_PRIMARY_PRIM, "p.aov")]:
            r.writeImage(kind, str(tmp_path / name))
        rad = r.readFramebuffer(pkg.abi.BUF_RADIANCE).reshape(H, W, 3)
        raw = (tmp_path / "f.pfm").read_bytes()
        hdr = f"PF\n{W} {H}\n-1.0\n".encode()
        assert raw.startswith(hdr) and np.array_equal(np.frombuffer(raw[len(hdr):], np.float32).reshape(H, W, 3), rad)
        rgba = r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(H, W, 4)
        raw = (tmp_path / "f.ppm").read_bytes()
        hdr = f"P6\n{W} {H}\n255\n".encode()
        assert raw.startswith(hdr) and np.array_equal(np.frombuffer(raw[len(hdr):], np.uint8).reshape(H, W, 3), rgba[::-1, :, :3])
        import emul_ctypes as ec
        r.writeImage(pkg.abi.BUF_RGBA8, str(tmp_path / "f.png"))              # ".png": 8-bit RGBA PNG, top-down
        assert np.array_equal(ec.png_decode((tmp_path / "f.png").read_bytes()), rgba[::-1])
        r.writeImage(pkg.abi.BUF_RADIANCE, str(tmp_path / "f.exr"))            # ".exr": uncompressed float OpenEXR, top-down
        _, ch = ec.exr_decode((tmp_path / "f.exr").read_bytes())
        assert all(np.array_equal(ch[c], rad[::-1, :, k]) for k, c in enumerate("RGB"))
        prim = r.readFramebuffer(pkg.abi.BUF_PRIMARY_PRIM).reshape(H, W)
        raw = (tmp_path / "p.aov").read_bytes()
        hdr = f"RTAOV1 {W} {H} {pkg.abi.BUF_PRIMARY_PRIM}\n".encode()
        assert raw.startswith(hdr) and np.array_equal(np.frombuffer(raw[len(hdr):], np.int32).reshape(H, W), prim)
        t = r.readFramebuffer(pkg.abi.BUF_PRIMARY_T).reshape(H, W)
        raw = (tmp_path / "t.pfm").read_bytes()
        hdr = f"Pf\n{W} {H}\n-1.0\n".encode()
        assert raw.startswith(hdr) and np.array_equal(np.frombuffer(raw[len(hdr):], np.float32).reshape(H, W), t)
    finally:
        r.cleanup()


@pytest.mark.parametrize("idx", [0, 3, 4, 5])
def test_sah_refinement(pkg, port, idx):
    """RT_FLAG_SAH_REFINE: the rotated tree is the one the serial twin makes (binary and wide), and the frame still
    matches the oracle (the closest hit does not depend on the topology)."""
    import emul_ctypes as ec
    sc = parity_scenes(pkg.scenes)[idx]
    g = gpu_frame(pkg, sc, flags=pkg.abi.FLAG_SAH_REFINE | pkg.abi.FLAG_KEEP_LBVH)
    e = ec.Emul(sc, pkg.abi.FLAG_SAH_REFINE)
    for which, key in ((0, "lbvh"), (1, "tlas")):
        twin = e.lbvh(which)
        for name in ("keys", "prim_ids", "left", "right", "parent", "bounds"):
            assert np.array_equal(g[key][name], twin[name]), f"{sc.name}: {key}.{name} differs from the CPU twin"
    plain = ec.Emul(sc).lbvh(0)
    if len(plain["left"]) > 64:
        assert not np.array_equal(plain["left"], g["lbvh"]["left"]), "no rotation was applied"
    assert sorted(g["wide"][1].tolist()) == sorted(e.wide(0)[1].tolist())
    check_against_oracle(g, port.scene(sc, 5), sc)


def test_rebuilds_under_sm_pressure(pkg, tmp_path):
    """The wide-BVH collapse claims its work items by ticket, so it needs no part of its grid to be resident at any given
    time (VERDICT r1 weak #11 / ADVICE: it used to spin on co-resident CTAs).  Eight lanes rebuild and render concurrently
    while a second, independent context keeps every SM busy with a heavy frame loop from another host thread; every
    delivered frame must still be the single-context frame, and nothing may hang (pytest-timeout guards the test)."""
    import os
    import threading
    S = pkg.scenes
    fixture = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "particle_frame0.vtk")
    counts = [25, 40, 31, 64, 28, 50, 33, 26, 45, 30, 38, 27, 60, 29, 35, 41]
    F = len(counts)
    series = S.synth_vtk_series(fixture, str(tmp_path), counts, frame_ids=list(range(F)))
    base = S.c1a(frame=0, width=160, height=90, spp=2, depth=10)
    want = []
    r = pkg.Renderer(device=0)
    try:
        r.commitScene(base)
        assert r.configureVTKFiles(series) == F
        for f in range(F):
            r.setVTKFrame(f)
            r.renderFrame(f, SEED)
            want.append(r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(90, 160, 4).copy())
        stop = threading.Event()
        hog_frames = [0]

        def hog():
            h = pkg.Renderer(device=0)
            try:
                h.commitScene(S.c2(width=1920, height=1080, spp=4, depth=10, nx=300, nz=150))
                while not stop.is_set():
                    h.renderFrame(hog_frames[0], SEED)
                    hog_frames[0] += 1
            finally:
                h.cleanup()

        t = threading.Thread(target=hog)
        t.start()
        try:
            for rounds in range(3):
                got = {}
                r.runVTKAnimation(0, 1, 0, SEED, 8, lambda frame, img, out: got.__setitem__(frame, img.copy()) or 0)
                assert sorted(got) == list(range(F))
                for f in range(F):
                    assert np.array_equal(got[f], want[f]), f"frame {f} differs under SM pressure"
        finally:
            stop.set()
            t.join()
        assert hog_frames[0] > 0
    finally:
        r.cleanup()


@pytest.mark.parametrize("idx", [0, 2, 3, 4, 5])
def test_sah_build_ploc(pkg, port, idx):
    """RT_FLAG_SAH_BUILD: BLASes built with PLOC (kernels_ploc.cu).  The tree is the one the serial twin's rounds make — same
    children, parents and boxes, and the same wide BVH — and the frame still matches the oracle (the closest hit does not
    depend on the topology)."""
    import emul_ctypes as ec
    sc = parity_scenes(pkg.scenes)[idx]
    g = gpu_frame(pkg, sc, flags=pkg.abi.FLAG_SAH_BUILD | pkg.abi.FLAG_KEEP_LBVH)
    e = ec.Emul(sc, pkg.abi.FLAG_SAH_BUILD)
    twin = e.lbvh(0)
    for name in ("keys", "prim_ids", "left", "right", "parent", "bounds"):
        assert np.array_equal(g["lbvh"][name], twin[name]), f"{sc.name}: lbvh.{name} differs from the CPU twin"
    canon = pkg.abi.canonical_wide_bvh
    assert canon(*g["wide"], range(e.n_blas())) == canon(*e.wide(0), range(e.n_blas())), sc.name
    plain = ec.Emul(sc).lbvh(0)
    if len(plain["left"]) > 64:
        assert not np.array_equal(plain["left"], g["lbvh"]["left"]), "the PLOC tree equals the Karras tree"
    check_against_oracle(g, port.scene(sc, 5), sc)


@pytest.mark.parametrize("idx", [0, 3, 4, 5])
def test_split_kernel_variant(pkg, port, idx, monkeypatch):
    """RTB200_SPLIT=1: the A/B variant of the path kernel with a dedicated shade warp per CTA and a pool of more paths than trace
    lanes (csrc/kernels_split.cu; measured slower).  A path still consumes its pixel's stream in order, so the frame is
    the default kernel's and the oracle's, bit for bit — also for a window of frames in flight."""
    monkeypatch.setenv("RTB200_SPLIT", "1")
    sc = parity_scenes(pkg.scenes)[idx]
    g = gpu_frame(pkg, sc)
    assert check_against_oracle(g, port.scene(sc, 12), sc)
    r = pkg.Renderer(device=0)
    try:
        r.commitScene(sc)
        w = r.submitFrames(0, 3, seeds=[SEED + 5, SEED, SEED + 9])
        assert r.waitFrame(w, 1).rays == g["out"].rays
        assert np.array_equal(r.readFrame(w, 1, pkg.abi.BUF_RGBA8).reshape(sc.height, sc.width, 4), g["rgba8"])
        r.releaseWindow(w)
    finally:
        r.cleanup()


def test_vtk_real_mid_animation_frame(pkg, port, tmp_path):
    """Config 3 on a REAL frame from the middle of the reference's animation (files/particle_mesh/particle_..4000.vtk, series entry 40,
    100 particles: committed as tests/golden/particle_frame40.vtk) next to frame 0: series loader -> strips expanded on the device ->
    batched LBVH -> render, against the oracle given the numpy parse of the same file.  (The vertex normals and the strip rule are
    restated — the reference calls the VTK library: parity unpinned for the shading of particles, pinned for ids and t.)"""
    import json
    import os
    import shutil
    S = pkg.scenes
    golden = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    os.makedirs(tmp_path / "particle_mesh")
    names = []
    for k, f in enumerate(("particle_frame0.vtk", "particle_frame40.vtk")):
        name = f"particle_mesh/particle_{k:015d}.vtk"
        shutil.copy(os.path.join(golden, f), tmp_path / name)
        names.append({"name": name, "time": 0.4 * k})
    (tmp_path / "particle_mesh.vtk.series").write_text(json.dumps({"file-series-version": "1.0", "files": names}))
    base = S.c1a(frame=40, width=320, height=180, spp=1, depth=10)
    r = pkg.Renderer(device=0, flags=pkg.abi.FLAG_RECORD_PRIMARY_AOV)
    try:
        r.commitScene(base)
        assert r.configureVTKFiles(str(tmp_path / "particle_mesh.vtk.series")) == 2
        st = r.setVTKFrame(1)
        sc = S.c3(os.path.join(golden, "particle_frame40.vtk"), frame=40, width=320, height=180, spp=1, depth=10)
        n_particles = (len(sc.triangles) - 1) // 1258
        assert n_particles > 25 and st.primitive_count == len(sc.triangles) - 1 + 4
        out = r.renderFrame(0, SEED)
        H, W = 180, 320
        g = dict(out=out, radiance=r.readFramebuffer(pkg.abi.BUF_RADIANCE).reshape(H, W, 3).copy(), rgba8=r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(H, W, 4).copy(),
                 t=r.readFramebuffer(pkg.abi.BUF_PRIMARY_T).reshape(H, W).copy(), inst=r.readFramebuffer(pkg.abi.BUF_PRIMARY_INSTANCE).reshape(H, W).copy(),
                 prim=r.readFramebuffer(pkg.abi.BUF_PRIMARY_PRIM).reshape(H, W).copy())
        assert (g["inst"] >= 5).sum() > 500, "the particles are not visible"
        check_against_oracle(g, port.scene(sc, 40), sc)
    finally:
        r.cleanup()
