"""GPU tests at BASELINE.json's FULL sizes (configs 2, 4, 5), through the C-ABI.  The oracle cannot render these frames
in seconds, so parity is checked (a) against the oracle on sampled scan-lines of the full-size frame and (b) through
size-independent properties: the schedule-independent modes must agree bit for bit, tile shards must reassemble to the
single-GPU frame, the leaf order of the wide BVH must be a permutation of the primitives, a frame is a pure function
of (scene, seed)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SEED = 0x5EED


def frame(pkg, sc, flags=0, seed=SEED, shard=None, want_bvh=False):
    r = pkg.Renderer(device=0, flags=flags | pkg.abi.FLAG_RECORD_PRIMARY_AOV)
    try:
        st = r.commitScene(sc)
        if shard:
            r.setTileShard(shard[0], shard[1], shard[2])
        out = r.renderFrame(0, seed)
        W = sc.width
        rad = r.readFramebuffer(pkg.abi.BUF_RADIANCE).reshape(-1, W, 3).copy()
        res = dict(out=out, stats=st, radiance=rad,
                   rgba8=r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(-1, W, 4).copy(),
                   t=r.readFramebuffer(pkg.abi.BUF_PRIMARY_T).reshape(-1, W).copy(),
                   inst=r.readFramebuffer(pkg.abi.BUF_PRIMARY_INSTANCE).reshape(-1, W).copy(),
                   prim=r.readFramebuffer(pkg.abi.BUF_PRIMARY_PRIM).reshape(-1, W).copy())
        if want_bvh:
            res["wide"] = r.readWideBVH(0)
    finally:
        r.cleanup()
    return res


def check_lines_against_oracle(g, port, sc, lines, ids=True, rgba8=True):
    """primary-hit ids >= 99.99 %, t within 1e-5 relative, radiance RMSE <= 1e-3 on the sampled scan-lines; the RGBA8
    bytes of those lines equal the oracle's wherever the primary hit agrees (the scenes with epsilon-ties, Q1, may keep
    another of the tied primitives on < 0.01 % of the pixels)"""
    h = port.scene(sc, 7)
    same_n, px_n, se, n = 0, 0, 0.0, 0
    for y in lines:
        img, rad, _ = h.render(SEED, y, y + 1, rgba8=rgba8)
        if not ids:                                     # radiance / bytes only (the primary AOV is the FIRST sample's)
            # bit-exact per pixel against the oracle; a pixel with an epsilon-tie on one of its paths (Q1: the reference keeps
            # the hit its randomly built BVH tests last, src/AS/BLAS.cu:84-89) must equal one of its build-seed outcomes
            bad = (g["radiance"][y] != rad[y]).any(axis=-1)
            if rgba8:
                bad |= (g["rgba8"][y] != img[y]).any(axis=-1)
            assert bad.sum() <= 3, f"{sc.name} line {y}: {int(bad.sum())} pixels differ from the oracle"
            for build_seed in (8, 9, 10, 11):
                if not bad.any():
                    break
                h2 = port.scene(sc, build_seed)
                img2, rad2, _ = h2.render(SEED, y, y + 1, rgba8=rgba8)
                h2.close()
                bad &= (g["radiance"][y] != rad2[y]).any(axis=-1) | ((g["rgba8"][y] != img2[y]).any(axis=-1) if rgba8 else False)
            # what is left differs under every build seed tried: it must be an epsilon-tie the reference resolves the same way in
            # all of them (tests/tie_check.py replays the pixel and finds the segment)
            from tie_check import is_epsilon_tie
            for x in np.nonzero(bad)[0].tolist():
                assert is_epsilon_tie(sc, h, int(x), int(y), SEED), f"{sc.name} pixel ({x},{y}) differs from the oracle and not through an epsilon-tie"
            continue
        t, inst, prim = h.primary(SEED, y, y + 1)
        same = (g["inst"][y] == inst[y]) & (g["prim"][y] == prim[y])
        same_n += int(same.sum())
        px_n += same.size
        hit = same & (inst[y] >= 0)
        if hit.any():
            rel = np.abs(g["t"][y][hit] - t[y][hit]) / np.abs(t[y][hit])
            assert rel.max() <= 1e-5, f"{sc.name} line {y}: hit t relative error {rel.max():.3e}"
        if rgba8:
            assert np.array_equal(g["rgba8"][y][same], img[y][same]), f"{sc.name} line {y}: RGBA8 differs from the oracle"
        se += float(((g["radiance"][y] - rad[y]) ** 2).sum())
        n += rad[y].size
    h.close()
    if not ids:
        return
    agree, rmse = same_n / px_n, (se / n) ** 0.5
    print(f"{sc.name}: {len(lines)} lines, id-agreement={agree:.6f} rmse={rmse:.3e}")
    assert agree >= 0.9999 and rmse <= 1e-3


def test_c2_full_size(pkg, port):
    """Config 2 geometry and resolution (1 000 000 triangles, 1920x1080), 4 spp to bound the run."""
    sc = pkg.scenes.c2(spp=4)
    g = frame(pkg, sc, want_bvh=True)
    assert g["out"].stack_overflow == 0
    nodes, prims = g["wide"]
    assert len(prims) == 1_000_000 and np.array_equal(np.sort(prims), np.arange(1_000_000, dtype=np.uint32))
    assert 1_000_000 / 8 <= len(nodes) <= 1_000_000 / 2
    # the schedule must not change a single bit: wave loop and one-thread-per-pixel kernel against the path kernel
    for flags in (pkg.abi.FLAG_WAVEFRONT, pkg.abi.FLAG_MEGAKERNEL):
        o = frame(pkg, sc, flags=flags)
        assert o["out"].rays == g["out"].rays
        assert np.array_equal(o["radiance"], g["radiance"]) and np.array_equal(o["rgba8"], g["rgba8"])
        assert np.array_equal(o["prim"], g["prim"]) and np.array_equal(o["t"], g["t"])
    # pure function of (scene, seed)
    again = frame(pkg, sc)
    assert np.array_equal(again["radiance"], g["radiance"])
    other = frame(pkg, sc, seed=SEED + 1)
    assert not np.array_equal(other["radiance"], g["radiance"])
    check_lines_against_oracle(g, port, sc, [5, 300, 541, 777, 1000, 1079])


def test_c2_headline_samples(pkg, port):
    """Config 2 as the bench runs it (1920x1080, 64 spp = 8x8 strata, depth 10): two scan-lines of the full frame equal
    the oracle's bit for bit, so sample indices 4..63 of the stratified loop (reference src/Global/Kernel.cu:119-143)
    are pinned at full width too."""
    sc = pkg.scenes.c2(spp=64)
    r = pkg.Renderer(device=0)
    try:
        r.commitScene(sc)
        out = r.renderFrame(0, SEED)
        g = dict(radiance=r.readFramebuffer(pkg.abi.BUF_RADIANCE).reshape(-1, sc.width, 3).copy(),
                 rgba8=r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(-1, sc.width, 4).copy())
    finally:
        r.cleanup()
    assert out.stack_overflow == 0
    check_lines_against_oracle(g, port, sc, [417, 902], ids=False)


@pytest.mark.parametrize("world,tile_rows", [(2, 8), (8, 8), (3, 16)])
def test_c2_tile_shards_reassemble(pkg, world, tile_rows):
    """Image-tile sharding at 1080p: the union of the ranks' planes is the single-GPU frame, bit for bit (the pixel
    streams are keyed by the GLOBAL pixel index)."""
    sc = pkg.scenes.c2(spp=1)
    full = frame(pkg, sc)
    H, W = sc.height, sc.width
    got = np.zeros((H, W, 3), np.float32)
    rays = 0
    for rank in range(world):
        s = frame(pkg, sc, shard=(rank, world, tile_rows))
        rays += s["out"].rays
        rows = pkg.sharding.owned_rows(H, rank, world, tile_rows)
        got[rows] = s["radiance"][:len(rows)]
    assert rays == full["out"].rays
    assert np.array_equal(got, full["radiance"])


def test_c5_instancing_stress(pkg, port):
    """Config 5: 10 000 instances of 16 meshes, incoherent diffuse bounces; 1080p at 1 spp, depth 8."""
    sc = pkg.scenes.c5(spp=1)
    g = frame(pkg, sc)
    assert g["out"].stack_overflow == 0
    assert len(np.unique(g["inst"][g["inst"] >= 0])) > 1000          # thousands of distinct instances are visible
    o = frame(pkg, sc, flags=pkg.abi.FLAG_MEGAKERNEL)
    assert o["out"].rays == g["out"].rays and np.array_equal(o["radiance"], g["radiance"])
    # interpenetrating instances meet in epsilon-ties (the hit the reference keeps there depends on ITS traversal order, Q1):
    # enough lines that the 99.99 % bar is a statistic, not one pixel
    check_lines_against_oracle(g, port, sc, list(range(13, 1080, 27)))


def test_c4_ten_million_triangles(pkg, port):
    """Config 4 geometry (10 000 000 triangles) at 3840x2160, 1 spp: the build holds at this size and the two
    schedules agree."""
    sc = pkg.scenes.c4(spp=1)
    g = frame(pkg, sc, want_bvh=True)
    assert g["out"].stack_overflow == 0
    # against the oracle: scan-lines whose pixel indices lie below 2^21, between 2^21 and 2^22 and above 2^22 (3840 px per
    # line: 2^21 starts in line 546, 2^22 in line 1092) pin the RNG jump tables, the 10 M-primitive Morton codes and the
    # quantised nodes at this extent together
    check_lines_against_oracle(g, port, sc, [3, 500, 547, 1000, 1093, 1500, 2100])
    nodes, prims = g["wide"]
    assert len(prims) == 10_000_000
    seen = np.zeros(10_000_000, np.uint8)
    seen[prims] = 1
    assert seen.all()
    assert (g["inst"] >= 0).mean() > 0.2
    o = frame(pkg, sc, flags=pkg.abi.FLAG_WAVEFRONT)
    assert o["out"].rays == g["out"].rays and np.array_equal(o["radiance"], g["radiance"])
    print(f"C4: build {g['stats'].blas_ms:.2f} ms for 10M triangles, {len(nodes)} wide nodes, frame {g['out'].render_ms:.1f} ms, {g['out'].rays} rays")


def test_pixel_indices_beyond_2_21(pkg, port):
    """Frames of more than 2 M pixels (config 4 is 8.3 M) take extra jump-table products for the top bits of the pixel
    index in the per-pixel RNG initialisation.  A 32 x 140 000 frame of the enclosed demo scene (every path scatters, so
    every pixel consumes its stream) puts the scene's rows just above index 2^21; the sampled scan-lines must equal the
    oracle's bit for bit (profiles/tools/check_high_pixel_index.py is the same check as a script)."""
    sc = pkg.scenes.c1a(frame=0, width=32, height=140000, spp=1, depth=4)
    lines = [65535, 65536, 69900, 69990, 70000, 70040]            # pixel index = 32 * y + x: 2^21 starts at y = 65536
    h = port.scene(sc, 7)
    want = {y: h.render(SEED, y, y + 1, rgba8=False)[1][y] for y in lines}
    h.close()
    r = pkg.Renderer(device=0)
    try:
        r.commitScene(sc)
        out = r.renderFrame(0, SEED)
        rad = r.readFramebuffer(pkg.abi.BUF_RADIANCE).reshape(sc.height, sc.width, 3)
    finally:
        r.cleanup()
    assert out.stack_overflow == 0
    assert len({float(want[y].sum()) for y in lines}) > 3           # the lines do differ: the streams are being consumed
    for y in lines:
        assert np.array_equal(rad[y], want[y]), f"scan-line {y} (pixel indices from {32 * y}) differs from the oracle"
