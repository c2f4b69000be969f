"""`not gpu`: the C-ABI library loads, exports every symbol include/rtb200.h declares, agrees with the Python mirror on
struct sizes, and fails loudly (no CPU fallback) without a CUDA device.  No compute calls."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "rtb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", src)) - {"rt_update_cb"})


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.abi.load_library()
    names = header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rtb200.h but not exported by librtb200.so"
    assert sorted(pkg.abi.EXPORTS) == names, "abi.EXPORTS out of sync with include/rtb200.h"


def test_struct_sizes(pkg):
    lib = pkg.abi.load_library()
    assert lib.rt_struct_size(b"rt_triangle") == pkg.abi.triangle_dt.itemsize == 88
    assert lib.rt_struct_size(b"rt_sphere") == 24 and lib.rt_struct_size(b"rt_instance_xform") == 36
    assert lib.rt_struct_size(b"nope") == 0
    assert b"sm_100a" in lib.rt_version()


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(pkg.RtError) as e:
        pkg.Renderer(device=0)
    assert e.value.code == pkg.abi.RT_ERR_CUDA


def test_missing_extension_fails_loudly(pkg, tmp_path):
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.abi.load_library(str(tmp_path / "librtb200.so"))


def test_product_does_not_link_the_oracle(pkg):
    import subprocess
    out = subprocess.run(["ldd", pkg.abi.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "emul" not in out
    syms = subprocess.run(["nm", "-D", "--defined-only", pkg.abi.LIB_PATH], capture_output=True, text=True).stdout
    assert "port_render" not in syms and "ref_render" not in syms and "emul_render" not in syms


def _np_camera_step(center, target, up, dx, dy, keys, sens, limit, speed):
    """Independent float32 restatement of SDL_OpenGLWindow::calculateNewPosition (reference
    src/Global/SDL_OpenGLWindow.cu:182-256, Vec3::rotate include/Basic/Vec3.cuh:140-160) for the test below."""
    f = np.float32

    def dot(a, b):
        s = f(a[0] * b[0]); s = f(s + f(a[1] * b[1])); s = f(s + f(a[2] * b[2])); return s

    def unit(a):
        k = f(f(1.0) / np.sqrt(dot(a, a), dtype=f)); return (a * k).astype(f)

    def cross(a, b):
        return np.array([f(a[1] * b[2]) - f(a[2] * b[1]), f(a[2] * b[0]) - f(a[0] * b[2]), f(a[0] * b[1]) - f(a[1] * b[0])], f)

    def rot(v, axis, ang):
        k = unit(axis); c, s = f(np.cos(f(ang))), f(np.sin(f(ang)))
        return ((v * c).astype(f) + (cross(k, v) * s).astype(f) + ((k * dot(k, v)).astype(f) * f(f(1.0) - c)).astype(f)).astype(f)

    center, target, up = (np.asarray(x, f) for x in (center, target, up))
    W = unit(target - center); U = unit(cross(W, up)); V = unit(cross(U, W))
    nc, nt, moved = center.copy(), target.copy(), False
    if dx or dy:
        moved = True
        view = target - center
        w = rot(unit(W), unit(V), f(-f(dx) * f(sens)))
        w = rot(w, unit(U), f(-f(dy) * f(sens)))
        pitch = f(np.arcsin(w[1]))
        if abs(pitch) > f(limit):
            pitch = f(np.sign(pitch) * f(limit))
            h = unit(np.array([w[0], 0, w[2]], f))
            w = (h * f(np.cos(pitch)) + np.array([0, f(np.sin(pitch)), 0], f)).astype(f)
        nt = (center + w * np.sqrt(dot(view, view), dtype=f)).astype(f)
    d = np.zeros(3, f)
    fwd = unit(np.array([W[0], 0, W[2]], f))
    for bit, vec in ((1, fwd), (4, -fwd), (8, U), (2, -U), (16, up), (32, -up)):
        if keys & bit:
            d = (d + vec).astype(f)
    if dot(d, d) > 0:
        moved = True
        t = (unit(d) * f(speed)).astype(f)
        nc, nt = (nc + t).astype(f), (nt + t).astype(f)
    return nc, nt, moved


def test_camera_step_scripted_input(pkg):
    """rt_camera_step: the reference's WASD / mouse controller as a pure host function (no GPU needed)."""
    abi = pkg.abi
    sc = pkg.scenes.c1a(frame=0, width=64, height=36, spp=1, depth=4)
    cam = sc.camera_input()
    limit = float(np.float32(np.pi) / np.float32(np.deg2rad(np.float32(80.0))))
    vec = lambda v: np.array([v.x, v.y, v.z], np.float32)
    rng = np.random.RandomState(11)
    cur = cam
    for step in range(40):
        dx, dy = (int(rng.randint(-200, 200)), int(rng.randint(-120, 120))) if step % 3 else (0, 0)
        keys = int(rng.randint(0, 64)) if step % 2 else 0
        want_c, want_t, want_moved = _np_camera_step(vec(cur.center), vec(cur.target), vec(cur.up), dx, dy, keys, 0.001, limit, 0.1)
        nxt, moved = pkg.cameraStep(cur, dx, dy, keys)
        assert moved == want_moved
        assert np.allclose(vec(nxt.center), want_c, rtol=0, atol=2e-6) and np.allclose(vec(nxt.target), want_t, rtol=0, atol=2e-5)
        # everything but center / target is carried over
        assert nxt.fov == cur.fov and nxt.sample_count == cur.sample_count and vec(nxt.up).tolist() == vec(cur.up).tolist()
        cur = nxt
    # mouse motion keeps the eye and the view distance, keys keep the view direction
    a, m = pkg.cameraStep(cam, dx=300, dy=-80)
    assert m and vec(a.center).tolist() == vec(cam.center).tolist()
    d0, d1 = np.linalg.norm(vec(cam.target) - vec(cam.center)), np.linalg.norm(vec(a.target) - vec(a.center))
    assert abs(d0 - d1) < 1e-4 * d0
    b, m = pkg.cameraStep(cam, keys=abi.KEY_W | abi.KEY_D, move_speed=0.5)
    step = vec(b.center) - vec(cam.center)
    assert m and abs(np.linalg.norm(step) - 0.5) < 1e-5 and abs(step[1]) < 1e-6          # diagonal is not faster, stays horizontal
    assert np.allclose(vec(b.target) - vec(b.center), vec(cam.target) - vec(cam.center), atol=1e-5)
    c, m = pkg.cameraStep(cam, keys=abi.KEY_W | abi.KEY_S)
    assert not m and vec(c.center).tolist() == vec(cam.center).tolist()
    # an explicit pitch limit clamps (the reference's own value, 2.25 rad, never does)
    for dy in (-1200, 1200):
        tilted, _ = pkg.cameraStep(cam, dy=dy, pitch_limit=0.5)
        w = vec(tilted.target) - vec(tilted.center)
        assert abs(abs(np.arcsin(w[1] / np.linalg.norm(w))) - 0.5) < 1e-4
        free, _ = pkg.cameraStep(cam, dy=dy)
        w = vec(free.target) - vec(free.center)
        assert abs(abs(np.arcsin(w[1] / np.linalg.norm(w))) - 1.2) < 1e-3


def test_header_is_plain_c(tmp_path):
    """include/rtb200.h is the drop-in boundary: it must compile as C99 (what a cgo / JNI / ctypes-side binding sees)
    and as C++, with no torch or CUDA types in it."""
    import subprocess
    src = tmp_path / "hdr.c"
    src.write_text('#include "include/rtb200.h"\nint main(void) { rt_camera_motion m; rt_animation_stats s; rt_frame_out o; (void) m; (void) s; (void) o; return 0; }\n')
    subprocess.check_call(["gcc", "-I", ROOT, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", str(src)])
    subprocess.check_call(["g++", "-I", ROOT, "-std=c++17", "-Wall", "-Werror", "-fsyntax-only", "-x", "c++", str(src)])
    text = open(os.path.join(ROOT, "include", "rtb200.h")).read()
    includes = re.findall(r"^\s*#\s*include\s*[<\"]([^>\"]+)", text, flags=re.M)
    assert includes and all(h in ("stddef.h", "stdint.h") for h in includes), includes
