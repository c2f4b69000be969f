"""ctypes loader of tests/emul/libemul.so — the g++ build of the product's __host__ __device__ cores.
TEST INFRASTRUCTURE ONLY (see tests/emul/emul.cpp)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import oracle_ctypes as oc  # noqa: E402

abi = oc.abi
LIB = os.path.join(HERE, "emul", "libemul.so")


def load(variant: str = ""):
    """variant "disc": the twin with the test-only fourth primitive type of tests/ext/ compiled in"""
    path = LIB.replace("libemul.so", f"libemul_{variant}.so") if variant else LIB
    if not os.path.exists(path):
        subprocess.check_call(["make", "-C", os.path.join(HERE, "emul")] + ([variant] if variant else []))
    lib = C.CDLL(path)
    vp, u64, i32, sz = C.c_void_p, C.c_uint64, C.c_int, C.c_size_t
    lib.emul_create.argtypes = [C.POINTER(oc.OracleScene), C.c_uint32]
    lib.emul_create.restype = vp
    lib.emul_destroy.argtypes = [vp]
    lib.emul_render.argtypes = [vp, u64, i32, i32, i32, vp, vp, C.POINTER(u64), vp, vp, vp]
    lib.emul_render.restype = C.c_uint32
    lib.emul_instances.argtypes = [vp, vp]
    lib.emul_camera.argtypes = [vp, vp]
    lib.emul_lbvh.argtypes = [vp, i32, C.POINTER(sz), vp, vp, vp, vp, vp, vp]
    lib.emul_wide.argtypes = [vp, i32, C.POINTER(sz), C.POINTER(sz), vp, vp]
    lib.emul_vtk_frame.argtypes = [C.c_char_p, i32, C.POINTER(u64), C.POINTER(u64), vp, vp, vp, vp]
    lib.emul_png_encode.argtypes = [vp, C.c_uint32, C.c_uint32, i32, vp, u64]
    lib.emul_png_encode.restype = u64
    lib.emul_exr_encode.argtypes = [vp, C.c_uint32, C.c_uint32, i32, i32, vp, u64]
    lib.emul_exr_encode.restype = u64
    lib.emul_vtk_series.argtypes = [C.c_char_p, C.c_char_p, u64, vp, u64]
    lib.emul_vtk_series.restype = C.c_int64
    return lib


def vtk_series(path: str):
    """The product's .vtk.series parser: returns ([names], times float32 array); raises RuntimeError with its message."""
    lib = load()
    buf = C.create_string_buffer(1 << 16)
    times = np.zeros(4096, np.float32)
    n = lib.emul_vtk_series(path.encode(), buf, len(buf), times.ctypes.data, len(times))
    if n < 0:
        raise RuntimeError(buf.value.decode())
    return [x for x in buf.value.decode().split("\n") if x], times[:n].copy()


def exr_encode(plane: np.ndarray, bottom_up: bool = True) -> bytes:
    """The product's EXR encoder (csrc/image_io.h) on an [H, W, 3] or [H, W] float32 array."""
    lib = load()
    plane = np.ascontiguousarray(plane, np.float32)
    h, w = plane.shape[:2]
    nc = 3 if plane.ndim == 3 else 1
    n = lib.emul_exr_encode(plane.ctypes.data, w, h, nc, int(bottom_up), None, 0)
    out = np.zeros(n, np.uint8)
    lib.emul_exr_encode(plane.ctypes.data, w, h, nc, int(bottom_up), out.ctypes.data, n)
    return out.tobytes()


def exr_decode(data: bytes):
    """Minimal OpenEXR reader for the tests (single-part, uncompressed, FLOAT channels): returns (attributes, {channel
    name: [H, W] float32 array, top-down})."""
    import struct
    magic, version = struct.unpack_from("<ii", data, 0)
    assert magic == 20000630 and version == 2
    pos, attrs = 8, {}

    def cstr(p):
        e = data.index(b"\0", p)
        return data[p:e].decode(), e + 1

    while data[pos] != 0:
        name, pos = cstr(pos)
        typ, pos = cstr(pos)
        (size,) = struct.unpack_from("<i", data, pos)
        attrs[name] = (typ, data[pos + 4:pos + 4 + size])
        pos += 4 + size
    pos += 1
    assert attrs["compression"] == ("compression", b"\0") and attrs["lineOrder"] == ("lineOrder", b"\0")
    xmin, ymin, xmax, ymax = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = xmax - xmin + 1, ymax - ymin + 1
    chans, p, body = [], 0, attrs["channels"][1]
    while body[p] != 0:
        e = body.index(b"\0", p)
        nm = body[p:e].decode()
        ptype, _plinear, xs, ys = struct.unpack_from("<iIii", body, e + 1)
        assert ptype == 2 and xs == 1 and ys == 1
        chans.append(nm)
        p = e + 1 + 16
    assert chans == sorted(chans)
    offsets = struct.unpack_from(f"<{h}Q", data, pos)
    out = {c: np.zeros((h, w), np.float32) for c in chans}
    for y in range(h):
        yy, nbytes = struct.unpack_from("<ii", data, offsets[y])
        assert yy == ymin + y and nbytes == w * 4 * len(chans)
        q = offsets[y] + 8
        for c in chans:
            out[c][y] = np.frombuffer(data, "<f4", w, q)
            q += w * 4
    assert offsets[-1] + 8 + w * 4 * len(chans) == len(data)
    return attrs, out


def png_encode(rgba: np.ndarray, bottom_up: bool = True) -> bytes:
    """The product's PNG encoder (csrc/image_io.h) on an [H, W, 4] uint8 array."""
    lib = load()
    rgba = np.ascontiguousarray(rgba, np.uint8)
    h, w = rgba.shape[:2]
    n = lib.emul_png_encode(rgba.ctypes.data, w, h, int(bottom_up), None, 0)
    out = np.zeros(n, np.uint8)
    lib.emul_png_encode(rgba.ctypes.data, w, h, int(bottom_up), out.ctypes.data, n)
    return out.tobytes()


def png_decode(data: bytes) -> np.ndarray:
    """Minimal PNG reader for the tests (8-bit RGBA, filter 0 only — what png_encode_rgba8 writes): checks the
    signature, every chunk CRC and the zlib stream; returns [H, W, 4] top-down."""
    import struct
    import zlib
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
        assert zlib.crc32(typ + body) == crc, f"bad CRC in {typ}"
        if typ == b"IHDR":
            w, h, depth, colour, comp, flt, lace = struct.unpack(">IIBBBBB", body)
            assert (depth, colour, comp, flt, lace) == (8, 6, 0, 0, 0)
        elif typ == b"IDAT":
            idat += body
        pos += 12 + n
    assert typ == b"IEND"
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, w * 4 + 1)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 4).copy()


def vtk_frame(path: str, mode: int):
    """The product's C++ reader + conversion of one .vtk frame on the CPU.  mode 0: host conversion (rt_triangle ->
    commit), mode 1: compact conversion + the device's per-triangle expansion function.  Returns (tri_verts [n, 9],
    tri_shade [n, 12] as uint32 bit patterns, instance descs, transforms)."""
    lib = load()
    nt, ni = C.c_uint64(), C.c_uint64()
    rc = lib.emul_vtk_frame(path.encode(), mode, C.byref(nt), C.byref(ni), None, None, None, None)
    if rc != 0:
        raise RuntimeError(f"emul_vtk_frame({path}) failed: {rc}")
    verts = np.zeros((nt.value, 9), np.float32)
    shade = np.zeros((nt.value, 12), np.uint32)
    inst = np.zeros(ni.value, abi.instance_desc_dt)
    xf = np.zeros(ni.value, abi.instance_xform_dt)
    rc = lib.emul_vtk_frame(path.encode(), mode, None, None, verts.ctypes.data, shade.ctypes.data, inst.ctypes.data, xf.ctypes.data)
    assert rc == 0
    return verts, shade, inst, xf


class Emul:
    def __init__(self, sc, flags: int = 0, variant: str = ""):
        self.lib = load(variant)
        self.sc = sc
        s = oc.OracleScene(abi.ptr(sc.spheres), len(sc.spheres), abi.ptr(sc.parallelograms), len(sc.parallelograms),
                           abi.ptr(sc.triangles), len(sc.triangles), abi.ptr(sc.roughs), len(sc.roughs),
                           abi.ptr(sc.metals), len(sc.metals), abi.ptr(sc.instances), abi.ptr(sc.xforms), len(sc.instances),
                           sc.window(), sc.camera_input(), 0, *oc._dielectrics(sc))
        self.h = self.lib.emul_create(C.byref(s), flags)
        assert self.h, "emul_create failed"
        self.W, self.H = sc.width, sc.height

    def render(self, seed, threads=None):
        img = np.zeros((self.H, self.W, 4), np.uint8)
        rad = np.zeros((self.H, self.W, 3), np.float32)
        t = np.zeros((self.H, self.W), np.float32)
        inst = np.zeros((self.H, self.W), np.int32)
        prim = np.zeros((self.H, self.W), np.int32)
        rays = C.c_uint64(0)
        ovf = self.lib.emul_render(self.h, seed, 0, self.H, threads or os.cpu_count() or 1, abi.ptr(img), abi.ptr(rad), C.byref(rays),
                                   abi.ptr(t), abi.ptr(inst), abi.ptr(prim))
        return dict(rgba8=img, radiance=rad, rays=rays.value, t=t, inst=inst, prim=prim, overflow=ovf)

    def count(self, seed, y0=0, y1=None, threads=None):
        """traversal work of rows [y0, y1): dict(rays, nodes, prims, insts) — what RT_FLAG_COUNT_TRAVERSAL reports on the device"""
        out = np.zeros(4, np.uint64)
        self.lib.emul_count.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p]
        self.lib.emul_count.restype = None
        self.lib.emul_count(self.h, seed, y0, self.H if y1 is None else y1, threads or os.cpu_count() or 1, abi.ptr(out))
        return dict(rays=int(out[0]), nodes=int(out[1]), prims=int(out[2]), insts=int(out[3]))

    def single_check(self, seed, y0=0, y1=None, threads=None):
        """every ray of rows [y0, y1) through the two-level walk AND the single-level walk of the path kernel's SINGLE
        instantiation (one-instance scenes): dict(rays, mismatches, nodes_two_level, nodes_single, rays_missing_top_box,
        prims_two_level, prims_single)"""
        out = np.zeros(7, np.uint64)
        self.lib.emul_single_check.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p]
        self.lib.emul_single_check.restype = C.c_int
        rc = self.lib.emul_single_check(self.h, seed, y0, self.H if y1 is None else y1, threads or os.cpu_count() or 1, abi.ptr(out))
        assert rc == 0, "emul_single_check: the scene is not exactly one instance"
        keys = ("rays", "mismatches", "nodes_two_level", "nodes_single", "rays_missing_top_box", "prims_two_level", "prims_single")
        return {k: int(v) for k, v in zip(keys, out)}

    def simt_tokens(self, seed, y0, y1, threads=None, two_level=False):
        """token streams (node / leaf / shade steps of the path kernel's SINGLE instantiation) of every pixel of rows [y0, y1),
        kept in the handle for simt_run; returns the number of tokens"""
        self.lib.emul_simt_tokens.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int]
        self.lib.emul_simt_tokens.restype = C.c_longlong
        n = self.lib.emul_simt_tokens(self.h, seed, y0, y1, threads or os.cpu_count() or 1, 1 if two_level else 0)
        assert n >= 0, "emul_simt_tokens: one-instance scenes, rows in multiples of 4"
        return int(n)

    def simt_run(self, n_warps, node_min=18, shade_min=10, variant=0):
        """the path kernel's warp votes replayed on the token streams (tests/emul/emul.cpp emul_simt_run)"""
        out = np.zeros(10, np.uint64)
        self.lib.emul_simt_run.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        self.lib.emul_simt_run.restype = C.c_int
        assert self.lib.emul_simt_run(self.h, n_warps, node_min, shade_min, variant, abi.ptr(out)) == 0
        o = [int(v) for v in out]
        return dict(node_steps=o[0], leaf_steps=o[1], shade_steps=o[2], node_lanes=o[3], leaf_lanes=o[4], shade_lanes=o[5],
                    leaf_rounds=o[6], leaf_lane_rounds=o[7], iterations=o[8], tokens=o[9])

    def pixel_rays(self, seed, y0, y1, threads=None):
        """rays traced by every pixel of rows [y0, y1): uint32 [y1 - y0, W]"""
        out = np.zeros((y1 - y0, self.W), np.uint32)
        self.lib.emul_pixel_rays.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p]
        self.lib.emul_pixel_rays.restype = None
        self.lib.emul_pixel_rays(self.h, seed, y0, y1, threads or os.cpu_count() or 1, abi.ptr(out))
        return out

    def instances(self):
        out = np.zeros((len(self.sc.instances), 33), np.float32)
        self.lib.emul_instances(self.h, abi.ptr(out))
        return out

    def camera(self):
        out = np.zeros(20, np.float32)
        self.lib.emul_camera(self.h, abi.ptr(out))
        return out

    def lbvh(self, which=0):
        nn = C.c_size_t()
        self.lib.emul_lbvh(self.h, which, C.byref(nn), None, None, None, None, None, None)
        n = nn.value
        m = max(n - 1, 0)
        keys, vals = np.zeros(n, np.uint64), np.zeros(n, np.uint32)
        left, right, parent = np.zeros(m, np.int32), np.zeros(m, np.int32), np.zeros(m, np.int32)
        bounds = np.zeros((m, 6), np.float32)
        self.lib.emul_lbvh(self.h, which, C.byref(nn), abi.ptr(keys), abi.ptr(vals), abi.ptr(left), abi.ptr(right), abi.ptr(parent), abi.ptr(bounds))
        return dict(keys=keys, prim_ids=vals, left=left, right=right, parent=parent, bounds=bounds)

    def wide(self, which=0):
        nn, npr = C.c_size_t(), C.c_size_t()
        self.lib.emul_wide(self.h, which, C.byref(nn), C.byref(npr), None, None)
        nodes = np.zeros((nn.value, 20), np.uint32)
        prims = np.zeros(npr.value, np.uint32)
        self.lib.emul_wide(self.h, which, C.byref(nn), C.byref(npr), abi.ptr(nodes), abi.ptr(prims))
        return nodes, prims

    def n_blas(self):
        return len({(int(i["type"]), int(i["first_prim"]), max(1, int(i["prim_count"]))) for i in self.sc.instances})

    def close(self):
        if self.h:
            self.lib.emul_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()
