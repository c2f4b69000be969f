"""ctypes / numpy mirror of include/rtb200.h (the C-ABI of the hot path).

Struct layouts are numpy structured dtypes (bulk scene data is generated with numpy) and are
checked against the library's own ``rt_struct_size`` at load time, so a drifted header fails loudly.
The product path has NO fallback: if ``librtb200.so`` (the CUDA extension) is missing, loading raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
# RTB200_LIB selects another in-tree build of the same sources (tuning variants made by csrc/Makefile VARIANT=...)
LIB_PATH = os.environ.get("RTB200_LIB") or os.path.join(PKG_DIR, "librtb200.so")

# enums (values identical to the reference: include/Basic/BasicTypes.cuh:9-16)
SPHERE, PARALLELOGRAM, TRIANGLE = 0, 1, 2
ROUGH, METAL, DIELECTRIC = 0, 1, 2

FLAG_FIX_PARALLELOGRAM_BBOX = 1 << 0
FLAG_RECORD_PRIMARY_AOV = 1 << 1
FLAG_KEEP_LBVH = 1 << 2
FLAG_COUNT_TRAVERSAL = 1 << 3
FLAG_TRAVERSE_BVH2 = 1 << 4
FLAG_MEGAKERNEL = 1 << 5
FLAG_WAVEFRONT = 1 << 6
FLAG_SAH_REFINE = 1 << 7
FLAG_SAH_BUILD = 1 << 8

BUF_RGBA8, BUF_RADIANCE, BUF_PRIMARY_T, BUF_PRIMARY_INSTANCE, BUF_PRIMARY_PRIM = range(5)

RT_OK, RT_ERR_IO, RT_ERR_ARGUMENT, RT_ERR_STATE, RT_ERR_CUDA, RT_ERR_NCCL = 0, -1, -2, -3, -200, -300

_f3 = (np.float32, 3)

sphere_dt = np.dtype([("center", *_f3), ("radius", np.float32), ("material_type", np.uint32), ("material_index", np.uint32)])
parallelogram_dt = np.dtype([("q", *_f3), ("u", *_f3), ("v", *_f3), ("material_type", np.uint32), ("material_index", np.uint32)])
triangle_dt = np.dtype([("v", np.float32, (3, 3)), ("n", np.float32, (3, 3)), ("material_type", np.uint32),
                        ("material_index", np.uint32), ("has_normals", np.uint32), ("_pad", np.uint32)])
rough_dt = np.dtype([("albedo", *_f3)])
metal_dt = np.dtype([("albedo", *_f3), ("fuzz", np.float32)])
dielectric_dt = np.dtype([("albedo", *_f3), ("ior", np.float32)])
instance_desc_dt = np.dtype([("type", np.uint32), ("has_bounds", np.uint32), ("first_prim", np.uint64), ("prim_count", np.uint64),
                             ("bounds_min", *_f3), ("bounds_max", *_f3), ("centroid", *_f3), ("_pad", np.uint32)])
instance_xform_dt = np.dtype([("shift", *_f3), ("rotate_deg", *_f3), ("scale", *_f3)])


class Window(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32)]


class Vec3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]

    @classmethod
    def of(cls, v):
        return cls(float(v[0]), float(v[1]), float(v[2]))


class CameraInput(C.Structure):
    _fields_ = [("background", Vec3), ("center", Vec3), ("target", Vec3), ("fov", C.c_float), ("up", Vec3),
                ("focus_disk_radius", C.c_float), ("sample_range", C.c_float),
                ("sample_count", C.c_uint64), ("ray_trace_depth", C.c_uint64)]


class CameraMotion(C.Structure):
    """One step of scripted input for rt_camera_step (the reference's WASD / mouse controller)."""
    _fields_ = [("dx", C.c_int32), ("dy", C.c_int32), ("keys", C.c_uint32), ("mouse_sensitivity", C.c_float),
                ("pitch_limit", C.c_float), ("move_speed", C.c_float)]


KEY_W, KEY_A, KEY_S, KEY_D, KEY_SPACE, KEY_LSHIFT = 1, 2, 4, 8, 16, 32


class FrameOut(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("primary_rays", C.c_uint64), ("render_ms", C.c_float), ("tlas_ms", C.c_float),
                ("rng_ms", C.c_float), ("trace_ms", C.c_float), ("kernel_launches", C.c_uint32), ("stack_overflow", C.c_uint32),
                ("waves", C.c_uint32), ("_pad", C.c_uint32), ("node_fetches", C.c_uint64), ("prim_fetches", C.c_uint64), ("instance_fetches", C.c_uint64)]


class BuildStats(C.Structure):
    _fields_ = [("blas_ms", C.c_float), ("tlas_ms", C.c_float), ("blas_count", C.c_uint64), ("primitive_count", C.c_uint64),
                ("lbvh_nodes", C.c_uint64), ("wide_nodes", C.c_uint64), ("node_bytes", C.c_uint64), ("prim_bytes", C.c_uint64),
                ("kernel_launches", C.c_uint32), ("_pad", C.c_uint32)]


class AnimationStats(C.Structure):
    _fields_ = [("frames", C.c_uint64), ("rays", C.c_uint64), ("triangles", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("wall_ms", C.c_double), ("device_ms", C.c_double), ("slowest_rebuild_ms", C.c_double),
                ("parse_ms", C.c_double), ("set_frame_ms", C.c_double), ("render_wall_ms", C.c_double), ("readback_ms", C.c_double),
                ("sink_ms", C.c_double)]


class WindowStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("frames", C.c_uint32), ("kernel_launches", C.c_uint32), ("device_ms", C.c_float),
                ("prepare_ms", C.c_float), ("stack_overflow", C.c_uint32), ("_pad", C.c_uint32)]


class LoopStats(C.Structure):
    _fields_ = [("frames", C.c_uint64), ("rays", C.c_uint64), ("kernel_launches", C.c_uint64), ("wall_ms", C.c_double),
                ("device_ms", C.c_double), ("span_ms", C.c_double), ("max_latency_ms", C.c_double)]


MAX_FRAMES_IN_FLIGHT = int(os.environ.get("RTB200_MAX_WINDOW", "8"))      # RT_MAX_FRAMES_IN_FLIGHT of the loaded build (8 unless a variant build says otherwise)
IPC_HANDLE_BYTES = 64

FRAME_SINK = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(FrameOut))
UPDATE_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p)

_STRUCT_SIZES = {
    "rt_sphere": sphere_dt.itemsize, "rt_parallelogram": parallelogram_dt.itemsize, "rt_triangle": triangle_dt.itemsize,
    "rt_rough": rough_dt.itemsize, "rt_metal": metal_dt.itemsize, "rt_dielectric": dielectric_dt.itemsize,
    "rt_instance_desc": instance_desc_dt.itemsize, "rt_instance_xform": instance_xform_dt.itemsize,
    "rt_window": C.sizeof(Window), "rt_camera_input": C.sizeof(CameraInput), "rt_frame_out": C.sizeof(FrameOut),
    "rt_build_stats": C.sizeof(BuildStats), "rt_animation_stats": C.sizeof(AnimationStats),
    "rt_camera_motion": C.sizeof(CameraMotion), "rt_window_stats": C.sizeof(WindowStats), "rt_loop_stats": C.sizeof(LoopStats),
}

# every symbol include/rtb200.h declares
EXPORTS = [
    "rt_create", "rt_cleanup", "rt_last_error", "rt_version", "rt_struct_size",
    "rt_commit_geometry", "rt_commit_materials", "rt_commit_dielectrics", "rt_configure_instances",
    "rt_build_acceleration_structure", "rt_get_build_stats", "rt_configure_camera",
    "rt_render_frame", "rt_set_instance_xforms", "rt_read_framebuffer", "rt_framebuffer_device_ptr", "rt_write_ppm", "rt_write_image",
    "rt_set_tile_shard", "rt_read_lbvh", "rt_read_wide_bvh", "rt_load_vtk_series", "rt_set_vtk_frame", "rt_prefetch_vtk_frame",
    "rt_probe_read_bandwidth", "rt_run_vtk_animation", "rt_camera_step",
    "rt_submit_frames", "rt_set_window_cameras", "rt_wait_frame", "rt_read_frame", "rt_release_window", "rt_run_frames",
    "rt_commit_extension_primitives", "rt_primitive_type_count",
    "rt_frame_ring_create", "rt_frame_ring_open", "rt_frame_ring_attach", "rt_frame_ring_close",
    "rt_build_tlas_no_alloc",
]

_lib = None


def load_library(path: str | None = None) -> C.CDLL:
    """Load librtb200.so.  Raises (no CPU fallback) when the CUDA extension has not been built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(f"CUDA extension missing: {p} (run `python -c 'import __graft_entry__ as g; g.build()'` "
                           f"or `make -C {PKG_DIR}/csrc`); there is no CPU fallback")
    lib = C.CDLL(p)
    vp, sz, u64, u32, i32 = C.c_void_p, C.c_size_t, C.c_uint64, C.c_uint32, C.c_int
    lib.rt_create.argtypes = [C.POINTER(vp), i32, u32]
    lib.rt_cleanup.argtypes = [vp]
    lib.rt_last_error.argtypes = [vp]
    lib.rt_last_error.restype = C.c_char_p
    lib.rt_version.restype = C.c_char_p
    lib.rt_struct_size.argtypes = [C.c_char_p]
    lib.rt_struct_size.restype = sz
    lib.rt_commit_geometry.argtypes = [vp, vp, sz, vp, sz, vp, sz]
    lib.rt_commit_materials.argtypes = [vp, vp, sz, vp, sz]
    lib.rt_commit_dielectrics.argtypes = [vp, vp, sz]
    lib.rt_configure_instances.argtypes = [vp, vp, sz, vp, vp, vp]
    lib.rt_build_acceleration_structure.argtypes = [vp]
    lib.rt_get_build_stats.argtypes = [vp, C.POINTER(BuildStats)]
    lib.rt_configure_camera.argtypes = [vp, C.POINTER(Window), C.POINTER(CameraInput)]
    lib.rt_render_frame.argtypes = [vp, u64, u64, C.POINTER(FrameOut)]
    lib.rt_set_instance_xforms.argtypes = [vp, vp, sz]
    lib.rt_read_framebuffer.argtypes = [vp, i32, vp, sz]
    lib.rt_framebuffer_device_ptr.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(sz)]
    lib.rt_write_ppm.argtypes = [vp, C.c_char_p]
    lib.rt_write_image.argtypes = [vp, i32, C.c_char_p]
    lib.rt_set_tile_shard.argtypes = [vp, u32, u32, u32, vp]
    lib.rt_read_lbvh.argtypes = [vp, sz, C.POINTER(sz), vp, vp, vp, vp, vp, vp]
    lib.rt_read_wide_bvh.argtypes = [vp, sz, C.POINTER(sz), C.POINTER(sz), vp, vp]
    lib.rt_load_vtk_series.argtypes = [vp, C.c_char_p, C.POINTER(sz)]
    lib.rt_set_vtk_frame.argtypes = [vp, sz]
    lib.rt_prefetch_vtk_frame.argtypes = [vp, sz]
    lib.rt_run_vtk_animation.argtypes = [vp, sz, sz, sz, u64, i32, FRAME_SINK, vp, C.POINTER(AnimationStats)]
    lib.rt_camera_step.argtypes = [C.POINTER(CameraInput), C.POINTER(CameraMotion), C.POINTER(CameraInput), C.POINTER(C.c_int)]
    lib.rt_probe_read_bandwidth.argtypes = [vp, sz, i32, C.POINTER(C.c_double)]
    # no rt_ctx, no allocation: device pointers of the current device + an optional error-text buffer (include/rtb200.h)
    lib.rt_build_tlas_no_alloc.argtypes = [vp, u32, vp, vp, vp, vp, C.c_char_p, sz]
    lib.rt_submit_frames.argtypes = [vp, u64, u32, vp, u64, C.POINTER(u32)]
    lib.rt_set_window_cameras.argtypes = [vp, vp, u32]
    lib.rt_wait_frame.argtypes = [vp, u32, u32, C.POINTER(FrameOut)]
    lib.rt_read_frame.argtypes = [vp, u32, u32, i32, vp, sz]
    lib.rt_release_window.argtypes = [vp, u32, C.POINTER(WindowStats)]
    lib.rt_run_frames.argtypes = [vp, u64, sz, u64, u32, FRAME_SINK, vp, C.POINTER(LoopStats)]
    lib.rt_commit_extension_primitives.argtypes = [vp, vp, sz, sz]
    lib.rt_primitive_type_count.argtypes = []
    lib.rt_frame_ring_create.argtypes = [vp, vp]
    lib.rt_frame_ring_open.argtypes = [vp, vp]
    lib.rt_frame_ring_attach.argtypes = [vp, vp]
    lib.rt_frame_ring_close.argtypes = [vp]
    for name in EXPORTS:
        fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
        if name not in ("rt_last_error", "rt_version", "rt_struct_size"):
            fn.restype = C.c_int
    for name, size in _STRUCT_SIZES.items():
        got = lib.rt_struct_size(name.encode())
        if got != size:
            raise RuntimeError(f"ABI drift: sizeof({name}) is {got} in librtb200.so but {size} in abi.py")
    if path is None:
        _lib = lib
    return lib


def ptr(a: np.ndarray | None):
    if a is None or a.size == 0:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def canonical_wide_bvh(nodes: np.ndarray, leaf_prims: np.ndarray, roots) -> list:
    """Layout-independent form of a compressed 8-wide BVH (csrc/rt_types.h BvhNode8): a pre-order walk from `roots`
    listing, per node, the 18 words that do not depend on memory placement (origin, exponents/imask, meta, quantised
    boxes) and the primitive ids of its leaf children in slot order.  nodes: (N, 20) uint32."""
    out = []
    stack = list(reversed(list(roots)))
    while stack:
        i = stack.pop()
        n = nodes[i]
        child_base, prim_base = int(n[4]), int(n[5])
        meta = [(int(n[6 + (s >> 2)]) >> (8 * (s & 3))) & 0xFF for s in range(8)]
        imask = int(n[3]) >> 24
        prims, kids = [], []
        for s in range(8):
            m = meta[s]
            if m == 0:
                continue
            if imask & (1 << s):
                kids.append(child_base + bin(imask & ((1 << s) - 1)).count("1"))
            else:
                cnt = bin(m >> 5).count("1")
                off = m & 0x1F
                prims.append(tuple(int(v) for v in leaf_prims[prim_base + off: prim_base + off + cnt]))
        out.append((tuple(int(v) for v in n[[0, 1, 2, 3, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19]]), tuple(prims)))
        stack.extend(reversed(kids))
    return out
