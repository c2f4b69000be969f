"""ctypes loader for the CPU checkers (oracle/oracle_api.h).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
from __future__ import annotations

import ctypes as C
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
pkg = importlib.import_module("real-time-gpu-ray-tracer_b200.abi")
abi = pkg

PORT_PATH = os.path.join(HERE, "liboracle_port.so")
REF_PATH = os.path.join(HERE, "_ref", "libref_oracle.so")


class OracleScene(C.Structure):
    _fields_ = [("spheres", C.c_void_p), ("sphere_count", C.c_size_t),
                ("parallelograms", C.c_void_p), ("parallelogram_count", C.c_size_t),
                ("triangles", C.c_void_p), ("triangle_count", C.c_size_t),
                ("roughs", C.c_void_p), ("rough_count", C.c_size_t),
                ("metals", C.c_void_p), ("metal_count", C.c_size_t),
                ("instances", C.c_void_p), ("xforms", C.c_void_p), ("instance_count", C.c_size_t),
                ("window", abi.Window), ("camera", abi.CameraInput), ("build_seed", C.c_uint64),
                ("dielectrics", C.c_void_p), ("dielectric_count", C.c_size_t),      # extension, parity unpinned
                ("ext_records", C.c_void_p), ("ext_record_bytes", C.c_size_t), ("ext_count", C.c_size_t)]   # test-only 4th primitive type


class OracleTiming(C.Structure):
    _fields_ = [("blas_build_ms", C.c_double), ("xform_ms", C.c_double), ("tlas_build_ms", C.c_double),
                ("blas_nodes", C.c_uint64), ("tlas_nodes", C.c_uint64)]


class Oracle:
    """One of the two checkers: kind = 'port' (restatement) or 'ref' (reference sources compiled as C++)."""

    def __init__(self, kind: str = "port"):
        # "port_disc": the port compiled with the test-only fourth primitive type of tests/ext/ (make -C oracle port_disc)
        path = PORT_PATH if kind == "port" else PORT_PATH.replace("liboracle_port.so", "liboracle_port_disc.so") if kind == "port_disc" else REF_PATH
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} not built (make -C oracle {kind})")
        self.kind = "port" if kind == "port_disc" else kind
        self.prefix = "ref_" if kind == "ref" else "port_"
        self.lib = C.CDLL(path)
        vp, sz, u64, i32 = C.c_void_p, C.c_size_t, C.c_uint64, C.c_int
        self._fn("create", [C.POINTER(OracleScene), C.POINTER(OracleTiming)], vp)
        self._fn("destroy", [vp], None)
        self._fn("render", [vp, u64, i32, i32, i32, vp, vp, C.POINTER(u64)], None)
        self._fn("primary", [vp, u64, i32, i32, i32, vp, vp, vp], None)
        self._fn("trace", [vp, vp, sz, C.c_float, vp, vp, vp, vp], None)
        self._fn("camera", [vp, vp], None)
        self._fn("instances", [vp, vp], None)
        self._fn("uniforms", [u64, u64, sz, vp], None)
        self._fn("describe", [], C.c_char_p)
        if kind == "ref":
            self._fn("render_kernel", [vp, u64, i32, vp], None)

    def _fn(self, name, argtypes, restype):
        f = getattr(self.lib, self.prefix + name)
        f.argtypes = argtypes
        f.restype = restype
        setattr(self, "_" + name, f)

    def describe(self) -> str:
        return self._describe().decode()

    def uniforms(self, seed: int, subsequence: int, n: int) -> np.ndarray:
        out = np.zeros(n, np.float32)
        self._uniforms(seed, subsequence, n, abi.ptr(out))
        return out

    def scene(self, sc, build_seed: int = 1234) -> "OracleSceneHandle":
        return OracleSceneHandle(self, sc, build_seed)


def _dielectrics(sc):
    """trailing fields of oracle_scene: dielectrics, then the records of a test-only extension primitive type"""
    d = getattr(sc, "dielectrics", None)
    e = getattr(sc, "ext_records", None)
    return ((abi.ptr(d), len(d)) if d is not None and len(d) else (None, 0)) + \
        ((abi.ptr(e), e.dtype.itemsize, len(e)) if e is not None and len(e) else (None, 0, 0))


class OracleSceneHandle:
    def __init__(self, oracle: Oracle, sc, build_seed: int):
        self.o = oracle
        self.sc = sc  # keeps the numpy arrays alive
        s = OracleScene(abi.ptr(sc.spheres), len(sc.spheres), abi.ptr(sc.parallelograms), len(sc.parallelograms),
                        abi.ptr(sc.triangles), len(sc.triangles), abi.ptr(sc.roughs), len(sc.roughs),
                        abi.ptr(sc.metals), len(sc.metals), abi.ptr(sc.instances), abi.ptr(sc.xforms), len(sc.instances),
                        sc.window(), sc.camera_input(), build_seed, *_dielectrics(sc))
        if len(getattr(sc, "dielectrics", ())) and oracle.kind != "port":
            raise ValueError("the compiled reference has no dielectric material (extension: port only)")
        self.timing = OracleTiming()
        self.h = oracle._create(C.byref(s), C.byref(self.timing))
        self.W, self.H = sc.width, sc.height

    def close(self):
        if self.h:
            self.o._destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def render(self, seed: int, y0: int = 0, y1: int | None = None, threads: int | None = None, rgba8=True, radiance=True):
        y1 = self.H if y1 is None else y1
        threads = threads or os.cpu_count() or 1
        img = np.zeros((self.H, self.W, 4), np.uint8) if rgba8 else None
        rad = np.zeros((self.H, self.W, 3), np.float32) if radiance else None
        rays = C.c_uint64(0)
        self.o._render(self.h, seed, y0, y1, threads, abi.ptr(img), abi.ptr(rad), C.byref(rays))
        return img, rad, rays.value

    def render_kernel(self, seed: int, threads: int | None = None):
        img = np.zeros((self.H, self.W, 4), np.uint8)
        self.o._render_kernel(self.h, seed, threads or os.cpu_count() or 1, abi.ptr(img))
        return img

    def primary(self, seed: int, y0: int = 0, y1: int | None = None, threads: int | None = None):
        y1 = self.H if y1 is None else y1
        t = np.full((self.H, self.W), np.inf, np.float32)
        inst = np.full((self.H, self.W), -1, np.int32)
        prim = np.full((self.H, self.W), -1, np.int32)
        self.o._primary(self.h, seed, y0, y1, threads or os.cpu_count() or 1, abi.ptr(t), abi.ptr(inst), abi.ptr(prim))
        return t, inst, prim

    def trace(self, rays: np.ndarray, tmin: float = 0.001):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = len(rays)
        t = np.zeros(n, np.float32)
        inst = np.zeros(n, np.int32)
        prim = np.zeros(n, np.int32)
        nrm = np.zeros((n, 3), np.float32)
        self.o._trace(self.h, abi.ptr(rays), n, tmin, abi.ptr(t), abi.ptr(inst), abi.ptr(prim), abi.ptr(nrm))
        return t, inst, prim, nrm

    def camera(self) -> np.ndarray:
        out = np.zeros(20, np.float32)
        self.o._camera(self.h, abi.ptr(out))
        return out

    def instances(self) -> np.ndarray:
        out = np.zeros((len(self.sc.instances), 33), np.float32)
        self.o._instances(self.h, abi.ptr(out))
        return out


# ---- the reference's own CUDA build (baseline timed beside ours on the GPU box; oracle/ref_cuda_driver.cu) -----------
REF_CUDA_PATH = os.path.join(HERE, "_ref", "libref_cuda.so")


def ref_cuda_bench(sc, warmup: int = 1, frames: int = 2):
    """Runs the reference's `render` kernel (unmodified sources, nvcc sm_100a) on scene `sc`.
    Returns (frame_ms list, host build ms dict, rgba8 image of the last frame)."""
    if not os.path.exists(REF_CUDA_PATH):
        raise FileNotFoundError(f"{REF_CUDA_PATH} not built (make -C oracle refcuda, needs /root/reference)")
    lib = C.CDLL(REF_CUDA_PATH)
    lib.refcuda_bench.argtypes = [C.POINTER(OracleScene), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.refcuda_bench.restype = C.c_int
    s = OracleScene(abi.ptr(sc.spheres), len(sc.spheres), abi.ptr(sc.parallelograms), len(sc.parallelograms),
                    abi.ptr(sc.triangles), len(sc.triangles), abi.ptr(sc.roughs), len(sc.roughs),
                    abi.ptr(sc.metals), len(sc.metals), abi.ptr(sc.instances), abi.ptr(sc.xforms), len(sc.instances),
                    sc.window(), sc.camera_input(), 0)
    ms = np.zeros(frames, np.float32)
    host = np.zeros(3, np.float64)
    img = np.zeros((sc.height, sc.width, 4), np.uint8)
    rc = lib.refcuda_bench(C.byref(s), warmup, frames, abi.ptr(ms), abi.ptr(host), abi.ptr(img))
    if rc != 0:
        raise RuntimeError(f"refcuda_bench failed with {rc}")
    return ms.tolist(), {"blas_build_ms": host[0], "xform_ms": host[1], "tlas_build_ms": host[2]}, img
