"""Host-side mirror of the reference's public API (renderer::Renderer, include/Global/Renderer.cuh:109-146)
over the C-ABI of librtb200.so.  Method names follow the reference; startRender's loop body is
``renderFrame`` (one iteration, src/Global/Renderer.cu:232-338, headless)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import abi


class RtError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"rtb200 error {code}: {msg}")
        self.code = code


def cameraStep(camera: abi.CameraInput, dx: int = 0, dy: int = 0, keys: int = 0, mouse_sensitivity: float = 0.001,
               pitch_limit: float = float(np.float32(np.pi) / np.float32(np.deg2rad(np.float32(80.0)))), move_speed: float = 0.1):
    """One step of the reference's camera controller (SDL_OpenGLWindow::calculateNewPosition) as scripted input:
    returns (next CameraInput, moved).  Defaults are the reference's (Renderer.cu:224-225)."""
    lib = abi.load_library()
    nxt, moved = abi.CameraInput(), C.c_int(0)
    m = abi.CameraMotion(dx, dy, keys, mouse_sensitivity, pitch_limit, move_speed)
    rc = lib.rt_camera_step(C.byref(camera), C.byref(m), C.byref(nxt), C.byref(moved))
    if rc != 0:
        raise RtError(rc, "rt_camera_step failed")
    return nxt, bool(moved.value)


class Renderer:
    def __init__(self, device: int = 0, flags: int = 0, lib_path: str | None = None):
        # lib_path: another in-tree build of the same sources (csrc/Makefile VARIANT=...), e.g. the test variant with a fourth primitive type
        self.lib = abi.load_library(lib_path)
        self.ctx = C.c_void_p()
        rc = self.lib.rt_create(C.byref(self.ctx), device, flags)
        if rc != 0:
            raise RtError(rc, "rt_create failed (no CUDA device?)")
        self._keep = []
        self.width = self.height = 0

    # -- plumbing ------------------------------------------------------------------------------
    def _check(self, rc: int):
        if rc != 0:
            raise RtError(rc, (self.lib.rt_last_error(self.ctx) or b"").decode())

    # -- Renderer::commitGeometryData ------------------------------------------------------------
    def commitGeometryData(self, spheres=None, parallelograms=None, triangles=None):
        n = lambda a: 0 if a is None else len(a)
        self._check(self.lib.rt_commit_geometry(self.ctx, abi.ptr(spheres), n(spheres), abi.ptr(parallelograms), n(parallelograms),
                                                abi.ptr(triangles), n(triangles)))

    def commitExtensionPrimitives(self, records: np.ndarray):
        """Records of the build's compile-time extension primitive type (csrc/rt_prims.h); RtError(-2) in the standard build."""
        self._check(self.lib.rt_commit_extension_primitives(self.ctx, abi.ptr(records), records.dtype.itemsize, len(records)))

    # -- Renderer::commitMaterialData ------------------------------------------------------------
    def commitMaterialData(self, roughs=None, metals=None, dielectrics=None):
        n = lambda a: 0 if a is None else len(a)
        self._check(self.lib.rt_commit_materials(self.ctx, abi.ptr(roughs), n(roughs), abi.ptr(metals), n(metals)))
        if dielectrics is not None and len(dielectrics):
            self._check(self.lib.rt_commit_dielectrics(self.ctx, abi.ptr(dielectrics), len(dielectrics)))

    # -- Renderer::configureInstances ------------------------------------------------------------
    def configureInstances(self, instances: np.ndarray, xforms: np.ndarray | None = None, update=None):
        """update(xforms_view: np.ndarray, frame: int) mutates the transform arguments in place
        (the reference's updateInstances callback, include/Global/Renderer.cuh:94)."""
        cb = None
        if update is not None:
            n_inst = len(instances)

            def _cb(xf_ptr, n, frame, _user):
                buf = (C.c_char * (n * abi.instance_xform_dt.itemsize)).from_address(xf_ptr)
                update(np.frombuffer(buf, abi.instance_xform_dt, count=n), int(frame))
            cb = abi.UPDATE_CB(_cb)
            self._keep.append(cb)
            assert n_inst >= 0
        self._check(self.lib.rt_configure_instances(self.ctx, abi.ptr(instances), len(instances), abi.ptr(xforms),
                                                    C.cast(cb, C.c_void_p) if cb else None, None))

    # -- Renderer::buildAccelerationStructure ------------------------------------------------------
    def buildAccelerationStructure(self) -> abi.BuildStats:
        self._check(self.lib.rt_build_acceleration_structure(self.ctx))
        st = abi.BuildStats()
        self._check(self.lib.rt_get_build_stats(self.ctx, C.byref(st)))
        return st

    # -- Renderer::configureCamera ---------------------------------------------------------------
    def configureCamera(self, window: abi.Window, camera: abi.CameraInput):
        self.width, self.height = window.width, window.height
        self._check(self.lib.rt_configure_camera(self.ctx, C.byref(window), C.byref(camera)))

    def commitScene(self, sc, update=None):
        """Convenience: the whole start-up sequence of src/Global/Main.cu:111-123 for a scenes.Scene."""
        self.commitGeometryData(sc.spheres, sc.parallelograms, sc.triangles)
        if getattr(sc, "ext_records", None) is not None and len(sc.ext_records):
            self.commitExtensionPrimitives(sc.ext_records)
        self.commitMaterialData(sc.roughs, sc.metals, getattr(sc, "dielectrics", None))
        self.configureInstances(sc.instances, sc.xforms, update)
        st = self.buildAccelerationStructure()
        self.configureCamera(sc.window(), sc.camera_input())
        return st

    # -- one iteration of Renderer::startRender ----------------------------------------------------
    def renderFrame(self, frame: int = 0, seed: int = 0x5EED) -> abi.FrameOut:
        out = abi.FrameOut()
        self._check(self.lib.rt_render_frame(self.ctx, frame, seed, C.byref(out)))
        return out

    def setInstanceTransforms(self, xforms: np.ndarray):
        self._check(self.lib.rt_set_instance_xforms(self.ctx, abi.ptr(xforms), len(xforms)))

    def setTileShard(self, rank: int, world: int, tile_rows: int = 8, peer_rgba8: int | None = None):
        self._check(self.lib.rt_set_tile_shard(self.ctx, rank, world, tile_rows, peer_rgba8))

    def devicePointer(self, kind: int):
        p, n = C.c_void_p(), C.c_size_t()
        self._check(self.lib.rt_framebuffer_device_ptr(self.ctx, kind, C.byref(p), C.byref(n)))
        return p.value, n.value

    def readFramebuffer(self, kind: int = abi.BUF_RGBA8, out: np.ndarray | None = None) -> np.ndarray:
        _, nbytes = self.devicePointer(kind)
        if out is None:
            out = np.empty(nbytes, np.uint8)
        self._check(self.lib.rt_read_framebuffer(self.ctx, kind, out.ctypes.data_as(C.c_void_p), out.nbytes))
        dt = {abi.BUF_RGBA8: np.uint8, abi.BUF_RADIANCE: np.float32, abi.BUF_PRIMARY_T: np.float32,
              abi.BUF_PRIMARY_INSTANCE: np.int32, abi.BUF_PRIMARY_PRIM: np.int32}[kind]
        return out.view(dt)

    def writePPM(self, path: str):
        self._check(self.lib.rt_write_ppm(self.ctx, path.encode()))

    def writeImage(self, kind: int, path: str):
        """Headless output stage: PPM (RGBA8), PFM (radiance, primary t) or raw int32 AOV dump of a plane."""
        self._check(self.lib.rt_write_image(self.ctx, kind, path.encode()))

    def probeReadBandwidth(self, nbytes: int, repeats: int = 20) -> float:
        """Measurement aid: GB/s of streaming an `nbytes` device buffer `repeats` times (L2-resident when it fits L2)."""
        g = C.c_double(0.0)
        self._check(self.lib.rt_probe_read_bandwidth(self.ctx, nbytes, repeats, C.byref(g)))
        return g.value

    def readLBVH(self, blas_index: int = 0):
        n = C.c_size_t()
        self._check(self.lib.rt_read_lbvh(self.ctx, blas_index, C.byref(n), None, None, None, None, None, None))
        n = n.value
        keys = np.zeros(n, np.uint64)
        ids = np.zeros(n, np.uint32)
        m = max(n - 1, 0)
        left, right, parent = np.zeros(m, np.int32), np.zeros(m, np.int32), np.zeros(m, np.int32)
        bounds = np.zeros((m, 6), np.float32)
        nn = C.c_size_t()
        self._check(self.lib.rt_read_lbvh(self.ctx, blas_index, C.byref(nn), abi.ptr(keys), abi.ptr(ids), abi.ptr(left), abi.ptr(right),
                                          abi.ptr(parent), abi.ptr(bounds)))
        return dict(keys=keys, prim_ids=ids, left=left, right=right, parent=parent, bounds=bounds)

    def readWideBVH(self, which: int = 0):
        nn, npr = C.c_size_t(), C.c_size_t()
        self._check(self.lib.rt_read_wide_bvh(self.ctx, which, C.byref(nn), C.byref(npr), None, None))
        nodes = np.zeros((nn.value, 20), np.uint32)
        prims = np.zeros(npr.value, np.uint32)
        self._check(self.lib.rt_read_wide_bvh(self.ctx, which, C.byref(nn), C.byref(npr), abi.ptr(nodes), abi.ptr(prims)))
        return nodes, prims

    # -- Renderer::configureVTKFiles (src/Global/Renderer.cu:394-443) + the per-frame particle rebuild -----------
    def configureVTKFiles(self, series_path: str) -> int:
        n = C.c_size_t()
        self._check(self.lib.rt_load_vtk_series(self.ctx, series_path.encode(), C.byref(n)))
        return n.value

    def setVTKFrame(self, frame: int) -> abi.BuildStats:
        self._check(self.lib.rt_set_vtk_frame(self.ctx, frame))
        st = abi.BuildStats()
        self._check(self.lib.rt_get_build_stats(self.ctx, C.byref(st)))
        return st

    def prefetchVTKFrame(self, frame: int):
        """Read + convert a series frame on a host thread while the GPU renders; the next setVTKFrame(frame) picks it up."""
        self._check(self.lib.rt_prefetch_vtk_frame(self.ctx, frame))

    def runVTKAnimation(self, first_frame: int = 0, frame_stride: int = 1, max_frames: int = 0, seed: int = 0x5EED, lanes: int = 2,
                        sink=None) -> abi.AnimationStats:
        """The headless frame loop (Renderer::startRender's loop) in the library: series frames first_frame,
        first_frame + frame_stride, ... on `lanes` renderer contexts of this GPU.  sink(frame, rgba8 [rows, W, 4] view,
        FrameOut) is called once per finished frame (frames of different lanes arrive out of order; copy what you keep)."""
        st = abi.AnimationStats()
        W = self.width

        def _sink(_user, frame, ptr, nbytes, out):
            try:
                buf = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(nbytes,))
                img = buf.reshape(-1, W, 4) if W else buf
                return int(sink(int(frame), img, out.contents) or 0)
            except Exception:                      # an exception must not unwind through the C frames
                import traceback
                traceback.print_exc()
                return 1

        cb = abi.FRAME_SINK(_sink) if sink is not None else abi.FRAME_SINK()
        self._check(self.lib.rt_run_vtk_animation(self.ctx, first_frame, frame_stride, max_frames, seed, lanes, cb, None, C.byref(st)))
        return st

    # -- frames in flight (rt_submit_frames .. rt_run_frames) and the multi-GPU frame ring ------------------------------
    def submitFrames(self, first_frame: int, n_frames: int, seed: int = 0x5EED, seeds=None) -> int:
        """Prepare `n_frames` consecutive frames and serve them with ONE launch of the path kernel; returns the window handle."""
        w = C.c_uint32()
        sp = None
        if seeds is not None:
            self._seeds = np.ascontiguousarray(seeds, np.uint64)
            assert len(self._seeds) == n_frames
            sp = abi.ptr(self._seeds)
        self._check(self.lib.rt_submit_frames(self.ctx, first_frame, n_frames, sp, seed, C.byref(w)))
        return w.value

    def setWindowCameras(self, cameras):
        """Cameras (abi.CameraInput) of the frames of the next submitted window, one per frame."""
        arr = (abi.CameraInput * len(cameras))(*cameras)
        self._check(self.lib.rt_set_window_cameras(self.ctx, C.cast(arr, C.c_void_p), len(cameras)))

    def waitFrame(self, window: int, k: int) -> abi.FrameOut:
        out = abi.FrameOut()
        self._check(self.lib.rt_wait_frame(self.ctx, window, k, C.byref(out)))
        return out

    def readFrame(self, window: int, k: int, kind: int = abi.BUF_RGBA8, out: np.ndarray | None = None, nbytes: int | None = None) -> np.ndarray:
        if nbytes is None:
            _, nbytes = self.devicePointer(kind)
            if kind == abi.BUF_RGBA8 and self.has_ring:
                nbytes = self.width * self.height * 4
        if out is None:
            out = np.empty(nbytes, np.uint8)
        self._check(self.lib.rt_read_frame(self.ctx, window, k, kind, out.ctypes.data_as(C.c_void_p), out.nbytes))
        dt = {abi.BUF_RGBA8: np.uint8, abi.BUF_RADIANCE: np.float32, abi.BUF_PRIMARY_T: np.float32,
              abi.BUF_PRIMARY_INSTANCE: np.int32, abi.BUF_PRIMARY_PRIM: np.int32}[kind]
        return out.view(dt)

    def releaseWindow(self, window: int) -> abi.WindowStats:
        st = abi.WindowStats()
        self._check(self.lib.rt_release_window(self.ctx, window, C.byref(st)))
        return st

    def runFrames(self, first_frame: int, n_frames: int, seed: int = 0x5EED, frames_per_window: int = abi.MAX_FRAMES_IN_FLIGHT, sink=None) -> abi.LoopStats:
        """The headless frame loop with frames in flight; sink(frame, rgba8 [rows, W, 4] view or None, FrameOut) per frame, in order."""
        st = abi.LoopStats()
        W = self.width

        def _sink(_user, frame, ptr, nbytes, out):
            try:
                img = None
                if ptr and nbytes:
                    img = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(nbytes,)).reshape(-1, W, 4)
                return int(sink(int(frame), img, out.contents) or 0)
            except Exception:                      # an exception must not unwind through the C frames
                import traceback
                traceback.print_exc()
                return 1

        cb = abi.FRAME_SINK(_sink) if sink is not None else abi.FRAME_SINK()
        self._check(self.lib.rt_run_frames(self.ctx, first_frame, n_frames, seed, frames_per_window, cb, None, C.byref(st)))
        return st

    has_ring = False

    def frameRingCreate(self) -> bytes:
        """Rank 0 of a tile shard: allocate the ring of full frames the ranks store into; returns the CUDA IPC handle bytes."""
        h = (C.c_ubyte * abi.IPC_HANDLE_BYTES)()
        self._check(self.lib.rt_frame_ring_create(self.ctx, C.cast(h, C.c_void_p)))
        self.has_ring = True
        return bytes(h)

    def frameRingOpen(self, handle: bytes):
        h = (C.c_ubyte * abi.IPC_HANDLE_BYTES).from_buffer_copy(handle)
        self._check(self.lib.rt_frame_ring_open(self.ctx, C.cast(h, C.c_void_p)))
        self.has_ring = True

    def frameRingAttach(self, owner: "Renderer"):
        self._check(self.lib.rt_frame_ring_attach(self.ctx, owner.ctx))
        self.has_ring = True

    def frameRingClose(self):
        self._check(self.lib.rt_frame_ring_close(self.ctx))
        self.has_ring = False

    # -- Renderer::cleanup -----------------------------------------------------------------------
    def cleanup(self):
        if self.ctx:
            self.lib.rt_cleanup(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.cleanup()
        except Exception:
            pass
