#!/usr/bin/env python
"""bench.py — the reference's headline metric (Mrays/s, ms/frame at 1080p fixed spp; LBVH build ms) on BASELINE.json's
config 2: synthetic 1 000 000-triangle mesh scene, 1920x1080, 64 spp, depth 10, rough/metal path tracing.

  python bench.py --gpus 1 --steps K --warmup W            # our arm (CUDA, through the C-ABI)
  torchrun ... bench.py --gpus N ...                       # N ranks: image tiles sharded, framebuffer gathered with NCCL
  python bench.py --impl reference ...                     # reference arm: the reference's own CPU code on the host cores

A step = one frame of the hot path (= one iteration of Renderer::startRender's loop, headless): per-frame instance
update + TLAS rebuild + per-pixel RNG init + trace/shade + framebuffer.  Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import ctypes as C
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

SEED = 0x5EED
METRIC = "Mrays/s (1M-triangle scene, 1920x1080, 64 spp, depth 10)"


def workload(args):
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    if args.config == "c4":      # configs[3]: 10M triangles, 3840x2160, 256 spp, image tiles over 2/4/8 GPUs
        sc = pkg.scenes.c4(**({"spp": args.spp} if args.spp != 64 else {}))
        cfg = {"workload": f"configs[3]: synthetic {len(sc.triangles)}-triangle height-field, {sc.width}x{sc.height}, {sc.sample_count} spp, depth {sc.ray_trace_depth}, "
                           f"frame seed 0x5EED", "triangles": int(len(sc.triangles)), "width": sc.width, "height": sc.height, "spp": sc.sample_count,
               "depth": sc.ray_trace_depth, "l2": "BVH + triangle records (~1 GB) exceed L2"}
        return pkg, sc, cfg
    if args.config == "c5":      # configs[4]: TLAS with 10k instanced BLAS meshes, incoherent diffuse bounces
        sc = pkg.scenes.c5(**({"spp": args.spp} if args.spp != 64 else {}))
        cfg = {"workload": f"configs[4]: {len(sc.instances)} instances of 16 particle meshes ({len(sc.triangles)} triangles), all rough, {sc.width}x{sc.height}, "
                           f"{sc.sample_count} spp, depth {sc.ray_trace_depth}, frame seed 0x5EED", "instances": int(len(sc.instances)),
               "width": sc.width, "height": sc.height, "spp": sc.sample_count, "depth": sc.ray_trace_depth, "l2": "working set L2-resident; rays incoherent"}
        return pkg, sc, cfg
    sc = pkg.scenes.c2(width=args.width, height=args.height, spp=args.spp, depth=args.depth, nx=args.nx, nz=args.nz)
    cfg = {"workload": f"configs[1]: synthetic {len(sc.triangles)}-triangle height-field, {sc.width}x{sc.height}, {sc.sample_count} spp, "
                       f"depth {sc.ray_trace_depth}, rough+metal, camera (0,6,9)->(0,0,0) fov 60, frame seed 0x5EED, XORWOW per-pixel streams",
           "triangles": int(len(sc.triangles)), "width": sc.width, "height": sc.height, "spp": sc.sample_count, "depth": sc.ray_trace_depth,
           "l2": "BVH + triangle records (~100 MB) exceed/approach L2 and every step re-traverses with 2M independent pixel streams; no L2 flush needed between steps (inputs larger than L2 slices touched per step)"}
    return pkg, sc, cfg


# ---------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""

    def __init__(self, gpu_index: int):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for k, nme in enumerate(names):
                    if r[3 + k].lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


class DevPtr:
    """Zero-copy view of a librtb200 device plane for torch (plumbing for the NCCL gather)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


# ---------------------------------------------------------------------------------------------------------------
def cpu_reference_sample(sc, rows, kind="ref"):
    """Time the reference's own CPU implementation (oracle/_ref, else the port) on `rows` scan-lines of the
    workload spread evenly over the image, one host thread per row.  Returns (rays, seconds, cores, build_ms, kind)."""
    import oracle_ctypes as oc
    if kind == "ref" and not os.path.exists(oc.REF_PATH):
        kind = "port"
    o = oc.Oracle(kind)
    t0 = time.perf_counter()
    h = o.scene(sc, 1234)
    build_s = time.perf_counter() - t0
    cores = os.cpu_count() or 1
    ys = sorted(set(int(round(v)) for v in np.linspace(0, sc.height - 1, rows)))
    rays_total = [0]

    def one(y):
        r = C.c_uint64(0)
        o._render(h.h, SEED, y, y + 1, 1, None, None, C.byref(r))
        return r.value
    t0 = time.perf_counter()
    with cf.ThreadPoolExecutor(max_workers=cores) as ex:
        for v in ex.map(one, ys):
            rays_total[0] += v
    dt = time.perf_counter() - t0
    timing = h.timing
    info = {"blas_build_ms_1core": timing.blas_build_ms, "tlas_build_ms_1core": timing.tlas_build_ms, "scene_setup_s": build_s}
    h.close()
    return rays_total[0], dt, min(cores, len(ys)), info, ("reference" if kind == "ref" else "port"), len(ys)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pkg, sc, cfg = workload(args)
    cores = os.cpu_count() or 1
    budget = 150.0 / max(1, args.steps + args.warmup)           # seconds per step
    rows = int(max(1, min(2 * cores, round(cores * budget / 5.0))))
    import oracle_ctypes as oc
    kind = "ref" if os.path.exists(oc.REF_PATH) else "port"
    o = oc.Oracle(kind)
    h = o.scene(sc, 1234)
    ys_all = np.linspace(0, sc.height - 1, rows * (args.steps + args.warmup)).round().astype(int)

    def step(k):
        ys = ys_all[k * rows:(k + 1) * rows]
        total = 0

        def one(y):
            r = C.c_uint64(0)
            o._render(h.h, SEED, int(y), int(y) + 1, 1, None, None, C.byref(r))
            return r.value
        t0 = time.perf_counter()
        with cf.ThreadPoolExecutor(max_workers=cores) as ex:
            for v in ex.map(one, ys):
                total += v
        return total, time.perf_counter() - t0
    for k in range(args.warmup):
        step(k)
    rays, secs = 0, 0.0
    for k in range(args.warmup, args.warmup + args.steps):
        r, s = step(k)
        rays += r
        secs += s
    value = rays / secs / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": min(cores, rows), "kind": "reference" if kind == "ref" else "port",
                             "sample": f"{rows} full scan-lines ({sc.width} px x {sc.sample_count} spp, all bounces) per step, spread evenly over the frame, one host thread per line; "
                                       f"host BLAS build {h.timing.blas_build_ms:.0f} ms on 1 core (not in the timed region)"},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------------
def reference_cuda_baseline(pkg, args, device, builds=3, budget_s=45.0):
    """north_star's first baseline: the reference's OWN CUDA build (unmodified sources compiled for sm_100a by
    oracle/Makefile `refcuda`, headless surface) on the same scene, camera and sample count as the bench line, timed in
    the same run.  The reference's host BVH builder picks its split axes at random (src/AS/BLAS.cu:84-89), so the frame
    is timed on `builds` independent builds and min / median / max are reported; the rays of that frame are counted by
    our renderer on the identical configuration (the reference kernel does not count, its seeds come from clock64()).
    Bounded: if one reference frame takes so long that `builds` of them would not fit `budget_s`, fewer are run."""
    import oracle_ctypes as oc
    if not os.path.exists(oc.REF_CUDA_PATH):
        return {"unavailable": "oracle/_ref/libref_cuda.so not built (needs /root/reference at build time)"}
    sc = pkg.scenes.c2(width=args.width, height=args.height, spp=args.spp, depth=args.depth, nx=args.nx, nz=args.nz)
    frames_ms, host_ms = [], []
    t_start = time.perf_counter()
    try:
        for b in range(builds):
            ms, host, _ = oc.ref_cuda_bench(sc, 0, 1)
            frames_ms.append(ms[0])
            host_ms.append(host["blas_build_ms"])
            spent = time.perf_counter() - t_start
            if spent + spent / (b + 1) > budget_s:
                break
    except Exception as e:   # noqa: BLE001
        if not frames_ms:
            return {"unavailable": f"{type(e).__name__}: {e}"}
    r = pkg.Renderer(device=device)
    r.commitScene(sc)
    out = r.renderFrame(0, SEED)
    out = r.renderFrame(1, SEED)
    r.cleanup()
    mr = sorted(out.rays / (m * 1e-3) / 1e6 for m in frames_ms)
    ours = out.rays / (out.render_ms * 1e-3) / 1e6
    return {"value": statistics.median(mr), "unit": "Mrays/s", "min": mr[0], "median": statistics.median(mr), "max": mr[-1], "builds": len(mr),
            "ms_per_frame": sorted(frames_ms), "spp": int(sc.sample_count), "rays": int(out.rays),
            "ours_same_config_ms": out.render_ms, "ours_same_config_mrays": ours, "speedup_vs_median": ours / statistics.median(mr),
            "host_blas_build_ms_1core": host_ms,
            "what": "reference render kernel (src/Global/Kernel.cu:105-147, 16x16 blocks) over the reference's host-built BVH on one B200, CUDA events, "
                    "one frame per independent (randomly split) reference build"}


def c2_dielectric_line(pkg, args, device, fpw):
    """The EXTENDED config 2 of SURVEY 8d: the same 1 M-triangle scene with the metal third of the quads made of glass (index of
    refraction 1.5) — BASELINE.json configs[1] names 'diffuse/metal/dielectric'.  The dielectric is an extension without a
    reference counterpart (parity unpinned; the CUDA path agrees bit for bit with the oracle port's independent restatement,
    tests/test_gpu_parity.py::test_dielectric_matches_port), so this line is reported beside the headline, never as it."""
    sc = pkg.scenes.c2_dielectric(width=args.width, height=args.height, spp=args.spp, depth=args.depth, nx=args.nx, nz=args.nz)
    r = pkg.Renderer(device=device)
    try:
        r.commitScene(sc)
        sync = r.renderFrame(0, SEED)
        r.runFrames(0, 2 * fpw, SEED, fpw, None)
        n = max(fpw, min(args.steps, 2 * fpw))
        st = r.runFrames(0, n, SEED, fpw, None)
    finally:
        r.cleanup()
    return {"value": st.rays / (st.span_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": st.span_ms / n, "steps": n, "rays_per_frame": st.rays / n,
            "synchronous_frame_ms": sync.render_ms, "materials": "rough + rough + dielectric (ior 1.5) by quad, as configs[1] with glass in place of metal",
            "parity": "unpinned: the reference has no dielectric (include/Basic/BasicTypes.cuh:14-16); GPU == oracle port restatement bit for bit"}


def c3_rebuild(pkg, device):
    """Config 3's per-frame work on its largest frame shape: 500 particles x 1258 strip triangles = 629 000 triangles
    (synthesised from the committed frame-0 fixture in the reference's on-disk layout): parse + upload + rebuild of all
    500 particle BLASes (one batched LBVH) + TLAS, then a 1080p 1 spp depth-10 frame."""
    import tempfile
    S = pkg.scenes
    fixture = os.path.join(ROOT, "tests", "golden", "particle_frame0.vtk")
    with tempfile.TemporaryDirectory() as tmp:
        series = S.synth_vtk_series(fixture, tmp, [500] * 6)
        r = pkg.Renderer(device=device)
        r.commitScene(S.c1a(frame=0, width=1920, height=1080, spp=1, depth=10))
        r.configureVTKFiles(series)
        blas, tlas, wall, frame_ms, loop_ms, rays = [], [], [], [], [], 0
        r.prefetchVTKFrame(0)
        for k in range(6):
            t0 = time.perf_counter()
            st = r.setVTKFrame(k)            # takes the frame read ahead by the host thread; commit + upload + rebuild
            wall.append(1e3 * (time.perf_counter() - t0))
            if k + 1 < 6:
                r.prefetchVTKFrame(k + 1)    # the next frame is parsed while this one renders
            blas.append(st.blas_ms)
            tlas.append(st.tlas_ms)
            out = r.renderFrame(k, SEED)
            frame_ms.append(out.render_ms)
            loop_ms.append(1e3 * (time.perf_counter() - t0))
            rays = out.rays
        r.cleanup()
    return {"triangles": int(st.primitive_count), "blas_count": int(st.blas_count), "lbvh_rebuild_ms": min(blas[1:]), "tlas_ms": min(tlas[1:]),
            "build_kernel_launches": int(st.kernel_launches), "achieved_gbs": 280 * st.primitive_count / min(blas[1:]) / 1e6,
            "commit_upload_rebuild_wall_ms": min(wall[1:]), "render_ms_1080p_1spp": min(frame_ms[1:]), "rays": int(rays),
            "animation_loop_wall_ms_per_frame": statistics.median(loop_ms[2:]),
            "parity": "particle vertex normals and the strip -> triangle rule are restated (the reference calls the VTK library, absent here): unpinned; primary-hit ids and t are not affected",
            "what": "rt_set_vtk_frame on 500-particle frames with rt_prefetch_vtk_frame reading the next one ahead; lbvh_rebuild_ms = CUDA events, "
                    "triangles resident -> wide BVH ready; animation_loop = set frame + rebuild + render 1080p, host wall per frame"}


def run_c3(args):
    """configs[2]: the particle_mesh.vtk.series animation, frames sharded i mod N over the ranks.  One
    step = one pass over the whole animation; every rank reads ahead, commits, rebuilds (all particle BLASes + TLAS)
    and renders its own frames at 1080p / 1 spp / depth 10 (Main.cu:88-89); nothing is exchanged per frame.
    `value` is device time (CUDA events: rebuild + render of every frame, one renderer per GPU, frame by frame through
    rt_set_vtk_frame / rt_render_frame); `e2e` is the host wall clock of the library's own loop (rt_run_vtk_animation)
    with --renderers-per-gpu R contexts per GPU (frames of the rank dealt round-robin): one context's commit + H2D
    overlaps the other's render, and every frame lands in pinned host memory."""
    import tempfile
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    distributed = world > 1
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (our arm) needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    cores = os.cpu_count() or 1
    # a lane keeps two host threads busy (its own + its read-ahead): leave each rank its share of the host's cores
    R = max(1, args.renderers_per_gpu) if args.renderers_per_gpu > 0 else min(3, max(1, cores // (2 * world)))
    os.environ.setdefault("RTB200_HOST_THREADS", str(max(1, min(8, cores // (world * R)))))     # the ranks share the host's cores
    os.environ.setdefault("RTB200_BLOCKING_WAIT", "1" if world * R * 4 > cores else "0")        # spin only when cores are plentiful
    if distributed:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    abi, S, sh = pkg.abi, pkg.scenes, pkg.sharding
    F = args.frames
    counts = S.c3_particle_counts(F)
    mine = sh.frames_of_rank(F, rank, world)
    fixture = os.path.join(ROOT, "tests", "golden", "particle_frame0.vtk")
    W, H = 1920, 1080

    class Lane:
        """one renderer context with its own copy of the frames it will render"""
        def __init__(self, tmp, frames, tag):
            self.frames = frames
            self.host_fb = torch.empty(W * H * 4, dtype=torch.uint8).pin_memory()
            series = S.synth_vtk_series(fixture, os.path.join(tmp, tag), [counts[f] for f in frames], frame_ids=frames)
            d = os.path.join(tmp, tag, "particle_mesh")
            self.bytes = sum(os.path.getsize(os.path.join(d, n)) for n in os.listdir(d))
            self.r = pkg.Renderer(device=local)
            self.r.commitScene(S.c1a(frame=0, width=W, height=H, spp=1, depth=10))
            assert self.r.configureVTKFiles(series) == len(frames)

        def animation(self, read_back):
            """this lane's frames once; returns (rays, device ms = rebuild + render, launches, triangles, rebuild ms list)"""
            r = self.r
            rays, dev, launches, tris, reb = 0, 0.0, 0, 0, []
            if self.frames:
                r.prefetchVTKFrame(0)
            for j, f in enumerate(self.frames):
                st = r.setVTKFrame(j)                    # takes the frame read ahead; commit + upload + rebuild
                if j + 1 < len(self.frames):
                    r.prefetchVTKFrame(j + 1)            # the next frame is parsed while this one renders
                out = r.renderFrame(f, SEED)
                rays += out.rays
                dev += st.blas_ms + st.tlas_ms + out.render_ms
                reb.append(st.blas_ms + st.tlas_ms)
                launches += st.kernel_launches + out.kernel_launches
                tris += st.primitive_count
                if read_back:
                    r.lib.rt_read_framebuffer(r.ctx, abi.BUF_RGBA8, C.c_void_p(self.host_fb.data_ptr()), self.host_fb.numel())
            return rays, dev, launches, tris, reb

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    with tempfile.TemporaryDirectory() as tmp:
        # ---- value: one renderer per GPU, device time ---------------------------------------------------------------
        one = Lane(tmp, mine, "v")
        for _ in range(args.warmup):
            one.animation(False)
        sampler = ClockSampler(local)
        barrier()
        if rank == 0:
            sampler.start()
        rays, dev_ms, launches, tris, rebuilds = 0, 0.0, 0, 0, []
        t0 = time.perf_counter()
        for _ in range(args.steps):
            n, ms, l, t, reb = one.animation(False)
            rays += n; dev_ms += ms; launches += l; tris += t; rebuilds += reb
        barrier()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        clocks = sampler.stop() if rank == 0 else None
        series_bytes = one.bytes
        # ---- e2e: the library's own frame loop (rt_run_vtk_animation), R renderer contexts per GPU, frames delivered
        #      to a host sink in pinned memory --------------------------------------------------------------------------
        delivered = []

        def sink(frame, img, out):
            delivered.append(frame)
            return 0

        for _ in range(max(1, args.warmup - 1)):
            one.r.runVTKAnimation(0, 1, 0, SEED, R, sink)
        barrier()
        e_rays = 0
        delivered.clear()
        host_ms = {"parse": 0.0, "set_frame": 0.0, "render": 0.0, "readback": 0.0, "sink": 0.0}
        t0 = time.perf_counter()
        for _ in range(args.steps):
            a = one.r.runVTKAnimation(0, 1, 0, SEED, R, sink)
            e_rays += a.rays
            for k, v in (("parse", a.parse_ms), ("set_frame", a.set_frame_ms), ("render", a.render_wall_ms), ("readback", a.readback_ms), ("sink", a.sink_ms)):
                host_ms[k] += v / max(1, a.frames) / args.steps
        barrier()
        e2e_s = time.perf_counter() - t0
        assert len(delivered) == len(mine) * args.steps
        one.r.cleanup()
    tot = torch.tensor([float(rays), float(e_rays), float(tris), float(launches), float(series_bytes)], dtype=torch.float64, device=f"cuda:{local}")
    mx = torch.tensor([dev_ms, e2e_s, wall_ms, max(rebuilds) if rebuilds else 0.0], dtype=torch.float64, device=f"cuda:{local}")
    if distributed:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    tot_rays, tot_e_rays, tot_tris, tot_launches, tot_bytes = tot.tolist()
    max_dev_ms, max_e2e_s, max_wall_ms, max_rebuild = mx.tolist()
    if rank == 0:
        line = {"metric": "Mrays/s (configs[2]: particle_mesh.vtk.series animation, per-frame LBVH rebuild + render, frames sharded over the GPUs)",
                "value": tot_rays / (max_dev_ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": max_dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"configs[2]: {F}-frame stand-in for files/particle_mesh.vtk.series ({counts[0]}..{counts[-1]} particles x 1258 strip triangles, "
                                       "the reference's on-disk layout, synthesised from the committed frame-0 fixture) + the 5 Main.cu instances, 1920x1080, 1 spp, depth 10",
                           "frames": F, "width": W, "height": H, "spp": 1, "depth": 10,
                           "l2": "every frame is a new scene (new BVH, new triangle records): nothing is reused between timed iterations"},
                "parallelism": f"frames i mod {world}, scene rebuilt per frame on the owning GPU, no per-frame exchange" if world > 1 else "single GPU",
                "parity": "particle vertex normals and the strip -> triangle rule are restated (the reference calls the VTK library, absent here): unpinned",
                "frames_per_s": F * args.steps / (max_dev_ms * 1e-3), "wall_ms_per_step": max_wall_ms / args.steps,
                "wall_frames_per_s_one_renderer": F * args.steps / (max_wall_ms * 1e-3),
                "triangles_rebuilt_per_step": tot_tris / args.steps, "slowest_rebuild_ms": max_rebuild,
                "e2e": {"value": tot_e_rays / max_e2e_s / 1e6, "unit": "Mrays/s", "frames_per_s": F * args.steps / max_e2e_s,
                        "renderers_per_gpu": R, "host_threads_per_renderer": int(os.environ["RTB200_HOST_THREADS"]), "host_cores": cores,
                        "h2d_bytes_per_step": int(tot_tris / args.steps * 36), "d2h_bytes_per_step": int(F * (W * H * 4 + 64)),
                        "ms_per_step": 1e3 * max_e2e_s / args.steps, "vtk_bytes_read_per_step": int(tot_bytes),
                        "rank0_host_ms_per_frame": {k: round(v, 3) for k, v in host_ms.items()},
                        "what": "rt_run_vtk_animation: per frame disk -> host (read ahead), commit, H2D, rebuild, render, RGBA8 -> pinned host, sink call; host wall clock"},
                "gpu_launches": int(tot_launches), "clocks": clocks}
        print(json.dumps(line))
    if distributed:
        dist.destroy_process_group()


def measure_synchronous(pkg, r, sc, args, torch, dist, rank, world, local, tile_rows):
    """The comparison modes (and the fallback when no frame ring can be mapped): rt_render_frame one frame at a time; at
    N > 1 the packed planes are gathered with NCCL and de-interleaved by torch ops."""
    abi = pkg.abi
    distributed = world > 1
    W, H = sc.width, sc.height
    ptr, nbytes = r.devicePointer(abi.BUF_RGBA8)
    if distributed:      # every rank's plane is allocated for the largest shard: gather equal-sized buffers
        nbytes = pkg.sharding.padded_plane_bytes(H, W * 4, world, tile_rows)
    local_plane = torch.as_tensor(DevPtr(ptr, nbytes), device=f"cuda:{local}")
    gathered = [torch.empty_like(local_plane) for _ in range(world)] if (distributed and rank == 0) else None
    host_fb = torch.empty(W * H * 4, dtype=torch.uint8).pin_memory()
    xf = sc.xforms.copy()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def frame(k, read_back):
        if read_back:
            r.setInstanceTransforms(xf)
        out = r.renderFrame(k, SEED)
        ms = out.render_ms
        if distributed:
            ev0.record()
            dist.gather(local_plane, gathered, dst=0)
            ev1.record()
            ev1.synchronize()
            ms += ev0.elapsed_time(ev1)
        if read_back and rank == 0:
            if distributed:
                full = pkg.sharding.assemble(gathered, H, W * 4, world, tile_rows)
                host_fb.copy_(full.view(-1), non_blocking=True)
                torch.cuda.synchronize()
            else:
                r.lib.rt_read_framebuffer(r.ctx, abi.BUF_RGBA8, C.c_void_p(host_fb.data_ptr()), host_fb.numel())
        return out.rays, ms, out

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    for k in range(args.warmup):
        frame(k, False)
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    rays, dev_ms, launches = 0, 0.0, 0
    trace_ms, rng_ms, tlas_ms = [], [], []
    t0 = time.perf_counter()
    for k in range(args.steps):
        n, ms, out = frame(args.warmup + k, False)
        rays += n
        dev_ms += ms
        trace_ms.append(out.trace_ms); rng_ms.append(out.rng_ms); tlas_ms.append(out.tlas_ms)
        launches += out.kernel_launches
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    clocks = sampler.stop() if rank == 0 else None
    frame(args.warmup, True)
    barrier()
    e_rays = 0
    t0 = time.perf_counter()
    for k in range(args.steps):
        n, _, _ = frame(args.warmup + k, True)
        e_rays += n
    barrier()
    e2e_s = time.perf_counter() - t0
    tot = torch.tensor([float(rays), float(e_rays), float(launches)], dtype=torch.float64, device=f"cuda:{local}")
    mx = torch.tensor([dev_ms, e2e_s, wall_ms], dtype=torch.float64, device=f"cuda:{local}")
    if distributed:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    tot_rays, tot_e_rays, launches = tot.tolist()
    max_dev_ms, max_e2e_s, max_wall_ms = mx.tolist()
    return {"value": tot_rays / (max_dev_ms * 1e-3) / 1e6, "ms_per_step": max_dev_ms / args.steps, "rays_per_frame": tot_rays / args.steps,
            "e2e_value": tot_e_rays / max_e2e_s / 1e6, "e2e_ms_per_step": 1e3 * max_e2e_s / args.steps, "wall_ms_per_step": max_wall_ms / args.steps,
            "launches": int(launches), "clocks": clocks,
            "stage_ms": {"tlas_rebuild": statistics.mean(tlas_ms), "rng_init": statistics.mean(rng_ms), "trace": statistics.mean(trace_ms)}}


def exchange_ring(r, dist, torch, rank, world, local):
    """Frame ring across the torchrun ranks: rank 0 allocates it and ships the 64-byte CUDA IPC handle (plumbing: one
    broadcast); the others map it.  Returns None on success, else the reason (the caller falls back to an NCCL gather)."""
    err = None
    h = torch.zeros(64, dtype=torch.uint8, device=f"cuda:{local}")
    try:
        if rank == 0:
            h = torch.frombuffer(bytearray(r.frameRingCreate()), dtype=torch.uint8).to(f"cuda:{local}")
    except Exception as e:   # noqa: BLE001
        err = f"rank 0: {e}"
    dist.broadcast(h, src=0)
    if rank != 0 and err is None:
        try:
            r.frameRingOpen(bytes(h.cpu().numpy().tobytes()))
        except Exception as e:   # noqa: BLE001
            err = f"rank {rank}: {e}"
    bad = torch.tensor([1.0 if err else 0.0], device=f"cuda:{local}")
    dist.all_reduce(bad)
    if bad.item() > 0:
        if r.has_ring:
            try:
                r.frameRingClose()
            except Exception:   # noqa: BLE001
                pass
        return err or "another rank could not map the ring"
    return None


def measure_pipelined(pkg, r, sc, args, torch, dist, rank, world, local, tile_rows, fpw, sync_frames=None, warm_frames=None):
    """value / e2e of one workload through the library's frame loop (rt_run_frames: windows of `fpw` frames in flight, two
    windows overlapping).  value: device clock of the loop, frames stay on the device.  e2e: host wall clock of the same
    loop with the per-frame transform arguments coming from host memory (H2D inside the loop) and every finished frame
    copied to pinned host memory (rank 0 reads the ring's assembled frame when the ranks share one)."""
    abi = pkg.abi
    distributed = world > 1

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    xf = sc.xforms.copy()
    # untimed: synchronous frames (stage times, single-frame latency) + one pipelined pass (allocations, tile-cost history)
    sync_ms, stage = [], None
    for k in range(sync_frames or max(3, args.warmup)):
        out = r.renderFrame(k, SEED)
        sync_ms.append(out.render_ms)
        stage = {"tlas_rebuild": out.tlas_ms, "rng_init": out.rng_ms, "trace": out.trace_ms}
    sync_rays = out.rays
    # both windows once: each has its own stream, buffers and tile-cost history (a launch hands the dear tiles out first,
    # from the costs the window's previous launch measured), and the first use of a window allocates them
    r.runFrames(0, warm_frames or 2 * fpw, SEED, fpw, None)
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    st = r.runFrames(args.warmup, args.steps, SEED, fpw, None)
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    clocks = sampler.stop() if rank == 0 else None
    # e2e
    seen = [0]

    def sink(frame, img, out):
        seen[0] += 1
        return 0
    r.setInstanceTransforms(xf)
    r.runFrames(0, min(2 * fpw, max(fpw, args.steps)), SEED, fpw, sink)        # untimed: first use of the read-back path
    barrier()
    seen[0] = 0
    t0 = time.perf_counter()
    r.setInstanceTransforms(xf)          # per-frame host input: the loop uploads the transform arguments of every frame (reference: Renderer.cu:269-291)
    es = r.runFrames(args.warmup, args.steps, SEED, fpw, sink)
    barrier()
    e2e_s = time.perf_counter() - t0
    assert seen[0] == args.steps
    tot = torch.tensor([float(st.rays), float(es.rays), float(st.kernel_launches)], dtype=torch.float64, device=f"cuda:{local}")
    mx = torch.tensor([st.span_ms, e2e_s, wall_ms, st.max_latency_ms, statistics.median(sync_ms)], dtype=torch.float64, device=f"cuda:{local}")
    if distributed:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    tot_rays, tot_e_rays, launches = tot.tolist()
    span_ms, e2e_s, wall_ms, max_lat, sync_med = mx.tolist()
    return {"value": tot_rays / (span_ms * 1e-3) / 1e6, "ms_per_step": span_ms / args.steps, "rays_per_frame": tot_rays / args.steps,
            "e2e_value": tot_e_rays / e2e_s / 1e6, "e2e_ms_per_step": 1e3 * e2e_s / args.steps, "wall_ms_per_step": wall_ms / args.steps,
            "launches": int(launches), "clocks": clocks, "stage_ms": stage, "max_latency_ms": max_lat,
            "sync_ms_per_frame": sync_med, "sync_rays": sync_rays, "span_ms": span_ms}


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    distributed = world > 1
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (our arm) needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")      # stdout carries exactly one JSON line
    if distributed:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg, sc, cfg = workload(args)
    abi = pkg.abi
    comparison = args.bvh2 or args.megakernel or args.wavefront or args.sync
    fpw = max(1, min(abi.MAX_FRAMES_IN_FLIGHT, args.frames_per_window, args.steps))

    base_flags = (abi.FLAG_TRAVERSE_BVH2 if args.bvh2 else 0) | (abi.FLAG_MEGAKERNEL if args.megakernel else 0) | (abi.FLAG_WAVEFRONT if args.wavefront else 0) | (abi.FLAG_SAH_REFINE if args.sah else 0)
    r = pkg.Renderer(device=local, flags=base_flags)
    t0 = time.perf_counter()
    st = r.commitScene(sc)
    commit_s = time.perf_counter() - t0
    # the same build again, twice: what a per-frame rebuild costs once the builder's buffers exist (the first build of a
    # context also allocates them)
    rebuilds = [r.buildAccelerationStructure() for _ in range(2)]
    warm_blas_ms = min(x.blas_ms for x in rebuilds)
    tile_rows = args.tile_rows
    exchange = "none (single GPU)"
    if distributed:
        r.setTileShard(rank, world, tile_rows)
        why = exchange_ring(r, dist, torch, rank, world, local) if not comparison else "comparison mode"
        if why is None:
            exchange = "frame ring: every rank's path kernel stores finished texels into rank 0's full frame through a CUDA IPC peer mapping over NVLink (fused gather)"
        else:
            print(f"[bench] frame ring unavailable ({why}); falling back to the synchronous loop with an NCCL gather", file=sys.stderr)
            comparison = True
            exchange = f"NCCL gather of packed planes (frame ring unavailable: {why})"
    elif args.emulate_shard > 1:        # tuning aid: this GPU renders rank 0's shard of an N-way split (not a bench line)
        r.setTileShard(0, args.emulate_shard, tile_rows)
    W, H = sc.width, sc.height

    if not comparison:
        m = measure_pipelined(pkg, r, sc, args, torch, dist, rank, world, local, tile_rows, fpw)
        value, e2e = m["value"], m["e2e_value"]
        ms_per_step, e2e_ms, wall_ms_per_step = m["ms_per_step"], m["e2e_ms_per_step"], m["wall_ms_per_step"]
        launches, clocks, stage_ms = m["launches"], m["clocks"], m["stage_ms"]
        rays_per_frame = m["rays_per_frame"]
        t_kernel_frame_ms = max(1e-6, ms_per_step - (stage_ms["tlas_rebuild"] + stage_ms["rng_init"]))
        loop = {"frames_per_window": fpw, "windows_in_flight": 2, "max_frame_latency_ms": m["max_latency_ms"],
                "synchronous_frame_ms": m["sync_ms_per_frame"], "synchronous_mrays": m["sync_rays"] / (m["sync_ms_per_frame"] * 1e-3) / 1e6,
                "what": "rt_run_frames: windows of frames in flight served by one launch of the path kernel each (a lane out of pixels of frame k starts on "
                        "frame k+1); ms_per_step = device clock of the loop / frames; synchronous_* = rt_render_frame, one frame at a time (its latency)"}
    else:
        m = measure_synchronous(pkg, r, sc, args, torch, dist, rank, world, local, tile_rows)
        value, e2e = m["value"], m["e2e_value"]
        ms_per_step, e2e_ms, wall_ms_per_step = m["ms_per_step"], m["e2e_ms_per_step"], m["wall_ms_per_step"]
        launches, clocks, stage_ms = m["launches"], m["clocks"], m["stage_ms"]
        rays_per_frame = m["rays_per_frame"]
        t_kernel_frame_ms = stage_ms["trace"]
        loop = {"frames_per_window": 1, "what": "rt_render_frame, one frame at a time"}

    if rank == 0:
        # ---- roofline of the dominant kernel (trace/shade): algorithmic bytes from an instrumented pass ----------
        roof = None
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            peak, peak_src = float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        if True:      # at N > 1: rank 0's kernel on its 1/N shard (tiles interleaved, so the whole-frame per-ray counts apply)
            rc = pkg.Renderer(device=local, flags=abi.FLAG_COUNT_TRAVERSAL | base_flags)
            rc.commitScene(sc)
            co = rc.renderFrame(0, SEED)
            rc.cleanup()
            node_b, prim_b, inst_b, io_b = (64 if args.bvh2 else 80), 48, 64, 24 + 16 + 12
            alg = (co.node_fetches * node_b + co.prim_fetches * prim_b + co.instance_fetches * inst_b + W * H * io_b) / world
            t_kernel = t_kernel_frame_ms * 1e-3
            achieved = alg / t_kernel / 1e9
            # DRAM traffic of one launch: from the committed ncu --set full capture of THIS workload, and only while the
            # kernel's sources still hash as they did when it was captured (profiles/tools/kernel_hash.py); else null
            traffic, traffic_src = None, None
            tp = os.path.join(ROOT, "profiles", "traffic_latest.json")
            if os.path.exists(tp) and world == 1 and args.emulate_shard == 1:
                try:
                    sys.path.insert(0, os.path.join(ROOT, "profiles", "tools"))
                    from kernel_hash import kernel_source_hash
                    rec = json.load(open(tp)).get(args.config if args.spp == 64 or args.config != "c2" else "-")
                    if rec and rec.get("kernel_source_sha256") == kernel_source_hash():
                        traffic, traffic_src = rec.get("dram_bytes_per_launch"), rec.get("from")
                        if rec.get("scale_to_spp") and rec["scale_to_spp"] != sc.sample_count:      # captured at fewer samples per pixel
                            traffic = traffic * sc.sample_count / rec["scale_to_spp"]
                            traffic_src += f" (captured at {rec['scale_to_spp']} spp, scaled to {sc.sample_count})"
                    elif rec:
                        traffic_src = "stale: kernel sources changed since the capture"
                except Exception:
                    traffic = None
            kname = ("render_pixels_kernel (one thread per pixel)" if args.megakernel else
                     "wave_trace kernels + wave_shade (wave loop)" if (args.wavefront or args.bvh2) else
                     "persist8_kernel (trace + shade; one launch per window of frames in flight)" if not comparison else
                     "persist8_kernel (trace + shade, one launch per frame)")
            l2 = None
            try:
                l2_gbs = r.probeReadBandwidth(32 << 20, 40)
                hbm_read_gbs = r.probeReadBandwidth(2 << 30, 2)
                l2 = {"peak": l2_gbs, "unit": "GB/s", "frac": achieved / l2_gbs, "hbm_read_probe_gbs": hbm_read_gbs,
                      "how": "rt_probe_read_bandwidth: 128-bit loads from every SM over a 32 MiB (L2-resident) / 2 GiB buffer, CUDA events"}
            except Exception as e:   # noqa: BLE001
                l2 = {"unavailable": str(e)}
            # What can bind: the node/primitive records are fetched through L2.  When they fit in it (126 MB; C1, C2, C3, C5) the
            # HBM sees almost none of the algorithmic bytes and the L2 -> SM read bandwidth, probed live here, is the roofline
            # (SURVEY 8d); when they do not (C4, ~1 GB) it is HBM.  Both fractions are reported either way.
            working_set = float(st.node_bytes + st.prim_bytes)
            l2_resident = working_set < 0.8 * 126e6 or (traffic is not None and traffic < 0.1 * alg)
            hbm = {"peak": peak, "unit": "GB/s", "peak_source": peak_src, "frac_algorithmic": achieved / peak,
                   "traffic": traffic, "traffic_gbs": (traffic / t_kernel / 1e9) if traffic else None,
                   "frac_traffic": (traffic / t_kernel / 1e9 / peak) if traffic else None, "traffic_source": traffic_src}
            if l2_resident and l2 and "peak" in l2:
                roof = {"bound": "l2", "achieved": achieved, "peak": l2["peak"], "unit": "GB/s", "frac": achieved / l2["peak"],
                        "peak_source": "measured live: " + l2["how"]}
            else:
                roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src}
            roof.update({"kernel": kname, "traffic": traffic, "algorithmic_bytes_per_launch": alg, "kernel_ms": t_kernel * 1e3,
                         "working_set_bytes": working_set, "hbm": hbm, "l2": l2,
                         "per_ray": {"node_fetches": co.node_fetches / co.rays, "prim_fetches": co.prim_fetches / co.rays,
                                     "instance_fetches": co.instance_fetches / co.rays, "bytes": alg / co.rays},
                         "record_bytes": {"node": node_b, "triangle": prim_b, "instance": inst_b, "pixel_io": io_b},
                         "note": "achieved = algorithmic bytes (instrumented fetch counts x record sizes) / kernel time; traffic = DRAM bytes of one launch "
                                 "from the committed ncu --set full capture, used only when the kernel sources hash as they did at capture time"})
        # ---- CPU baseline: the reference's own code on this box's host cores, bounded sample ---------------------
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            c_rays, c_s, c_cores, info, kind, nrows = cpu_reference_sample(sc, rows=max(4, min(4 * cores, 256)))
            cpu = {"value": c_rays / c_s / 1e6, "unit": "Mrays/s", "cores": c_cores, "kind": kind,
                   "sample": f"{nrows} full scan-lines ({W} px x {sc.sample_count} spp, all bounces) spread evenly over the frame, one host thread per line, {c_s:.1f} s",
                   "host_bvh_build_ms_1core": info["blas_build_ms_1core"]}
        refcuda = None
        if world == 1 and not args.no_ref_cuda:
            refcuda = reference_cuda_baseline(pkg, args, local)
        c3 = None
        if world == 1 and not args.no_c3:
            c3 = c3_rebuild(pkg, local)
        c2d = None
        if world == 1 and not args.no_c3 and args.config == "c2" and not comparison:
            try:
                c2d = c2_dielectric_line(pkg, args, local, fpw)
            except Exception as e:   # noqa: BLE001
                c2d = {"unavailable": f"{type(e).__name__}: {e}"}
        metric = METRIC if args.config == "c2" else f"Mrays/s ({cfg['workload'].split(',')[0]})"
        line = {"metric": metric, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": cfg,
                "parallelism": f"image tiles ({tile_rows} rows) x{world}, scene replicated; exchange: {exchange}" if world > 1 else "single GPU",
                "frame_loop": loop,
                "rays_per_frame": rays_per_frame, "wall_ms_per_step": wall_ms_per_step,
                "stage_ms": stage_ms,
                "lbvh_build_ms": {"blas_1M_tris": st.blas_ms, "blas_rebuild": warm_blas_ms, "tlas": st.tlas_ms, "kernel_launches": st.kernel_launches,
                                  "bytes_per_prim_algorithmic": 280, "achieved_gbs": 280 * st.primitive_count / max(warm_blas_ms, 1e-6) / 1e6,
                                  "what": "blas_1M_tris = first build of the context (allocates the builder's buffers), blas_rebuild = best of two more builds of the same geometry; CUDA events, primitives resident -> wide BVH ready"},
                "commit_s": commit_s,
                "e2e": {"value": e2e, "unit": "Mrays/s", "h2d_bytes_per_step": int(len(sc.instances) * 92), "d2h_bytes_per_step": int(W * H * 4 + 64),
                        "ms_per_step": e2e_ms,
                        "what": "rt_set_instance_xforms(host) + rt_run_frames with a sink: per frame the transform arguments go host -> device and the finished "
                                "RGBA8 frame device -> pinned host (rank 0: the assembled full frame), host wall clock" if not comparison else
                                "rt_set_instance_xforms(host) + rt_render_frame + rt_read_framebuffer(RGBA8 -> pinned host), host wall clock"},
                "gpu_launches": int(launches), "clocks": clocks}
        if c3:
            line["c3_particle_rebuild"] = c3
        if c2d:
            line["c2_dielectric"] = c2d
        if refcuda:
            line["reference_cuda"] = refcuda
            if "speedup_vs_median" in refcuda:
                line["vs_reference_cuda"] = refcuda["speedup_vs_median"]
        if roof:
            line["roofline"] = roof
        if cpu:
            line["cpu_baseline"] = cpu
    r.cleanup()
    # ---- config 4 beside the headline line (BASELINE.json names it for 2/4/8 GPUs): 1 warm-up + 2 steps ------------------
    c4 = None
    if args.config == "c2" and not args.no_c4 and not comparison and args.emulate_shard == 1:
        try:
            c4 = config4_line(pkg, args, torch, dist, rank, world, local, tile_rows)
        except Exception as e:   # noqa: BLE001
            c4 = {"unavailable": f"{type(e).__name__}: {e}"}
    if rank == 0:
        if c4:
            line["c4"] = c4
        print(json.dumps(line))
    if distributed:
        dist.destroy_process_group()


def config4_line(pkg, args, torch, dist, rank, world, local, tile_rows):
    """configs[3] (10 M triangles, 3840x2160, 256 spp, image tiles over the GPUs) through the same frame loop: 1 untimed
    window + 2 timed frames, so that the driver's scaling record carries the configuration BASELINE.json names for 2/4/8 GPUs."""
    import copy
    a = copy.copy(args)
    a.config, a.steps, a.warmup, a.spp = "c4", 2, 1, 64
    _, sc, cfg = workload(a)
    r = pkg.Renderer(device=local)
    try:
        st = r.commitScene(sc)
        if world > 1:
            r.setTileShard(rank, world, tile_rows)
            why = exchange_ring(r, dist, torch, rank, world, local)
            if why is not None:
                return {"unavailable": f"frame ring: {why}"}
        m = measure_pipelined(pkg, r, sc, a, torch, dist, rank, world, local, tile_rows, 1, sync_frames=1, warm_frames=2)
    finally:
        r.cleanup()
    return {"metric": f"Mrays/s ({cfg['workload'].split(',')[0]})", "value": m["value"], "unit": "Mrays/s", "ms_per_step": m["ms_per_step"],
            "steps": a.steps, "warmup": a.warmup, "e2e": {"value": m["e2e_value"], "ms_per_step": m["e2e_ms_per_step"]},
            "rays_per_frame": m["rays_per_frame"], "synchronous_frame_ms": m["sync_ms_per_frame"], "config": cfg,
            "lbvh_build_ms": {"blas_10M_tris": st.blas_ms, "tlas": st.tlas_ms, "achieved_gbs": 280 * st.primitive_count / max(st.blas_ms, 1e-6) / 1e6}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c2", choices=["c2", "c3", "c4", "c5"], help="c2 = the bench line (BASELINE configs[1]); c3 (animation, frames sharded) / c4 / c5 = the other GPU configs, extra lines")
    ap.add_argument("--frames", type=int, default=64, help="--config c3: frames of the synthesised particle animation")
    ap.add_argument("--renderers-per-gpu", type=int, default=0, help="--config c3, e2e pass: renderer contexts (lanes of rt_run_vtk_animation) per GPU; 0 = min(3, host cores / (2 x ranks))")
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--depth", type=int, default=10)
    ap.add_argument("--nx", type=int, default=1000)
    ap.add_argument("--nz", type=int, default=500)
    ap.add_argument("--tile-rows", type=int, default=8)
    ap.add_argument("--emulate-shard", type=int, default=1, help="tuning aid: render only rank 0's tiles of an N-way split on one GPU")
    ap.add_argument("--frames-per-window", type=int, default=8, help="frames in flight per launch of the path kernel (1..8)")
    ap.add_argument("--sync", action="store_true", help="comparison mode: rt_render_frame one frame at a time (at N > 1 with an NCCL gather) instead of the frame loop")
    ap.add_argument("--no-c4", action="store_true", help="skip the config-4 sub-line")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true", help="skip the reference-CUDA-build baseline")
    ap.add_argument("--no-c3", action="store_true", help="skip the config-3 particle rebuild measurement")
    ap.add_argument("--megakernel", action="store_true", help="comparison mode: one thread per pixel (the reference's kernel structure)")
    ap.add_argument("--wavefront", action="store_true", help="comparison mode: wave loop (shade + trace launch per bounce) instead of the persistent path kernel")
    ap.add_argument("--sah", action="store_true", help="optional SAH refinement of the LBVH (tree rotations in the refit)")
    ap.add_argument("--bvh2", action="store_true", help="comparison mode: traverse the binary LBVH instead of the wide BVH")
    args = ap.parse_args()
    if args.config != "c2":
        args.no_ref_cuda = args.no_c3 = args.no_cpu_baseline = args.no_c4 = True
    if args.impl == "reference":
        run_reference(args)
    elif args.config == "c3":
        run_c3(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
