"""Scene builders for the BASELINE.json configs, as numpy arrays in the C-ABI layouts.

C1a  the reference's own demo scene, src/Global/Main.cu:48-106 + updateInstance (Main.cu:6-42)
C1b  ~100 analytic primitives (ground quad + 99 spheres)
C2   exactly 1 000 000-triangle displaced height-field (C4: 10 M, same generator)
C5   TLAS stress: many instances of a few shared triangle meshes
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import abi

f32 = np.float32


@dataclass
class Scene:
    spheres: np.ndarray = field(default_factory=lambda: np.zeros(0, abi.sphere_dt))
    parallelograms: np.ndarray = field(default_factory=lambda: np.zeros(0, abi.parallelogram_dt))
    triangles: np.ndarray = field(default_factory=lambda: np.zeros(0, abi.triangle_dt))
    roughs: np.ndarray = field(default_factory=lambda: np.zeros(0, abi.rough_dt))
    metals: np.ndarray = field(default_factory=lambda: np.zeros(0, abi.metal_dt))
    instances: np.ndarray = field(default_factory=lambda: np.zeros(0, abi.instance_desc_dt))
    xforms: np.ndarray = field(default_factory=lambda: np.zeros(0, abi.instance_xform_dt))
    width: int = 640
    height: int = 360
    background: tuple = (0.7, 0.8, 0.9)
    center: tuple = (0.0, 2.0, 10.0)
    target: tuple = (0.0, 2.0, 0.0)
    fov: float = 90.0
    up: tuple = (0.0, 1.0, 0.0)
    focus_disk_radius: float = 0.0
    sample_range: float = 0.5
    sample_count: int = 1
    ray_trace_depth: int = 4
    name: str = "scene"

    def camera_input(self) -> abi.CameraInput:
        return abi.CameraInput(abi.Vec3.of(self.background), abi.Vec3.of(self.center), abi.Vec3.of(self.target), self.fov,
                               abi.Vec3.of(self.up), self.focus_disk_radius, self.sample_range, self.sample_count, self.ray_trace_depth)

    def window(self) -> abi.Window:
        return abi.Window(self.width, self.height)


def identity_xforms(n: int) -> np.ndarray:
    x = np.zeros(n, abi.instance_xform_dt)
    x["scale"] = 1.0
    return x


def single_instances(kinds) -> np.ndarray:
    """kinds: list of (primitive_type, index) — the reference's insMapArray (Main.cu:93-99)."""
    ins = np.zeros(len(kinds), abi.instance_desc_dt)
    for k, (t, i) in enumerate(kinds):
        ins[k]["type"] = t
        ins[k]["first_prim"] = i
        ins[k]["prim_count"] = 1
    return ins


def main_cu_xforms(frame: int) -> np.ndarray:
    """updateInstance of the reference demo (src/Global/Main.cu:6-42), float32 arithmetic."""
    radius, speed = f32(2.0), f32(0.02)
    angle = f32(frame) * speed
    c, s = f32(np.cos(angle)), f32(np.sin(angle))
    nc = np.array([f32(0.0) + radius * c * f32(1.5), f32(2.0) + radius * s * c, f32(0.0) + radius * s * f32(1.5)], f32)
    x = identity_xforms(5)
    x[0]["shift"] = (0.0, -1000.0, 0.0)
    x[1]["shift"] = nc
    x[2]["shift"] = (-5.0, 0.0, 0.0)
    x[3]["shift"] = (-nc[0], nc[1], -nc[2])
    x[3]["rotate_deg"] = f32(frame) * f32(0.4)
    x[3]["scale"] = 3.0
    x[4]["shift"] = (-nc[0], nc[1] + f32(5.0), nc[2])
    return x


def c1a(frame: int = 0, width: int = 640, height: int = 360, spp: int = 1, depth: int = 4) -> Scene:
    """The reference's demo scene without VTK particles (src/Global/Main.cu:48-106)."""
    sc = Scene(name="C1a", width=width, height=height, sample_count=spp, ray_trace_depth=depth)
    sc.spheres = np.zeros(2, abi.sphere_dt)
    sc.spheres[0] = ((0, 0, 0), 1000.0, abi.ROUGH, 3)
    sc.spheres[1] = ((0, 0, 0), 2.0, abi.ROUGH, 0)
    sc.parallelograms = np.zeros(1, abi.parallelogram_dt)
    sc.parallelograms[0] = ((0, 0, 0), (1, 0, 1), (0, 4, 0), abi.ROUGH, 1)
    sc.triangles = np.zeros(1, abi.triangle_dt)
    sc.triangles[0]["v"] = [(0, 0, 0), (1, 0, 1), (0, 1, 0)]
    sc.triangles[0]["material_type"] = abi.ROUGH
    sc.triangles[0]["material_index"] = 2
    sc.roughs = np.zeros(4, abi.rough_dt)
    sc.roughs["albedo"] = [(.65, .05, .05), (.73, .73, .73), (.12, .45, .15), (.70, .60, .50)]
    sc.metals = np.zeros(1, abi.metal_dt)
    sc.metals[0] = ((0.8, 0.85, 0.88), 0.0)
    sc.instances = single_instances([(abi.SPHERE, 0), (abi.SPHERE, 1), (abi.PARALLELOGRAM, 0), (abi.TRIANGLE, 0), (abi.SPHERE, 1)])
    sc.xforms = main_cu_xforms(frame)
    return sc


def c1b(width: int = 640, height: int = 360, spp: int = 1, depth: int = 4, seed: int = 1) -> Scene:
    """Ground parallelogram + 99 spheres on a jittered 11x9 grid."""
    rs = np.random.RandomState(seed)
    sc = Scene(name="C1b", width=width, height=height, sample_count=spp, ray_trace_depth=depth)
    sc.parallelograms = np.zeros(1, abi.parallelogram_dt)
    sc.parallelograms[0] = ((-50, 0, -50), (100, 0, 0), (0, 0, 100), abi.ROUGH, 0)
    roughs = [(0.5, 0.5, 0.5)]
    metals = []
    sc.spheres = np.zeros(99, abi.sphere_dt)
    k = 0
    for gz in range(9):
        for gx in range(11):
            r = f32(rs.uniform(0.2, 0.5))
            cx = f32((gx - 5) * 1.6 + rs.uniform(-0.3, 0.3))
            cz = f32((gz - 4) * 1.6 - 2.0 + rs.uniform(-0.3, 0.3))
            if rs.uniform() < 0.7:
                roughs.append(tuple(rs.uniform(0, 1, 3)))
                sc.spheres[k] = ((cx, r, cz), r, abi.ROUGH, len(roughs) - 1)
            else:
                metals.append((tuple(rs.uniform(0.5, 1, 3)), rs.uniform(0, 0.3)))
                sc.spheres[k] = ((cx, r, cz), r, abi.METAL, len(metals) - 1)
            k += 1
    sc.roughs = np.zeros(len(roughs), abi.rough_dt)
    sc.roughs["albedo"] = roughs
    sc.metals = np.zeros(max(len(metals), 1), abi.metal_dt)
    for i, (a, fz) in enumerate(metals):
        sc.metals[i] = (a, fz)
    sc.instances = single_instances([(abi.PARALLELOGRAM, 0)] + [(abi.SPHERE, i) for i in range(99)])
    sc.xforms = identity_xforms(100)
    return sc


def heightfield_triangles(nx: int, nz: int, x_range, z_range, seed: int, material_tile: int = 25) -> np.ndarray:
    """Displaced height-field of nx*nz quads (2 triangles each), SURVEY §8d C2.

    y = 0.6 sin(0.9x) cos(1.3z) + 0.15 sin(5.1x + 2.3z) + 0.02 U(-1,1); vertices row-major (z rows, x columns);
    quad (i, j) split along (i,j)->(i+1,j+1); triangle order row-major, two per quad; flat normals
    (has_normals = 0 -> Triangle's 3-vertex constructor); material by quad (i/tile + j/tile) mod 3 ->
    rough#0 / rough#1 / metal#0.
    """
    xs = np.linspace(x_range[0], x_range[1], nx + 1, dtype=np.float64)
    zs = np.linspace(z_range[0], z_range[1], nz + 1, dtype=np.float64)
    X, Z = np.meshgrid(xs, zs)  # shape (nz+1, nx+1): row j (z), column i (x)
    noise = np.random.RandomState(seed).uniform(-1.0, 1.0, size=X.shape)
    Y = 0.6 * np.sin(0.9 * X) * np.cos(1.3 * Z) + 0.15 * np.sin(5.1 * X + 2.3 * Z) + 0.02 * noise
    P = np.stack([X, Y, Z], axis=-1).astype(f32)  # (nz+1, nx+1, 3)
    p00 = P[:-1, :-1]
    p10 = P[:-1, 1:]   # (i+1, j)
    p01 = P[1:, :-1]   # (i, j+1)
    p11 = P[1:, 1:]
    tris = np.zeros((nz, nx, 2), abi.triangle_dt)
    # wind so that e1 x e2 points to +y
    tris["v"][:, :, 0, 0] = p00
    tris["v"][:, :, 0, 1] = p11
    tris["v"][:, :, 0, 2] = p10
    tris["v"][:, :, 1, 0] = p00
    tris["v"][:, :, 1, 1] = p01
    tris["v"][:, :, 1, 2] = p11
    ii, jj = np.meshgrid(np.arange(nx), np.arange(nz))
    m = ((ii // material_tile) + (jj // material_tile)) % 3
    mtype = np.where(m == 2, abi.METAL, abi.ROUGH).astype(np.uint32)
    mindex = np.where(m == 2, 0, m).astype(np.uint32)
    tris["material_type"] = mtype[:, :, None]
    tris["material_index"] = mindex[:, :, None]
    return np.ascontiguousarray(tris.reshape(-1))


def _mesh_scene(tris: np.ndarray, name: str, width, height, spp, depth, center, target, fov) -> Scene:
    sc = Scene(name=name, width=width, height=height, sample_count=spp, ray_trace_depth=depth, center=center, target=target, fov=fov)
    sc.triangles = tris
    sc.roughs = np.zeros(2, abi.rough_dt)
    sc.roughs["albedo"] = [(.73, .73, .73), (.65, .05, .05)]
    sc.metals = np.zeros(1, abi.metal_dt)
    sc.metals[0] = ((0.8, 0.85, 0.88), 0.1)
    ins = np.zeros(1, abi.instance_desc_dt)
    ins[0]["type"] = abi.TRIANGLE
    ins[0]["first_prim"] = 0
    ins[0]["prim_count"] = len(tris)
    sc.instances = ins
    sc.xforms = identity_xforms(1)
    return sc


def c2(width: int = 1920, height: int = 1080, spp: int = 64, depth: int = 10, nx: int = 1000, nz: int = 500) -> Scene:
    """1 M-triangle height-field, camera (0,6,9)->(0,0,0), fov 60.  nx/nz shrink it for tests."""
    tris = heightfield_triangles(nx, nz, (-10.0, 10.0), (-5.0, 5.0), seed=2)
    return _mesh_scene(tris, f"C2[{len(tris)} tris]", width, height, spp, depth, (0.0, 6.0, 9.0), (0.0, 0.0, 0.0), 60.0)


def c4(width: int = 3840, height: int = 2160, spp: int = 256, depth: int = 10, nx: int = 5000, nz: int = 1000) -> Scene:
    """10 M-triangle height-field, camera (0,12,18)."""
    tris = heightfield_triangles(nx, nz, (-50.0, 50.0), (-10.0, 10.0), seed=4)
    return _mesh_scene(tris, f"C4[{len(tris)} tris]", width, height, spp, depth, (0.0, 12.0, 18.0), (0.0, 0.0, 0.0), 60.0)


def _icosphere(subdiv: int) -> tuple[np.ndarray, np.ndarray]:
    t = (1.0 + 5 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t), (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    v = [np.array(p, np.float64) / np.linalg.norm(p) for p in v]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6), (7, 1, 8),
         (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    for _ in range(subdiv):
        cache, nf = {}, []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = v[a] + v[b]
                v.append(m / np.linalg.norm(m))
                cache[key] = len(v) - 1
            return cache[key]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    return np.array(v), np.array(f)


def c5(n_instances: int = 10000, n_meshes: int = 16, width: int = 1920, height: int = 1080, spp: int = 16, depth: int = 8,
       subdiv: int = 3, seed: int = 5) -> Scene:
    """TLAS stress: n_instances instances of n_meshes shared ~1280-triangle blobs, all rough."""
    rs = np.random.RandomState(seed)
    V, F = _icosphere(subdiv)
    tris_all = []
    for m in range(n_meshes):
        k = rs.uniform(1.0, 4.0, 3)
        ph = rs.uniform(0, 6.28, 3)
        rad = 1.0 + 0.25 * np.sin(k[0] * V[:, 0] + ph[0]) * np.sin(k[1] * V[:, 1] + ph[1]) * np.sin(k[2] * V[:, 2] + ph[2])
        Pm = (V * rad[:, None]).astype(f32)
        nrm = Pm / np.linalg.norm(Pm, axis=1, keepdims=True)
        t = np.zeros(len(F), abi.triangle_dt)
        t["v"] = Pm[F]
        t["n"] = nrm[F].astype(f32)
        t["has_normals"] = 1
        t["material_type"] = abi.ROUGH
        t["material_index"] = m % 8
        tris_all.append(t)
    per = len(F)
    sc = Scene(name=f"C5[{n_instances} inst x {n_meshes} meshes]", width=width, height=height, sample_count=spp, ray_trace_depth=depth,
               center=(0.0, 30.0, 120.0), target=(0.0, 0.0, 0.0), fov=60.0)
    sc.triangles = np.ascontiguousarray(np.concatenate(tris_all))
    sc.roughs = np.zeros(8, abi.rough_dt)
    sc.roughs["albedo"] = rs.uniform(0.2, 0.9, (8, 3)).astype(f32)
    sc.metals = np.zeros(1, abi.metal_dt)
    ins = np.zeros(n_instances, abi.instance_desc_dt)
    mesh_of = rs.randint(0, n_meshes, n_instances)
    ins["type"] = abi.TRIANGLE
    ins["first_prim"] = mesh_of * per
    ins["prim_count"] = per
    sc.instances = ins
    x = np.zeros(n_instances, abi.instance_xform_dt)
    x["shift"] = rs.uniform(-50, 50, (n_instances, 3)).astype(f32)
    x["rotate_deg"] = rs.uniform(0, 360, (n_instances, 3)).astype(f32)
    x["scale"] = rs.uniform(0.5, 2.0, (n_instances, 1)).astype(f32)
    sc.xforms = x
    return sc
