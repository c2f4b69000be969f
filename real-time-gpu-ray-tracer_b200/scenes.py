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
    dielectrics: np.ndarray = field(default_factory=lambda: np.zeros(0, abi.dielectric_dt))   # extension, parity unpinned
    ext_records: np.ndarray | None = None      # records of a build's extension primitive type (csrc/rt_prims.h); tests only
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


def c1_mixed(width: int = 160, height: int = 90, spp: int = 4, depth: int = 8, seed: int = 3) -> Scene:
    """All three materials on all three primitive types (the dielectric is an EXTENSION with no reference counterpart —
    parity unpinned): C1b's layout with a third of the spheres made of glass (index of refraction 1.5, one of 1 / 1.5 = an
    air bubble), a glass parallelogram pane, a glass triangle and a rotated, non-uniformly scaled glass sphere instance."""
    rs = np.random.RandomState(seed)
    sc = Scene(name="C1-mixed-materials", width=width, height=height, sample_count=spp, ray_trace_depth=depth,
               center=(0.0, 3.0, 9.0), target=(0.0, 0.5, 0.0), fov=70.0)
    sc.parallelograms = np.zeros(2, abi.parallelogram_dt)
    sc.parallelograms[0] = ((-50, 0, -50), (100, 0, 0), (0, 0, 100), abi.ROUGH, 0)
    sc.parallelograms[1] = ((-2.0, 0.2, 3.0), (4.0, 0.0, 0.0), (0.0, 2.5, -0.4), abi.DIELECTRIC, 1)
    sc.triangles = np.zeros(1, abi.triangle_dt)
    sc.triangles[0]["v"] = [(-4.0, 0.3, 1.0), (-2.5, 0.3, 2.0), (-3.2, 2.4, 1.2)]
    sc.triangles[0]["material_type"] = abi.DIELECTRIC
    sc.triangles[0]["material_index"] = 0
    roughs, metals = [(0.5, 0.5, 0.5)], []
    diel = [((1.0, 1.0, 1.0), 1.5), ((0.9, 0.95, 1.0), 1.33), ((1.0, 1.0, 1.0), 1.0 / 1.5)]
    n = 48
    sc.spheres = np.zeros(n + 1, abi.sphere_dt)
    for k in range(n):
        gx, gz = k % 8, k // 8
        r = f32(rs.uniform(0.25, 0.55))
        cx = f32((gx - 3.5) * 1.7 + rs.uniform(-0.3, 0.3))
        cz = f32((gz - 3) * 1.7 - 1.0 + rs.uniform(-0.3, 0.3))
        u = rs.uniform()
        if u < 0.34:
            roughs.append(tuple(rs.uniform(0, 1, 3)))
            sc.spheres[k] = ((cx, r, cz), r, abi.ROUGH, len(roughs) - 1)
        elif u < 0.67:
            metals.append((tuple(rs.uniform(0.5, 1, 3)), rs.uniform(0, 0.3)))
            sc.spheres[k] = ((cx, r, cz), r, abi.METAL, len(metals) - 1)
        else:
            sc.spheres[k] = ((cx, r, cz), r, abi.DIELECTRIC, int(rs.randint(0, 3)))
    sc.spheres[n] = ((0.0, 0.0, 0.0), 1.0, abi.DIELECTRIC, 0)
    sc.roughs = np.zeros(len(roughs), abi.rough_dt)
    sc.roughs["albedo"] = roughs
    sc.metals = np.zeros(max(len(metals), 1), abi.metal_dt)
    for i, (a, fz) in enumerate(metals):
        sc.metals[i] = (a, fz)
    sc.dielectrics = np.zeros(len(diel), abi.dielectric_dt)
    for i, (a, ior) in enumerate(diel):
        sc.dielectrics[i] = (a, ior)
    sc.instances = single_instances([(abi.PARALLELOGRAM, 0), (abi.PARALLELOGRAM, 1), (abi.TRIANGLE, 0)] + [(abi.SPHERE, i) for i in range(n + 1)])
    sc.xforms = identity_xforms(len(sc.instances))
    sc.xforms[-1]["shift"] = (3.0, 1.6, 3.5)
    sc.xforms[-1]["rotate_deg"] = (20.0, 35.0, 10.0)
    sc.xforms[-1]["scale"] = (1.3, 0.8, 1.0)
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


def c2_dielectric(**kw) -> Scene:
    """The EXTENDED config 2 of SURVEY 8d: the same height-field with the metal third of the quads made of glass
    (index of refraction 1.5) instead — 'diffuse/metal/dielectric' of BASELINE.json configs[1].  The dielectric is an
    extension with no reference counterpart: parity unpinned, reported separately from the headline line."""
    sc = c2(**kw)
    glass = sc.triangles["material_type"] == abi.METAL
    sc.triangles["material_type"][glass] = abi.DIELECTRIC
    sc.dielectrics = np.zeros(1, abi.dielectric_dt)
    sc.dielectrics[0] = ((1.0, 1.0, 1.0), 1.5)
    sc.name = sc.name.replace("C2", "C2-dielectric")
    return sc


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


# ---- config 3: particle_mesh.vtk.series ---------------------------------------------------------------------------
def read_vtk_frame(path: str):
    """numpy restatement of the legacy-binary POLYDATA layout: returns (points float64 [N,3],
    list of strips as int arrays).  Independent of csrc/vtk_reader.cpp so that each checks the other."""
    raw = open(path, "rb").read()
    pos = raw.index(b"POINTS")
    eol = raw.index(b"\n", pos)
    _, n, typ = raw[pos:eol].split()
    n = int(n)
    dt = ">f8" if typ == b"double" else ">f4"
    pts = np.frombuffer(raw, dt, n * 3, eol + 1).astype(np.float64).reshape(n, 3)
    pos = raw.index(b"TRIANGLE_STRIPS", eol + 1 + n * 3 * np.dtype(dt).itemsize)
    eol = raw.index(b"\n", pos)
    _, cells, size = raw[pos:eol].split()
    conn = np.frombuffer(raw, ">i4", int(size), eol + 1).astype(np.int64)
    strips, k = [], 0
    for _ in range(int(cells)):
        ln = int(conn[k])
        strips.append(conn[k + 1:k + 1 + ln])
        k += 1 + ln
    return pts, strips


def write_vtk_frame(path: str, pts: np.ndarray, strips) -> None:
    """Writes a frame in the same on-disk layout (used to synthesise larger frames from a committed one)."""
    p = len(strips)
    size = sum(len(s) + 1 for s in strips)
    with open(path, "wb") as f:
        f.write(b"# vtk DataFile Version 2.0\nGenerated by rtb200 tests\nBINARY \nDATASET POLYDATA\n")
        f.write(b"POINTS %d double\n" % len(pts))
        f.write(np.ascontiguousarray(pts, ">f8").tobytes())
        f.write(b"\nTRIANGLE_STRIPS %d %d\n" % (p, size))
        for s in strips:
            f.write(np.concatenate([[len(s)], s]).astype(">i4").tobytes())
        f.write(b"\nCELL_DATA %d\nSCALARS id int 1\nLOOKUP_TABLE default\n" % p)
        f.write(np.arange(p).astype(">i4").tobytes())
        f.write(b"\nVECTORS vel double\n")
        f.write(np.zeros((p, 3), ">f8").tobytes())
        f.write(b"\n")


def vtk_particles(pts: np.ndarray, strips):
    """VTKReader::convertToRendererData (reference src/Global/VTKReader.cu:167-220) + the restated point normals
    (vtkPolyDataNormals stand-in, parity unpinned): triangles, instance descs, transforms."""
    nrm = np.zeros_like(pts)
    for s in strips:
        c = pts[np.unique(s)].mean(axis=0)
        a, b, d = s[:-2], s[1:-1], s[2:]
        ok = (a != b) & (b != d) & (a != d)
        a, b, d = a[ok], b[ok], d[ok]
        n = np.cross(pts[b] - pts[a], pts[d] - pts[a])
        ln = np.sqrt((n * n).sum(1))
        keep = ln > 0
        a, b, d, n = a[keep], b[keep], d[keep], n[keep] / ln[keep, None]
        out = (n * ((pts[a] + pts[b] + pts[d]) / 3.0 - c)).sum(1)
        n = np.where(out[:, None] < 0, -n, n)
        for ids in (a, b, d):
            np.add.at(nrm, ids, n)
    ln = np.sqrt((nrm * nrm).sum(1))
    nrm = np.where(ln[:, None] > 0, nrm / np.maximum(ln, 1e-300)[:, None], nrm)
    tris, ins = [], np.zeros(len(strips), abi.instance_desc_dt)
    first = 0
    for k, s in enumerate(strips):
        m = len(s) - 2
        j = np.arange(m)
        i0, i1, i2 = s[j], np.where(j & 1, s[j + 2], s[j + 1]), np.where(j & 1, s[j + 1], s[j + 2])
        t = np.zeros(m, abi.triangle_dt)
        t["v"] = np.stack([pts[i0], pts[i1], pts[i2]], 1).astype(f32)
        t["n"] = np.stack([nrm[i0], nrm[i1], nrm[i2]], 1).astype(f32)
        t["material_type"], t["material_index"], t["has_normals"] = abi.METAL, 0, 1
        tris.append(t)
        ins[k]["type"], ins[k]["has_bounds"], ins[k]["first_prim"], ins[k]["prim_count"] = abi.TRIANGLE, 1, first, m
        ins[k]["bounds_min"], ins[k]["bounds_max"] = pts[s].min(0).astype(f32), pts[s].max(0).astype(f32)
        ins[k]["centroid"] = (pts[s].sum(0) / float(len(s))).astype(f32)
        first += m
    xf = identity_xforms(len(strips))
    xf["shift"], xf["rotate_deg"], xf["scale"] = (0.0, 4.0, 0.0), (90.0, 0.0, 0.0), 3.0
    return np.concatenate(tris), ins, xf


def c3(vtk_path: str, frame: int = 0, width: int = 1920, height: int = 1080, spp: int = 1, depth: int = 10) -> "Scene":
    """Config 3, one frame: the reference demo scene (Main.cu:48-106) + the particles of one .vtk file; VTK triangles
    sit at the head of the triangle array and VTK instances after the user's (Renderer.cu:17,106-116)."""
    sc = c1a(frame=frame, width=width, height=height, spp=spp, depth=depth)
    tris, ins, xf = vtk_particles(*read_vtk_frame(vtk_path))
    user = sc.instances.copy()
    user["first_prim"] = np.where(user["type"] == abi.TRIANGLE, user["first_prim"] + len(tris), user["first_prim"])
    sc.triangles = np.concatenate([tris, sc.triangles])
    sc.instances = np.concatenate([user, ins])
    sc.xforms = np.concatenate([sc.xforms, xf])
    sc.name = f"C3-frame{frame}"
    return sc


def synth_vtk_series(src_vtk: str, out_dir: str, particle_counts, seed: int = 3, frame_ids=None) -> str:
    """Synthesises a .vtk.series whose frame k holds particle_counts[k] rigid copies of the particles of `src_vtk`
    (the reference's own data is 1.2 GB and cannot travel to the GPU box; frame 0 of it is committed as a fixture).
    With `frame_ids` (frame sharding: a rank writes only its own frames) entry k is animation frame frame_ids[k]; its
    content then depends on (seed, frame id) alone, so every rank's copy of a frame is the same file."""
    import os
    pts, strips = read_vtk_frame(src_vtk)
    rng = np.random.RandomState(seed)
    os.makedirs(os.path.join(out_dir, "particle_mesh"), exist_ok=True)
    names = []
    for k, count in enumerate(particle_counts):
        if frame_ids is not None:
            rng = np.random.RandomState((seed * 1000003 + int(frame_ids[k])) & 0x7FFFFFFF)
        P, S, base = [], [], 0
        for q in range(count):
            s = strips[q % len(strips)]
            ids, inv = np.unique(s, return_inverse=True)
            shift = np.zeros(3) if q < len(strips) else rng.uniform(-1.5, 1.5, 3) * np.array([1.0, 1.0, 0.3])
            P.append(pts[ids] + shift)
            S.append(inv + base)
            base += len(ids)
        fid = k if frame_ids is None else int(frame_ids[k])
        name = f"particle_mesh/particle_{fid * 100:015d}.vtk"
        write_vtk_frame(os.path.join(out_dir, name), np.concatenate(P), S)
        names.append(name)
    series = os.path.join(out_dir, "particle_mesh.vtk.series")
    with open(series, "w") as f:
        f.write('{\n  "file-series-version": "1.0",\n  "files": [\n')
        f.write(",\n".join('    {"name": "%s", "time": %e}' % (n, 0.01 * (k if frame_ids is None else int(frame_ids[k]))) for k, n in enumerate(names)))
        f.write("\n  ]\n}\n")
    return series


def c3_particle_counts(n_frames: int, first: int = 25, last: int = 500):
    """Particle count of each frame of an n_frames-long stand-in for files/particle_mesh.vtk.series (the reference's
    401 frames grow from 25 particles at t = 0 to 500 at t = 4.0)."""
    if n_frames == 1:
        return [last]
    return [int(round(first + (last - first) * i / (n_frames - 1))) for i in range(n_frames)]
