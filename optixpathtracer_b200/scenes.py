"""Deterministic procedural scenes for the BASELINE.json workloads (the reference ships no assets: every
OBJ/HDR path in sv4/main.cpp:188-202 is an absolute Windows path).  SURVEY.md §8d table:

  C1 cornell()            ~32 triangles, authored as OBJ+MTL text and read back through loadOBJ
  C2 sponza_class()       ~262 k triangles: atrium with colonnades, arches, cloth height fields, floor tiles
  C3 lost_empire_class()  ~225 k triangles: voxel terrain of addBox-style boxes (12 triangles each)
  C5 city_class()         N-million triangles: grid of atrium-like cells, flattened (no instancing in the reference)

All generators use numpy's PCG64 with a fixed seed and emit float32 geometry through Model/TriangleMesh.
"""
import math
import os
import tempfile

import numpy as np

from .model import Model, TriangleMesh, Texture, loadOBJ
from .structs import Material

# ------------------------------------------------------------------------------------------ materials
def material_mix(rng, kind=None):
    """70 % default-constructed sv4 Material with random colour, 20 % diffuse preset, 10 % glass."""
    r = rng.random() if kind is None else {"default": 0.0, "diffuse": 0.75, "glass": 0.95}[kind]
    col = tuple(float(c) for c in rng.uniform(0.2, 0.9, 3))
    if r < 0.7:
        return Material(color=col, emission=(0, 0, 0))
    if r < 0.9:
        return Material.orig_defaults(color=col)
    return Material(color=col, emission=(0, 0, 0), transmission=1.0, metallic=0.0, roughness=0.05, eta=1.5)


def checker_texture(rng, n=512):
    """512^2 RGBA8 procedural texture (two-colour checker with per-texel noise)."""
    a, b = rng.integers(40, 255, 3), rng.integers(40, 255, 3)
    yy, xx = np.mgrid[0:n, 0:n]
    mask = (((xx // 32) + (yy // 32)) & 1).astype(bool)
    rgb = np.where(mask[..., None], a[None, None, :], b[None, None, :]).astype(np.int32)
    rgb = np.clip(rgb + rng.integers(-12, 13, (n, n, 1)), 0, 255).astype(np.uint32)
    px = rgb[..., 0] | (rgb[..., 1] << 8) | (rgb[..., 2] << 16) | np.uint32(255 << 24)
    return Texture(np.ascontiguousarray(px, np.uint32), (n, n))


def _mesh(vertex, index, material, texcoord=None, tex_id=-1):
    m = TriangleMesh()
    m.vertex = np.ascontiguousarray(vertex, np.float32).reshape(-1, 3)
    m.index = np.ascontiguousarray(index, np.uint32).reshape(-1, 3)
    m.material = material
    if texcoord is not None:
        m.texcoord = np.ascontiguousarray(texcoord, np.float32).reshape(-1, 2)
    m.diffuseTextureID = tex_id
    return m


# ------------------------------------------------------------------------------------------ C1 Cornell
_CORNELL_MTL = """newmtl white
Kd 0.73 0.73 0.73
Ke 0 0 0
newmtl red
Kd 0.65 0.05 0.05
Ke 0 0 0
newmtl green
Kd 0.12 0.45 0.15
Ke 0 0 0
newmtl light
Kd 0.78 0.78 0.78
Ke 12 12 12
newmtl boxes
Kd 0.6 0.6 0.7
Ke 0 0 0
"""


def _quad(v, a, b, c, d):
    return [f"f {a} {b} {c}", f"f {a} {c} {d}"]


def cornell_obj_text():
    """Cornell box, inward-facing walls, open towards +z.  5 walls x2 + light x2 + two 5-face boxes x10 = 32."""
    V, F = [], []

    def add(p):
        V.append(p)
        return len(V)

    def quad(mat, p0, p1, p2, p3):
        i = [add(p) for p in (p0, p1, p2, p3)]
        F.append((mat, f"f {i[0]} {i[1]} {i[2]}"))
        F.append((mat, f"f {i[0]} {i[2]} {i[3]}"))

    quad("white", (-1, 0, 1), (1, 0, 1), (1, 0, -1), (-1, 0, -1))            # floor, normal +y
    quad("white", (-1, 2, -1), (1, 2, -1), (1, 2, 1), (-1, 2, 1))            # ceiling, normal -y
    quad("white", (-1, 0, -1), (1, 0, -1), (1, 2, -1), (-1, 2, -1))          # back, normal +z
    quad("red", (-1, 0, 1), (-1, 0, -1), (-1, 2, -1), (-1, 2, 1))            # left, normal +x
    quad("green", (1, 0, -1), (1, 0, 1), (1, 2, 1), (1, 2, -1))              # right, normal -x
    quad("light", (-0.3, 1.98, -0.3), (0.3, 1.98, -0.3), (0.3, 1.98, 0.3), (-0.3, 1.98, 0.3))   # light, normal -y

    def box(cx, cz, half, height, angle):
        ca, sa = math.cos(angle), math.sin(angle)
        c = [(cx + ca * x - sa * z, cz + sa * x + ca * z) for x, z in ((-half, half), (half, half), (half, -half), (-half, -half))]
        lo = [(x, 0.0, z) for x, z in c]
        hi = [(x, height, z) for x, z in c]
        for k in range(4):                                                   # four sides, outward normals
            n = (k + 1) % 4
            quad("boxes", lo[k], lo[n], hi[n], hi[k])
        quad("boxes", hi[0], hi[1], hi[2], hi[3])                            # top, normal +y

    box(0.35, 0.35, 0.3, 0.6, 0.3)
    box(-0.35, -0.3, 0.3, 1.2, -0.35)
    lines = ["mtllib cornell.mtl", "o cornell"]
    lines += [f"v {p[0]:.6f} {p[1]:.6f} {p[2]:.6f}" for p in V]
    cur = None
    for mat, f in F:
        if mat != cur:
            lines.append(f"usemtl {mat}")
            cur = mat
        lines.append(f)
    return "\n".join(lines) + "\n", _CORNELL_MTL


def cornell(variant="disney", directory=None):
    """C1.  variant: 'disney' = Material() defaults as loadOBJ leaves them (40 % transmissive, half metallic);
    'diffuse' = the orig/sv3 defaults (transmission 0, metallic 0, specular 0.5)."""
    directory = directory or tempfile.mkdtemp(prefix="b200pt_cornell_")
    obj, mtl = cornell_obj_text()
    with open(os.path.join(directory, "cornell.obj"), "w") as fh:
        fh.write(obj)
    with open(os.path.join(directory, "cornell.mtl"), "w") as fh:
        fh.write(mtl)
    model = loadOBJ(directory.rstrip("/") + "/cornell.obj")
    if variant == "diffuse":
        for m in model.meshes:
            c, e = m.material.color.tolist(), m.material.emission.tolist()
            m.material = Material.orig_defaults(color=tuple(c), emission=tuple(e))
    return model


# ------------------------------------------------------------------------------------------ primitives
def _grid(nu, nv):
    """indices of a (nu x nv)-cell grid over (nu+1)*(nv+1) vertices, two triangles per cell."""
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    a = (i * (nv + 1) + j).ravel()
    b, c, d = a + (nv + 1), a + (nv + 1) + 1, a + 1
    return np.stack([np.stack([a, b, c], 1), np.stack([a, c, d], 1)], 1).reshape(-1, 3)


def _cylinder(cx, cz, y0, y1, radius, seg, rings):
    th = np.linspace(0, 2 * np.pi, seg + 1)
    ys = np.linspace(y0, y1, rings + 1)
    T, Y = np.meshgrid(th, ys, indexing="ij")
    v = np.stack([cx + radius * np.cos(T), Y, cz + radius * np.sin(T)], -1).reshape(-1, 3)
    return v, _grid(seg, rings)[:, ::-1]                                     # outward winding


def _height_field(x0, x1, z0, z1, nu, nv, fn):
    xs, zs = np.linspace(x0, x1, nu + 1), np.linspace(z0, z1, nv + 1)
    X, Z = np.meshgrid(xs, zs, indexing="ij")
    v = np.stack([X, fn(X, Z), Z], -1).reshape(-1, 3)
    uv = np.stack(np.meshgrid(np.linspace(0, 4, nu + 1), np.linspace(0, 4, nv + 1), indexing="ij"), -1).reshape(-1, 2)
    return v, _grid(nu, nv)[:, ::-1], uv                                     # normal +y


_BOX_TRIS = np.array([[0, 1, 2], [0, 2, 3], [4, 7, 6], [4, 6, 5], [4, 0, 3], [4, 3, 7],
                      [1, 5, 6], [1, 6, 2], [3, 2, 6], [3, 6, 7], [4, 5, 1], [4, 1, 0]], np.uint32)


def _boxes(centers, half):
    """Vectorised addBox: 8 shared corners + 12 outward triangles per box (corner order A..H of Model.cpp:220-229)."""
    centers = np.asarray(centers, np.float32).reshape(-1, 3)
    half = np.broadcast_to(np.asarray(half, np.float32), centers.shape)
    sg = np.array([[-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1], [-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1]], np.float32)
    v = centers[:, None, :] + sg[None, :, :] * half[:, None, :]
    idx = _BOX_TRIS[None, :, :] + (8 * np.arange(len(centers), dtype=np.uint32))[:, None, None]
    return v.reshape(-1, 3), idx.reshape(-1, 3)


def _merge(parts):
    vs, is_, off = [], [], 0
    for v, i in parts:
        vs.append(np.asarray(v, np.float32).reshape(-1, 3))
        is_.append(np.asarray(i, np.int64).reshape(-1, 3) + off)
        off += len(vs[-1])
    return np.concatenate(vs), np.concatenate(is_).astype(np.uint32)


# ------------------------------------------------------------------------------------------ C2 Sponza-class
def _atrium_parts(rng, x0, x1, z0, z1, height, detail):
    """One atrium cell: tiled floor, two colonnades with arches, a roof ring and hanging cloth."""
    parts = {"floor": [], "columns": [], "arches": [], "cloth": [], "walls": []}
    nt = max(2, int(24 * detail))
    v, i, uv = _height_field(x0, x1, z0, z1, nt * 2, nt, lambda X, Z: 0.0 * X)
    parts["floor"].append((v, i, uv))
    ncol = 8
    seg, rings = max(6, int(20 * detail)), max(2, int(24 * detail))
    zc = [z0 + (z1 - z0) * 0.25, z0 + (z1 - z0) * 0.75]
    xs = np.linspace(x0 + (x1 - x0) * 0.08, x1 - (x1 - x0) * 0.08, ncol)
    r = (x1 - x0) * 0.012
    for z in zc:
        for x in xs:
            parts["columns"].append(_cylinder(x, z, 0.0, height * 0.55, r, seg, rings))
            c = np.array([[x, height * 0.55 + r, z]], np.float32)
            parts["columns"].append(_boxes(c, np.array([r * 1.6, r, r * 1.6], np.float32)))
        for xa, xb in zip(xs[:-1], xs[1:]):                                  # arches: half tori between columns
            na, nb = max(6, int(18 * detail)), max(4, int(8 * detail))
            A, B = np.meshgrid(np.linspace(0, np.pi, na + 1), np.linspace(0, 2 * np.pi, nb + 1), indexing="ij")
            R, rr = (xb - xa) * 0.5, r * 0.8
            vx = (xa + xb) * 0.5 + (R + rr * np.cos(B)) * np.cos(A)
            vy = height * 0.55 + 2 * r + (R * 0.6 + rr * np.cos(B)) * np.sin(A)
            vz = z + rr * np.sin(B)
            parts["arches"].append((np.stack([vx, vy, vz], -1).reshape(-1, 3), _grid(na, nb)))
    nc = max(4, int(40 * detail))
    for k in range(4):                                                       # draped cloth between the colonnades
        xa = x0 + (x1 - x0) * (0.15 + 0.2 * k)
        ph = rng.uniform(0, 6.28)
        amp = height * 0.03
        v, i, uv = _height_field(xa, xa + (x1 - x0) * 0.12, zc[0], zc[1], nc, nc * 2,
                                 lambda X, Z, ph=ph, amp=amp: height * 0.7 - amp * 4 * np.sin((Z - zc[0]) / (zc[1] - zc[0]) * np.pi)
                                 + amp * np.sin(X * 0.05 + ph) * np.cos(Z * 0.07 + ph))
        parts["cloth"].append((v, i, uv))
    t = (x1 - x0) * 0.01                                                      # outer walls + roof ring (boxes)
    cx, cz = (x0 + x1) / 2, (z0 + z1) / 2
    walls = np.array([[cx, height / 2, z0], [cx, height / 2, z1], [x0, height / 2, cz], [x1, height / 2, cz],
                      [cx, height, z0 + (z1 - z0) * 0.12], [cx, height, z1 - (z1 - z0) * 0.12]], np.float32)
    halves = np.array([[(x1 - x0) / 2, height / 2, t], [(x1 - x0) / 2, height / 2, t], [t, height / 2, (z1 - z0) / 2],
                       [t, height / 2, (z1 - z0) / 2], [(x1 - x0) / 2, t, (z1 - z0) * 0.12], [(x1 - x0) / 2, t, (z1 - z0) * 0.12]], np.float32)
    parts["walls"].append(_boxes(walls, halves))
    return parts


def _parts_to_model(model, rng, parts, textured=("floor", "cloth")):
    for name, lst in parts.items():
        if not lst:
            continue
        has_uv = len(lst[0]) == 3
        v, i = _merge([(p[0], p[1]) for p in lst])
        uv = np.concatenate([np.asarray(p[2], np.float32).reshape(-1, 2) for p in lst]) if has_uv else None
        tex_id = -1
        if has_uv and name in textured and (name == "floor" or rng.random() < 0.5):
            model.textures.append(checker_texture(rng))
            tex_id = len(model.textures) - 1
        kind = {"floor": "diffuse", "walls": "default", "columns": "default", "arches": "default", "cloth": None}[name]
        model.meshes.append(_mesh(v, i, material_mix(rng, kind), uv if tex_id >= 0 else None, tex_id))


def sponza_class(target_tris=262144, seed=0x5702A):
    """C2: procedural atrium scaled to Crytek Sponza's ~±1900 extent; camera preset 'sponza' looks down its axis."""
    detail, model = 1.0, None
    for _ in range(12):                                                      # fixed-point on the detail knob
        rng = np.random.default_rng(seed)
        model = Model()
        _parts_to_model(model, rng, _atrium_parts(rng, -1900.0, 1900.0, -800.0, 800.0, 1400.0, detail))
        n = model.num_triangles()
        if abs(n - target_tris) <= 0.01 * target_tris:
            break
        detail *= math.sqrt(target_tris / n)
    return model


# ------------------------------------------------------------------------------------------ C3 Lost-Empire-class
def lost_empire_class(target_tris=225000, seed=0x10573):
    """C3: voxel terrain (fBm height map, each column's exposed voxels as boxes of 12 triangles)."""
    nb = target_tris // 12
    ratio = 1.55                                                             # boxes per column, refined below
    for _ in range(8):
        rng = np.random.default_rng(seed)
        n = int(math.sqrt(nb / ratio))
        xs = np.arange(n) - n / 2
        X, Z = np.meshgrid(xs, xs, indexing="ij")
        h = np.zeros_like(X, dtype=np.float64)
        for o in range(5):
            fx, fz, ph = rng.uniform(0.5, 1.5, 2) * (2 ** o) * 0.035, rng.uniform(0.5, 1.5, 2) * (2 ** o) * 0.035, rng.uniform(0, 6.28, 2)
            h += (np.sin(X * fx[0] + ph[0]) * np.cos(Z * fz[0] + ph[1]) + np.sin((X + Z) * fx[1] * 0.7)) * 9.0 / (2 ** o)
        hq = np.floor(h).astype(np.int32)
        # a column shows its top voxel plus the voxels above its lowest neighbour (cliff faces)
        pad = np.pad(hq, 1, mode="edge")
        low = np.minimum.reduce([pad[:-2, 1:-1], pad[2:, 1:-1], pad[1:-1, :-2], pad[1:-1, 2:]])
        depth = np.clip(hq - low, 1, 6)
        total = int(depth.sum())
        if nb <= total <= 1.01 * nb:
            break
        ratio *= (total / nb) ** 0.5 if total < nb else (total / (1.005 * nb)) ** 0.5
        ratio = total / (n * n) * (1.0 if total >= nb else 0.995)
    centers, layer = [], []
    for d in range(6):
        m = depth > d
        centers.append(np.stack([X[m] * 1.0, (hq[m] - d) * 1.0, Z[m] * 1.0], -1))
        layer.append(np.full(m.sum(), d))
    centers = np.concatenate(centers)[:nb]
    band = (centers[:, 1] > np.percentile(centers[:, 1], 66)).astype(int) + (centers[:, 1] > np.percentile(centers[:, 1], 90)).astype(int)
    model = Model()
    for b, kind in enumerate(("diffuse", "default", "default")):
        sel = centers[band == b]
        if len(sel):
            v, i = _boxes(sel, 0.5)
            model.meshes.append(_mesh(v, i, material_mix(rng, kind)))
    water = _boxes(np.array([[0.0, float(np.percentile(hq, 20)) + 0.4, 0.0]], np.float32), np.array([n / 2, 0.05, n / 2], np.float32))
    model.meshes.append(_mesh(water[0], water[1], material_mix(rng, "glass")))
    return model


# ------------------------------------------------------------------------------------------ C5 20M-triangle class
def city_class(target_tris=20_000_000, seed=0x20000000, cell_tris=None):
    """C5: an (n x n) grid of atrium cells, each a distinct copy (flattened; the reference has no instancing)."""
    rng = np.random.default_rng(seed)
    cell_tris = cell_tris or min(262144, max(2000, target_tris // 64))
    n = max(1, int(round(math.sqrt(target_tris / cell_tris))))
    per = target_tris / (n * n)
    base = sponza_class(int(per), seed=seed ^ 0x5A5A)
    model = Model()
    model.textures = base.textures
    span_x, span_z = 4000.0, 1800.0
    for m in base.meshes:
        vs, is_, uvs = [], [], []
        nv = len(m.vertex)
        for gx in range(n):
            for gz in range(n):
                jitter = rng.uniform(-20, 20, 3) * np.array([1, 0, 1])
                off = np.array([(gx - (n - 1) / 2) * span_x, 0.0, (gz - (n - 1) / 2) * span_z]) + jitter
                scale = np.array([1.0, rng.uniform(0.8, 1.25), 1.0])
                vs.append(m.vertex * scale.astype(np.float32) + off.astype(np.float32))
                is_.append(m.index.astype(np.int64) + nv * (gx * n + gz))
                if len(m.texcoord):
                    uvs.append(m.texcoord)
        mm = _mesh(np.concatenate(vs), np.concatenate(is_), m.material, np.concatenate(uvs) if uvs else None, m.diffuseTextureID)
        model.meshes.append(mm)
    return model


def by_name(name, **kw):
    return {"cornell": cornell, "sponza": sponza_class, "lost_empire": lost_empire_class, "city": city_class}[name](**kw)


def add_smooth_normals(model):
    """Area-weighted per-vertex normals for every mesh (the Whitted integrator interpolates them, 04HelloRaytracing/
    deviceProgram.cu:86-90; the path tracer never reads normals).  Returns the model."""
    for mesh in model.meshes:
        v = np.asarray(mesh.vertex, np.float64)
        idx = np.asarray(mesh.index, np.int64)
        fn = np.cross(v[idx[:, 1]] - v[idx[:, 0]], v[idx[:, 2]] - v[idx[:, 0]])
        n = np.zeros_like(v)
        for k in range(3):
            np.add.at(n, idx[:, k], fn)
        length = np.linalg.norm(n, axis=1, keepdims=True)
        n = np.where(length > 0, n / np.where(length > 0, length, 1.0), np.array([0.0, 1.0, 0.0]))
        mesh.normal = np.ascontiguousarray(n, np.float32)
    return model
