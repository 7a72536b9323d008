"""Scene input types of the reference, host side: TriangleMesh / Texture / Model, loadOBJ, addBox.

reference: HelloPathtracing_sv4_vmv23/Model.h:10-46, Model.cpp:137-216 (loadOBJ), :218-290 (addBox).
Arrays are numpy (float32 / uint32) so they can be handed to the C-ABI (include/b200pt.h) without copies.
"""
import os
import struct
import zlib

import numpy as np

from .structs import Material


class TriangleMesh:
    """Model.h:10-20 — an indexed triangle mesh with one material."""

    def __init__(self):
        self.vertex = np.zeros((0, 3), np.float32)
        self.normal = np.zeros((0, 3), np.float32)
        self.texcoord = np.zeros((0, 2), np.float32)
        self.index = np.zeros((0, 3), np.uint32)
        self.material = Material()
        self.diffuseTextureID = -1


class Texture:
    """Model.h:22-30 — RGBA8 pixels, R in the low byte, row 0 = bottom row after loadOBJ's vertical flip."""

    def __init__(self, pixel=None, resolution=(-1, -1)):
        self.pixel = pixel
        self.resolution = tuple(resolution)


class Model:
    """Model.h:32-42."""

    def __init__(self):
        self.meshes = []
        self.textures = []

    def num_triangles(self):
        return int(sum(len(m.index) for m in self.meshes))


def addBox(model, mat, pos, extend):
    """Model.cpp:218-290 — 12 triangles, 36 un-shared vertices, face order front/back/left/right/top/bottom."""
    px, py, pz = (float(v) for v in pos)
    ex, ey, ez = (float(v) for v in extend)
    f = np.float32
    A = (f(-ex) + f(px), f(-ey) + f(py), f(ez) + f(pz))
    B = (f(ex) + f(px), f(-ey) + f(py), f(ez) + f(pz))
    Cc = (f(ex) + f(px), f(ey) + f(py), f(ez) + f(pz))
    D = (f(-ex) + f(px), f(ey) + f(py), f(ez) + f(pz))
    E = (f(-ex) + f(px), f(-ey) + f(py), f(-ez) + f(pz))
    F = (f(ex) + f(px), f(-ey) + f(py), f(-ez) + f(pz))
    G = (f(ex) + f(px), f(ey) + f(py), f(-ez) + f(pz))
    H = (f(-ex) + f(px), f(ey) + f(py), f(-ez) + f(pz))
    verts = [A, B, Cc, A, Cc, D,  E, H, G, E, G, F,  E, A, D, E, D, H,
             B, F, G, B, G, Cc,  D, Cc, G, D, G, H,  E, A, B, E, B, F]
    normals = [(0, 0, 1)] * 6 + [(0, 0, -1)] * 6 + [(-1, 0, 0)] * 6 + [(1, 0, 0)] * 6 + [(0, 1, 0)] * 6 + [(0, -1, 0)] * 6
    mesh = TriangleMesh()
    mesh.vertex = np.array(verts, np.float32)
    mesh.normal = np.array(normals, np.float32)
    mesh.texcoord = np.zeros((36, 2), np.float32)
    mesh.index = np.arange(36, dtype=np.uint32).reshape(12, 3)
    m = Material()
    import ctypes
    ctypes.memmove(ctypes.byref(m), ctypes.byref(mat), ctypes.sizeof(Material))
    mesh.material = m
    model.meshes.append(mesh)
    return mesh


# ----------------------------------------------------------------------------------------------- images
def _read_png(path):
    """Minimal PNG reader (8-bit gray/RGB/RGBA/palette, non-interlaced) -> (H, W, 4) uint8."""
    with open(path, "rb") as fh:
        data = fh.read()
    if data[:8] != b"\x89PNG\r\n\x1a\n":
        raise ValueError("not a PNG")
    pos, idat, plte, trns = 8, b"", None, None
    while pos < len(data):
        ln, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + ln]
        pos += 12 + ln
        if typ == b"IHDR":
            w, h, depth, ctype, _, _, interlace = struct.unpack(">IIBBBBB", body)
        elif typ == b"IDAT":
            idat += body
        elif typ == b"PLTE":
            plte = np.frombuffer(body, np.uint8).reshape(-1, 3)
        elif typ == b"tRNS":
            trns = np.frombuffer(body, np.uint8)
        elif typ == b"IEND":
            break
    if depth != 8 or interlace:
        raise ValueError("unsupported PNG (need 8-bit, non-interlaced)")
    ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    raw = zlib.decompress(idat)
    stride = w * ch
    out = np.zeros((h, stride), np.uint8)
    prev = np.zeros(stride, np.int32)
    p = 0
    for y in range(h):
        ft = raw[p]
        line = np.frombuffer(raw, np.uint8, stride, p + 1).astype(np.int32)
        p += stride + 1
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 255
        else:
            cur = np.zeros(stride, np.int32)
            for i in range(stride):
                a = cur[i - ch] if i >= ch else 0
                b = prev[i]
                c = prev[i - ch] if i >= ch else 0
                if ft == 1:
                    pr = a
                elif ft == 3:
                    pr = (a + b) >> 1
                else:
                    pa, pb, pc = abs(b - c), abs(a - c), abs(a + b - 2 * c)
                    pr = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                cur[i] = (line[i] + pr) & 255
        out[y] = cur
        prev = cur
    px = out.reshape(h, w, ch)
    rgba = np.full((h, w, 4), 255, np.uint8)
    if ctype == 3:
        rgba[..., :3] = plte[px[..., 0]]
        if trns is not None:
            a = np.full(256, 255, np.uint8)
            a[:len(trns)] = trns
            rgba[..., 3] = a[px[..., 0]]
    elif ch == 1:
        rgba[..., :3] = px
    elif ch == 2:
        rgba[..., :3] = px[..., :1]
        rgba[..., 3] = px[..., 1]
    else:
        rgba[..., :ch] = px
    return rgba


def _read_ppm(path):
    with open(path, "rb") as fh:
        data = fh.read()
    toks, pos = [], 0
    while len(toks) < 4:
        while data[pos:pos + 1].isspace():
            pos += 1
        if data[pos:pos + 1] == b"#":
            pos = data.index(b"\n", pos) + 1
            continue
        end = pos
        while not data[end:end + 1].isspace():
            end += 1
        toks.append(data[pos:end])
        pos = end
    if toks[0] != b"P6" or int(toks[3]) != 255:
        raise ValueError("unsupported PPM")
    w, h = int(toks[1]), int(toks[2])
    rgb = np.frombuffer(data, np.uint8, w * h * 3, pos + 1).reshape(h, w, 3)
    return np.concatenate([rgb, np.full((h, w, 1), 255, np.uint8)], axis=2)


def load_image_rgba8(path):
    """stbi_load(..., STBI_rgb_alpha) stand-in for the formats we can decode without third-party code."""
    ext = os.path.splitext(path)[1].lower()
    return _read_png(path) if ext == ".png" else _read_ppm(path)


def _load_texture(model, known, name, model_dir):
    """Model.cpp:86-135: cache by name, fix backslashes, flip vertically, -1 when unreadable."""
    if name == "":
        return -1
    if name in known:
        return known[name]
    path = os.path.join(model_dir, name.replace("\\", "/"))
    tex_id = -1
    try:
        rgba = load_image_rgba8(path)
        rgba = rgba[::-1].copy()                       # Model.cpp:118-125
        tex = Texture(rgba.view(np.uint32).reshape(rgba.shape[0], rgba.shape[1]).copy(), (rgba.shape[1], rgba.shape[0]))
        tex_id = len(model.textures)
        model.textures.append(tex)
    except Exception:
        print(f"Could not load texture from {path}!")
    known[name] = tex_id
    return tex_id


# ----------------------------------------------------------------------------------------------- OBJ
def _parse_mtl(path):
    mats, cur = [], None
    with open(path, "r", errors="replace") as fh:
        for line in fh:
            t = line.split()
            if not t or t[0].startswith("#"):
                continue
            if t[0] == "newmtl":
                cur = dict(name=" ".join(t[1:]), diffuse=(0.6, 0.6, 0.6), emission=(0.0, 0.0, 0.0), diffuse_texname="")
                mats.append(cur)
            elif cur is None:
                continue
            elif t[0] == "Kd":
                cur["diffuse"] = tuple(float(x) for x in t[1:4])
            elif t[0] == "Ke":
                cur["emission"] = tuple(float(x) for x in t[1:4])
            elif t[0] == "map_Kd":
                cur["diffuse_texname"] = t[-1]
    return mats


def loadOBJ(objFile):
    """Model.cpp:137-216.  One TriangleMesh per (shape, material id); vertices de-duplicated by their
    (v, vn, vt) index triple; material.color <- Kd, material.emission <- Ke, every other Material field keeps
    its constructor default (Model.cpp:189-190); diffuse texture from map_Kd.

    Deliberate deviation: the reference shares one `knownVertices` map across the material-split meshes of a
    shape (Model.cpp:176), so a vertex used by two materials gets the FIRST mesh's index in the second mesh;
    here the map is per mesh.  Polygons are fan-triangulated like tinyobjloader's simple path.
    """
    if not os.path.exists(objFile):
        raise RuntimeError(f"Could not read OBJ model from {objFile} : file not found")
    model_dir = objFile[:objFile.rfind("/") + 1]
    V, VN, VT = [], [], []
    materials, mat_index = [], {}
    shapes = []                                     # list of dict(faces=[(i0,i1,i2)...], mats=[...])
    cur = dict(faces=[], mats=[])
    cur_mat = -1

    def fix(i, n):
        i = int(i)
        return i - 1 if i > 0 else n + i

    with open(objFile, "r", errors="replace") as fh:
        for line in fh:
            t = line.split()
            if not t or t[0].startswith("#"):
                continue
            k = t[0]
            if k == "v":
                V.append((float(t[1]), float(t[2]), float(t[3])))
            elif k == "vn":
                VN.append((float(t[1]), float(t[2]), float(t[3])))
            elif k == "vt":
                VT.append((float(t[1]), float(t[2]) if len(t) > 2 else 0.0))
            elif k == "f":
                corners = []
                for c in t[1:]:
                    p = c.split("/")
                    vi = fix(p[0], len(V))
                    ti = fix(p[1], len(VT)) if len(p) > 1 and p[1] else -1
                    ni = fix(p[2], len(VN)) if len(p) > 2 and p[2] else -1
                    corners.append((vi, ni, ti))
                for j in range(1, len(corners) - 1):
                    cur["faces"].append((corners[0], corners[j], corners[j + 1]))
                    cur["mats"].append(cur_mat)
            elif k in ("o", "g"):
                if cur["faces"]:
                    shapes.append(cur)
                cur = dict(faces=[], mats=[])
            elif k == "usemtl":
                cur_mat = mat_index.get(" ".join(t[1:]), -1)
            elif k == "mtllib":
                try:
                    for m in _parse_mtl(os.path.join(model_dir, " ".join(t[1:]))):
                        mat_index.setdefault(m["name"], len(materials))
                        materials.append(m)
                except OSError as e:
                    raise RuntimeError(f"Could not read OBJ model from {objFile} : {e}")
    if cur["faces"]:
        shapes.append(cur)

    V = np.array(V, np.float32).reshape(-1, 3)
    VN = np.array(VN, np.float32).reshape(-1, 3)
    VT = np.array(VT, np.float32).reshape(-1, 2)
    model = Model()
    for shape in shapes:
        known_tex = {}
        for mid in sorted(set(shape["mats"])):
            known, verts, norms, tcs, tris = {}, [], [], [], []
            for face, fm in zip(shape["faces"], shape["mats"]):
                if fm != mid:
                    continue
                tri = []
                for key in face:
                    if key not in known:
                        known[key] = len(verts)
                        verts.append(V[key[0]])
                        if key[1] >= 0:
                            norms.append(VN[key[1]])
                        if key[2] >= 0:
                            tcs.append(VT[key[2]])
                    tri.append(known[key])
                tris.append(tri)
            if not verts:
                continue
            mesh = TriangleMesh()
            mesh.vertex = np.array(verts, np.float32)
            mesh.normal = np.array(norms, np.float32).reshape(-1, 3) if len(norms) == len(verts) else np.zeros((0, 3), np.float32)
            mesh.texcoord = np.array(tcs, np.float32).reshape(-1, 2) if len(tcs) == len(verts) else np.zeros((0, 2), np.float32)
            mesh.index = np.array(tris, np.uint32)
            if mid >= 0:
                m = materials[mid]
                mesh.material.color.x, mesh.material.color.y, mesh.material.color.z = m["diffuse"]
                mesh.material.emission.x, mesh.material.emission.y, mesh.material.emission.z = m["emission"]
                mesh.diffuseTextureID = _load_texture(model, known_tex, m["diffuse_texname"], model_dir)
            model.meshes.append(mesh)
    return model
