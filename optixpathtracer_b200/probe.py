"""Environment probe, host side: ProbeData + BuildCDF, loadColor, analytic sky.

reference: HelloPathtracing_sv4_vmv23/Probe.h:7-89 (ProbeData::BuildCDF), main.cpp:154-180 (loadProbe/loadColor).
BuildCDF is a *sequential* fp32 accumulation in the reference and that order defines the bits of the tables
(SURVEY.md §8a row a17); numpy's float32 cumsum accumulates in the same left-to-right order, and
tests/test_oracle_kat.py checks the result bit-for-bit against the reference's own BuildCDF.
"""
import numpy as np


class ProbeData:
    def __init__(self, data=None, offset=(0.0, 0.0, 0.0)):
        self.data = None if data is None else np.ascontiguousarray(data, np.float32)   # (H, W, 4)
        self.offset = tuple(offset)
        self.valid = False
        self.pdfX = self.cdfX = self.pdfY = self.cdfY = None

    @property
    def width(self):
        return self.data.shape[1]

    @property
    def height(self):
        return self.data.shape[0]

    def BuildCDF(self):
        d = self.data
        f = np.float32
        lum = (d[..., 0] * f(0.3) + d[..., 1] * f(0.6)) + d[..., 2] * f(0.1)          # Luminance, maths.h:165-168
        cx = np.cumsum(lum, axis=1, dtype=np.float32)                              # running row weight (Probe.h:42-51)
        tot = cx[:, -1].copy()
        inv = f(1.0) / tot                                                         # Probe.h:53
        self.pdfX = np.ascontiguousarray(lum * inv[:, None], np.float32)
        self.cdfX = np.ascontiguousarray(cx * inv[:, None], np.float32)
        cy = np.cumsum(tot, dtype=np.float32)                                      # Probe.h:62-66
        total = cy[-1]
        self.cdfY = np.ascontiguousarray(cy / total, np.float32)                   # Probe.h:70-74
        self.pdfY = np.ascontiguousarray(tot / total, np.float32)
        self.valid = True
        return self


def loadColor(color, width=3840, height=2160):
    """main.cpp:167-180 — constant-colour environment (the reference hard-codes 3840x2160)."""
    data = np.empty((height, width, 4), np.float32)
    data[..., 0], data[..., 1], data[..., 2], data[..., 3] = color[0], color[1], color[2], 1.0
    return ProbeData(data).BuildCDF()


def analytic_sky(width=2048, height=1024, sun_dir=(0.35, 0.8, 0.45), sun_strength=40.0, seed=7):
    """Synthetic HDR sky (the reference ships no .hdr): gradient sky + ground + a broad sun lobe.
    Stands in for garden_nook_8k.hdr (main.cpp:201-202) in the C2-C5 workloads."""
    v = (np.arange(height, dtype=np.float32) + 0.5) / height
    u = (np.arange(width, dtype=np.float32) + 0.5) / width
    theta = v[:, None] * np.float32(np.pi)
    phi = u[None, :] * np.float32(2 * np.pi)
    x = -np.sin(theta) * np.cos(phi)
    y = np.cos(theta) * np.ones_like(phi)
    z = -np.sin(theta) * np.sin(phi)
    s = np.asarray(sun_dir, np.float32)
    s = s / np.linalg.norm(s)
    cosang = np.clip(x * s[0] + y * s[1] + z * s[2], -1, 1)
    up = np.clip(y, 0, 1)
    sky = np.stack([0.35 + 0.25 * (1 - up), 0.5 + 0.2 * (1 - up), 0.9 - 0.1 * (1 - up)], -1)
    ground = np.array([0.18, 0.16, 0.14], np.float32)
    base = np.where((y > 0)[..., None], sky, ground[None, None, :])
    sun = (sun_strength * np.exp((cosang - 1.0) * 400.0) + 2.0 * np.exp((cosang - 1.0) * 12.0))[..., None] * np.array([1.0, 0.93, 0.8], np.float32)
    data = np.concatenate([base + sun, np.ones((height, width, 1), np.float32)], -1).astype(np.float32)
    return ProbeData(data).BuildCDF()


# ---- Radiance .hdr (RGBE) files: loadProbe (main.cpp:154-165) ------------------------------------------------------------
# The reference reads its environment with stbi_loadf(file, &w, &h, &channels, 4).  stb_image is a vendored third-party file of
# the reference and is not copied; this is our own reader of the published Radiance picture format with stbi_loadf's conventions
# (top row first, no flip; 4 channels with alpha 1; value = mantissa * 2^(exponent - 136), exponent 0 -> 0; "-Y h +X w" layout
# only; new-style RLE scanlines for widths 8..32767, otherwise flat RGBE quads).  tests/golden/hdr.npz pins it against the
# reference's own stb_image on RLE and flat files.
def _rgbe_to_float(rgbe):
    rgbe = np.asarray(rgbe, np.uint8).reshape(-1, 4)
    e = rgbe[:, 3].astype(np.int32)
    scale = np.ldexp(np.float32(1.0), e - 136).astype(np.float32)
    out = np.empty((len(rgbe), 4), np.float32)
    out[:, :3] = rgbe[:, :3].astype(np.float32) * scale[:, None]
    out[e == 0, :3] = 0.0
    out[:, 3] = 1.0
    return out


def read_hdr(path_or_bytes):
    """-> (H, W, 4) float32, like stbi_loadf(..., 4) on a Radiance .hdr file."""
    try:
        return _read_hdr(path_or_bytes)
    except (IndexError, ValueError) as e:
        raise RuntimeError("Could not load HDR probe: truncated or corrupt data") from e


def _read_hdr(path_or_bytes):
    if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
        buf = bytes(path_or_bytes)
    else:
        with open(path_or_bytes, "rb") as f:
            buf = f.read()
    pos = 0

    def line():
        nonlocal pos
        end = buf.find(b"\n", pos)
        if end < 0:
            raise RuntimeError("Could not load HDR probe: truncated header")
        s = buf[pos:end]
        pos = end + 1
        return s.decode("latin-1")

    first = line()
    if first not in ("#?RADIANCE", "#?RGBE"):
        raise RuntimeError("Could not load HDR probe: not a Radiance picture")
    valid = False
    while True:
        s = line()
        if s == "":
            break
        if s == "FORMAT=32-bit_rle_rgbe":
            valid = True
    if not valid:
        raise RuntimeError("Could not load HDR probe: unsupported format")
    tok = line().split()
    if len(tok) != 4 or tok[0] != "-Y" or tok[2] != "+X":
        raise RuntimeError("Could not load HDR probe: unsupported data layout")
    h, w = int(tok[1]), int(tok[3])
    if w <= 0 or h <= 0:
        raise RuntimeError("Could not load HDR probe: bad size")
    raw = np.frombuffer(buf, np.uint8, offset=pos)
    flat = w < 8 or w >= 32768
    if not flat and len(raw) >= 4 and (raw[0] != 2 or raw[1] != 2 or (raw[2] & 0x80)):
        flat = True                                   # stb: the first scanline decides for the whole file
    if flat:
        if len(raw) < 4 * w * h:
            raise RuntimeError("Could not load HDR probe: truncated data")
        return _rgbe_to_float(raw[:4 * w * h]).reshape(h, w, 4)
    out = np.empty((h, w, 4), np.float32)
    scan = np.empty((4, w), np.uint8)
    p = 0
    data = raw.tobytes()
    for j in range(h):
        if p + 4 > len(data) or data[p] != 2 or data[p + 1] != 2 or (data[p + 2] & 0x80):
            raise RuntimeError("Could not load HDR probe: invalid decoded scanline length")
        if ((data[p + 2] << 8) | data[p + 3]) != w:
            raise RuntimeError("Could not load HDR probe: invalid decoded scanline length")
        p += 4
        for k in range(4):
            i = 0
            while i < w:
                count = data[p]
                p += 1
                if count > 128:
                    count -= 128
                    if count > w - i:
                        raise RuntimeError("Could not load HDR probe: corrupt RLE data")
                    scan[k, i:i + count] = data[p]
                    p += 1
                else:
                    if count == 0 or count > w - i:
                        raise RuntimeError("Could not load HDR probe: corrupt RLE data")
                    scan[k, i:i + count] = np.frombuffer(data, np.uint8, count, p)
                    p += count
                i += count
        out[j] = _rgbe_to_float(scan.T)
    return out


def write_hdr(path, rgb, rle=True):
    """Writes (H, W, >=3) float32 as a Radiance .hdr file (test assets; the reference has no writer).  Returns the RGBE bytes."""
    rgb = np.asarray(rgb, np.float32)[..., :3]
    h, w = rgb.shape[:2]
    m = rgb.max(axis=-1)
    mant, expo = np.frexp(m)
    rgbe = np.zeros((h, w, 4), np.uint8)
    ok = m > 1e-32
    scale = np.where(ok, mant * 256.0 / np.where(ok, m, 1.0), 0.0).astype(np.float32)
    rgbe[..., :3] = np.clip(rgb * scale[..., None], 0, 255).astype(np.uint8)
    rgbe[..., 3] = np.where(ok, expo + 128, 0).astype(np.uint8)
    rgbe[~ok] = 0
    body = bytearray()
    if rle and 8 <= w < 32768:
        for j in range(h):
            body += bytes([2, 2, w >> 8, w & 255])
            for k in range(4):
                row = rgbe[j, :, k]
                i = 0
                while i < w:
                    run = 1
                    while i + run < w and run < 127 and row[i + run] == row[i]:
                        run += 1
                    if run >= 3:
                        body += bytes([128 + run, int(row[i])])
                        i += run
                    else:
                        start = i
                        while i < w and i - start < 128:
                            if i + 2 < w and row[i] == row[i + 1] == row[i + 2]:
                                break
                            i += 1
                        body += bytes([i - start]) + row[start:i].tobytes()
    else:
        body += rgbe.tobytes()
    blob = b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n" + f"-Y {h} +X {w}\n".encode() + bytes(body)
    if path is not None:
        with open(path, "wb") as f:
            f.write(blob)
    return blob


def loadProbe(hdrFile):
    """main.cpp:154-165: stbi_loadf + BuildCDF."""
    return ProbeData(read_hdr(hdrFile)).BuildCDF()
