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
