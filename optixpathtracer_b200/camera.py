"""Pinhole camera of the reference's sutil library.

reference: sutil/Camera.h:37-76, sutil/Camera.cpp:34-45 (UVWFrame); presets from sv4/main.cpp:211-240.
"""
import numpy as np

f32 = np.float32


def _norm(v):
    inv = f32(1.0) / np.sqrt(f32(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]), dtype=f32)
    return (v * inv).astype(f32)


def _cross(a, b):
    return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]], f32)


class Camera:
    def __init__(self, eye=(1, 1, 1), lookat=(0, 0, 0), up=(0, 1, 0), fovY=35.0, aspectRatio=1.0):
        self.eye = np.asarray(eye, f32)
        self.lookat = np.asarray(lookat, f32)
        self.up = np.asarray(up, f32)
        self.fovY = f32(fovY)
        self.aspectRatio = f32(aspectRatio)

    def setAspectRatio(self, a):
        self.aspectRatio = f32(a)

    def UVWFrame(self):
        """W is NOT normalised (it carries the focal length); U, V are scaled by the half extents."""
        W = (self.lookat - self.eye).astype(f32)
        wlen = np.sqrt(f32(W[0] * W[0] + W[1] * W[1] + W[2] * W[2]), dtype=f32)
        U = _norm(_cross(W, self.up))
        V = _norm(_cross(U, W))
        vlen = f32(wlen * np.tan(f32(f32(0.5) * self.fovY * f32(3.14159265358979323846) / f32(180.0)), dtype=f32))
        V = (V * vlen).astype(f32)
        ulen = f32(vlen * self.aspectRatio)
        U = (U * ulen).astype(f32)
        return U, V, W


PRESETS = {
    "sponza": dict(eye=(-1293.07, 154.681, 1.0), lookat=(0.0, 200.0, 0.0), up=(0, 1, 0), fovY=35.0),      # main.cpp:213-217
    "san_miguel": dict(eye=(26, 8, -2), lookat=(0, 0, 0), up=(0, 1, 0), fovY=35.0),                        # :220-224
    "lost_empire": dict(eye=(-70, 40, 100), lookat=(0, 0, 0), up=(0, 1, 0), fovY=45.0),                    # :227-231
    "cornell": dict(eye=(0.0, 1.0, 3.2), lookat=(0.0, 1.0, 0.0), up=(0, 1, 0), fovY=40.0),
}


def preset_camera(name):
    return Camera(**PRESETS[name])
