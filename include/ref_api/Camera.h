// Pinhole camera with the semantics of sutil::Camera (reference: sutil/Camera.h, sutil/Camera.cpp:34-45).
#pragma once
#include <cmath>
#include <vector_types.h>

namespace sutil {
class Camera {
public:
    Camera() : m_eye{1.f, 1.f, 1.f}, m_lookat{0.f, 0.f, 0.f}, m_up{0.f, 1.f, 0.f}, m_fovY(35.f), m_aspectRatio(1.f) {}
    Camera(const float3& eye, const float3& lookat, const float3& up, float fovY, float aspectRatio)
        : m_eye(eye), m_lookat(lookat), m_up(up), m_fovY(fovY), m_aspectRatio(aspectRatio) {}
    const float3& eye() const { return m_eye; }
    const float3& lookat() const { return m_lookat; }
    const float3& up() const { return m_up; }
    float fovY() const { return m_fovY; }
    float aspectRatio() const { return m_aspectRatio; }
    void setEye(const float3& v) { m_eye = v; }
    void setLookat(const float3& v) { m_lookat = v; }
    void setUp(const float3& v) { m_up = v; }
    void setFovY(float v) { m_fovY = v; }
    void setAspectRatio(float v) { m_aspectRatio = v; }

    // W = lookat - eye (NOT normalised: it carries the focal length); U, V scaled to the half extents of the image plane.
    void UVWFrame(float3& U, float3& V, float3& W) const {
        W = sub(m_lookat, m_eye);
        const float wlen = sqrtf(dot(W, W));
        U = normalize(cross(W, m_up));
        V = normalize(cross(U, W));
        const float vlen = wlen * tanf(0.5f * m_fovY * 3.14159265358979323846f / 180.0f);
        V = scale(V, vlen);
        const float ulen = vlen * m_aspectRatio;
        U = scale(U, ulen);
    }

private:
    static float3 sub(const float3& a, const float3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
    static float dot(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
    static float3 cross(const float3& a, const float3& b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
    static float3 scale(const float3& a, float s) { return {a.x * s, a.y * s, a.z * s}; }
    static float3 normalize(const float3& v) { const float inv = 1.0f / sqrtf(dot(v, v));  return scale(v, inv); }
    float3 m_eye, m_lookat, m_up;
    float m_fovY, m_aspectRatio;
};
}  // namespace sutil
