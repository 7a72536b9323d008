// Surface material record, layout-compatible with the reference's `struct Material`
// (reference: HelloPathtracing_sv4_vmv23/Material.h:11-70; 104 bytes, offsets in SURVEY.md §8b).
// Re-typed for the B200 path tracer; the device kernels read this struct verbatim from the
// per-mesh material table, so the field order below IS the ABI.
#pragma once
#include <vector_types.h>
#include <cstddef>
#include <cmath>

static const int MATERIAL_FLAG_NONE = 0;
static const int MATERIAL_FLAG_SHADOW_CATCHER = 1 << 0;

struct Material {
    // --- ABI fields (do not reorder) -------------------------------------------------------
    float3 emission;        // @0
    float3 color;           // @12
    float3 absorption;      // @24
    float  eta;             // @36   0 => derive the IOR from `specular`
    float  metallic;        // @40
    float  subsurface;      // @44
    float  specular;        // @48
    float  roughness;       // @52
    float  specularTint;    // @56
    float  anisotropic;     // @60   (unused by the path, kept for layout)
    float  sheen;           // @64   (unused)
    float  sheenTint;       // @68   (unused)
    float  clearcoat;       // @72
    float  clearcoatGloss;  // @76
    float  transmission;    // @80
    float  bump;            // @84   (unused)
    float3 bumpTile;        // @88   (unused)
    int    flags;           // @100

    // Default = the sv4_vmv23 constructor values (Material.h:13-38): a 40 %-transmissive,
    // half-metallic surface with unit emission.  loadOBJ only overwrites color + emission.
    Material() { *this = sv4Defaults(); }
    struct NoInit {};
    explicit Material(NoInit) {}

    static Material sv4Defaults() {
        Material m{NoInit{}};
        m.emission = {1.f, 1.f, 1.f};  m.color = {1.f, 0.f, 0.f};  m.absorption = {1.f, 1.f, 1.f};
        m.eta = 1.4f;  m.metallic = 0.5f;  m.subsurface = 0.f;  m.specular = 1.f;  m.roughness = 1.f;
        m.specularTint = 1.f;  m.anisotropic = 0.f;  m.sheen = 0.f;  m.sheenTint = 0.f;
        m.clearcoat = 0.f;  m.clearcoatGloss = 1.f;  m.transmission = 0.4f;  m.bump = 0.f;
        m.bumpTile = {1.f, 1.f, 1.f};  m.flags = 0;
        return m;
    }
    // HelloPathtracing_original/Material.h:13-36 (the plain diffuse-ish Disney default).
    static Material sv3Defaults() {                       // HelloPathtracing_sv3/Material.h:13-37
        Material m = origDefaults();
        m.color = {0.6f, 1.6f, 0.6f};  m.eta = 1.4f;
        return m;
    }
    static Material origDefaults() {
        Material m{NoInit{}};
        m.emission = {0.f, 0.f, 0.f};  m.color = {0.6f, 0.6f, 0.6f};  m.absorption = {0.f, 0.f, 0.f};
        m.eta = 0.f;  m.metallic = 0.f;  m.subsurface = 0.f;  m.specular = 0.5f;  m.roughness = 1.f;
        m.specularTint = 0.f;  m.anisotropic = 0.f;  m.sheen = 0.f;  m.sheenTint = 0.f;
        m.clearcoat = 0.f;  m.clearcoatGloss = 1.f;  m.transmission = 0.f;  m.bump = 0.f;
        m.bumpTile = {10.f, 10.f, 10.f};  m.flags = 0;
        return m;
    }

    // Material.h:40-46
    float GetIndexOfRefraction() const {
        return eta == 0.0f ? 2.0f / (1.0f - std::sqrt(0.08f * specular)) - 1.0f : eta;
    }
};

static_assert(sizeof(Material) == 104, "Material must stay 104 bytes (reference ABI)");
static_assert(offsetof(Material, color) == 12 && offsetof(Material, eta) == 36 &&
              offsetof(Material, transmission) == 80 && offsetof(Material, bumpTile) == 88 &&
              offsetof(Material, flags) == 100, "Material field offsets changed");
