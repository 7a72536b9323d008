// Scene input types with the reference's names and ownership rules
// (reference: HelloPathtracing_sv4_vmv23/Model.h:10-46, Model.cpp:137-290).  Implemented in host/Model.cpp.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include <vector_types.h>
#include "Material.h"

// one indexed triangle mesh = one OptiX build input / SBT record in the reference
struct TriangleMesh {
    std::vector<float3> vertex;
    std::vector<float3> normal;      // uploaded by the reference, never read by its shaders
    std::vector<float2> texcoord;
    std::vector<uint3>  index;
    Material material;
    int diffuseTextureID{-1};
};

struct Texture {
    ~Texture() { delete[] pixel; }
    uint32_t* pixel{nullptr};        // RGBA8, R in the low byte, flipped vertically on load (Model.cpp:118-125)
    int2 resolution{-1, -1};
};

struct Model {                       // owns its meshes and textures (Model.h:31-36); the renderer only borrows it
    ~Model() { for (auto m : meshes) delete m;  for (auto t : textures) delete t; }
    std::vector<TriangleMesh*> meshes;
    std::vector<Texture*> textures;
    size_t numTriangles() const { size_t n = 0;  for (auto m : meshes) n += m->index.size();  return n; }
};

// Heap Model the caller deletes; throws std::runtime_error when the file cannot be read or parsed (Model.cpp:158-160).
// One TriangleMesh per (shape, material id); vertices de-duplicated by their (v, vn, vt) index triple;
// material.color <- Kd, material.emission <- Ke, every other field keeps the Material() default (Model.cpp:189-190).
Model* loadOBJ(const std::string& objFile);

// 12-triangle axis-aligned box, 36 un-shared vertices (Model.cpp:218-290).
void addBox(Model* model, Material& mat, const float3& pos, const float3& extend);
