// loadOBJ / addBox with the reference's semantics (HelloPathtracing_sv4_vmv23/Model.cpp:137-290), written against our own
// small Wavefront OBJ/MTL reader (the reference uses the vendored tinyobjloader + stb_image, neither of which is copied).
//  * one TriangleMesh per (shape, material id) (Model.cpp:162-213); polygons fan-triangulated;
//  * vertices de-duplicated by their (v, vn, vt) index triple (Model.cpp:51-84).  Deliberate deviation, as in the Python
//    mirror: the map is per mesh (the reference shares one map across the material-split meshes of a shape, Model.cpp:176,
//    which hands the second mesh indices into the first mesh's arrays);
//  * material.color <- Kd, material.emission <- Ke; every other Material field keeps its constructor default (:189-190);
//  * map_Kd textures: binary PPM (P6) / PGM (P5) only here, flipped vertically like the reference (:118-125); a texture that
//    cannot be read leaves diffuseTextureID = -1 (the reference prints a warning and does the same, :127-129).
#include "../include/ref_api/Model.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <tuple>

namespace {

struct MtlEntry { std::string name;  float kd[3] = {0.6f, 0.6f, 0.6f};  float ke[3] = {0.f, 0.f, 0.f};  std::string mapKd; };

int fixIndex(int i, size_t n) { return i > 0 ? i - 1 : (int)n + i; }

std::vector<MtlEntry> parseMtl(const std::string& path) {
    std::ifstream in(path);
    if (!in) throw std::runtime_error("cannot open material library " + path);
    std::vector<MtlEntry> out;
    std::string line;
    while (std::getline(in, line)) {
        std::istringstream ss(line);
        std::string k;
        if (!(ss >> k) || k[0] == '#') continue;
        if (k == "newmtl") { out.emplace_back();  ss >> out.back().name; }
        else if (out.empty()) continue;
        else if (k == "Kd") ss >> out.back().kd[0] >> out.back().kd[1] >> out.back().kd[2];
        else if (k == "Ke") ss >> out.back().ke[0] >> out.back().ke[1] >> out.back().ke[2];
        else if (k == "map_Kd") { std::string tok;  while (ss >> tok) out.back().mapKd = tok; }
    }
    return out;
}

// binary PPM/PGM -> RGBA8 (R in the low byte), flipped vertically
bool readPnm(const std::string& path, std::vector<uint32_t>& px, int& w, int& h) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    char magic[3] = {0, 0, 0};
    int maxv = 0;
    auto token = [&](int& v) {
        int c = fgetc(f);
        while (c == '#' || isspace(c)) { if (c == '#') while (c != '\n' && c != EOF) c = fgetc(f);  c = fgetc(f); }
        v = 0;
        while (c >= '0' && c <= '9') { v = v * 10 + (c - '0');  c = fgetc(f); }
    };
    if (fread(magic, 1, 2, f) != 2 || magic[0] != 'P' || (magic[1] != '6' && magic[1] != '5')) { fclose(f);  return false; }
    token(w);  token(h);  token(maxv);
    if (w <= 0 || h <= 0 || maxv != 255) { fclose(f);  return false; }
    const int ch = magic[1] == '6' ? 3 : 1;
    std::vector<unsigned char> raw((size_t)w * h * ch);
    const bool ok = fread(raw.data(), 1, raw.size(), f) == raw.size();
    fclose(f);
    if (!ok) return false;
    px.resize((size_t)w * h);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const unsigned char* p = &raw[((size_t)(h - 1 - y) * w + x) * ch];
            const uint32_t r = p[0], g = ch == 3 ? p[1] : p[0], b = ch == 3 ? p[2] : p[0];
            px[(size_t)y * w + x] = r | (g << 8) | (b << 16) | (255u << 24);
        }
    return true;
}

int loadTexture(Model* model, std::map<std::string, int>& known, const std::string& name, const std::string& dir) {
    if (name.empty()) return -1;
    auto it = known.find(name);
    if (it != known.end()) return it->second;
    std::string file = dir + name;
    for (auto& c : file) if (c == '\\') c = '/';
    std::vector<uint32_t> px;
    int w = 0, h = 0, id = -1;
    if (readPnm(file, px, w, h)) {
        Texture* t = new Texture;
        t->resolution = {w, h};
        t->pixel = new uint32_t[px.size()];
        memcpy(t->pixel, px.data(), px.size() * sizeof(uint32_t));
        id = (int)model->textures.size();
        model->textures.push_back(t);
    } else fprintf(stderr, "Could not load texture from %s!\n", file.c_str());
    known[name] = id;
    return id;
}

struct Corner { int v, n, t;  bool operator<(const Corner& o) const { return std::tie(v, n, t) < std::tie(o.v, o.n, o.t); } };
struct Face { Corner c[3];  int mat; };
struct Shape { std::vector<Face> faces; };

}  // namespace

Model* loadOBJ(const std::string& objFile) {
    std::ifstream in(objFile);
    if (!in) throw std::runtime_error("Could not read OBJ model from " + objFile + " : file not found");
    const std::string dir = objFile.substr(0, objFile.rfind('/') + 1);
    std::vector<float3> V, VN;
    std::vector<float2> VT;
    std::vector<MtlEntry> materials;
    std::map<std::string, int> matIndex;
    std::vector<Shape> shapes;
    Shape cur;
    int curMat = -1;
    std::string line;
    while (std::getline(in, line)) {
        std::istringstream ss(line);
        std::string k;
        if (!(ss >> k) || k[0] == '#') continue;
        if (k == "v") { float3 p{0, 0, 0};  ss >> p.x >> p.y >> p.z;  V.push_back(p); }
        else if (k == "vn") { float3 p{0, 0, 0};  ss >> p.x >> p.y >> p.z;  VN.push_back(p); }
        else if (k == "vt") { float2 p{0, 0};  ss >> p.x >> p.y;  VT.push_back(p); }
        else if (k == "f") {
            std::vector<Corner> cs;
            std::string tok;
            while (ss >> tok) {
                Corner c{0, -1, -1};
                int part = 0;
                size_t pos = 0;
                while (pos <= tok.size() && part < 3) {
                    const size_t slash = tok.find('/', pos);
                    const std::string s = tok.substr(pos, slash == std::string::npos ? std::string::npos : slash - pos);
                    if (!s.empty()) {
                        const int i = atoi(s.c_str());
                        if (part == 0) c.v = fixIndex(i, V.size());
                        else if (part == 1) c.t = fixIndex(i, VT.size());
                        else c.n = fixIndex(i, VN.size());
                    }
                    if (slash == std::string::npos) break;
                    pos = slash + 1;  ++part;
                }
                if (c.v < 0 || c.v >= (int)V.size()) throw std::runtime_error("Could not read OBJ model from " + objFile + " : vertex index out of range");
                cs.push_back(c);
            }
            for (size_t j = 1; j + 1 < cs.size(); ++j) cur.faces.push_back(Face{{cs[0], cs[j], cs[j + 1]}, curMat});
        } else if (k == "o" || k == "g") {
            if (!cur.faces.empty()) shapes.push_back(cur);
            cur = Shape();
        } else if (k == "usemtl") {
            std::string name;  ss >> name;
            auto it = matIndex.find(name);
            curMat = it == matIndex.end() ? -1 : it->second;
        } else if (k == "mtllib") {
            std::string name;  ss >> name;
            try {
                for (auto& m : parseMtl(dir + name)) { matIndex.emplace(m.name, (int)materials.size());  materials.push_back(m); }
            } catch (const std::exception& e) { throw std::runtime_error("Could not read OBJ model from " + objFile + " : " + e.what()); }
        }
    }
    if (!cur.faces.empty()) shapes.push_back(cur);

    Model* model = new Model;
    try {
        for (const Shape& shape : shapes) {
            std::map<std::string, int> knownTextures;
            std::map<int, int> matIds;
            for (const Face& f : shape.faces) matIds[f.mat] = 1;
            for (const auto& mi : matIds) {
                const int mid = mi.first;
                std::map<Corner, int> known;
                TriangleMesh* mesh = new TriangleMesh;
                bool allN = true, allT = true;
                for (const Face& f : shape.faces) {
                    if (f.mat != mid) continue;
                    uint3 tri{0, 0, 0};
                    unsigned* tp = &tri.x;
                    for (int c = 0; c < 3; ++c) {
                        const Corner& key = f.c[c];
                        auto it = known.find(key);
                        if (it == known.end()) {
                            it = known.emplace(key, (int)mesh->vertex.size()).first;
                            mesh->vertex.push_back(V[key.v]);
                            if (key.n >= 0 && key.n < (int)VN.size()) mesh->normal.push_back(VN[key.n]); else allN = false;
                            if (key.t >= 0 && key.t < (int)VT.size()) mesh->texcoord.push_back(VT[key.t]); else allT = false;
                        }
                        tp[c] = (unsigned)it->second;
                    }
                    mesh->index.push_back(tri);
                }
                if (!allN) mesh->normal.clear();
                if (!allT) mesh->texcoord.clear();
                if (mesh->vertex.empty()) { delete mesh;  continue; }
                if (mid >= 0) {
                    const MtlEntry& m = materials[mid];
                    mesh->material.color = {m.kd[0], m.kd[1], m.kd[2]};
                    mesh->material.emission = {m.ke[0], m.ke[1], m.ke[2]};
                    mesh->diffuseTextureID = loadTexture(model, knownTextures, m.mapKd, dir);
                }
                model->meshes.push_back(mesh);
            }
        }
    } catch (...) { delete model;  throw; }
    return model;
}

void addBox(Model* model, Material& mat, const float3& pos, const float3& e) {
    const float3 A{-e.x + pos.x, -e.y + pos.y, e.z + pos.z}, B{e.x + pos.x, -e.y + pos.y, e.z + pos.z};
    const float3 C{e.x + pos.x, e.y + pos.y, e.z + pos.z}, D{-e.x + pos.x, e.y + pos.y, e.z + pos.z};
    const float3 E{-e.x + pos.x, -e.y + pos.y, -e.z + pos.z}, F{e.x + pos.x, -e.y + pos.y, -e.z + pos.z};
    const float3 G{e.x + pos.x, e.y + pos.y, -e.z + pos.z}, H{-e.x + pos.x, e.y + pos.y, -e.z + pos.z};
    const float3 verts[36] = {A, B, C, A, C, D,  E, H, G, E, G, F,  E, A, D, E, D, H,
                              B, F, G, B, G, C,  D, C, G, D, G, H,  E, A, B, E, B, F};
    const float3 normals[6] = {{0, 0, 1}, {0, 0, -1}, {-1, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, -1, 0}};
    TriangleMesh* mesh = new TriangleMesh;
    for (int i = 0; i < 36; ++i) {
        mesh->vertex.push_back(verts[i]);
        mesh->normal.push_back(normals[i / 6]);
        mesh->texcoord.push_back(float2{0.f, 0.f});
    }
    for (unsigned t = 0; t < 12; ++t) mesh->index.push_back(uint3{3 * t, 3 * t + 1, 3 * t + 2});
    mesh->material = mat;
    model->meshes.push_back(mesh);
}
