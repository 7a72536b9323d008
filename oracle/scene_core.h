// ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// Scene container + the oracle's statement of optixTrace semantics:
//   closest hit  = argmin t over ALL triangles with tmin < t < tmax, ties -> lowest global triangle id
//                  (sv4/deviceProgram.cu:194-220, RAY_FLAG_NONE, empty any-hit);
//   occlusion    = exists a triangle hit in (tmin,tmax); with CULL_BACK only triangles FRONT-facing the
//                  ray count, front = dot(cross(v1-v0, v2-v0), dir) < 0  (sv4/deviceProgram.cu:222-246,282-298).
// The ray/triangle test is a fixed fp32 Moeller-Trumbore sequence WITHOUT fused multiply-add; the CUDA
// traversal kernel evaluates the identical sequence (with __fmul_rn/__fadd_rn), so the selected triangle
// and (t,u,v) are BVH-independent and reproducible.  OptiX's own arithmetic is closed: this IS the contract.
// Included after a shading backend (shading_port.h or ref_backend.h) which supplies float3 algebra.
#pragma once
#include <vector>
#include <algorithm>
#include <cstring>
#include <cfloat>
#include <cstdint>

struct OTexture { int w = 0, h = 0; std::vector<uint32_t> px; };

struct OMesh {
    std::vector<float3> vertex;
    std::vector<float2> texcoord;      // may be empty
    std::vector<float3> normal;        // may be empty (read by the Whitted restatement only)
    std::vector<uint3>  index;
    Material material;
    int textureID = -1;
    size_t firstGid = 0;               // global id of this mesh's triangle 0
};

struct OTri { float3 v0, e1, e2; };

struct OBvhNode { float lo[3], hi[3]; int left, right, first, count; };   // count>0 => leaf [first,first+count)

struct OScene {
    std::vector<OMesh> meshes;
    std::vector<OTexture> textures;
    std::vector<OTri> tris;            // by global id (mesh order, then primitive order)
    std::vector<int>  meshOf;          // gid -> mesh
    std::vector<OBvhNode> nodes;
    std::vector<int> bvhOrder;         // leaf slots -> gid
    bool bruteForce = false;
    Probe probe;                        // host pointers
    std::vector<float> probeStore[5];
};

struct OHit { float t, u, v; int gid; };

// ---- the contract ray/triangle test -----------------------------------------------------------------
// returns true and fills (t,u,v) when the ray hits with tmin < t < tmax.  det>0 <=> front-facing.
static inline bool oracleTriTest(const OTri& T, const float3& o, const float3& d, float tmin, float tmax,
                                 bool cullBack, float& t, float& u, float& v) {
    const float px = d.y * T.e2.z - d.z * T.e2.y;
    const float py = d.z * T.e2.x - d.x * T.e2.z;
    const float pz = d.x * T.e2.y - d.y * T.e2.x;
    const float det = (T.e1.x * px + T.e1.y * py) + T.e1.z * pz;
    if (det == 0.0f) return false;
    if (cullBack && det < 0.0f) return false;
    const float sx = o.x - T.v0.x, sy = o.y - T.v0.y, sz = o.z - T.v0.z;
    const float U = (sx * px + sy * py) + sz * pz;
    const float qx = sy * T.e1.z - sz * T.e1.y;
    const float qy = sz * T.e1.x - sx * T.e1.z;
    const float qz = sx * T.e1.y - sy * T.e1.x;
    const float V = (d.x * qx + d.y * qy) + d.z * qz;
    if (det > 0.0f) { if (U < 0.0f || V < 0.0f || U + V > det) return false; }
    else            { if (U > 0.0f || V > 0.0f || U + V < det) return false; }
    const float Tn = (T.e2.x * qx + T.e2.y * qy) + T.e2.z * qz;
    const float tt = Tn / det;
    if (!(tt > tmin && tt < tmax)) return false;
    t = tt;  u = U / det;  v = V / det;
    return true;
}

// ---- conservative slab test (shared statement with the CUDA kernel: far side widened by 2 ulp) --------
static inline bool oracleBoxTest(const float* lo, const float* hi, const float3& o, const float3& inv,
                                 float tmin, float tmax) {
    float t0 = tmin, t1 = tmax;
    const float oo[3] = {o.x, o.y, o.z}, ii[3] = {inv.x, inv.y, inv.z};
    for (int a = 0; a < 3; ++a) {
        float ta = (lo[a] - oo[a]) * ii[a], tb = (hi[a] - oo[a]) * ii[a];
        float tn = fminf(ta, tb), tf = fmaxf(ta, tb);    // NaN (0*inf) is ignored by fmin/fmax
        t0 = fmaxf(t0, tn);  t1 = fminf(t1, tf);
    }
    return t0 <= t1 * 1.0000004f;
}

static void oracleFlatten(OScene& s) {
    s.tris.clear();  s.meshOf.clear();
    for (size_t m = 0; m < s.meshes.size(); ++m) {
        OMesh& M = s.meshes[m];
        M.firstGid = s.tris.size();
        for (const uint3& id : M.index) {
            OTri T;
            T.v0 = M.vertex[id.x];
            T.e1 = M.vertex[id.y] - M.vertex[id.x];
            T.e2 = M.vertex[id.z] - M.vertex[id.x];
            s.tris.push_back(T);
            s.meshOf.push_back((int)m);
        }
    }
}

// ---- host BVH: top-down binned SAH (16 bins), leaves <= 4.  Only an accelerator for the contract above.
struct OBuildRef { float lo[3], hi[3], c[3]; int gid; };

static int oracleBuildRec(OScene& s, std::vector<OBuildRef>& refs, int first, int count, int depth) {
    OBvhNode node;
    float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int a = 0; a < 3; ++a) { node.lo[a] = FLT_MAX; node.hi[a] = -FLT_MAX; }
    for (int i = first; i < first + count; ++i)
        for (int a = 0; a < 3; ++a) {
            node.lo[a] = fminf(node.lo[a], refs[i].lo[a]);  node.hi[a] = fmaxf(node.hi[a], refs[i].hi[a]);
            clo[a] = fminf(clo[a], refs[i].c[a]);  chi[a] = fmaxf(chi[a], refs[i].c[a]);
        }
    node.left = node.right = -1;  node.first = first;  node.count = count;
    const int me = (int)s.nodes.size();
    s.nodes.push_back(node);
    if (count <= 4 || depth > 60) return me;
    const int NB = 16;
    int bestAxis = -1, bestSplit = -1;  float bestCost = FLT_MAX;
    for (int a = 0; a < 3; ++a) {
        float ext = chi[a] - clo[a];
        if (!(ext > 0.0f)) continue;
        int cnt[NB] = {0};  float blo[NB][3], bhi[NB][3];
        for (int b = 0; b < NB; ++b) for (int k = 0; k < 3; ++k) { blo[b][k] = FLT_MAX; bhi[b][k] = -FLT_MAX; }
        for (int i = first; i < first + count; ++i) {
            int b = std::min(NB - 1, (int)((refs[i].c[a] - clo[a]) / ext * NB));
            cnt[b]++;
            for (int k = 0; k < 3; ++k) { blo[b][k] = fminf(blo[b][k], refs[i].lo[k]); bhi[b][k] = fmaxf(bhi[b][k], refs[i].hi[k]); }
        }
        float rArea[NB];  int rCnt[NB];
        float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};  int c = 0;
        for (int b = NB - 1; b > 0; --b) {
            for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], blo[b][k]); hi[k] = fmaxf(hi[k], bhi[b][k]); }
            c += cnt[b];
            float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
            rArea[b] = c ? 2.f * (dx * dy + dy * dz + dz * dx) : 0.f;  rCnt[b] = c;
        }
        for (int k = 0; k < 3; ++k) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; }
        c = 0;
        for (int b = 0; b < NB - 1; ++b) {
            for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], blo[b][k]); hi[k] = fmaxf(hi[k], bhi[b][k]); }
            c += cnt[b];
            if (c == 0 || rCnt[b + 1] == 0) continue;
            float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
            float cost = 2.f * (dx * dy + dy * dz + dz * dx) * c + rArea[b + 1] * rCnt[b + 1];
            if (cost < bestCost) { bestCost = cost; bestAxis = a; bestSplit = b; }
        }
    }
    int mid;
    if (bestAxis < 0) {
        mid = first + count / 2;            // all centroids coincide: split by order
    } else {
        float ext = chi[bestAxis] - clo[bestAxis], base = clo[bestAxis];
        auto it = std::partition(refs.begin() + first, refs.begin() + first + count, [&](const OBuildRef& r) {
            int b = std::min(NB - 1, (int)((r.c[bestAxis] - base) / ext * NB));
            return b <= bestSplit;
        });
        mid = (int)(it - refs.begin());
        if (mid == first || mid == first + count) mid = first + count / 2;
    }
    int l = oracleBuildRec(s, refs, first, mid - first, depth + 1);
    int r = oracleBuildRec(s, refs, mid, first + count - mid, depth + 1);
    s.nodes[me].left = l;  s.nodes[me].right = r;  s.nodes[me].count = 0;
    return me;
}

static void oracleBuild(OScene& s) {
    oracleFlatten(s);
    s.nodes.clear();  s.bvhOrder.clear();
    s.bruteForce = s.tris.size() <= 64;
    if (s.tris.empty()) return;
    std::vector<OBuildRef> refs(s.tris.size());
    for (size_t i = 0; i < s.tris.size(); ++i) {
        const OTri& T = s.tris[i];
        const float3 a = T.v0, b = T.v0 + T.e1, c = T.v0 + T.e2;
        // NB: v0+e1 may differ from the source vertex by 1 ulp; boxes are built from the source vertices.
        const OMesh& M = s.meshes[s.meshOf[i]];
        const uint3 id = M.index[i - M.firstGid];
        const float3 p[3] = {M.vertex[id.x], M.vertex[id.y], M.vertex[id.z]};
        (void)a; (void)b; (void)c;
        OBuildRef& R = refs[i];
        const float xs[3] = {p[0].x, p[1].x, p[2].x}, ys[3] = {p[0].y, p[1].y, p[2].y}, zs[3] = {p[0].z, p[1].z, p[2].z};
        R.lo[0] = std::min({xs[0], xs[1], xs[2]});  R.hi[0] = std::max({xs[0], xs[1], xs[2]});
        R.lo[1] = std::min({ys[0], ys[1], ys[2]});  R.hi[1] = std::max({ys[0], ys[1], ys[2]});
        R.lo[2] = std::min({zs[0], zs[1], zs[2]});  R.hi[2] = std::max({zs[0], zs[1], zs[2]});
        for (int k = 0; k < 3; ++k) R.c[k] = 0.5f * (R.lo[k] + R.hi[k]);
        R.gid = (int)i;
    }
    s.nodes.reserve(refs.size());
    oracleBuildRec(s, refs, 0, (int)refs.size(), 0);
    s.bvhOrder.resize(refs.size());
    for (size_t i = 0; i < refs.size(); ++i) s.bvhOrder[i] = refs[i].gid;
}

// NB on box conservativeness: the MT hit point can lie a few ulp outside the exact vertex box; the
// 2-ulp widening of the far slab plus an absolute pad applied at traversal time covers that.
static inline void oraclePad(const OBvhNode& n, float* lo, float* hi) {
    for (int a = 0; a < 3; ++a) {
        float m = fmaxf(fabsf(n.lo[a]), fabsf(n.hi[a])) * 4e-7f + 1e-30f;
        lo[a] = n.lo[a] - m;  hi[a] = n.hi[a] + m;
    }
}

static bool oracleClosest(const OScene& s, const float3& o, const float3& d, float tmin, float tmax, OHit& best) {
    best.t = tmax;  best.gid = -1;  best.u = best.v = 0.f;
    float t, u, v;
    if (s.bruteForce || s.nodes.empty()) {
        for (size_t g = 0; g < s.tris.size(); ++g)
            if (oracleTriTest(s.tris[g], o, d, tmin, tmax, false, t, u, v))
                if (t < best.t || (t == best.t && (int)g < best.gid)) { best.t = t; best.u = u; best.v = v; best.gid = (int)g; }
        return best.gid >= 0;
    }
    const float3 inv = make_float3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    int stack[128], sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const OBvhNode& n = s.nodes[stack[--sp]];
        float lo[3], hi[3];
        oraclePad(n, lo, hi);
        if (!oracleBoxTest(lo, hi, o, inv, tmin, best.t)) continue;
        if (n.count > 0) {
            for (int i = n.first; i < n.first + n.count; ++i) {
                const int g = s.bvhOrder[i];
                if (oracleTriTest(s.tris[g], o, d, tmin, tmax, false, t, u, v))
                    if (t < best.t || (t == best.t && g < best.gid)) { best.t = t; best.u = u; best.v = v; best.gid = g; }
            }
        } else { stack[sp++] = n.left;  stack[sp++] = n.right; }
    }
    return best.gid >= 0;
}

static bool oracleAnyHit(const OScene& s, const float3& o, const float3& d, float tmin, float tmax, bool cullBack) {
    float t, u, v;
    if (s.bruteForce || s.nodes.empty()) {
        for (size_t g = 0; g < s.tris.size(); ++g)
            if (oracleTriTest(s.tris[g], o, d, tmin, tmax, cullBack, t, u, v)) return true;
        return false;
    }
    const float3 inv = make_float3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    int stack[128], sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const OBvhNode& n = s.nodes[stack[--sp]];
        float lo[3], hi[3];
        oraclePad(n, lo, hi);
        if (!oracleBoxTest(lo, hi, o, inv, tmin, tmax)) continue;
        if (n.count > 0) {
            for (int i = n.first; i < n.first + n.count; ++i)
                if (oracleTriTest(s.tris[s.bvhOrder[i]], o, d, tmin, tmax, cullBack, t, u, v)) return true;
        } else { stack[sp++] = n.left;  stack[sp++] = n.right; }
    }
    return false;
}
