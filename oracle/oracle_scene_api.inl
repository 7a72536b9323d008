// ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Scene-level C entry points (included by oracle_api.cpp).

// Host reference for the integer BVH-build stages (bit-exact contract with csrc/bvh_build.cu):
//   scene box   = min/max over every triangle's three vertices
//   centroid    = 0.5f * (triLo + triHi)                 (triangle AABB centre)
//   q           = (centroid - sceneLo) * (1.0f / extent) * 2097152.0f, extent==0 -> scale 0
//   cell        = min(max((int)q, 0), 2097151)            (21 bits per axis)
//   code        = interleave(x)<<2 | interleave(y)<<1 | interleave(z)   (63-bit Morton)
//   order       = stable sort of triangle ids by code
static inline uint64_t oracleExpand21(uint32_t v) {
    uint64_t x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

extern "C" {

OScene* oracle_scene_create() { OScene* s = new OScene();  memset(&s->probe, 0, sizeof(Probe));  return s; }
void oracle_scene_destroy(OScene* s) { delete s; }
int oracle_scene_add_texture(OScene* s, const uint32_t* rgba8, int w, int h) {
    OTexture t;  t.w = w;  t.h = h;  t.px.assign(rgba8, rgba8 + (size_t)w * h);
    s->textures.push_back(std::move(t));
    return (int)s->textures.size() - 1;
}
int oracle_scene_add_mesh(OScene* s, const float* vtx, size_t nv, const float* texcoord, const uint32_t* idx, size_t ntri,
                          const void* material104, int textureID) {
    OMesh m;
    m.vertex.resize(nv);
    for (size_t i = 0; i < nv; ++i) m.vertex[i] = make_float3(vtx[3 * i], vtx[3 * i + 1], vtx[3 * i + 2]);
    if (texcoord) { m.texcoord.resize(nv);  for (size_t i = 0; i < nv; ++i) m.texcoord[i] = make_float2(texcoord[2 * i], texcoord[2 * i + 1]); }
    m.index.resize(ntri);
    for (size_t i = 0; i < ntri; ++i) m.index[i] = make_uint3(idx[3 * i], idx[3 * i + 1], idx[3 * i + 2]);
    memcpy(&m.material, material104, sizeof(Material));
    m.textureID = textureID;
    s->meshes.push_back(std::move(m));
    return (int)s->meshes.size() - 1;
}
void oracle_scene_set_mesh_normals(OScene* s, int meshID, const float* normal) {
    OMesh& m = s->meshes[meshID];
    m.normal.resize(m.vertex.size());
    for (size_t i = 0; i < m.vertex.size(); ++i) m.normal[i] = make_float3(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]);
}
void oracle_scene_build(OScene* s, int force_brute_force) {
    oracleBuild(*s);
    if (force_brute_force) s->bruteForce = true;
}
void oracle_scene_set_probe(OScene* s, int w, int h, const float* data, const float* pdfX, const float* cdfX,
                            const float* pdfY, const float* cdfY) {
    const size_t n = (size_t)w * h;
    s->probeStore[0].assign(data, data + 4 * n);   s->probeStore[1].assign(pdfX, pdfX + n);
    s->probeStore[2].assign(cdfX, cdfX + n);       s->probeStore[3].assign(pdfY, pdfY + h);
    s->probeStore[4].assign(cdfY, cdfY + h);
    s->probe = mkProbe(w, h, s->probeStore[0].data(), s->probeStore[1].data(), s->probeStore[2].data(),
                       s->probeStore[3].data(), s->probeStore[4].data());
}
uint64_t oracle_scene_num_triangles(OScene* s) { return s->tris.size(); }

void oracle_morton(OScene* s, uint64_t* codes, uint32_t* order, float* bounds6) {
    const size_t n = s->tris.size();
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    std::vector<float> cen(3 * n);
    for (size_t g = 0; g < n; ++g) {
        const OMesh& M = s->meshes[s->meshOf[g]];
        const uint3 id = M.index[g - M.firstGid];
        const float3 p[3] = {M.vertex[id.x], M.vertex[id.y], M.vertex[id.z]};
        const float c[3][3] = {{p[0].x, p[1].x, p[2].x}, {p[0].y, p[1].y, p[2].y}, {p[0].z, p[1].z, p[2].z}};
        for (int a = 0; a < 3; ++a) {
            float tl = fminf(fminf(c[a][0], c[a][1]), c[a][2]), th = fmaxf(fmaxf(c[a][0], c[a][1]), c[a][2]);
            lo[a] = fminf(lo[a], tl);  hi[a] = fmaxf(hi[a], th);
            cen[3 * g + a] = 0.5f * (tl + th);
        }
    }
    float scale[3];
    for (int a = 0; a < 3; ++a) { float e = hi[a] - lo[a];  scale[a] = e > 0.f ? 1.0f / e : 0.f;  bounds6[a] = lo[a];  bounds6[3 + a] = hi[a]; }
    for (size_t g = 0; g < n; ++g) {
        uint32_t cell[3];
        for (int a = 0; a < 3; ++a) {
            float q = (cen[3 * g + a] - lo[a]) * scale[a] * 2097152.0f;
            int qi = (int)q;
            cell[a] = (uint32_t)(qi < 0 ? 0 : (qi > 2097151 ? 2097151 : qi));
        }
        codes[g] = (oracleExpand21(cell[0]) << 2) | (oracleExpand21(cell[1]) << 1) | oracleExpand21(cell[2]);
    }
    for (size_t g = 0; g < n; ++g) order[g] = (uint32_t)g;
    std::stable_sort(order, order + n, [&](uint32_t a, uint32_t b) { return codes[a] < codes[b]; });
}

// SampleRenderer::render() on host buffers.  counters_out: samples, radiance rays, shadow rays.  Returns seconds.
double oracle_render(OScene* s, LaunchParams* params, const b200pt_config* cfg, float* accum, uint32_t* frame,
                     float* normal, float* albedo, float* color, uint64_t* counters_out) {
    OFrame F{(float4*)accum, frame, (float4*)normal, (float4*)albedo, (float4*)color};
    OCounters total;
    auto t0 = std::chrono::steady_clock::now();
    oracleRenderFrame(*s, *params, *cfg, F, total);
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (counters_out) { counters_out[0] = total.samples;  counters_out[1] = total.radianceRays;  counters_out[2] = total.shadowRays; }
    return sec;
}

// Bounded-sample variant for the CPU baseline: renders only launch rows [row0, row1) of a uniform launch.
double oracle_render_rows(OScene* s, const LaunchParams* params, const b200pt_config* cfg, float* accum, uint32_t* frame,
                          int row0, int row1, uint64_t* counters_out) {
    OFrame F{(float4*)accum, frame, nullptr, nullptr, nullptr};
    OCounters total;
    const int W = params->frame.size.x;
    auto t0 = std::chrono::steady_clock::now();
    #pragma omp parallel
    {
        OCounters local;
        OTraceCtx C{s, cfg, &local};
        #pragma omp for schedule(dynamic, 1)
        for (int ly = row0; ly < row1; ++ly)
            for (int lx = 0; lx < W; ++lx) oracleRaygen(C, *params, F, (unsigned)lx, (unsigned)ly, false, nullptr);
        #pragma omp critical
        { total.samples += local.samples;  total.radianceRays += local.radianceRays;  total.shadowRays += local.shadowRays; }
    }
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (counters_out) { counters_out[0] = total.samples;  counters_out[1] = total.radianceRays;  counters_out[2] = total.shadowRays; }
    return sec;
}

// First radiance optixTrace of sample 0 for every launch index of a (launchW x launchH) launch.
void oracle_trace_primary(OScene* s, const LaunchParams* params, const b200pt_config* cfg, int launchW, int launchH,
                          int32_t* prim, int32_t* mesh, float* t, float* u, float* v) {
    #pragma omp parallel
    {
        OCounters local;
        OTraceCtx C{s, cfg, &local};
        OFrame F{nullptr, nullptr, nullptr, nullptr, nullptr};
        #pragma omp for schedule(dynamic, 4)
        for (int ly = 0; ly < launchH; ++ly)
            for (int lx = 0; lx < launchW; ++lx) {
                const size_t i = (size_t)ly * launchW + lx;
                OHit h;  h.gid = -1;  h.t = 0.f;  h.u = h.v = 0.f;
                bool live = oracleRaygen(C, *params, F, (unsigned)lx, (unsigned)ly, true, &h);
                if (!live || h.gid < 0) { prim[i] = -1;  mesh[i] = live ? -1 : -2;  t[i] = 0.f;  u[i] = v[i] = 0.f;  continue; }
                const int m = s->meshOf[h.gid];
                mesh[i] = m;  prim[i] = h.gid - (int)s->meshes[m].firstGid;  t[i] = h.t;  u[i] = h.u;  v[i] = h.v;
            }
    }
}

void oracle_trace_rays(OScene* s, const float* org, const float* dir, size_t n, float tmin, float tmax, int any_hit,
                       int shadow_mode, int32_t* prim, int32_t* mesh, float* t, float* u, float* v) {
    #pragma omp parallel for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; ++i) {
        const float3 o = ld3(org, (int)i), d = ld3(dir, (int)i);
        if (any_hit) {
            prim[i] = oracleAnyHit(*s, o, d, tmin, tmax, shadow_mode == B200PT_SHADOW_CULL_BACK) ? 1 : 0;
            continue;
        }
        OHit h;
        if (oracleClosest(*s, o, d, tmin, tmax, h)) {
            const int m = s->meshOf[h.gid];
            mesh[i] = m;  prim[i] = h.gid - (int)s->meshes[m].firstGid;  t[i] = h.t;  u[i] = h.u;  v[i] = h.v;
        } else { prim[i] = -1;  mesh[i] = -1;  t[i] = 0.f;  u[i] = v[i] = 0.f; }
    }
}


// CPU restatement of 04HelloRaytracing/deviceProgram.cu:66-243 (raygen, closest hit with its shadow trace, miss, shadow any-hit)
// over this oracle's traversal contract.  cam = position, direction, horizontal, vertical (SampleRenderer.cpp:125-138).
void oracle_render_whitted(OScene* sp, int W, int H, const float* cam12, const float* light3, uint32_t* colorBuffer) {
    const OScene& S = *sp;
    const float3 position = ld3(cam12, 0), direction = ld3(cam12, 1), horizontal = ld3(cam12, 2), vertical = ld3(cam12, 3);
    const float3 lightPos = ld3(light3, 0);
    #pragma omp parallel for schedule(dynamic, 4)
    for (int iy = 0; iy < H; ++iy)
        for (int ix = 0; ix < W; ++ix) {
            float3 prd = make_float3(0.f, 0.f, 0.f);
            const float2 screen = make_float2((ix + .5f) / (float)W, (iy + .5f) / (float)H);                 // :213-214
            const float3 rayDir = normalize(direction + (screen.x - 0.5f) * horizontal + (screen.y - 0.5f) * vertical);
            OHit hit;
            if (oracleClosest(S, position, rayDir, 0.f, 1e20f, hit)) {                                     // :221-232
                const int meshID = S.meshOf[hit.gid];
                const OMesh& mesh = S.meshes[meshID];
                const uint3 index = mesh.index[hit.gid - (int)mesh.firstGid];
                const float u = hit.u, v = hit.v;
                const float3 A = mesh.vertex[index.x], B = mesh.vertex[index.y], C = mesh.vertex[index.z];
                float3 Ng = cross(B - A, C - A);                                                           // :85
                float3 Ns = !mesh.normal.empty()
                    ? ((1.f - u - v) * mesh.normal[index.x] + u * mesh.normal[index.y] + v * mesh.normal[index.z]) : Ng;
                if (dot(rayDir, Ng) > 0.f) Ng = -Ng;                                                       // :97-102
                Ng = normalize(Ng);
                if (dot(Ng, Ns) < 0.f) Ns = Ns - (2.f * dot(Ng, Ns)) * Ng;
                Ns = normalize(Ns);
                float3 diffuseColor = mesh.material.color;                                                 // :108
                if (mesh.textureID >= 0 && !mesh.texcoord.empty()) {
                    const float2 tc = (1.f - u - v) * mesh.texcoord[index.x] + u * mesh.texcoord[index.y] + v * mesh.texcoord[index.z];
                    diffuseColor *= make_float3(oracleTex2D(S.textures[mesh.textureID], tc.x, tc.y));
                }
                const float3 surfPos = (1.f - u - v) * A + u * B + v * C;                                  // :121-124
                const float3 lightDir = lightPos - surfPos;
                // shadow ray: first hit terminates, no culling (:131-153); any hit -> 0, miss -> 1 (:177-191)
                const bool occluded = oracleAnyHit(S, surfPos + 1e-3f * Ng, lightDir, 1e-3f, 1.f - 1e-3f, false);
                const float3 lightVisibility = occluded ? make_float3(0.f) : make_float3(1.f);
                const float cosDN = 0.1f + .8f * fabsf(dot(rayDir, Ns));                                   // :158-160
                const float3 lv = .8f * lightVisibility;
                const float3 shade = make_float3(.1f + (.2f + lv.x) * cosDN, .1f + (.2f + lv.y) * cosDN, .1f + (.2f + lv.z) * cosDN);
                prd = shade * diffuseColor;                                                                // :163
            }
            const int r = int(255.99f * prd.x), g = int(255.99f * prd.y), b = int(255.99f * prd.z);         // :232-242
            colorBuffer[ix + (size_t)iy * W] = 0xff000000u | (r << 0) | (g << 8) | (b << 16);
        }
}

}  // extern "C"
