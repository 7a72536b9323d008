// ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// __raygen__renderFrame (sv4/deviceProgram.cu:389-588, orig/deviceProgram.cu:339-475) and the host launch
// schedule (sv4/SimplePathtracer.cpp:77-217) restated for the CPU.  One call = one optixLaunch.
#pragma once
#include "render_core.h"

// One launch index of one launch.  Returns false when the ring test culled it.
static inline bool oracleRaygen(const OTraceCtx& C, const LaunchParams& params, const OFrame& F,
                                unsigned lx, unsigned ly, bool primaryOnly, OHit* primaryOut) {
    const b200pt_config& cfg = *C.cfg;
    const int w = params.frame.size.x, h = params.frame.size.y;
    const float3 eye = params.camera.eye, U = params.camera.U, V = params.camera.V, W = params.camera.W;
    uint3 idx = make_uint3(lx, ly, 0);
    const unsigned int subframe_index = params.frame.subframe_index;
    int samples_per_launch = params.samples_per_launch;
    int i = samples_per_launch;

    unsigned int seed = tea<4>(idx.y * w + idx.x, subframe_index);               // :407 (unscaled index)
    float3 result = make_float3(0.0f);

    idx = make_uint3(idx.x * params.frame.factor.x + params.frame.offset.x,        // :419
                     idx.y * params.frame.factor.y + params.frame.offset.y,
                     idx.z * params.frame.factor.z + 0u);
    const float3 cf = make_float3((float)params.frame.c.x, (float)params.frame.c.y, 0.0f);
    float range = length(make_float3((float)idx.x, (float)idx.y, (float)idx.z) - cf);          // :421
    if (range < params.frame.r_inner || range > params.frame.r_outer) return false;        // :423

    float3 normal = make_float3(0.f), albedo = make_float3(0.f), alpha = make_float3(0.f);
    float3 backplate = make_float3(0.f);
    do {
        float3 directLight = make_float3(0.0f), indirectLight = make_float3(0.0f);
        RadiancePRD prd;
        prd.radiance = make_float3(0.f);
        prd.alpha = make_float3(0.f);
        prd.seed = seed;
        prd.rand = Random(seed);
        prd.rayEta = 1.0f;
        prd.pathThroughput = make_float3(1.f);
        prd.rayAbsorption = make_float3(0.f);
        prd.bsdfPdf = 1.0f;
        prd.normal = make_float3(0.0f);
        prd.albedo = make_float3(0.0f);
        prd.stateFlags = 0;
        prd.depth = 0;
        prd.rayType = eReflected;
        C.counters->samples++;

        const float jx = rnd(seed);                                              // :459 (x first, then y)
        const float jy = rnd(seed);
        const float dx = 2.0f * ((static_cast<float>(idx.x) + jx) / static_cast<float>(w)) - 1.0f;
        const float dy = 2.0f * ((static_cast<float>(idx.y) + jy) / static_cast<float>(h)) - 1.0f;
        float3 ray_direction = normalize(dx * U + dy * V + W);
        float3 ray_origin = eye;

        if (primaryOnly) {                       // parity hook: the first radiance trace of sample 0
            C.counters->radianceRays++;
            oracleClosest(*C.scene, ray_origin, ray_direction, cfg.tmin, cfg.tmax, *primaryOut);
            return true;
        }
        backplate = make_float3(ProbeEval(C.scene->probe, ProbeDirToUV(ray_direction)));      // :475

        for (;;) {
            prd.radiance = make_float3(0.f);
            OHit hit;
            C.counters->radianceRays++;
            if (oracleClosest(*C.scene, ray_origin, ray_direction, cfg.tmin, cfg.tmax, hit))
                oracleClosestHitProgram(C, hit, ray_origin, ray_direction, &prd);
            else {
                oracleMissProgram(C, ray_direction, &prd);
                // the reference's loop drops a DONE segment's radiance (:495); the opt-in miss term is kept, also for the last
                // segment (depth == max_depth), which is the MIS partner of the NEE sample taken at depth max_depth - 1
                if (cfg.mis_on_miss && prd.depth > 0) indirectLight += prd.radiance;
            }
            if (prd.depth == 0.f) { normal += prd.normal;  albedo += prd.albedo; }
            if ((prd.stateFlags & RAY_STATE_FLAGS_DONE) || prd.depth >= cfg.max_depth) break;   // :495
            if (prd.depth == 0) directLight += prd.radiance;
            else indirectLight += prd.radiance;
            if (prd.stateFlags & RAY_STATE_FLAGS_RR_KILLED) break;                              // opt-in Russian roulette
            ++prd.depth;
            ray_origin = prd.origin;
            ray_direction = prd.direction;
        }
        result += directLight + indirectLight;
        alpha += prd.alpha;
    } while (--i);

    normal /= static_cast<float>(samples_per_launch);
    albedo /= static_cast<float>(samples_per_launch);
    alpha /= static_cast<float>(samples_per_launch);

    for (int fi = 0; fi < params.frame.fillSize; ++fi) {                          // :526-587
        for (int fj = 0; fj < params.frame.fillSize; ++fj) {
            unsigned int ix = lx * params.frame.factor.x + fi + params.frame.offset.x;
            unsigned int iy = ly * params.frame.factor.y + fj + params.frame.offset.y;
            ix = ix > (unsigned)(w - 1) ? (unsigned)(w - 1) : ix;                 // clamp(uint2, 0, (w-1,h-1))
            iy = iy > (unsigned)(h - 1) ? (unsigned)(h - 1) : iy;
            const unsigned int image_index = iy * w + ix;
            const float spl = static_cast<float>(params.samples_per_launch);
            float3 color = (backplate * spl) * (make_float3(1.0f) - alpha) + result;
            float3 accum_color = color / spl;
            const bool accumulate = subframe_index > 0 && !params.frame.redraw;   // sv4 :545 (orig :459 has no redraw; it is 0 there)
            if (cfg.accum_mode == B200PT_ACCUM_SUM) {
                // multi-GPU sample split: associative sum of per-pass clamped colours (pass 0 unclamped in the
                // reference's running mean, so the first pass is left unclamped here too).
                float3 c = subframe_index > 0 ? clamp(accum_color, make_float3(0.0f), make_float3(10.0f)) : accum_color;
                float4 prev = F.accum[image_index];
                F.accum[image_index] = make_float4(prev.x + c.x, prev.y + c.y, prev.z + c.z, prev.w + 1.0f);
                continue;
            }
            if (accumulate) {
                accum_color = clamp(accum_color, make_float3(0.0f), make_float3(10.0f));
                const float a = 1.0f / static_cast<float>(subframe_index + 1);
                const float3 prev = make_float3(F.accum[image_index]);
                accum_color = lerp(prev, accum_color, a);
            }
            F.accum[image_index] = make_float4(accum_color.x, accum_color.y, accum_color.z, 1.0f);
            uchar4 px;
            if (cfg.display == B200PT_DISPLAY_REINHARD)
                px = make_color(reinhardToneMap(accum_color * cfg.exposure, 1.0f));       // :557,568
            else
                px = make_color(accum_color);                                             // orig :467
            F.frame[image_index] = (uint32_t)px.x | ((uint32_t)px.y << 8) | ((uint32_t)px.z << 16) | ((uint32_t)px.w << 24);
            if (cfg.write_aovs && F.normal) {                                             // orig :472-474
                F.normal[image_index] = make_float4(normal.x, normal.y, normal.z, 1.0f);
                F.color[image_index] = make_float4(accum_color.x, accum_color.y, accum_color.z, 1.0f);
                F.albedo[image_index] = make_float4(albedo.x, albedo.y, albedo.z, 1.0f);
            }
        }
    }
    return true;
}

// One optixLaunch(launchW, launchH, 1).  Rows are distributed over OpenMP threads.  Where several launch
// indices write the same pixel (border clamp, :534) the reference races; here the highest launch row/column
// wins deterministically only in the serial build — tests avoid such gaze positions.
static void oracleLaunch(const OScene& scene, const LaunchParams& params, const b200pt_config& cfg, const OFrame& F,
                         int launchW, int launchH, OCounters& total) {
    #pragma omp parallel
    {
        OCounters local;
        OTraceCtx C{&scene, &cfg, &local};
        #pragma omp for schedule(dynamic, 1)
        for (int ly = 0; ly < launchH; ++ly) {
            const int bands = cfg.tile_count > 1 ? cfg.tile_count : 1;
            if (bands > 1 && ((ly / 32) % bands) != cfg.tile_rank) continue;
            for (int lx = 0; lx < launchW; ++lx) oracleRaygen(C, params, F, (unsigned)lx, (unsigned)ly, false, nullptr);
        }
        #pragma omp critical
        { total.samples += local.samples;  total.radianceRays += local.radianceRays;  total.shadowRays += local.shadowRays; }
    }
}

// SampleRenderer::render() — mutates *params the way the reference leaves its public launchParams.
static void oracleRenderFrame(const OScene& scene, LaunchParams& P, const b200pt_config& cfg, const OFrame& F, OCounters& total) {
    if (P.frame.size.x == 0) return;                                             // :81
    const int W = P.frame.size.x, H = P.frame.size.y;
    if (cfg.schedule == B200PT_SCHEDULE_AS_GIVEN) {                               // orig/SimplePathtracer.cpp:73-97
        oracleLaunch(scene, P, cfg, F, cfg.launch_w ? cfg.launch_w : W, cfg.launch_h ? cfg.launch_h : H, total);
        if (cfg.increment_subframe) P.frame.subframe_index++;
        return;
    }
    if (cfg.schedule == B200PT_SCHEDULE_UNIFORM) {                                // :85-131 (fields set BEFORE the launch)
        if (!cfg.accumulate) P.frame.subframe_index = 0;
        P.frame.factor = make_uint3(1, 1, 1);
        P.frame.fillSize = 1;
        P.frame.r_outer = 1000000000;
        P.frame.r_inner = 0;
        if (cfg.uniform_spl > 0) P.samples_per_launch = cfg.uniform_spl;
        P.frame.offset = make_uint2(0, 0);
        P.frame.redraw = 0;
        P.viewportSize.x = W;  P.viewportSize.y = H;
        int temp_frame = P.frame.subframe_index;
        oracleLaunch(scene, P, cfg, F, W, H, total);
        P.frame.subframe_index = temp_frame;
        if (cfg.increment_subframe) P.frame.subframe_index++;
        return;
    }
    // foveated 3-pass schedule, :136-214
    const int inner_radius = cfg.inner_radius, outer_radius = cfg.outer_radius;
    P.frame.factor = make_uint3(4, 4, 1);
    P.frame.fillSize = 4;
    P.frame.r_outer = 1000000000;
    P.frame.r_inner = outer_radius;
    P.samples_per_launch = 1;
    P.frame.offset = make_uint2(0, 0);
    P.frame.redraw = 0;
    oracleLaunch(scene, P, cfg, F, W / 4, H / 4, total);
    int temp_frame = P.frame.subframe_index;

    P.frame.subframe_index = 0;
    P.frame.factor = make_uint3(2, 2, 1);
    P.frame.fillSize = 2;
    P.frame.r_outer = outer_radius + 2;
    P.frame.r_inner = inner_radius;
    P.samples_per_launch = 2;
    P.frame.offset = make_uint2(P.frame.c.x - (outer_radius + 2), P.frame.c.y - (outer_radius + 2));
    P.frame.redraw = 1;
    oracleLaunch(scene, P, cfg, F, (int)P.frame.r_outer, (int)P.frame.r_outer, total);

    P.frame.factor = make_uint3(1, 1, 1);
    P.frame.fillSize = 1;
    P.frame.r_outer = inner_radius + 1;
    P.frame.r_inner = 0;
    P.samples_per_launch = 8;
    P.frame.offset = make_uint2(P.frame.c.x - (inner_radius + 1), P.frame.c.y - (inner_radius + 1));
    P.frame.redraw = 1;
    oracleLaunch(scene, P, cfg, F, (int)(P.frame.r_outer * 2), (int)(P.frame.r_outer * 2), total);

    P.frame.subframe_index = temp_frame;
    if (cfg.increment_subframe) P.frame.subframe_index++;
}
