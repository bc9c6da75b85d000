// vrt_wave.cuh — the voxel_volume pass cut into ray waves (see WaveBuf in vrt_types.h). These are the per-ray /
// per-record stage functions, written for the device and compilable for the host (tests/host_harness runs them
// through a serial queue emulation against the oracle). Each follows the same lines of
// /root/reference/shader/voxel_volume.frag as the per-pixel schedule in vrt_pixel.cuh does; the float operation
// sequence of every ray and of the shading is identical, only the order BETWEEN rays differs.
#pragma once
#include <string.h>
#include "vrt_pixel.cuh"

namespace vrt {

VRT_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
VRT_HD RecQ mkq(V3 v, uint32_t w) { RecQ q; q.x = v.x; q.y = v.y; q.z = v.z; q.w = w; return q; }
VRT_HD V3 q3(const RecQ& q) { return mk3(q.x, q.y, q.z); }

// plane index of pixel (px,py): row-major frame, or tile-major compact shard (FrameParams / VRT_TILE)
VRT_HD size_t plane_index(const FrameParams& fp, int px, int py) {
    if (fp.world == 1) return (size_t)py * fp.cam.screenSize[0] + px;
    const int t = (py / VRT_TILE) * fp.tilesX + (px / VRT_TILE);
    return (size_t)(t / fp.world) * (VRT_TILE * VRT_TILE) + (py % VRT_TILE) * VRT_TILE + (px % VRT_TILE);
}

// where pixel (px,py) of the assembled frame lies in the gathered buffer (world shard buffers back to back, rank-major,
// each rankStridePixels long): the inverse of plane_index over all ranks (vrt_assemble / detile_kernel)
VRT_HD size_t gathered_index(int px, int py, int tilesX, int world, size_t rankStridePixels) {
    const int t = (py / VRT_TILE) * tilesX + (px / VRT_TILE);
    return (size_t)(t % world) * rankStridePixels + (size_t)(t / world) * (VRT_TILE * VRT_TILE) + (py % VRT_TILE) * VRT_TILE + (px % VRT_TILE);
}

enum { WAVE_PIXEL_FINAL = 0, WAVE_PIXEL_SHADE = 1, WAVE_PIXEL_SHADE_AND_BOUNCE = 2 };

// main :309-345 up to the first colorMainRay decision. Fills the G-buffer fields of `o`; `o.color` only when the
// pixel is final (sky, or albedo when SHADE is off).
template <int TRAVERSAL>
VRT_HD int wave_primary(const SceneDev& sc, const FrameParams& fp, int px, int py, PixelOut& o, V3& rayDirOut) {
    const vrt_camera& cam = fp.cam;
    const int W = cam.screenSize[0], H = cam.screenSize[1];
    const float vsx = ((float)px + 0.5f) / (float)W, vsy = ((float)py + 0.5f) / (float)H;
    const float spx = vsx * 2.0f - 1.0f, spy = vsy * 2.0f - 1.0f;
    const V3 U = mk3(cam.camRight[0], cam.camRight[1], cam.camRight[2]);
    const V3 Vv = mk3(cam.camUp[0], cam.camUp[1], cam.camUp[2]) * (float)H / (float)W;
    const V3 jit = mk3(cam.cameraJitter[0] / (float)W * -2.0f, cam.cameraJitter[1] / (float)H * 2.0f, 0.0f);
    const V3 rayDir = normalize3(normalize3(mk3(cam.camDir[0], cam.camDir[1], cam.camDir[2])) + spx * U + spy * Vv + jit);
    const V3 camPos = mk3(cam.camPos[0], cam.camPos[1], cam.camPos[2]);
    rayDirOut = rayDir;
    const TraceRes r = trace_ray<TRAVERSAL>(sc, camPos, rayDir, MAX_RAY_STEPS);
    o.rays = 1; o.fetches = r.fetches;
    o.hx = r.mx; o.hy = r.my; o.hz = r.mz; o.steps = r.steps; o.palette = r.material;
    o.pos = r.pos; o.normal = r.normal;
    if (r.material == 0) {                                   // :336-345
        o.color = sky_color(sc, rayDir);
        o.depth = 0.0f; o.mask = 0.0f; o.normal = mk3(0.0f, 0.0f, 0.0f);
        return WAVE_PIXEL_FINAL;
    }
    o.depth = length3(r.pos - camPos); o.mask = 0.9f;        // :327-335
    if (!(fp.pass_mask & VRT_PASS_SHADE)) {
        const vrt_material& m = sc.palette[r.material];
        o.color = mk3(m.diffuse[0], m.diffuse[1], m.diffuse[2]);
        return WAVE_PIXEL_FINAL;
    }
    if ((fp.pass_mask & VRT_PASS_REFLECT) && sc.palette[r.material].metallic > 0.0f) return WAVE_PIXEL_SHADE_AND_BOUNCE;   // :274
    return WAVE_PIXEL_SHADE;
}

// one reflection bounce :282-297 from the hit (pos, normal, dir)
template <int TRAVERSAL>
VRT_HD TraceRes wave_bounce(const SceneDev& sc, V3 pos, V3 normal, V3 dir, V3& rd, uint32_t roundCap = ROUND_CAP_SECONDARY) {
    rd = reflect3(dir, normal);
    const V3 rs = pos + normal * 0.01f;
    return trace_ray<TRAVERSAL>(sc, rs, rd, MAX_RAY_STEPS, true, roundCap);   // a miss carries no state (traceRay :180-195): see lin_round ANYHIT
}

// ray i of a shaded hit: i < aoN -> AO ray i (:214-223, 64 steps), i == aoN -> shadow ray (:230-233, 512 steps)
VRT_HD void wave_anyhit_ray(const SceneDev& sc, const FrameParams& fp, V3 pos, V3 normal, int px, int py, uint32_t depth, uint32_t i,
                            uint32_t aoN, V3& rs, V3& rd, uint32_t& maxSteps) {
    if (i < aoN) {
        const V3 dir = normal + random_dir(sc, fp.cam.frame, px, py, i + depth * aoN);
        rs = pos + dir * 0.01f; rd = dir; maxSteps = 64;
    } else {
        rd = mk3(fp.lightDir[0], fp.lightDir[1], fp.lightDir[2]);
        rs = pos + normal * 0.01f; maxSteps = MAX_RAY_STEPS;
    }
}
template <int TRAVERSAL>
VRT_HD bool wave_anyhit(const SceneDev& sc, const FrameParams& fp, V3 pos, V3 normal, int px, int py, uint32_t depth, uint32_t i,
                        uint32_t aoN, uint32_t& fetches) {
    V3 rs, rd; uint32_t maxSteps;
    wave_anyhit_ray(sc, fp, pos, normal, px, py, depth, i, aoN, rs, rd, maxSteps);
    const TraceRes r = trace_ray<TRAVERSAL>(sc, rs, rd, maxSteps, true, ROUND_CAP_SECONDARY);
    fetches = r.fetches;
    return r.material != 0;
}

// ambient of calcAmbient :205-227 from the number of AO rays that hit: `ambient += sampleFrac` once per hit
VRT_HD float wave_ambient(uint32_t aoN, uint32_t hits) {
    if (aoN == 0) return 1.0f;
    const float sampleFrac = 1.0f / (float)aoN;
    float ambient = 0.0f;
    for (uint32_t k = 0; k < hits; k++) ambient += sampleFrac;
    return ambient;
}

// skyColor of a surface normal. Normals are normalize(mask * -step): almost always an axis vector whose components are exactly
// +-0 or +-1 (the sign of a zero matters to atan2), so the resolve kernel tabulates skyColor for the 64 such vectors once per block
// (skyTab, 3 floats each, computed by sky_color itself) instead of two binary64 inverse-trig calls per shaded record. Any other
// normal (two axes stepped at once on a tie) goes through sky_color directly.
VRT_HD uint32_t sky_axis_code(float c) { const uint32_t b = f2u(c); return (b >> 31 << 1) | ((b & 0x7fffffffu) == 0x3f800000u ? 1u : 0u); }
VRT_HD bool sky_axis_ok(float c) { const uint32_t a = f2u(c) & 0x7fffffffu; return a == 0u || a == 0x3f800000u; }
VRT_HD V3 sky_of_normal(const SceneDev& sc, V3 n, const float* skyTab) {
    if (skyTab && sky_axis_ok(n.x) && sky_axis_ok(n.y) && sky_axis_ok(n.z)) {
        const float* t = skyTab + 3u * (sky_axis_code(n.x) | (sky_axis_code(n.y) << 2) | (sky_axis_code(n.z) << 4));
        return mk3(t[0], t[1], t[2]);
    }
    return sky_color(sc, n);
}
VRT_HD V3 sky_axis_vector(uint32_t code) {
    const float v[4] = {0.0f, 1.0f, -0.0f, -1.0f};
    return mk3(v[code & 3u], v[(code >> 2) & 3u], v[(code >> 4) & 3u]);
}

// colorHit :252-264 (+ color :236-248) of one record
VRT_HD V3 wave_shade(const SceneDev& sc, const FrameParams& fp, uint32_t material, V3 normal, V3 dir, float ambient, bool shadowed,
                     V3 reflection, uint32_t depth, const float* skyTab = nullptr) {
    if (material == 0) return sky_color(sc, dir);
    const V3 lightDir = mk3(fp.lightDir[0], fp.lightDir[1], fp.lightDir[2]);
    const V3 ambientVec = (ambient * fp.ambientIntensity) * sky_of_normal(sc, normal, skyTab);
    const vrt_material& mat = sc.palette[material];
    V3 diffuse = mk3(0.0f, 0.0f, 0.0f);
    if (!shadowed) {
        const float diff = fmaxf(dot3(normal, lightDir), 0.0f);
        diffuse = mk3(diff * fp.lightColor[0], diff * fp.lightColor[1], diff * fp.lightColor[2]) * fp.lightIntensity;
    }
    const V3 specular = reflection * mat.metallic;
    const V3 lit = ((diffuse + specular) + ambientVec) * mk3(mat.diffuse[0], mat.diffuse[1], mat.diffuse[2]);
    return (lit * 1.0f) / (float)(depth + 1u);
}

// shading inputs of record `idx` from the any-hit result bytes
VRT_HD void wave_results(const WaveBuf& wb, const FrameParams& fp, uint32_t idx, uint32_t aoN, float& ambient, bool& shadowed) {
    uint32_t hits = 0;
    for (uint32_t i = 0; i < aoN; i++) hits += wb.res[(size_t)i * wb.recCap + idx] ? 1u : 0u;
    ambient = wave_ambient(aoN, hits);
    shadowed = (fp.pass_mask & VRT_PASS_SHADOW) ? wb.res[(size_t)aoN * wb.recCap + idx] != 0 : false;
}

// colorMainRay :298-306 for the primary record `root`: walk the reflection chain from its deepest record up
VRT_HD V3 wave_resolve(const SceneDev& sc, const FrameParams& fp, const WaveBuf& wb, uint32_t root, uint32_t aoN, const float* skyTab = nullptr) {
    const RecQ r1 = wb.rec[3 * (size_t)root + 1], r2 = wb.rec[3 * (size_t)root + 2];
    V3 reflection = mk3(0.0f, 0.0f, 0.0f);
    uint32_t k = r2.w;
    while (k != WAVE_NONE && k != root) {
        const RecQ b1 = wb.rec[3 * (size_t)k + 1], b2 = wb.rec[3 * (size_t)k + 2];
        const uint32_t material = b1.w & 0xffu, depth = b1.w >> 8;
        float ambient = 0.0f; bool shadowed = false;
        if (material != 0) wave_results(wb, fp, k, aoN, ambient, shadowed);
        reflection = reflection + wave_shade(sc, fp, material, q3(b1), q3(b2), ambient, shadowed, reflection, depth, skyTab);
        k = b2.w;
    }
    float ambient; bool shadowed;
    wave_results(wb, fp, root, aoN, ambient, shadowed);
    return wave_shade(sc, fp, r1.w & 0xffu, q3(r1), q3(r2), ambient, shadowed, reflection, 0u, skyTab);
}

}  // namespace vrt
