// vrt_pixel.cuh — one pixel of the voxel_volume pass: ray generation, the secondary-ray schedule and shading.
// Follows /root/reference/shader/voxel_volume.frag:80-105 (noise, sky), :205-307 (ambient, shadow, colour,
// reflection stack) and :309-346 (main). The shader's nested calls are flattened into a state machine with ONE
// trace_ray call site, so that lanes of a warp tracing different ray kinds (primary / bounce / AO / shadow)
// still execute the same DDA loop together. Every ray and every float operation of the reference is kept.
#pragma once
#include "vrt_core.cuh"

namespace vrt {

struct PixelOut {
    V3 color;            // before the UNORM8 store
    float depth, mask;
    V3 pos, normal;
    int hx, hy, hz;      // instrumentation: primary mapPos at exit
    uint32_t steps, palette;
    uint32_t rays, fetches;
};

// fragmentNoiseSeq :80-89. texel of gl_FragCoord.xy/512+0.5 under NEAREST/REPEAT = ((px+256)%512, (py+256)%512)
VRT_HD V3 fragment_noise_seq(const SceneDev& sc, uint32_t frame, int px, int py, uint32_t num) {
    const uint32_t offset = num * 32u + frame % 32u;
    // a = (1/g, 1/g^2, 1/g^3), g = 1.22074408460575947536, folded in binary64 then narrowed
    const float ax = 0x1.a36a94p-1f, ay = 0x1.579308p-1f, az = 0x1.197258p-1f;
    float nx = 0.0f, ny = 0.0f, nz = 0.0f;
    if (sc.noise) {
        const uint32_t tx = ((uint32_t)px + 256u) % 512u, ty = ((uint32_t)py + 256u) % 512u;
        const uint8_t* t = sc.noise + 4u * (ty * 512u + tx);
        nx = (float)t[0] / 255.0f; ny = (float)t[1] / 255.0f; nz = (float)t[2] / 255.0f;
    }
    const float fo = (float)offset;
    const float vx = nx + fo * ax, vy = ny + fo * ay, vz = nz + fo * az;
    return mk3(vx - floorf(vx), vy - floorf(vy), vz - floorf(vz));   // mod(v, 1.0)
}
VRT_HD V3 random_dir(const SceneDev& sc, uint32_t frame, int px, int py, uint32_t num) {   // :92-95
    V3 s = fragment_noise_seq(sc, frame, px, py, num);
    return normalize3(mk3(s.x * 2.0f - 1.0f, s.y * 2.0f - 1.0f, s.z * 2.0f - 1.0f));
}
// skyColor :97-105; atan/asin pinned to the correctly rounded binary32 value (evaluate in binary64, narrow)
VRT_HD V3 sky_color(const SceneDev& sc, V3 d) {
    if (!sc.sky) return mk3(0.0f, 0.0f, 0.0f);
    float u = (float)atan2((double)d.z, (double)d.x), v = (float)asin((double)(-d.y));
    u *= 0.1591f; v *= 0.3183f;
    u += 0.5f; v += 0.5f;
    const int W = (int)sc.skyW, H = (int)sc.skyH;
    int ix = f2i(floorf(u * (float)W)), iy = f2i(floorf(v * (float)H));
    ix %= W; if (ix < 0) ix += W;
    iy %= H; if (iy < 0) iy += H;
    const float* t = sc.sky + 4u * ((size_t)iy * W + ix);
    return mk3(t[0], t[1], t[2]);
}

enum { PH_PRIMARY = 0, PH_BOUNCE = 1, PH_AO = 2, PH_SHADOW = 3 };
enum { SUB_BEGIN_LEVEL = 0, SUB_AFTER_AO = 1, SUB_FINISH = 2, SUB_ACCUM = 3 };
constexpr uint32_t MAX_RAY_STEPS = 512;   // :68
constexpr int MAX_REFLECTIONS = 5;        // :69

template <int TRAVERSAL>
VRT_HD void render_pixel(const SceneDev& sc, const FrameParams& fp, int px, int py, PixelOut& o) {
    const vrt_camera& cam = fp.cam;
    const int W = cam.screenSize[0], H = cam.screenSize[1];
    // ---- main :309-325 ----
    const float vsx = ((float)px + 0.5f) / (float)W, vsy = ((float)py + 0.5f) / (float)H;   // screen_quad.vert uv at the pixel centre
    const float spx = vsx * 2.0f - 1.0f, spy = vsy * 2.0f - 1.0f;
    const V3 U = mk3(cam.camRight[0], cam.camRight[1], cam.camRight[2]);
    const V3 Vv = mk3(cam.camUp[0], cam.camUp[1], cam.camUp[2]) * (float)H / (float)W;
    const V3 jit = mk3(cam.cameraJitter[0] / (float)W * -2.0f, cam.cameraJitter[1] / (float)H * 2.0f, 0.0f);
    const V3 rayDir = normalize3(normalize3(mk3(cam.camDir[0], cam.camDir[1], cam.camDir[2])) + spx * U + spy * Vv + jit);
    const V3 camPos = mk3(cam.camPos[0], cam.camPos[1], cam.camPos[2]);
    const V3 lightDir = mk3(fp.lightDir[0], fp.lightDir[1], fp.lightDir[2]);
    const uint32_t aoN = (fp.pass_mask & VRT_PASS_AO) ? fp.aoSamples : 0u;

    RayHit hit0; hit0.material = 0; hit0.pos = hit0.normal = hit0.dir = mk3(0.0f, 0.0f, 0.0f);
    RayHit bounces[MAX_REFLECTIONS];
    RayHit H_; H_ = hit0;                 // hit being shaded
    int phase = PH_PRIMARY, sub = SUB_BEGIN_LEVEL;
    int nb = 0, level = -1;
    uint32_t aoI = 0;
    float ambient = 0.0f, sampleFrac = 0.0f;
    bool shadowed = false;
    V3 reflection = mk3(0.0f, 0.0f, 0.0f), ambientVec = reflection, c = reflection;
    V3 rs = camPos, rd = rayDir;
    uint32_t maxSteps = MAX_RAY_STEPS;
    o.rays = 0; o.fetches = 0;

    for (;;) {
        const TraceRes r = trace_ray<TRAVERSAL>(sc, rs, rd, maxSteps, phase != PH_PRIMARY, phase != PH_PRIMARY ? ROUND_CAP_SECONDARY : ROUND_CAP_PRIMARY);   // the only call site
        o.rays++; o.fetches += r.fetches;
        bool done = false, need_ray = false;

        if (phase == PH_PRIMARY) {
            hit0.material = r.material; hit0.pos = r.pos; hit0.normal = r.normal; hit0.dir = rd;
            o.hx = r.mx; o.hy = r.my; o.hz = r.mz; o.steps = r.steps; o.palette = r.material;
            if (r.material == 0) {                       // :336-345
                o.color = sky_color(sc, rayDir);
                o.depth = 0.0f; o.mask = 0.0f; o.pos = r.pos; o.normal = mk3(0.0f, 0.0f, 0.0f);
                return;
            }
            o.depth = length3(r.pos - camPos); o.mask = 0.9f; o.pos = r.pos; o.normal = r.normal;   // :327-335
            if (!(fp.pass_mask & VRT_PASS_SHADE)) {
                const vrt_material& m = sc.palette[r.material];
                o.color = mk3(m.diffuse[0], m.diffuse[1], m.diffuse[2]);
                return;
            }
            if ((fp.pass_mask & VRT_PASS_REFLECT) && sc.palette[r.material].metallic > 0.0f) {   // :274
                phase = PH_BOUNCE; nb = 0;
                rd = reflect3(hit0.dir, hit0.normal); rs = hit0.pos + hit0.normal * 0.01f; maxSteps = MAX_RAY_STEPS;
                continue;
            }
            level = -1; sub = SUB_BEGIN_LEVEL;
        } else if (phase == PH_BOUNCE) {                  // :282-297
            RayHit& bh = bounces[nb];
            bh.material = r.material; bh.pos = r.pos; bh.normal = r.normal; bh.dir = rd;
            if (r.material == 0 || sc.palette[r.material].metallic <= 0.0f) {
                level = nb; sub = SUB_BEGIN_LEVEL;       // lastIdx = nb
            } else if (++nb == MAX_REFLECTIONS) {
                level = -1; sub = SUB_BEGIN_LEVEL;       // lastIdx stays -1: the reflection sum is skipped
            } else {
                rd = reflect3(bh.dir, bh.normal); rs = bh.pos + bh.normal * 0.01f; maxSteps = MAX_RAY_STEPS;
                continue;
            }
        } else if (phase == PH_AO) {                      // :214-223
            if (r.material != 0) ambient += sampleFrac;
            aoI++;
            if (aoI < aoN) {
                const uint32_t depth = level < 0 ? 0u : (uint32_t)level;
                const V3 dir = H_.normal + random_dir(sc, cam.frame, px, py, aoI + depth * aoN);
                rs = H_.pos + dir * 0.01f; rd = dir; maxSteps = 64;
                continue;
            }
            sub = SUB_AFTER_AO;
        } else {                                          // PH_SHADOW :230-233
            shadowed = r.material != 0;
            sub = SUB_FINISH;
        }

        // advance the shading schedule until another ray is needed or the pixel is finished
        while (!need_ray && !done) {
            const uint32_t depth = level < 0 ? 0u : (uint32_t)level;
            if (sub == SUB_BEGIN_LEVEL) {                 // colorHit :252-264
                if (level < 0) H_ = hit0; else H_ = bounces[level];
                if (H_.material == 0) { c = sky_color(sc, H_.dir); sub = SUB_ACCUM; continue; }
                ambient = 0.0f; aoI = 0;
                if (aoN == 0) { ambient = 1.0f; sub = SUB_AFTER_AO; continue; }       // :209-210
                sampleFrac = 1.0f / (float)aoN;
                const V3 dir = H_.normal + random_dir(sc, cam.frame, px, py, 0u + depth * aoN);
                rs = H_.pos + dir * 0.01f; rd = dir; maxSteps = 64;
                phase = PH_AO; need_ray = true;
            } else if (sub == SUB_AFTER_AO) {             // :226, then isShadowed
                ambientVec = (ambient * fp.ambientIntensity) * sky_color(sc, H_.normal);
                if (fp.pass_mask & VRT_PASS_SHADOW) {
                    rs = H_.pos + H_.normal * 0.01f; rd = lightDir; maxSteps = MAX_RAY_STEPS;
                    phase = PH_SHADOW; need_ray = true;
                } else { shadowed = false; sub = SUB_FINISH; }
            } else if (sub == SUB_FINISH) {               // color :236-248 and the /(depth+1) of :258
                const vrt_material& mat = sc.palette[H_.material];
                V3 diffuse = mk3(0.0f, 0.0f, 0.0f);
                if (!shadowed) {
                    const float diff = fmaxf(dot3(H_.normal, lightDir), 0.0f);
                    diffuse = mk3(diff * fp.lightColor[0], diff * fp.lightColor[1], diff * fp.lightColor[2]) * fp.lightIntensity;
                }
                const V3 specular = reflection * mat.metallic;
                const V3 lit = ((diffuse + specular) + ambientVec) * mk3(mat.diffuse[0], mat.diffuse[1], mat.diffuse[2]);
                c = (lit * 1.0f) / (float)(depth + 1u);
                sub = SUB_ACCUM;
            } else {                                      // SUB_ACCUM: :300-306
                if (level >= 0) { reflection = reflection + c; level--; sub = SUB_BEGIN_LEVEL; }
                else { o.color = c; done = true; }
            }
        }
        if (done) return;
    }
}

}  // namespace vrt
