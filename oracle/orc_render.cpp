// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.h). CPU restatement of
// /root/reference/shader/voxel_volume.frag (all line numbers below refer to that file unless noted).
// PARITY UNPINNED by the reference (no golden vectors exist); pinned decisions: SURVEY.md F5,F8,F9.
#include "vrt_oracle.h"
#include "orc_math.h"
#include <omp.h>

using namespace orc;

namespace {

struct Material { vec4 diffuse; float metallic, pad1, pad2, pad3; };  // :24-31

struct RayHitInternal {  // :33-41 ; F5: every local is zero-initialised
    vec3 pos{0, 0, 0}, sideDist{0, 0, 0}, deltaDist{0, 0, 0};
    ivec3 rayStep{0, 0, 0};
    uint32_t material = 0;
    bvec3 mask{false, false, false};
    // instrumentation (not in the shader)
    ivec3 mapPos{0, 0, 0};
    uint32_t steps = 0, fetches = 0;
};

struct RayHit {  // :43-49
    uint32_t material = 0;
    vec3 pos{0, 0, 0}, normal{0, 0, 0}, dir{0, 0, 0};
};

struct Ctx {
    const orc_scene* sc;
    const orc_camera* cam;
    const orc_settings* st;
    vec3 bounds;          // vec3(pushConstants.volumeBounds)
    vec3 lightDir, lightColor;
    const Material* materials;
    // per-fragment
    int px, py;
    uint32_t rays, fetches;
};

const uint32_t MAX_RAY_STEPS = 512;    // :68
const int MAX_REFLECTIONS = 5;         // :69

inline vec3 bv(bvec3 m) { return {m.x ? 1.0f : 0.0f, m.y ? 1.0f : 0.0f, m.z ? 1.0f : 0.0f}; }
inline vec3 iv(ivec3 v) { return {(float)v.x, (float)v.y, (float)v.z}; }

// :73-77 getVoxel — pinned to an integer texel fetch
inline uint32_t getVoxel(Ctx& c, ivec3 p) {
    c.fetches++;
    const orc_scene* s = c.sc;
    return s->volume[(size_t)p.x + (size_t)p.y * s->w + (size_t)p.z * s->w * (size_t)s->h];
}

// :80-89 fragmentNoiseSeq. gl_FragCoord.xy/512 + 0.5 with NEAREST/REPEAT -> texel ((px+256)%512,(py+256)%512)
inline vec3 fragmentNoiseSeq(Ctx& c, uint32_t num) {
    uint32_t offset = num * 32u + c.cam->frame % 32u;
    // const float g = 1.22074408460575947536; a = (1/g, 1/(g*g), 1/(g*g*g)) folded in double, then narrowed
    const double g = 1.22074408460575947536;
    const vec3 a = {(float)(1.0 / g), (float)(1.0 / (g * g)), (float)(1.0 / (g * g * g))};
    uint32_t tx = ((uint32_t)c.px + 256u) % 512u, ty = ((uint32_t)c.py + 256u) % 512u;
    static const uint8_t zero[4] = {0, 0, 0, 0};
    const uint8_t* t = c.sc->noise ? c.sc->noise + 4 * ((size_t)ty * 512 + tx) : zero;
    vec3 noise = {from_unorm8(t[0]), from_unorm8(t[1]), from_unorm8(t[2])};
    vec3 v = noise + (float)offset * a;
    return {gmod1(v.x), gmod1(v.y), gmod1(v.z)};
}

// :92-95
inline vec3 randomDir(Ctx& c, uint32_t num) { return normalize(fragmentNoiseSeq(c, num) * 2.0f - v3(1.0f)); }

// :97-105 skyColor; NEAREST/REPEAT sampler (texture_2d.cpp:157-163)
inline vec4 skyColor(Ctx& c, vec3 rayDir) {
    vec2 uv = {atan2_cr(rayDir.z, rayDir.x), asin_cr(-rayDir.y)};
    uv.x *= 0.1591f; uv.y *= 0.3183f;
    uv.x += 0.5f; uv.y += 0.5f;
    const orc_scene* s = c.sc;
    if (!s->sky) return {0, 0, 0, 0};  // no skybox bound (headless configs without a sky): black
    int W = (int)s->skyW, H = (int)s->skyH;
    int ix = f2i(floorf(uv.x * (float)W)), iy = f2i(floorf(uv.y * (float)H));
    ix %= W; if (ix < 0) ix += W;
    iy %= H; if (iy < 0) iy += H;
    const float* t = s->sky + 4 * ((size_t)iy * W + ix);
    return {t[0], t[1], t[2], t[3]};
}

// :109-125
inline vec3 boxIntersection(Ctx& c, vec3 start, vec3 dir) {
    vec3 invDir = 1.0f / dir;
    vec3 t1 = (-start) * invDir;
    vec3 t2 = (c.bounds - start) * invDir;
    vec3 tminDir = gmin(t1, t2);
    vec3 tmaxDir = gmax(t1, t2);
    float tmin = gmax(tminDir.x, gmax(tminDir.y, tminDir.z));
    float tmax = gmin(tmaxDir.x, gmin(tmaxDir.y, tmaxDir.z));
    if (tmin >= 0 && tmax >= tmin) return start + (tmin + 0.1f) * dir;
    return start;
}

// :127-174
RayHitInternal traceRayInt(Ctx& c, vec3 start, vec3 dir, uint32_t maxSteps) {
    RayHitInternal r;
    c.rays++;
    r.pos = boxIntersection(c, start, dir);
    vec3 fl = gfloor(r.pos);
    ivec3 mapPos = {f2i(fl.x), f2i(fl.y), f2i(fl.z)};
    r.deltaDist = gabs(1.0f / dir);
    vec3 sg = gsign(dir);
    r.rayStep = {f2i(sg.x), f2i(sg.y), f2i(sg.z)};
    r.sideDist = (sg * (iv(mapPos) - r.pos) + (sg * 0.5f) + 0.5f) * r.deltaDist;
    const int bx = (int)c.cam->volumeBounds[0], by = (int)c.cam->volumeBounds[1], bz = (int)c.cam->volumeBounds[2];
    uint32_t i = 0, f0 = c.fetches;
    for (; i < maxSteps; i++) {
        if (mapPos.x < 0 || mapPos.x >= bx || mapPos.y < 0 || mapPos.y >= by || mapPos.z < 0 || mapPos.z >= bz) break;
        r.material = getVoxel(c, mapPos);
        if (r.material != 0) break;
        vec3 sd = r.sideDist;
        vec3 mn = gmin(v3(sd.y, sd.z, sd.x), v3(sd.z, sd.x, sd.y));
        r.mask = {sd.x <= mn.x, sd.y <= mn.y, sd.z <= mn.z};
        r.sideDist = r.sideDist + bv(r.mask) * r.deltaDist;
        vec3 mf = bv(r.mask);
        mapPos = {mapPos.x + f2i(mf.x) * r.rayStep.x, mapPos.y + f2i(mf.y) * r.rayStep.y, mapPos.z + f2i(mf.z) * r.rayStep.z};
    }
    r.mapPos = mapPos; r.steps = i; r.fetches = c.fetches - f0;
    return r;
}

// :176-196
RayHit traceRay(Ctx& c, vec3 start, vec3 dir, uint32_t maxSteps, RayHitInternal* internalOut = nullptr) {
    RayHitInternal in = traceRayInt(c, start, dir, maxSteps);
    if (internalOut) *internalOut = in;
    RayHit r;
    r.material = in.material;
    r.dir = dir;
    if (r.material != 0) {
        r.normal = normalize(bv(in.mask) * -iv(in.rayStep));
        float d = length(bv(in.mask) * (in.sideDist - in.deltaDist));
        r.pos = in.pos + d * dir;
    }
    return r;
}

// :198-202
inline bool traceRayHit(Ctx& c, vec3 start, vec3 dir, uint32_t maxSteps) {
    return traceRayInt(c, start, dir, maxSteps).material != 0;
}

// :205-227. pass bit AO off == aoSamples 0.
vec3 calcAmbient(Ctx& c, const RayHit& hit, uint32_t depth) {
    float ambient = 0.0f;
    uint32_t aoSamples = (c.st->pass_mask & ORC_PASS_AO) ? c.st->aoSamples : 0u;
    if (aoSamples == 0) {
        ambient = 1.0f;
    } else {
        float sampleFrac = 1.0f / (float)aoSamples;
        for (uint32_t i = 0; i < aoSamples; i++) {
            vec3 dir = hit.normal + randomDir(c, i + depth * aoSamples);
            bool occluded = traceRayHit(c, hit.pos + dir * 0.01f, dir, 64);  // :219 (GLSL initialiser sees the parameter `hit`)
            if (occluded) ambient += sampleFrac;
        }
    }
    vec4 sky = skyColor(c, hit.normal);
    return (ambient * c.st->ambientIntensity) * v3(sky.x, sky.y, sky.z);
}

// :230-233. pass bit SHADOW off == never shadowed, no ray.
inline bool isShadowed(Ctx& c, const RayHit& hit) {
    if (!(c.st->pass_mask & ORC_PASS_SHADOW)) return false;
    return traceRayHit(c, hit.pos + hit.normal * 0.01f, c.lightDir, MAX_RAY_STEPS);
}

// :236-248
inline vec3 color(Ctx& c, vec3 normal, const Material& mat, vec3 ambient, vec3 reflection, bool shadowed) {
    vec3 diffuse = v3(0.0f);
    if (!shadowed) {
        float diff = gmax(dot(normal, c.lightDir), 0.0f);
        diffuse = diff * c.lightColor * c.st->lightIntensity;
    }
    vec3 specular = reflection * mat.metallic;
    return (diffuse + specular + ambient) * v3(mat.diffuse.x, mat.diffuse.y, mat.diffuse.z);
}

// :252-264
vec3 colorHit(Ctx& c, const RayHit& hit, vec3 reflection, uint32_t depth) {
    if (hit.material != 0) {
        vec3 ambient = calcAmbient(c, hit, depth);
        bool shadowed = isShadowed(c, hit);
        return color(c, hit.normal, c.materials[hit.material], ambient, reflection, shadowed) * 1.0f / (float)(depth + 1);
    }
    vec4 s = skyColor(c, hit.dir);
    return {s.x, s.y, s.z};
}

// :267-307. pass bit REFLECT off == the metallic branch is skipped (reflection = 0).
vec3 colorMainRay(Ctx& c, const RayHit& hit) {
    Material mat = c.materials[hit.material];
    vec3 reflection = v3(0.0f);
    if ((c.st->pass_mask & ORC_PASS_REFLECT) && mat.metallic > 0) {
        RayHit bounces[MAX_REFLECTIONS];
        RayHit lastHit = hit;
        int lastIdx = -1;
        for (int i = 0; i < MAX_REFLECTIONS; i++) {
            vec3 reflectDir = reflect(lastHit.dir, lastHit.normal);
            RayHit reflectHit = traceRay(c, lastHit.pos + lastHit.normal * 0.01f, reflectDir, MAX_RAY_STEPS);
            bounces[i] = reflectHit;
            lastHit = reflectHit;
            if (lastHit.material == 0 || c.materials[lastHit.material].metallic <= 0) { lastIdx = i; break; }
        }
        for (int i = lastIdx; i >= 0; i--) reflection = reflection + colorHit(c, bounces[i], reflection, (uint32_t)i);
    }
    return colorHit(c, hit, reflection, 0);
}

}  // namespace

// main(), :309-346, for one fragment. vScreenPos from screen_quad.vert:16-32 (full-screen triangle uv)
// at the pixel centre: ((px+.5)/W, (py+.5)/H), py = 0 at the top row (viewport voxel_renderer.cpp:57-63).
static void shade_fragment(Ctx& c, const orc_planes* o) {
    const orc_camera* cam = c.cam;
    const int W = cam->screenSize[0], H = cam->screenSize[1];
    c.rays = 0; c.fetches = 0;
    vec2 vScreenPos = {((float)c.px + 0.5f) / (float)W, ((float)c.py + 0.5f) / (float)H};
    vec2 screenPos = {vScreenPos.x * 2.0f - 1.0f, vScreenPos.y * 2.0f - 1.0f};
    vec3 cameraPlaneU = {cam->camRight[0], cam->camRight[1], cam->camRight[2]};
    vec3 cameraPlaneV = v3(cam->camUp[0], cam->camUp[1], cam->camUp[2]) * (float)H / (float)W;
    vec2 jit = {cam->cameraJitter[0] / (float)W * -2.0f, cam->cameraJitter[1] / (float)H * 2.0f};
    vec3 rayDir = normalize(normalize(v3(cam->camDir[0], cam->camDir[1], cam->camDir[2])) + screenPos.x * cameraPlaneU +
                            screenPos.y * cameraPlaneV + v3(jit.x, jit.y, 0.0f));
    vec3 rayStart = {cam->camPos[0], cam->camPos[1], cam->camPos[2]};

    RayHitInternal prim;
    RayHit result = traceRay(c, rayStart, rayDir, MAX_RAY_STEPS, &prim);

    vec3 outColor, outPos, outNormal;
    float outDepth, outMask;
    if (result.material != 0) {
        if (c.st->pass_mask & ORC_PASS_SHADE) {
            outColor = colorMainRay(c, result);
        } else {  // unlit: palette albedo (new pass-mask behaviour, not in the reference)
            vec4 d = c.materials[result.material].diffuse;
            outColor = {d.x, d.y, d.z};
        }
        outDepth = length(result.pos - rayStart);
        outMask = 0.9f;
        outPos = result.pos;
        outNormal = result.normal;
    } else {
        vec4 s = skyColor(c, rayDir);
        outColor = {s.x, s.y, s.z};
        outDepth = 0.0f;
        outMask = 0.0f;
        outPos = result.pos;  // F5: zero-initialised
        outNormal = v3(0.0f);
    }

    size_t p = (size_t)c.py * W + c.px;
    if (o->color) { uint8_t* q = o->color + 4 * p; q[0] = to_unorm8(outColor.x); q[1] = to_unorm8(outColor.y); q[2] = to_unorm8(outColor.z); q[3] = 0; }
    if (o->depth) o->depth[p] = outDepth;
    if (o->motion) { o->motion[2 * p] = 0.0f; o->motion[2 * p + 1] = 0.0f; }
    if (o->mask) o->mask[p] = to_unorm8(outMask);
    if (o->position) { float* q = o->position + 4 * p; q[0] = outPos.x; q[1] = outPos.y; q[2] = outPos.z; q[3] = 0.0f; }
    if (o->normal) { int8_t* q = o->normal + 4 * p; q[0] = to_snorm8(outNormal.x); q[1] = to_snorm8(outNormal.y); q[2] = to_snorm8(outNormal.z); q[3] = 0; }
    if (o->hit) {
        int16_t* q = o->hit + 4 * p;
        auto sat = [](int v) { return (int16_t)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); };
        q[0] = sat(prim.mapPos.x); q[1] = sat(prim.mapPos.y); q[2] = sat(prim.mapPos.z); q[3] = (int16_t)prim.steps;
    }
    if (o->hit_palette) o->hit_palette[p] = (uint8_t)prim.material;
    if (o->color_f32) { float* q = o->color_f32 + 3 * p; q[0] = outColor.x; q[1] = outColor.y; q[2] = outColor.z; }
    if (o->rays) o->rays[p] = c.rays;
    if (o->fetches) o->fetches[p] = c.fetches;
}

static Ctx make_ctx(const orc_scene* sc, const orc_camera* cam, const orc_settings* st) {
    Ctx c{};
    c.sc = sc; c.cam = cam; c.st = st;
    c.bounds = {(float)cam->volumeBounds[0], (float)cam->volumeBounds[1], (float)cam->volumeBounds[2]};
    c.lightDir = {st->lightDir[0], st->lightDir[1], st->lightDir[2]};
    c.lightColor = {st->lightColor[0], st->lightColor[1], st->lightColor[2]};
    c.materials = reinterpret_cast<const Material*>(sc->palette);
    return c;
}

extern "C" int orc_num_threads(void) { return omp_get_max_threads(); }

extern "C" int orc_render(const orc_scene* sc, const orc_camera* cam, const orc_settings* st, const orc_planes* out,
                          int row_begin, int row_end, int row_step, int nthreads) {
    if (!sc || !cam || !st || !out || !sc->volume || !sc->palette) return -1;
    if (cam->volumeBounds[0] != sc->w || cam->volumeBounds[1] != sc->h || cam->volumeBounds[2] != sc->d) return -1;
    const int W = cam->screenSize[0], H = cam->screenSize[1];
    if (W <= 0 || H <= 0 || row_step <= 0) return -1;
    if (row_begin < 0) row_begin = 0;
    if (row_end > H) row_end = H;
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    const int nrows = row_end > row_begin ? (row_end - row_begin + row_step - 1) / row_step : 0;
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads)
    for (int r = 0; r < nrows; r++) {
        Ctx c = make_ctx(sc, cam, st);
        c.py = row_begin + r * row_step;
        for (int x = 0; x < W; x++) { c.px = x; shade_fragment(c, out); }
    }
    return 0;
}

extern "C" int orc_trace_one(const orc_scene* sc, const float start[3], const float dir[3], uint32_t maxSteps,
                             int32_t out[6], float fout[6]) {
    orc_camera cam{}; orc_settings st{};
    cam.volumeBounds[0] = sc->w; cam.volumeBounds[1] = sc->h; cam.volumeBounds[2] = sc->d;
    Ctx c = make_ctx(sc, &cam, &st);
    RayHitInternal in;
    RayHit h = traceRay(c, v3(start[0], start[1], start[2]), v3(dir[0], dir[1], dir[2]), maxSteps, &in);
    out[0] = in.mapPos.x; out[1] = in.mapPos.y; out[2] = in.mapPos.z; out[3] = (int32_t)in.steps;
    out[4] = (int32_t)in.material; out[5] = (int32_t)in.fetches;
    fout[0] = h.pos.x; fout[1] = h.pos.y; fout[2] = h.pos.z; fout[3] = h.normal.x; fout[4] = h.normal.y; fout[5] = h.normal.z;
    return 0;
}

// CameraController::updateDirectionVectors, source/voxels/resource/camera_controller.cpp:15-28
// (glm::radians(x) = x * 0.01745329251994329576923690768489f; cos/sin are the float overloads)
extern "C" void orc_camera_vectors(float yaw, float pitch, float focal, float dir[3], float right[3], float up[3]) {
    const float k = 0.01745329251994329576923690768489f;
    const vec3 worldUp = {0.0f, -1.0f, 0.0f};
    vec3 n;
    n.x = cosf(yaw * k) * cosf(pitch * k);
    n.y = sinf(pitch * k);
    n.z = sinf(yaw * k) * cosf(pitch * k);
    n = normalize(n);
    vec3 r = normalize(cross(n, worldUp));
    vec3 u = normalize(cross(r, n));
    vec3 d = n * focal;
    dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
    right[0] = r.x; right[1] = r.y; right[2] = r.z;
    up[0] = u.x; up[1] = u.y; up[2] = u.z;
}
