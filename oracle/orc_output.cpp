// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.h). CPU restatement of the two steps either side of the render pass
// (SURVEY.md §8f N2, N3), paths relative to /root/reference:
//   * the frame driver's jitter sequence: UpscalerStage::update (source/voxels/stages/upscaler_stage.cpp:59-70) on
//     ffxFsr2GetJitterPhaseCount / ffxFsr2GetJitterOffset. FSR2 is a third-party dependency (thirdparty/fsr2, v2.1.0 per
//     include/ffx_fsr2.h:33-43) whose implementation is NOT in the checkout (headers + Windows .lib LFS stubs only); its published
//     algorithm (FidelityFX-FSR2 v2.1.0, src/ffx-fsr2-api/ffx_fsr2.cpp) is restated here and pinned on the header's own table of
//     phase counts (ffx_fsr2.h:341-347: 1.5x -> 18, 1.7x -> 23, 2.0x -> 32, 3.0x -> 72) in tests/test_oracle.py;
//   * VoxelRenderSettings::renderResolution (source/voxels/voxel_render_settings.cpp:3-13);
//   * the blit: shader/blit.frag:14-22 sampled through RenderImage's sampler (LINEAR / CLAMP_TO_EDGE,
//     source/engine/resource/render_image.cpp:59-67) into the swap-chain image (vk-bootstrap default format selection,
//     source/engine/swapchain.cpp:15 = B8G8R8A8_SRGB: the store encodes linear -> sRGB, alpha stays linear).
// PARITY UNPINNED for the blit (no golden image in the reference; graphics APIs unavailable on this box). Pinned decisions:
// bilinear weights per the Vulkan spec's texel-filtering equations evaluated in binary32, products first, sums left to right;
// the sRGB transfer function evaluated in binary64 and rounded to nearest-even UNORM8.
#include <cmath>
#include <cstdint>
#include <algorithm>
#include "orc_math.h"
#include "vrt_oracle.h"
using namespace orc;

extern "C" int orc_jitter_phase_count(int renderWidth, int displayWidth) {
    const float basePhaseCount = 8.0f;
    return (int32_t)(basePhaseCount * powf((float)displayWidth / (float)renderWidth, 2.0f));
}

static float halton(int32_t index, int32_t base) {
    float f = 1.0f, result = 0.0f;
    for (int32_t cur = index; cur > 0;) {
        f /= (float)base;
        result = result + f * (float)(cur % base);
        cur = (int32_t)(uint32_t)floorf((float)cur / (float)base);
    }
    return result;
}

extern "C" void orc_jitter_offset(float* outX, float* outY, int index, int phaseCount) {
    *outX = halton((index % phaseCount) + 1, 2) - 0.5f;
    *outY = halton((index % phaseCount) + 1, 3) - 0.5f;
}

// UpscalerStage::update, upscaler_stage.cpp:59-70: one call per frame; state = frameCount. Returns the jitter of this frame and
// advances frameCount the way the reference does (it wraps only once it EXCEEDS the phase count, so index 0 comes up twice).
extern "C" void orc_upscaler_update(int renderWidth, int targetWidth, int* frameCount, float* jitterX, float* jitterY) {
    const int32_t phases = orc_jitter_phase_count(renderWidth, targetWidth);
    orc_jitter_offset(jitterX, jitterY, *frameCount % phases, phases);
    (*frameCount)++;
    if (*frameCount > phases) *frameCount = 0;
}

// voxel_render_settings.cpp:3-13; scaling = FsrScaling value (10 NONE, 15, 17, 20, 30), voxel_render_settings.hpp:6-13
extern "C" void orc_render_resolution(int fsrEnable, uint32_t scaling, const uint32_t target[2], uint32_t out[2]) {
    for (int k = 0; k < 2; k++) out[k] = fsrEnable ? (uint32_t)((10.0f / (float)scaling) * (float)target[k]) : target[k];
}

static inline uint8_t srgb8(float L) {
    if (!(L == L)) return 0;
    const double l = L < 0.0f ? 0.0 : (L > 1.0f ? 1.0 : (double)L);
    const double e = l <= 0.0031308 ? 12.92 * l : 1.055 * pow(l, 1.0 / 2.4) - 0.055;
    return (uint8_t)rint(e * 255.0);
}

// blit.frag:14-22 for every pixel of the target; src/dst RGBA8, row-major. flags: bit 0 = sRGB-encode rgb on store.
extern "C" int orc_blit(const uint8_t* src, int sw, int sh, uint8_t* dst, int tw, int th, int flags, int nthreads) {
    if (!src || !dst || sw <= 0 || sh <= 0 || tw <= 0 || th <= 0) return -1;
    (void)nthreads;
    const float fsw = (float)sw, fsh = (float)sh, ftw = (float)tw, fth = (float)th;
#pragma omp parallel for schedule(static)
    for (int py = 0; py < th; py++) for (int px = 0; px < tw; px++) {
        const vec2 vScreenPos = {((float)px + 0.5f) / ftw, ((float)py + 0.5f) / fth};
        const vec2 targetPos = {vScreenPos.x * ftw, vScreenPos.y * fth};
        const float scale = gmin(fsw / ftw, fsh / fth);
        const vec2 scaledTarget = {ftw * scale, fth * scale};
        const vec2 sourcePos = {targetPos.x * scale + (fsw - scaledTarget.x) / 2.0f, targetPos.y * scale + (fsh - scaledTarget.y) / 2.0f};
        const vec2 uv = {sourcePos.x / fsw, sourcePos.y / fsh};
        // texture(): LINEAR, CLAMP_TO_EDGE, one mip level
        const float u = uv.x * fsw - 0.5f, v = uv.y * fsh - 0.5f;
        const float fi = floorf(u), fj = floorf(v);
        const float a = u - fi, b = v - fj;
        const int i0 = std::clamp(f2i(fi), 0, sw - 1), i1 = std::clamp(f2i(fi) + 1, 0, sw - 1);
        const int j0 = std::clamp(f2i(fj), 0, sh - 1), j1 = std::clamp(f2i(fj) + 1, 0, sh - 1);
        const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
        const uint8_t* t00 = src + 4 * ((size_t)j0 * sw + i0); const uint8_t* t10 = src + 4 * ((size_t)j0 * sw + i1);
        const uint8_t* t01 = src + 4 * ((size_t)j1 * sw + i0); const uint8_t* t11 = src + 4 * ((size_t)j1 * sw + i1);
        uint8_t* o = dst + 4 * ((size_t)py * tw + px);
        for (int c = 0; c < 4; c++) {
            const float val = ((w00 * from_unorm8(t00[c]) + w10 * from_unorm8(t10[c])) + w01 * from_unorm8(t01[c])) + w11 * from_unorm8(t11[c]);
            o[c] = ((flags & 1) && c < 3) ? srgb8(val) : to_unorm8(val);
        }
    }
    return 0;
}
