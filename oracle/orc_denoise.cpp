// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.h). CPU restatement of
// /root/reference/shader/denoiser.frag:38-73 and the host parameters of
// /root/reference/source/voxels/stages/denoiser_stage.cpp:37-61 (offsets, weights), :143-154 (per-iteration
// phi / stepWidth), :203-257 (ping-pong).  PARITY UNPINNED by the reference (no golden vectors).
// Pinned: SURVEY F6 (UBO stride mismatch -> two modes), F7 (iteration 0 has phi = +inf).
#include "vrt_oracle.h"
#include "orc_math.h"
#include <omp.h>
#include <vector>

using namespace orc;

namespace {

struct Taps { float kernel[25]; vec2 offset[25]; };

// Host side: denoiser_stage.cpp:37-61. glm::gauss(Coord, 0, sd=2) = exp(-((x*x)/(2*sd*sd) + (y*y)/(2*sd*sd))).
void host_tables(float weights[9], vec2 offsets[9]) {
    int i = 0;
    for (int y = -1; y <= 1; y++)
        for (int x = -1; x <= 1; x++, i++) {
            offsets[i] = {(float)x, (float)y};
            float sx = ((float)x * (float)x) / (2.0f * 2.0f * 2.0f), sy = ((float)y * (float)y) / (2.0f * 2.0f * 2.0f);
            weights[i] = expf(-(sx + sy));
        }
}

// What the shader's `float kernel[9]; vec2 offset[9]` (std140: 16-byte array stride) see when the host
// uploads tightly packed arrays and the loop runs i < 25.
Taps make_taps(int mode) {
    float w[9]; vec2 o[9];
    host_tables(w, o);
    Taps t{};
    if (mode == ORC_DENOISE_AS_INTENDED) {
        for (int i = 0; i < 9; i++) { t.kernel[i] = w[i]; t.offset[i] = o[i]; }
        return t;  // taps 9..24: weight 0 (the intended 3x3 kernel)
    }
    // AS_EXECUTED: element i is read at byte 16*i; reads past the bound range give 0.
    const float* wf = w;                                    // 36 bytes
    const float* of = reinterpret_cast<const float*>(o);    // 72 bytes
    for (int i = 0; i < 25; i++) {
        int fi = 4 * i;  // float index of byte 16*i
        t.kernel[i] = fi < 9 ? wf[fi] : 0.0f;
        t.offset[i].x = fi < 18 ? of[fi] : 0.0f;
        t.offset[i].y = fi + 1 < 18 ? of[fi + 1] : 0.0f;
    }
    return t;
}

inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

}  // namespace

extern "C" int orc_denoise(const uint8_t* color_in, const int8_t* normal, const float* position, int w, int h,
                           const orc_denoise_settings* ds, uint8_t* color_out, int nthreads) {
    if (!color_in || !normal || !position || !ds || !color_out || w <= 0 || h <= 0) return -1;
    if (ds->iterations < 1 || ds->iterations > 10) return -1;  // MAX_DENOISER_PASSES, denoiser_stage.hpp:9
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    const Taps taps = make_taps(ds->mode);
    const size_t n = (size_t)w * h;
    std::vector<uint8_t> ping[2] = {std::vector<uint8_t>(4 * n), std::vector<uint8_t>(4 * n)};
    const uint8_t* src = color_in;
    int lastOutput = 0;
    for (int it = 0; it < ds->iterations; it++) {
        // denoiser_stage.cpp:147-151
        const float phiColor = 1.0f / (float)it * ds->phiColor0;
        const float phiNormal = 1.0f / (float)it * ds->phiNormal0;
        const float phiPos = 1.0f / (float)it * ds->phiPos0;
        const float stepWidth = (float)it * ds->stepWidth + 1.0f;
        uint8_t* dst = ping[it % 2].data();
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int py = 0; py < h; py++) {
            for (int px = 0; px < w; px++) {
                auto ldc = [&](int x, int y) { const uint8_t* q = src + 4 * ((size_t)y * w + x);
                    return vec4{from_unorm8(q[0]), from_unorm8(q[1]), from_unorm8(q[2]), from_unorm8(q[3])}; };
                auto ldn = [&](int x, int y) { const int8_t* q = normal + 4 * ((size_t)y * w + x);
                    return vec4{from_snorm8(q[0]), from_snorm8(q[1]), from_snorm8(q[2]), from_snorm8(q[3])}; };
                auto ldp = [&](int x, int y) { const float* q = position + 4 * ((size_t)y * w + x);
                    return vec4{q[0], q[1], q[2], q[3]}; };
                vec4 sum = {0, 0, 0, 0};
                vec4 sampleColor = ldc(px, py), sampleNormal = ldn(px, py), samplePos = ldp(px, py);
                float totalWeight = 0.0f;
                for (int i = 0; i < 25; i++) {
                    // uv = vScreenPos + offset*step*stepWidth sampled LINEAR/CLAMP (render_image.cpp:60-66):
                    // pinned to the texel containing the sample point (exact texel centre for integral stepWidth).
                    int tx = clampi(px + f2i(floorf(0.5f + taps.offset[i].x * stepWidth)), 0, w - 1);
                    int ty = clampi(py + f2i(floorf(0.5f + taps.offset[i].y * stepWidth)), 0, h - 1);
                    vec4 offsetColor = ldc(tx, ty);
                    vec4 t = sampleColor - offsetColor;
                    float dist2 = dot(t, t);
                    float colorWeight = gmin(expf(-(dist2) / phiColor), 1.0f);
                    vec4 offsetNormal = ldn(tx, ty);
                    t = sampleNormal - offsetNormal;
                    dist2 = gmax(dot(t, t) / (stepWidth * stepWidth), 0.0f);
                    float normalWeight = gmin(expf(-(dist2) / phiNormal), 1.0f);
                    vec4 offsetPos = ldp(tx, ty);
                    t = samplePos - offsetPos;
                    dist2 = dot(t, t);
                    float posWeight = gmin(expf(-(dist2) / phiPos), 1.0f);
                    float weight = colorWeight * normalWeight * posWeight;
                    sum = sum + offsetColor * weight * taps.kernel[i];
                    totalWeight += weight * taps.kernel[i];
                }
                vec4 o = sum / totalWeight;
                uint8_t* q = dst + 4 * ((size_t)py * w + px);
                q[0] = to_unorm8(o.x); q[1] = to_unorm8(o.y); q[2] = to_unorm8(o.z); q[3] = to_unorm8(o.w);
            }
        }
        src = dst;
        lastOutput = it % 2;
    }
    memcpy(color_out, ping[lastOutput].data(), 4 * n);
    return 0;
}
