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
                    // uv = vScreenPos + offset*step*stepWidth sampled LINEAR / CLAMP_TO_EDGE (render_image.cpp:60-66). In texel space the
                    // sample point is px + offset*stepWidth (+0.5 for the pixel centre, -0.5 for the texel centre): on a texel centre when
                    // offset*stepWidth is integral (the filter then returns that texel), between four texels otherwise (stepWidth is a
                    // float slider in the reference GUI, voxel_settings_gui.cpp:106). Pinned: the texel-space coordinate is evaluated as
                    // float(px) + offset*stepWidth in binary32; weights per the Vulkan texel-filtering equations in binary32, products
                    // first, sums left to right (as in orc_blit); all three inputs are filtered the same way, after their UNORM / SNORM decode.
                    const float ux = (float)px + taps.offset[i].x * stepWidth, uy = (float)py + taps.offset[i].y * stepWidth;
                    const float fi = floorf(ux), fj = floorf(uy);
                    const float a = ux - fi, b = uy - fj;
                    const int i0 = clampi(f2i(fi), 0, w - 1), i1 = clampi(f2i(fi) + 1, 0, w - 1);
                    const int j0 = clampi(f2i(fj), 0, h - 1), j1 = clampi(f2i(fj) + 1, 0, h - 1);
                    const bool onTexel = a == 0.0f && b == 0.0f;
                    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
                    auto mix4 = [&](const vec4& t00, const vec4& t10, const vec4& t01, const vec4& t11) {
                        return vec4{((w00 * t00.x + w10 * t10.x) + w01 * t01.x) + w11 * t11.x, ((w00 * t00.y + w10 * t10.y) + w01 * t01.y) + w11 * t11.y,
                                    ((w00 * t00.z + w10 * t10.z) + w01 * t01.z) + w11 * t11.z, ((w00 * t00.w + w10 * t10.w) + w01 * t01.w) + w11 * t11.w}; };
                    vec4 offsetColor = onTexel ? ldc(i0, j0) : mix4(ldc(i0, j0), ldc(i1, j0), ldc(i0, j1), ldc(i1, j1));
                    vec4 t = sampleColor - offsetColor;
                    float dist2 = dot(t, t);
                    float colorWeight = gmin(expf(-(dist2) / phiColor), 1.0f);
                    vec4 offsetNormal = onTexel ? ldn(i0, j0) : mix4(ldn(i0, j0), ldn(i1, j0), ldn(i0, j1), ldn(i1, j1));
                    t = sampleNormal - offsetNormal;
                    dist2 = gmax(dot(t, t) / (stepWidth * stepWidth), 0.0f);
                    float normalWeight = gmin(expf(-(dist2) / phiNormal), 1.0f);
                    vec4 offsetPos = onTexel ? ldp(i0, j0) : mix4(ldp(i0, j0), ldp(i1, j0), ldp(i0, j1), ldp(i1, j1));
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
