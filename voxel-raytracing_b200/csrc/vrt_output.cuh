// vrt_output.cuh — per-pixel arithmetic of the output stage (the step downstream of the render pass), written for
// the device and compilable for the host (tests/host_harness). Follows /root/reference/shader/blit.frag:14-22 sampled through
// RenderImage's LINEAR / CLAMP_TO_EDGE sampler (source/engine/resource/render_image.cpp:59-67). Same exactness rules as vrt_core.cuh.
#pragma once
#include "vrt_core.cuh"

namespace vrt {

VRT_HD int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

struct BlitParams { int sw, sh, tw, th; uint32_t flags; };

// The footprint is separable: column taps depend on px only, row taps on py only (blit.frag's vec2 arithmetic never mixes the axes
// except through the scalar `scale`). One axis: the two clamped texel coordinates and alpha (Vulkan texel filtering: i0 = floor(u - 0.5),
// alpha = frac(u - 0.5)).
struct BlitAxis { int i0, i1; float a; };

VRT_HD float blit_scale(const BlitParams& p) { return fminf((float)p.sw / (float)p.tw, (float)p.sh / (float)p.th); }   // blit.frag:17

VRT_HD BlitAxis blit_axis(int pix, int srcN, int tgtN, float scale) {
    const float fs = (float)srcN, ft = (float)tgtN;
    const float screen = ((float)pix + 0.5f) / ft;                // vScreenPos (screen_quad.vert:16-32)
    const float target = screen * ft;                             // blit.frag:16
    const float scaled = ft * scale;                              // :18
    const float source = target * scale + (fs - scaled) / 2.0f;   // :19
    const float u = (source / fs) * fs - 0.5f;                    // :20 + un-normalisation by the sampler
    const float fi = floorf(u);
    const int i = f2i(fi);
    BlitAxis t;
    t.a = u - fi; t.i0 = clampi(i, 0, srcN - 1); t.i1 = clampi(i + 1, 0, srcN - 1);
    return t;
}

// weights as products, applied in the order 00, 10, 01, 11
struct BlitTap { float w00, w10, w01, w11; };
VRT_HD BlitTap blit_weights(float a, float b) {
    BlitTap t;
    t.w00 = (1.0f - a) * (1.0f - b); t.w10 = a * (1.0f - b); t.w01 = (1.0f - a) * b; t.w11 = a * b;
    return t;
}

VRT_HD float blit_mix(const BlitTap& t, float c00, float c10, float c01, float c11) {
    return ((t.w00 * c00 + t.w10 * c10) + t.w01 * c01) + t.w11 * c11;
}

// sRGB store: thr[k] (k = 1..255) is the smallest binary32 value whose encoded UNORM8 code is >= k (built on the host by bisection over
// the exact transfer function, vrt_output.cu), so the code of L is the number of thresholds <= L: 8 compares instead of a powf whose
// last bit would differ between libraries. thr[0] is unused.
VRT_HD uint8_t srgb8_search(const float* thr, float L) {
    int code = 0;
    for (int step = 128; step > 0; step >>= 1) if (L >= thr[code + step]) code += step;
    return (uint8_t)code;
}
#if defined(__CUDACC__)
static __device__ __noinline__ uint8_t srgb8_search_cold(const float* thr, float L) { return srgb8_search(thr, L); }
#endif
VRT_HD uint8_t srgb8_lookup(const float* thr, float L) {
#if defined(__CUDA_ARCH__)
    // a guess from the fast transfer function (lg2.approx / ex2.approx, no branches), moved by at most one onto the exact code by the two
    // thresholds around it; the result is defined by thr[] alone: if the guess had to move, the new interval is checked and the search runs
    // when it does not hold. NaN clamps to 0 (fmaxf returns the other operand), like the NaN -> 0 rule of the UNORM store.
    const float l = fminf(fmaxf(L, 0.0f), 1.0f);
    float lg, p;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(l));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p) : "f"(lg * (1.0f / 2.4f)));
    const float e = l <= 0.0031308f ? 12.92f * l : 1.055f * p - 0.055f;
    const int guess = min(max(__float2int_rn(e * 255.0f), 1), 254);          // thr[guess - 1 .. guess + 2] are all in the table
    int code = guess + (l >= thr[guess + 1] ? 1 : 0) - (l < thr[guess] ? 1 : 0);
    if (code != guess) {
        const bool ok = (code == 0 || l >= thr[code]) && (code == 255 || l < thr[code + 1]);
        if (!ok) code = srgb8_search_cold(thr, l);
    }
    return (uint8_t)code;
#else
    if (!(L == L)) return 0;
    return srgb8_search(thr, L);
#endif
}

// lut[(word >> 8k) & 255] with the byte turned into a byte offset by one shift + mask (instead of extract, scale, add)
VRT_HD float lut_byte(const float* lut, uint32_t word, int k) {
    const uint32_t off = (k == 0 ? word << 2 : word >> (8 * k - 2)) & 0x3FCu;
    return *reinterpret_cast<const float*>(reinterpret_cast<const char*>(lut) + off);
}

// ---- sRGB thresholds (host). encode(L) = round-half-even(255 * E(L)), E = the sRGB transfer function in binary64; it is monotone in L,
// so the smallest binary32 L with encode(L) >= k is found by bisection over the (ordered) bit patterns of [0, 1].
inline int srgb_code(float L) {
    const double l = (double)L;
    const double e = l <= 0.0031308 ? 12.92 * l : 1.055 * pow(l, 1.0 / 2.4) - 0.055;
    return (int)rint(e * 255.0);
}
inline void srgb_thresholds(float thr[256]) {
    thr[0] = 0.0f;
    for (int k = 1; k < 256; k++) {
        uint32_t lo = 0u, hi = 0x3f800000u;                 // code(lo) < k <= code(hi); positive floats order like their bit patterns
        while (hi - lo > 1u) {
            const uint32_t mid = lo + (hi - lo) / 2u;
            float f; memcpy(&f, &mid, 4);
            if (srgb_code(f) >= k) hi = mid; else lo = mid;
        }
        memcpy(&thr[k], &hi, 4);
    }
}

}  // namespace vrt
