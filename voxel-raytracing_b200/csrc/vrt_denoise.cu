// vrt_denoise.cu — edge-avoiding a-trous filter: one iteration of shader/denoiser.frag:38-73 as a CUDA kernel.
// Host side (tap tables, per-iteration phi/stepWidth, ping-pong) is in vrt_api.cu and follows
// source/voxels/stages/denoiser_stage.cpp:37-61,143-154,203-257. HBM-bound: 24 B read + 4 B written per pixel
// per iteration algorithmically; taps of neighbouring pixels overlap and are served by L1/L2.
#include <cuda_runtime.h>
#include "vrt_core.cuh"
#include "vrt_launch.h"

#ifndef DENOISE_MIN_BLOCKS
#define DENOISE_MIN_BLOCKS 5                 // 48 registers for the general iteration (247 -> 238 us at 4K)
#endif

namespace vrt {

// UNORM8 / SNORM8 texel decode through per-block tables of the exactly divided values: the kernel was bound by its IEEE
// divisions (8 per tap for the two decodes alone), not by memory (ncu r1: 0.69 ms per iteration at 4K before, see profiles/)
__device__ __forceinline__ float4 ld_color(const uint8_t* c, size_t p, const float* lutU) {
    const uchar4 q = reinterpret_cast<const uchar4*>(c)[p];
    return make_float4(lutU[q.x], lutU[q.y], lutU[q.z], lutU[q.w]);
}
__device__ __forceinline__ float snorm(signed char v) { const float f = (float)v / 127.0f; return f < -1.0f ? -1.0f : f; }
__device__ __forceinline__ float4 ld_normal(const int8_t* n, size_t p, const float* lutS) {
    const uchar4 q = reinterpret_cast<const uchar4*>(n)[p];
    return make_float4(lutS[q.x], lutS[q.y], lutS[q.z], lutS[q.w]);
}
__device__ __forceinline__ float dot4d(float4 a, float4 b) {   // t = a - b; dot(t, t), summed left to right
    const float x = a.x - b.x, y = a.y - b.y, z = a.z - b.z, w = a.w - b.w;
    return ((x * x + y * y) + z * z) + w * w;
}

// a / b for a divisor that is a per-launch constant, rb = 1.0f / b (IEEE, computed once on the host): two Markstein corrections of a * rb.
// With rb the correctly rounded reciprocal, the first correction leaves the quotient within half an ulp and the second makes it the
// correctly rounded a / b — the value the oracle's division produces — in 5 FMA-pipe instructions and WITHOUT the slow path the
// compiler's IEEE division takes for a zero numerator (a tap on the pixel itself, or a neighbour with the same colour / normal /
// position: the common case; ncu r2p1 had 32 % of the kernel's instructions in those subroutines). Quotients that are not finite or
// so small that the corrections would work on denormals take the plain division (rare, and exact by construction).
__device__ __forceinline__ float div_const(float a, float b, float rb) {
    const float q0 = a * rb;
    const float q1 = __fmaf_rn(__fmaf_rn(-b, q0, a), rb, q0);
    const float q2 = __fmaf_rn(__fmaf_rn(-b, q1, a), rb, q1);
    // An infinite first estimate (infinite numerator, or a quotient that overflows) would turn into NaN in the corrections: it is the
    // answer. NaN stays NaN. A zero numerator gives a zero of either sign and a denormal quotient may be one denormal ulp off: both only
    // ever reach exp(), which maps them to the same 1.0f, or a UNORM8 store, which maps them to the same 0.
    return fabsf(q0) == INFINITY ? q0 : q2;
}

// BLIND: all three phi are +inf, which is what the reference's host code produces for iteration 0 (1.0f / float(0) * phi0,
// denoiser_stage.cpp:143-154). Then every edge-stopping weight is exactly 1 whatever the inputs are — x / inf is +-0 for finite x and
// NaN for x = inf / NaN, exp gives 1 or NaN, and min(NaN, 1) = 1 — so the iteration is a plain weighted blur of the colour plane:
// normals and positions are not read, no exp, no division (the IEEE divisions by inf all took the slow path). The sums are the same
// float operations in the same order ((c * 1) * k == c * k, 1 * k == k), so the bytes are those of the general kernel.
// LINEAR: some tap offset*stepWidth of this iteration is not integral (stepWidth is a float slider in the reference GUI,
// voxel_settings_gui.cpp:106): the inputs are then sampled between four texels, as the reference's LINEAR / CLAMP_TO_EDGE sampler does
// (render_image.cpp:60-66). Texel-space coordinate float(px) + offset*stepWidth, weights per the Vulkan texel-filtering equations in
// binary32, products first, sums left to right: the same float operations as oracle/orc_denoise.cpp. Integral offsets keep the
// one-texel path (the filter returns the texel itself there).
struct Bilin { int i0, i1, j0, j1; float w00, w10, w01, w11; bool onTexel; };
__device__ __forceinline__ Bilin bilin_setup(int px, int py, float ox, float oy, int W, int H) {
    Bilin q;
    const float ux = (float)px + ox, uy = (float)py + oy;
    const float fi = floorf(ux), fj = floorf(uy);
    const float a = ux - fi, b = uy - fj;
    q.i0 = min(max(f2i(fi), 0), W - 1); q.i1 = min(max(f2i(fi) + 1, 0), W - 1);
    q.j0 = min(max(f2i(fj), 0), H - 1); q.j1 = min(max(f2i(fj) + 1, 0), H - 1);
    q.onTexel = a == 0.0f && b == 0.0f;
    q.w00 = (1.0f - a) * (1.0f - b); q.w10 = a * (1.0f - b); q.w01 = (1.0f - a) * b; q.w11 = a * b;
    return q;
}
__device__ __forceinline__ float4 bilin_mix(const Bilin& q, float4 t00, float4 t10, float4 t01, float4 t11) {
    return make_float4(((q.w00 * t00.x + q.w10 * t10.x) + q.w01 * t01.x) + q.w11 * t11.x, ((q.w00 * t00.y + q.w10 * t10.y) + q.w01 * t01.y) + q.w11 * t11.y,
                       ((q.w00 * t00.z + q.w10 * t10.z) + q.w01 * t01.z) + q.w11 * t11.z, ((q.w00 * t00.w + q.w10 * t10.w) + q.w01 * t01.w) + q.w11 * t11.w);
}

template <bool BLIND, bool LINEAR>
__global__ void __launch_bounds__(256, DENOISE_MIN_BLOCKS) denoise_kernel(const uint8_t* __restrict__ color, const int8_t* __restrict__ normal,
                                                      const float* __restrict__ position, uint8_t* __restrict__ out, int W, int H,
                                                      const DenoiseTaps taps, const DenoiseIter it) {
    __shared__ float lutU[256], lutS[256];
    lutU[threadIdx.x] = (float)threadIdx.x / 255.0f;
    if (!BLIND) lutS[threadIdx.x] = snorm((signed char)threadIdx.x);          // indexed by the byte pattern
    __syncthreads();
    const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (px >= W || py >= H) return;
    const size_t p = (size_t)py * W + px;
    float4 sum = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    float total = 0.0f;
    if (BLIND) {
#pragma unroll 1
        for (int i = 0; i < taps.n; i++) {
            const int tx = min(max(px + taps.dx[i], 0), W - 1), ty = min(max(py + taps.dy[i], 0), H - 1);
            const float4 oc = ld_color(color, (size_t)ty * W + tx, lutU);
            const float k = taps.k[i];
            sum.x += oc.x * k; sum.y += oc.y * k; sum.z += oc.z * k; sum.w += oc.w * k;
            total += k;
        }
    } else {
        const float4 sc = ld_color(color, p, lutU), sn = ld_normal(normal, p, lutS), sp = reinterpret_cast<const float4*>(position)[p];
        const float sw2 = it.stepWidth * it.stepWidth;
        const float rPhiC = 1.0f / it.phiColor, rPhiN = 1.0f / it.phiNormal, rPhiP = 1.0f / it.phiPos, rSw2 = 1.0f / sw2;   // once per thread (IEEE)
#pragma unroll 1
        for (int i = 0; i < taps.n; i++) {
            // uv = vScreenPos + offset*step*stepWidth under LINEAR/CLAMP = the texel holding the sample point; the texel offsets
            // floor(0.5 + offset*stepWidth) are per-iteration constants (taps.dx / dy, computed on the host with the same float operations)
            float4 oc, on, op;
            if (LINEAR) {
                const Bilin b = bilin_setup(px, py, taps.fx[i], taps.fy[i], W, H);
                const size_t q00 = (size_t)b.j0 * W + b.i0, q10 = (size_t)b.j0 * W + b.i1, q01 = (size_t)b.j1 * W + b.i0, q11 = (size_t)b.j1 * W + b.i1;
                if (b.onTexel) {
                    oc = ld_color(color, q00, lutU); on = ld_normal(normal, q00, lutS); op = reinterpret_cast<const float4*>(position)[q00];
                } else {
                    oc = bilin_mix(b, ld_color(color, q00, lutU), ld_color(color, q10, lutU), ld_color(color, q01, lutU), ld_color(color, q11, lutU));
                    on = bilin_mix(b, ld_normal(normal, q00, lutS), ld_normal(normal, q10, lutS), ld_normal(normal, q01, lutS), ld_normal(normal, q11, lutS));
                    const float4* pp = reinterpret_cast<const float4*>(position);
                    op = bilin_mix(b, pp[q00], pp[q10], pp[q01], pp[q11]);
                }
            } else {
                const int tx = min(max(px + taps.dx[i], 0), W - 1), ty = min(max(py + taps.dy[i], 0), H - 1);
                const size_t q = (size_t)ty * W + tx;
                oc = ld_color(color, q, lutU); on = ld_normal(normal, q, lutS); op = reinterpret_cast<const float4*>(position)[q];
            }
            const float cw = fminf(expf(div_const(-(dot4d(sc, oc)), it.phiColor, rPhiC)), 1.0f);
            const float nw = fminf(expf(div_const(-(fmaxf(div_const(dot4d(sn, on), sw2, rSw2), 0.0f)), it.phiNormal, rPhiN)), 1.0f);
            const float pw = fminf(expf(div_const(-(dot4d(sp, op)), it.phiPos, rPhiP)), 1.0f);
            const float weight = (cw * nw) * pw;
            const float k = taps.k[i];
            sum.x += (oc.x * weight) * k; sum.y += (oc.y * weight) * k; sum.z += (oc.z * weight) * k; sum.w += (oc.w * weight) * k;
            total += weight * k;
        }
    }
    {
        // one IEEE reciprocal + corrected products instead of four divisions (the alpha sum is always zero: a slow-path division per pixel);
        // a total that is zero, denormal or not finite (all weights underflowed / NaN inputs) keeps the plain divisions
        const float rt = 1.0f / total;
        const bool plain = !(total >= 1.0e-30f && total <= 1.0e30f);
        reinterpret_cast<uchar4*>(out)[p] = plain ? make_uchar4(to_unorm8(sum.x / total), to_unorm8(sum.y / total), to_unorm8(sum.z / total), to_unorm8(sum.w / total))
                                                  : make_uchar4(to_unorm8(div_const(sum.x, total, rt)), to_unorm8(div_const(sum.y, total, rt)), to_unorm8(div_const(sum.z, total, rt)), to_unorm8(div_const(sum.w, total, rt)));
    }
}

cudaError_t launch_denoise(const uint8_t* color, const int8_t* normal, const float* position, uint8_t* out, int W, int H,
                           const DenoiseTaps& taps, const DenoiseIter& it, cudaStream_t stream) {
    DenoiseTaps t = taps;
    bool integral = true;
    for (int i = 0; i < t.n; i++) {   // the tap offsets of this iteration in texels: offset * stepWidth, the product the shader forms
        t.fx[i] = (float)t.ox[i] * it.stepWidth; t.fy[i] = (float)t.oy[i] * it.stepWidth;
        integral = integral && t.fx[i] == floorf(t.fx[i]) && t.fy[i] == floorf(t.fy[i]) && fabsf(t.fx[i]) < 1.0e6f && fabsf(t.fy[i]) < 1.0e6f;
        t.dx[i] = f2i(t.fx[i]); t.dy[i] = f2i(t.fy[i]);
    }
    const float inf = INFINITY;
    const dim3 grid((W + 31) / 32, (H + 7) / 8);
    const bool blind = it.phiColor == inf && it.phiNormal == inf && it.phiPos == inf;
    if (blind && integral) denoise_kernel<true, false><<<grid, 256, 0, stream>>>(color, normal, position, out, W, H, t, it);
    else if (integral) denoise_kernel<false, false><<<grid, 256, 0, stream>>>(color, normal, position, out, W, H, t, it);
    else denoise_kernel<false, true><<<grid, 256, 0, stream>>>(color, normal, position, out, W, H, t, it);   // (a blind iteration has stepWidth 1: always integral)
    return cudaGetLastError();
}

}  // namespace vrt

// ---- vrt_denoise_static: the denoiser without any allocation (include/vrt.h; no counterpart in the reference).
// The intermediate image of the ping-pong is this file-scope array. In HOST code `g_denoise_scratch` is only a
// stand-in for the symbol, not a device address: the address is resolved once per device with cudaGetSymbolAddress (that device
// current) and it is that pointer the kernels are given. All calls on a device share the array; see the header for the rules.
#include <mutex>
#include "vrt_internal.h"

static __device__ __align__(256) unsigned char g_denoise_scratch[(size_t)VRT_DENOISE_STATIC_MAX_PIXELS * 4];

static constexpr int kScratchMaxDevices = 64;
static std::mutex g_scratchMutex;
static void* g_scratchPtr[kScratchMaxDevices] = {};   // device address of g_denoise_scratch on each device, resolved on first use

// the current device must be `device`
static int denoise_scratch_address(int device, void** out) {
    if (device < 0 || device >= kScratchMaxDevices) return fail(VRT_E_UNSUPPORTED, "vrt_denoise_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(g_scratchMutex);
    if (!g_scratchPtr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_denoise_scratch);
        if (e != cudaSuccess || !p) return fail(VRT_E_CUDA, std::string("cudaGetSymbolAddress(g_denoise_scratch): ") + cudaGetErrorString(e));
        g_scratchPtr[device] = p;
    }
    *out = g_scratchPtr[device];
    return VRT_OK;
}

extern "C" int vrt_denoise_static(vrt_context* c, const void* d_color, const void* d_normal, const void* d_position, int w, int h,
                                  const vrt_denoiser_settings* ds, void* d_out, void* stream) {
    if (!c || !d_color || !d_normal || !d_position || !ds || !d_out || w <= 0 || h <= 0) return fail(VRT_E_INVALID, "null argument");
    if (d_out == d_color) return fail(VRT_E_INVALID, "d_out must not alias d_color (iteration 0 reads neighbours while it writes)");
    if ((long long)w * (long long)h > (long long)VRT_DENOISE_STATIC_MAX_PIXELS)
        return fail(VRT_E_INVALID, "vrt_denoise_static: w * h = " + std::to_string((long long)w * h) + " pixels exceeds VRT_DENOISE_STATIC_MAX_PIXELS = " +
                                       std::to_string((long long)VRT_DENOISE_STATIC_MAX_PIXELS) + " (the built-in scratch image); use vrt_denoise_on with a scratch image of your own");
    CU(cudaSetDevice(c->device));
    void* scratch = nullptr;
    if (int rc = denoise_scratch_address(c->device, &scratch)) return rc;
    void* ping[2] = {d_out, scratch};
    if (((ds->iterations - 1) % 2) == 1) { ping[0] = scratch; ping[1] = d_out; }   // the last iteration lands in d_out
    int last = 0; uint32_t n = 0;
    return run_denoise(c, stream ? (cudaStream_t)stream : c->stream, d_color, d_normal, d_position, w, h, ds, ping, &last, &n);
}
