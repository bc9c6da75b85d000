// vrt_upscale.cu — temporal upscaler over the planes the pass already writes: the stage the reference runs after the
// denoiser, UpscalerStage::record -> ffxFsr2ContextDispatch (source/voxels/stages/upscaler_stage.cpp:78-172).
//
// PARITY UNPINNED — OWN DEFINITION. FSR2 is closed in the reference checkout (headers + LFS stubs of a Windows .lib); there is nothing to
// restate and no oracle can exist. What is kept is the stage's INTERFACE (colour, depth, motion vectors and the reactive mask of the
// render resolution in, one image of the target resolution out, per-frame jitter from UpscalerStage::update, reset flag, motion-vector
// scale) and the class of algorithm (jittered temporal accumulation with neighbourhood rectification). Defined here as:
//   1. r = centre of the output pixel in render-pixel units. Pixel (i, j) of this frame was sampled at (i + 0.5 - jx, j + 0.5 + jy): the
//      shader adds cameraJitter / screenSize * (-2, 2) to the ray direction (voxel_volume.frag:322), i.e. -jx / +jy pixels.
//   2. current colour = tent-weighted mean of the (up to) four samples around r, the tent HALF an output pixel in radius: w = max(0,
//      1 - 2|dx| tw/rw) * max(0, 1 - 2|dy| th/rh), dx, dy in render pixels; W = sum of w (a frame whose samples miss the pixel contributes
//      nothing: W = 0). A wider tent, or a floor under alpha, was measured to converge to a blurred / noisy image (profiles/r2_upscaler.txt).
//      When nothing better exists (first frame, reset) the pixel starts from the plain bilinear value with a token weight.
//   3. history (target resolution, RGBA32F: colour + accumulated weight) is fetched at the output pixel moved back by the motion vector
//      of the nearest sample (x motionVectorScale x target/render), nearest texel; rejected when that falls outside the image.
//   4. the history colour is clamped to the min / max of the 3x3 render pixels around the nearest sample (rectification).
//   5. W > 0: alpha = max(W / (W + history weight), W * reactive mask of the nearest sample); out = history + (current - history) * alpha;
//      new weight = min(history weight + W, 15) (the cap keeps alpha >= W / (W + 15): the history never freezes). W = 0: the (rectified) history is kept. reset, or no valid history: out = current.
// Properties tested (tests/test_gpu.py): deterministic; exact pass-through at scale 1, jitter (0, 0); a static scene converges.
// Sharpening (the reference asks FSR2 for RCAS at 1.0) is not part of this definition.
#include <cuda_runtime.h>
#include <new>
#include "vrt_internal.h"
#include "vrt_core.cuh"

struct vrt_upscaler {
    vrt_context* ctx = nullptr;
    int rw = 0, rh = 0, tw = 0, th = 0;
    float4* hist[2] = {nullptr, nullptr};   // ping-pong: the pass reads neighbours of the previous history
    int cur = 0;
    bool valid = false;
};

namespace vrt {

struct UpscaleParams { int rw, rh, tw, th; float jx, jy, mvx, mvy; int useHistory; };

__device__ __forceinline__ float4 up_color(const uint8_t* c, int x, int y, int rw, const float* lut) {
    const uchar4 q = reinterpret_cast<const uchar4*>(c)[(size_t)y * rw + x];
    return make_float4(lut[q.x], lut[q.y], lut[q.z], lut[q.w]);
}

__global__ void __launch_bounds__(256) upscale_kernel(const uint8_t* __restrict__ color, const float* __restrict__ depth, const float* __restrict__ motion,
                                                      const uint8_t* __restrict__ mask, const float4* __restrict__ histIn, float4* __restrict__ histOut,
                                                      uint8_t* __restrict__ out, const UpscaleParams p) {
    __shared__ float lut[256];
    lut[threadIdx.x] = (float)threadIdx.x / 255.0f;
    __syncthreads();
    const int ox = blockIdx.x * 32 + (threadIdx.x & 31), oy = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (ox >= p.tw || oy >= p.th) return;
    (void)depth;                                            // part of the stage's interface (FSR2 uses it for disocclusion); not used by this definition
    // 1. output pixel centre in render-pixel units; sample (i, j) sits at (i + 0.5 - jx, j + 0.5 + jy)
    const float rx = ((float)ox + 0.5f) * ((float)p.rw / (float)p.tw), ry = ((float)oy + 0.5f) * ((float)p.rh / (float)p.th);   // scale 1: exactly ox + 0.5
    const float ux = rx - 0.5f + p.jx, uy = ry - 0.5f - p.jy;   // in units where sample (i, j) is at integer (i, j)
    const float fx = floorf(ux), fy = floorf(uy);
    const float ax = ux - fx, ay = uy - fy;
    const int i0 = f2i(fx), j0 = f2i(fy);
    // 2. tent-weighted current colour over the samples that exist
    const float kx = 2.0f * (float)p.tw / (float)p.rw, ky = 2.0f * (float)p.th / (float)p.rh;   // tent of half an output pixel in radius
    float4 cur = make_float4(0.0f, 0.0f, 0.0f, 0.0f), wide = cur;
    float W = 0.0f, Wwide = 0.0f;
#pragma unroll
    for (int dy = 0; dy < 2; dy++)
#pragma unroll
        for (int dx = 0; dx < 2; dx++) {
            const int i = i0 + dx, j = j0 + dy;
            if (i < 0 || j < 0 || i >= p.rw || j >= p.rh) continue;
            const float ddx = dx ? 1.0f - ax : ax, ddy = dy ? 1.0f - ay : ay;             // distance to the sample, render pixels
            const float wb = (1.0f - ddx) * (1.0f - ddy);
            const float w = fmaxf(1.0f - ddx * kx, 0.0f) * fmaxf(1.0f - ddy * ky, 0.0f);
            if (wb <= 0.0f) continue;
            const float4 c = up_color(color, i, j, p.rw, lut);
            wide.x += c.x * wb; wide.y += c.y * wb; wide.z += c.z * wb; wide.w += c.w * wb; Wwide += wb;
            if (w > 0.0f) { cur.x += c.x * w; cur.y += c.y * w; cur.z += c.z * w; cur.w += c.w * w; W += w; }
        }
    const int ni = min(max(f2i(floorf(ux + 0.5f)), 0), p.rw - 1), nj = min(max(f2i(floorf(uy + 0.5f)), 0), p.rh - 1);   // nearest sample
    const float4 nearest = up_color(color, ni, nj, p.rw, lut);
    if (Wwide > 0.0f) { wide.x /= Wwide; wide.y /= Wwide; wide.z /= Wwide; wide.w /= Wwide; } else wide = nearest;
    if (W > 0.0f) { cur.x /= W; cur.y /= W; cur.z /= W; cur.w /= W; } else cur = wide;
    float4 o = cur;
    float weight = W > 0.0f ? W : 0.015625f;                 // a pixel no sample reached starts from the bilinear value with a token weight
    if (p.useHistory) {
        // 3. history at the output pixel moved back by the nearest sample's motion vector
        const float2 mv = motion ? reinterpret_cast<const float2*>(motion)[(size_t)nj * p.rw + ni] : make_float2(0.0f, 0.0f);
        const int hx = f2i(floorf((float)ox + 0.5f - mv.x * p.mvx * (float)p.tw / (float)p.rw)), hy = f2i(floorf((float)oy + 0.5f - mv.y * p.mvy * (float)p.th / (float)p.rh));
        if (hx >= 0 && hy >= 0 && hx < p.tw && hy < p.th) {
            float4 h = histIn[(size_t)hy * p.tw + hx];
            if (h.w > 0.0f) {
                // 4. rectification: clamp to the 3x3 neighbourhood of the nearest sample
                float3 lo = make_float3(nearest.x, nearest.y, nearest.z), hi = lo;
                for (int dy = -1; dy <= 1; dy++)
                    for (int dx = -1; dx <= 1; dx++) {
                        const float4 c = up_color(color, min(max(ni + dx, 0), p.rw - 1), min(max(nj + dy, 0), p.rh - 1), p.rw, lut);
                        lo.x = fminf(lo.x, c.x); lo.y = fminf(lo.y, c.y); lo.z = fminf(lo.z, c.z);
                        hi.x = fmaxf(hi.x, c.x); hi.y = fmaxf(hi.y, c.y); hi.z = fmaxf(hi.z, c.z);
                    }
                h.x = fminf(fmaxf(h.x, lo.x), hi.x); h.y = fminf(fmaxf(h.y, lo.y), hi.y); h.z = fminf(fmaxf(h.z, lo.z), hi.z);
                // 5. blend
                const float reactive = mask ? lut[mask[(size_t)nj * p.rw + ni]] : 0.0f;
                const float alpha = W > 0.0f ? fmaxf(W / (W + h.w), W * reactive) : 0.0f;
                o.x = h.x + (cur.x - h.x) * alpha; o.y = h.y + (cur.y - h.y) * alpha; o.z = h.z + (cur.z - h.z) * alpha;
                weight = fminf(h.w + W, 15.0f);
            }
        }
    }
    histOut[(size_t)oy * p.tw + ox] = make_float4(o.x, o.y, o.z, weight);
    reinterpret_cast<uchar4*>(out)[(size_t)oy * p.tw + ox] = make_uchar4(to_unorm8(o.x), to_unorm8(o.y), to_unorm8(o.z), to_unorm8(cur.w));
}

}  // namespace vrt

// == UpscalerStage::UpscalerStage / ffxFsr2ContextCreate (upscaler_stage.cpp:18-57): sizes fixed at creation, history owned by the stage
extern "C" int vrt_upscaler_create(vrt_context* c, int renderW, int renderH, int targetW, int targetH, vrt_upscaler** out) {
    if (!c || !out || renderW <= 0 || renderH <= 0 || targetW <= 0 || targetH <= 0) return fail(VRT_E_INVALID, "bad argument");
    CU(cudaSetDevice(c->device));
    vrt_upscaler* u = new (std::nothrow) vrt_upscaler();
    if (!u) return fail(VRT_E_OOM, "host allocation failed");
    u->ctx = c; u->rw = renderW; u->rh = renderH; u->tw = targetW; u->th = targetH;
    for (int k = 0; k < 2; k++) {
        const cudaError_t e = cudaMalloc(&u->hist[k], (size_t)targetW * targetH * sizeof(float4));
        if (e != cudaSuccess) { cudaFree(u->hist[0]); delete u; return fail(e == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA, cudaGetErrorString(e)); }
    }
    *out = u;
    return VRT_OK;
}
extern "C" void vrt_upscaler_destroy(vrt_upscaler* u) {
    if (!u) return;
    cudaSetDevice(u->ctx->device);
    cudaFree(u->hist[0]); cudaFree(u->hist[1]);
    delete u;
}
// == UpscalerStage::record (upscaler_stage.cpp:78-172). Planes are row-major device images of the render resolution; d_motion / d_mask
// may be NULL (no motion, nothing reactive). Asynchronous on `stream` (NULL: the context stream).
extern "C" int vrt_upscaler_dispatch(vrt_upscaler* u, const void* d_color, const void* d_depth, const void* d_motion, const void* d_mask,
                                     float jitterX, float jitterY, float mvScaleX, float mvScaleY, int reset, void* d_out, void* stream) {
    if (!u || !d_color || !d_out) return fail(VRT_E_INVALID, "null argument");
    CU(cudaSetDevice(u->ctx->device));
    vrt::UpscaleParams p{u->rw, u->rh, u->tw, u->th, jitterX, jitterY, mvScaleX, mvScaleY, (!reset && u->valid) ? 1 : 0};
    const dim3 grid((u->tw + 31) / 32, (u->th + 7) / 8);
    vrt::upscale_kernel<<<grid, 256, 0, stream ? (cudaStream_t)stream : u->ctx->stream>>>((const uint8_t*)d_color, (const float*)d_depth, (const float*)d_motion, (const uint8_t*)d_mask,
                                                                                         u->hist[u->cur], u->hist[u->cur ^ 1], (uint8_t*)d_out, p);
    CU(cudaGetLastError());
    u->cur ^= 1; u->valid = true;
    return VRT_OK;
}
