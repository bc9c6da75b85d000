// vrt_denoise.cu — edge-avoiding a-trous filter: one iteration of shader/denoiser.frag:38-73 as a CUDA kernel.
// Host side (tap tables, per-iteration phi/stepWidth, ping-pong) is in vrt_api.cu and follows
// source/voxels/stages/denoiser_stage.cpp:37-61,143-154,203-257. HBM-bound: 24 B read + 4 B written per pixel
// per iteration algorithmically; taps of neighbouring pixels overlap and are served by L1/L2.
#include <cuda_runtime.h>
#include "vrt_core.cuh"
#include "vrt_launch.h"

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

__global__ void __launch_bounds__(256) denoise_kernel(const uint8_t* __restrict__ color, const int8_t* __restrict__ normal,
                                                      const float* __restrict__ position, uint8_t* __restrict__ out, int W, int H,
                                                      const DenoiseTaps taps, const DenoiseIter it) {
    __shared__ float lutU[256], lutS[256];
    lutU[threadIdx.x] = (float)threadIdx.x / 255.0f;
    lutS[threadIdx.x] = snorm((signed char)threadIdx.x);          // indexed by the byte pattern
    __syncthreads();
    const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (px >= W || py >= H) return;
    const size_t p = (size_t)py * W + px;
    const float4 sc = ld_color(color, p, lutU), sn = ld_normal(normal, p, lutS), sp = reinterpret_cast<const float4*>(position)[p];
    float4 sum = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    float total = 0.0f;
    const float sw2 = it.stepWidth * it.stepWidth;
#pragma unroll 1
    for (int i = 0; i < taps.n; i++) {
        // uv = vScreenPos + offset*step*stepWidth under LINEAR/CLAMP = the texel holding the sample point
        int tx = px + f2i(floorf(0.5f + (float)taps.ox[i] * it.stepWidth));
        int ty = py + f2i(floorf(0.5f + (float)taps.oy[i] * it.stepWidth));
        tx = min(max(tx, 0), W - 1); ty = min(max(ty, 0), H - 1);
        const size_t q = (size_t)ty * W + tx;
        const float4 oc = ld_color(color, q, lutU);
        const float cw = fminf(expf(-(dot4d(sc, oc)) / it.phiColor), 1.0f);
        const float4 on = ld_normal(normal, q, lutS);
        const float nw = fminf(expf(-(fmaxf(dot4d(sn, on) / sw2, 0.0f)) / it.phiNormal), 1.0f);
        const float4 op = reinterpret_cast<const float4*>(position)[q];
        const float pw = fminf(expf(-(dot4d(sp, op)) / it.phiPos), 1.0f);
        const float weight = (cw * nw) * pw;
        const float k = taps.k[i];
        sum.x += (oc.x * weight) * k; sum.y += (oc.y * weight) * k; sum.z += (oc.z * weight) * k; sum.w += (oc.w * weight) * k;
        total += weight * k;
    }
    reinterpret_cast<uchar4*>(out)[p] = make_uchar4(to_unorm8(sum.x / total), to_unorm8(sum.y / total), to_unorm8(sum.z / total), to_unorm8(sum.w / total));
}

cudaError_t launch_denoise(const uint8_t* color, const int8_t* normal, const float* position, uint8_t* out, int W, int H,
                           const DenoiseTaps& taps, const DenoiseIter& it, cudaStream_t stream) {
    denoise_kernel<<<dim3((W + 31) / 32, (H + 7) / 8), 256, 0, stream>>>(color, normal, position, out, W, H, taps, it);
    return cudaGetLastError();
}

}  // namespace vrt
