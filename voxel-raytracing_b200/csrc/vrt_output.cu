// vrt_output.cu — the output stage: BlitStage::record (source/voxels/stages/blit_stage.cpp:25-57) running shader/blit.frag as a CUDA
// kernel, headless: the "window" is a device buffer of the target size. HBM-streaming: algorithmically 4 B read per source pixel the
// footprints cover + 4 B written per target pixel; the four taps of neighbouring pixels overlap and are served by L1/L2.
#include <cuda_runtime.h>
#include <math.h>
#include <mutex>
#include "vrt_output.cuh"
#include "vrt_launch.h"

namespace vrt {

// One block = a 32 x 32 tile of the target, 256 threads, 4 rows per thread (rows ty, ty+8, ty+16, ty+24): the tables and the tile's
// 32 column + 32 row taps are set up once per 1024 pixels, and a thread keeps its column tap in registers.
#define VRT_BLIT_TILE 32
__global__ void __launch_bounds__(256) blit_kernel(const uchar4* __restrict__ src, uchar4* __restrict__ dst, const BlitParams p,
                                                   const float* __restrict__ srgbThr) {
    __shared__ float lutU[256], thr[257];
    __shared__ BlitAxis col[VRT_BLIT_TILE], row[VRT_BLIT_TILE];
    lutU[threadIdx.x] = (float)threadIdx.x / 255.0f;       // exact UNORM8 decode without a division per tap (as in denoise_kernel)
    thr[threadIdx.x] = srgbThr[threadIdx.x];
    // the tile's column and row taps, once: 4 IEEE divisions per tap instead of per pixel (warp 0: columns, warp 1: rows)
    const float scale = blit_scale(p);
    if (threadIdx.x < 32) col[threadIdx.x] = blit_axis(min((int)(blockIdx.x * VRT_BLIT_TILE + threadIdx.x), p.tw - 1), p.sw, p.tw, scale);
    else if (threadIdx.x < 64) row[threadIdx.x - 32] = blit_axis(min((int)(blockIdx.y * VRT_BLIT_TILE + threadIdx.x - 32), p.th - 1), p.sh, p.th, scale);
    __syncthreads();
    const int lx = threadIdx.x & 31, px = blockIdx.x * VRT_BLIT_TILE + lx;
    if (px >= p.tw) return;
    const BlitAxis cx = col[lx];
    const bool srgb = (p.flags & VRT_BLIT_SRGB) != 0, bgra = (p.flags & VRT_BLIT_BGRA) != 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int ly = (threadIdx.x >> 5) + 8 * k, py = blockIdx.y * VRT_BLIT_TILE + ly;
        if (py >= p.th) break;
        const BlitAxis ry = row[ly];
        const uchar4* r0 = src + (size_t)ry.i0 * p.sw; const uchar4* r1 = src + (size_t)ry.i1 * p.sw;
        const uint32_t* w0 = reinterpret_cast<const uint32_t*>(r0); const uint32_t* w1 = reinterpret_cast<const uint32_t*>(r1);
        const uint32_t c00 = __ldg(w0 + cx.i0);
        float r, g, b, a;
        if (__all_sync(__activemask(), cx.a == 0.0f && ry.a == 0.0f)) {
            // sample points on texel centres: the weights are (1, 0, 0, 0) and the mix returns c00's values bit for bit
            r = lut_byte(lutU, c00, 0); g = lut_byte(lutU, c00, 1); b = lut_byte(lutU, c00, 2); a = lut_byte(lutU, c00, 3);
        } else {
            const uint32_t c10 = __ldg(w0 + cx.i1), c01 = __ldg(w1 + cx.i0), c11 = __ldg(w1 + cx.i1);
            const BlitTap t = blit_weights(cx.a, ry.a);
            r = blit_mix(t, lut_byte(lutU, c00, 0), lut_byte(lutU, c10, 0), lut_byte(lutU, c01, 0), lut_byte(lutU, c11, 0));
            g = blit_mix(t, lut_byte(lutU, c00, 1), lut_byte(lutU, c10, 1), lut_byte(lutU, c01, 1), lut_byte(lutU, c11, 1));
            b = blit_mix(t, lut_byte(lutU, c00, 2), lut_byte(lutU, c10, 2), lut_byte(lutU, c01, 2), lut_byte(lutU, c11, 2));
            a = blit_mix(t, lut_byte(lutU, c00, 3), lut_byte(lutU, c10, 3), lut_byte(lutU, c01, 3), lut_byte(lutU, c11, 3));
        }
        uchar4 o;
        if (srgb) { o.x = srgb8_lookup(thr, r); o.y = srgb8_lookup(thr, g); o.z = srgb8_lookup(thr, b); }
        else { o.x = to_unorm8(r); o.y = to_unorm8(g); o.z = to_unorm8(b); }
        o.w = to_unorm8(a);                                      // alpha of an sRGB format is linear
        if (bgra) { const unsigned char sw = o.x; o.x = o.z; o.z = sw; }
        dst[(size_t)py * p.tw + px] = o;
    }
}

static std::mutex g_thrLock;
static float* g_thrDev[64] = {};

cudaError_t launch_blit(const void* src, int sw, int sh, void* dst, int tw, int th, uint32_t flags, cudaStream_t stream) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    {
        std::lock_guard<std::mutex> lk(g_thrLock);
        if (!g_thrDev[dev]) {
            float h[256]; srgb_thresholds(h);
            float* d = nullptr;
            if ((e = cudaMalloc(&d, sizeof(h))) != cudaSuccess) return e;
            if ((e = cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice)) != cudaSuccess) { cudaFree(d); return e; }
            g_thrDev[dev] = d;
        }
    }
    BlitParams p; p.sw = sw; p.sh = sh; p.tw = tw; p.th = th; p.flags = flags;
    dim3 grid((tw + VRT_BLIT_TILE - 1) / VRT_BLIT_TILE, (th + VRT_BLIT_TILE - 1) / VRT_BLIT_TILE);
    blit_kernel<<<grid, 256, 0, stream>>>((const uchar4*)src, (uchar4*)dst, p, g_thrDev[dev]);
    return cudaGetLastError();
}

}  // namespace vrt
