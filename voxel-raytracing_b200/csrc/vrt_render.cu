// vrt_render.cu — the voxel_volume pass as a CUDA kernel for sm_100a (replaces shader/voxel_volume.frag driven by
// GeometryStage::record, geometry_stage.cpp:106-174). One thread = one pixel = one fragment invocation; a warp
// covers an 8x4 pixel patch (ray coherence: at 4K adjacent primary rays walk the same bricks), a 128-thread
// block covers 16x8 pixels, 8 blocks cover one 32x32 shard tile.
#include <cuda_runtime.h>
#include "vrt_wave.cuh"
#include "vrt_launch.h"

namespace vrt {

#ifndef VRT_MIN_BLOCKS
#define VRT_MIN_BLOCKS 1
#endif

template <int TRAVERSAL>
__global__ void __launch_bounds__(128, VRT_MIN_BLOCKS) render_kernel(const SceneDev sc, const FrameParams fp, const PlanesDev pl, Counters* __restrict__ ctr) {
    const int W = fp.cam.screenSize[0], H = fp.cam.screenSize[1];
    const int j = blockIdx.x >> 3, sub = blockIdx.x & 7;           // local tile, 16x8 sub-block of the tile
    const int t = j * fp.world + fp.rank;                          // global tile (interleaved shard)
    const int tx = t % fp.tilesX, ty = t / fp.tilesX;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int lx = (sub & 1) * 16 + (warp & 1) * 8 + (lane & 7);    // pixel inside the tile
    const int ly = (sub >> 1) * 8 + (warp >> 1) * 4 + (lane >> 3);
    const int px = tx * VRT_TILE + lx, py = ty * VRT_TILE + ly;
    const bool active = px < W && py < H;

    unsigned rays = 0, fetches = 0;
    if (active) {
        PixelOut o;
        o.hx = o.hy = o.hz = 0; o.steps = 0; o.palette = 0;
        render_pixel<TRAVERSAL>(sc, fp, px, py, o);
        rays = o.rays; fetches = o.fetches;
        const size_t p = fp.world == 1 ? (size_t)py * W + px : (size_t)j * (VRT_TILE * VRT_TILE) + ly * VRT_TILE + lx;
        const uchar4 c4 = make_uchar4(to_unorm8(o.color.x), to_unorm8(o.color.y), to_unorm8(o.color.z), 0);
        const size_t xp = (size_t)py * W + px;                      // fused exchange (XPlanes): the pixel in rank 0's row-major planes
        reinterpret_cast<uchar4*>(pl.color)[p] = c4;
        if (fp.x.color) reinterpret_cast<uchar4*>(fp.x.color)[xp] = c4;
        if (fp.x.colorHost) reinterpret_cast<uchar4*>(fp.x.colorHost)[xp] = c4;
        if (fp.out_mask & VRT_OUT_GBUFFER) {
            const float4 pos4 = make_float4(o.pos.x, o.pos.y, o.pos.z, 0.0f);
            const char4 n4 = make_char4(to_snorm8(o.normal.x), to_snorm8(o.normal.y), to_snorm8(o.normal.z), 0);
            pl.depth[p] = o.depth;
            reinterpret_cast<float2*>(pl.motion)[p] = make_float2(0.0f, 0.0f);
            pl.mask[p] = to_unorm8(o.mask);
            reinterpret_cast<float4*>(pl.position)[p] = pos4;
            reinterpret_cast<char4*>(pl.normal)[p] = n4;
            if (fp.x.depth) fp.x.depth[xp] = o.depth;
            if (fp.x.motion) reinterpret_cast<float2*>(fp.x.motion)[xp] = make_float2(0.0f, 0.0f);
            if (fp.x.mask) fp.x.mask[xp] = to_unorm8(o.mask);
            if (fp.x.position) reinterpret_cast<float4*>(fp.x.position)[xp] = pos4;
            if (fp.x.normal) reinterpret_cast<char4*>(fp.x.normal)[xp] = n4;
        }
        if (fp.out_mask & VRT_OUT_INSTRUMENT) {
            auto sat = [](int v) { return (short)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); };
            reinterpret_cast<short4*>(pl.hit)[p] = make_short4(sat(o.hx), sat(o.hy), sat(o.hz), (short)o.steps);
            pl.hit_palette[p] = (uint8_t)o.palette;
            pl.color_f32[3 * p + 0] = o.color.x; pl.color_f32[3 * p + 1] = o.color.y; pl.color_f32[3 * p + 2] = o.color.z;
            pl.rays[p] = o.rays;
            pl.fetches[p] = o.fetches;
        }
    }
    // frame totals: one atomic per warp
    const unsigned r = __reduce_add_sync(0xffffffffu, rays);
    const unsigned f = __reduce_add_sync(0xffffffffu, fetches);
    const unsigned n = __popc(__ballot_sync(0xffffffffu, active));
    if (lane == 0) {
        atomicAdd(&ctr->rays, (unsigned long long)r); atomicAdd(&ctr->stageRays[0], (unsigned long long)r);
        atomicAdd(&ctr->fetches, (unsigned long long)f); atomicAdd(&ctr->stageFetches[0], (unsigned long long)f);
        atomicAdd(&ctr->pixels, (unsigned long long)n);
    }
}

cudaError_t launch_render(const SceneDev& sc, const FrameParams& fp, const PlanesDev& pl, Counters* ctr, cudaStream_t stream) {
    if (fp.localTiles <= 0) return cudaSuccess;
    const dim3 grid((unsigned)fp.localTiles * 8u), block(128);
    if (fp.traversal == VRT_TRAVERSAL_DENSE_U8) render_kernel<VRT_TRAVERSAL_DENSE_U8><<<grid, block, 0, stream>>>(sc, fp, pl, ctr);
    else if (fp.traversal == VRT_TRAVERSAL_FINE) render_kernel<VRT_TRAVERSAL_FINE><<<grid, block, 0, stream>>>(sc, fp, pl, ctr);
    else if (fp.traversal == VRT_TRAVERSAL_LINEAR) render_kernel<VRT_TRAVERSAL_LINEAR><<<grid, block, 0, stream>>>(sc, fp, pl, ctr);
    else if (fp.traversal == VRT_TRAVERSAL_OCTANT) render_kernel<VRT_TRAVERSAL_OCTANT><<<grid, block, 0, stream>>>(sc, fp, pl, ctr);
    else if (fp.traversal == VRT_TRAVERSAL_SPARSE) render_kernel<VRT_TRAVERSAL_SPARSE><<<grid, block, 0, stream>>>(sc, fp, pl, ctr);
    else render_kernel<VRT_TRAVERSAL_BRICK><<<grid, block, 0, stream>>>(sc, fp, pl, ctr);
    return cudaGetLastError();
}

// ---- de-tile: tile-major shard buffers (gathered rank-major) -> row-major frame on the assembling rank ----
__global__ void detile_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int W, int H, int tilesX, int world,
                              size_t rankStridePixels, int bpp) {
    const int px = blockIdx.x * blockDim.x + threadIdx.x, py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= W || py >= H) return;
    const size_t sp = gathered_index(px, py, tilesX, world, rankStridePixels);
    const size_t dp = (size_t)py * W + px;
    if (bpp == 4) reinterpret_cast<uint32_t*>(dst)[dp] = reinterpret_cast<const uint32_t*>(src)[sp];
    else if (bpp == 16) reinterpret_cast<uint4*>(dst)[dp] = reinterpret_cast<const uint4*>(src)[sp];
    else if (bpp == 8) reinterpret_cast<uint2*>(dst)[dp] = reinterpret_cast<const uint2*>(src)[sp];
    else for (int k = 0; k < bpp; k++) dst[dp * bpp + k] = src[sp * bpp + k];
}

cudaError_t launch_detile(const void* src, void* dst, int W, int H, int world, size_t rankStridePixels, int bpp, cudaStream_t stream) {
    const int tilesX = (W + VRT_TILE - 1) / VRT_TILE;
    dim3 block(32, 8), grid((W + 31) / 32, (H + 7) / 8);
    detile_kernel<<<grid, block, 0, stream>>>((const uint8_t*)src, (uint8_t*)dst, W, H, tilesX, world, rankStridePixels, bpp);
    return cudaGetLastError();
}

// ---- push: this rank's tile-major shard -> its pixels of a row-major frame that may live on ANOTHER GPU (peer memory over
// NVLink: plain stores, 128-byte rows of a tile are contiguous in both layouts). Replaces gather + de-tile on the assembling rank.
__global__ void push_tiles_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int W, int H, int tilesX, int rank, int world,
                                  int localTiles, int bpp) {
    const int j = blockIdx.x, lx = threadIdx.x, ly0 = threadIdx.y;
    if (j >= localTiles) return;
    const int t = rank + j * world, px = (t % tilesX) * VRT_TILE + lx;
    if (px >= W) return;
    for (int ly = ly0; ly < VRT_TILE; ly += blockDim.y) {
        const int py = (t / tilesX) * VRT_TILE + ly;
        if (py >= H) break;
        const size_t sp = (size_t)j * (VRT_TILE * VRT_TILE) + ly * VRT_TILE + lx, dp = (size_t)py * W + px;
        if (bpp == 4) reinterpret_cast<uint32_t*>(dst)[dp] = reinterpret_cast<const uint32_t*>(src)[sp];
        else if (bpp == 16) reinterpret_cast<uint4*>(dst)[dp] = reinterpret_cast<const uint4*>(src)[sp];
        else if (bpp == 8) reinterpret_cast<uint2*>(dst)[dp] = reinterpret_cast<const uint2*>(src)[sp];
        else for (int k = 0; k < bpp; k++) dst[dp * bpp + k] = src[sp * bpp + k];
    }
}

cudaError_t launch_push_tiles(const void* src, void* dst, int W, int H, int rank, int world, int localTiles, int bpp, cudaStream_t stream) {
    if (localTiles <= 0) return cudaSuccess;
    const int tilesX = (W + VRT_TILE - 1) / VRT_TILE;
    push_tiles_kernel<<<(unsigned)localTiles, dim3(VRT_TILE, 8), 0, stream>>>((const uint8_t*)src, (uint8_t*)dst, W, H, tilesX, rank, world, localTiles, bpp);
    return cudaGetLastError();
}

}  // namespace vrt
