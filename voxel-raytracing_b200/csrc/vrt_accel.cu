// vrt_accel.cu — builds the brick acceleration levels on the GPU from the dense palette-indexed volume
// (the volume itself is what VoxelScene uploads as an R8_UINT 3D texture: voxel_scene.cpp:122).
//   bmask: one 64-bit occupancy word per 4x4x4 brick.
//   bdist: per-brick Chebyshev distance (in bricks, capped) to the nearest brick that is non-empty, straddles the
//          volume bounds, or lies outside the grid. Computed by VRT_DIST_CAP rounds of a separable 3x3x3 min filter.
// These only let the traversal skip voxel FETCHES; every DDA step of the reference is still executed.
#include <cuda_runtime.h>
#include "vrt_launch.h"

namespace vrt {

__global__ void build_bmask_kernel(const uint8_t* __restrict__ vol, uint32_t W, uint32_t H, uint32_t D, uint32_t Bx, uint32_t By, uint32_t Bz,
                                   unsigned long long* __restrict__ bmask, uint8_t* __restrict__ bdist) {
    const uint32_t bx = blockIdx.x * blockDim.x + threadIdx.x, by = blockIdx.y * blockDim.y + threadIdx.y, bz = blockIdx.z;
    if (bx >= Bx || by >= By || bz >= Bz) return;
    const uint32_t x0 = bx * 4, y0 = by * 4, z0 = bz * 4;
    const bool inside = x0 + 4 <= W && y0 + 4 <= H && z0 + 4 <= D;
    unsigned long long m = 0ull;
    if (inside && (W & 3u) == 0) {
        for (uint32_t z = 0; z < 4; z++)
            for (uint32_t y = 0; y < 4; y++) {
                const uint32_t v = *reinterpret_cast<const uint32_t*>(vol + (size_t)x0 + (size_t)(y0 + y) * W + (size_t)(z0 + z) * W * H);
                const uint32_t bits = ((v & 0xffu) ? 1u : 0u) | ((v & 0xff00u) ? 2u : 0u) | ((v & 0xff0000u) ? 4u : 0u) | ((v & 0xff000000u) ? 8u : 0u);
                m |= (unsigned long long)bits << (y * 4 + z * 16);
            }
    } else {
        for (uint32_t z = 0; z < 4 && z0 + z < D; z++)
            for (uint32_t y = 0; y < 4 && y0 + y < H; y++)
                for (uint32_t x = 0; x < 4 && x0 + x < W; x++)
                    if (vol[(size_t)(x0 + x) + (size_t)(y0 + y) * W + (size_t)(z0 + z) * W * H]) m |= 1ull << (x + y * 4 + z * 16);
    }
    const size_t b = (size_t)bz * By * Bx + (size_t)by * Bx + bx;
    bmask[b] = m;
    bdist[b] = (m != 0ull || !inside) ? 0 : (uint8_t)VRT_DIST_CAP;
}

// AXIS 0/1: out = min3 along the axis (outside the grid = 0). AXIS 2: out = min(out, min3_z(in) + 1) in place.
template <int AXIS>
__global__ void dist_pass_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int Bx, int By, int Bz) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y, z = blockIdx.z;
    if (x >= Bx || y >= By || z >= Bz) return;
    const size_t b = (size_t)z * By * Bx + (size_t)y * Bx + x;
    const size_t stride = AXIS == 0 ? 1 : (AXIS == 1 ? (size_t)Bx : (size_t)Bx * By);
    const int c = AXIS == 0 ? x : (AXIS == 1 ? y : z), n = AXIS == 0 ? Bx : (AXIS == 1 ? By : Bz);
    const int lo = c > 0 ? in[b - stride] : 0, mid = in[b], hi = c + 1 < n ? in[b + stride] : 0;
    const int m = min(mid, min(lo, hi));
    if (AXIS == 2) { const int cur = out[b]; out[b] = (uint8_t)min(cur, m + 1); }
    else out[b] = (uint8_t)m;
}

cudaError_t launch_build_accel(const uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, unsigned long long* bmask, uint8_t* bdist,
                               uint8_t* tmpA, uint8_t* tmpB, cudaStream_t stream, int* launches) {
    const uint32_t Bx = (W + 3) / 4, By = (H + 3) / 4, Bz = (D + 3) / 4;
    const dim3 block(32, 8), grid((Bx + 31) / 32, (By + 7) / 8, Bz);
    build_bmask_kernel<<<grid, block, 0, stream>>>(vol, W, H, D, Bx, By, Bz, bmask, bdist);
    int n = 1;
    for (int it = 0; it < VRT_DIST_CAP; it++) {
        dist_pass_kernel<0><<<grid, block, 0, stream>>>(bdist, tmpA, (int)Bx, (int)By, (int)Bz);
        dist_pass_kernel<1><<<grid, block, 0, stream>>>(tmpA, tmpB, (int)Bx, (int)By, (int)Bz);
        dist_pass_kernel<2><<<grid, block, 0, stream>>>(tmpB, bdist, (int)Bx, (int)By, (int)Bz);
        n += 3;
    }
    if (launches) *launches = n;
    return cudaGetLastError();
}

}  // namespace vrt
