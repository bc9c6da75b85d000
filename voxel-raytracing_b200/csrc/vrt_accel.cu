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
    if (launches) *launches += n;
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// Per-voxel safe-step field (fdist). Fine part: Chebyshev distance to the nearest solid / out-of-bounds voxel,
// exact up to VRT_FINE_CAP, by VRT_FINE_CAP-1 rounds of a separable 3x3x3 min (+1) over the linear volume.
__global__ void fine_init_kernel(const uint8_t* __restrict__ vol, uint8_t* __restrict__ a, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = vol[i] ? 0 : (uint8_t)VRT_FINE_CAP;
}
template <int AXIS>
__global__ void fine_pass_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int W, int H, int D) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
    if (x >= W) return;
    const size_t v = (size_t)x + (size_t)y * W + (size_t)z * W * H;
    const size_t stride = AXIS == 0 ? 1 : (AXIS == 1 ? (size_t)W : (size_t)W * H);
    const int c = AXIS == 0 ? x : (AXIS == 1 ? y : z), n = AXIS == 0 ? W : (AXIS == 1 ? H : D);
    const int lo = c > 0 ? in[v - stride] : 0, mid = in[v], hi = c + 1 < n ? in[v + stride] : 0;
    const int m = min(mid, min(lo, hi));
    if (AXIS == 2) { const int cur = out[v]; out[v] = (uint8_t)min(cur, m + 1); }
    else out[v] = (uint8_t)m;
}
// same passes, 4 voxels along x per thread (W % 4 == 0): byte-wise SIMD min
template <int AXIS>
__global__ void fine_pass4_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int W4, int H, int D) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
    if (x >= W4) return;
    const size_t v = (size_t)x + (size_t)y * W4 + (size_t)z * W4 * H;
    uint32_t m;
    if (AXIS == 0) {
        const uint32_t mid = in[v], l = x > 0 ? in[v - 1] : 0u, r = x + 1 < W4 ? in[v + 1] : 0u;
        const uint32_t left = (mid << 8) | (l >> 24), right = (mid >> 8) | (r << 24);   // neighbours at x-1 / x+1 per byte lane
        m = __vminu4(mid, __vminu4(left, right));
    } else {
        const size_t stride = AXIS == 1 ? (size_t)W4 : (size_t)W4 * H;
        const int c = AXIS == 1 ? y : z, n = AXIS == 1 ? H : D;
        const uint32_t lo = c > 0 ? in[v - stride] : 0u, hi = c + 1 < n ? in[v + stride] : 0u;
        m = __vminu4(in[v], __vminu4(lo, hi));
    }
    if (AXIS == 2) out[v] = __vminu4(out[v], m + 0x01010101u);   // values <= VRT_FINE_CAP: no carry between bytes
    else out[v] = m;
}
// merge with the far-field bound of the brick level and store 4x4x4-bricked
__global__ void fine_merge_kernel(const uint8_t* __restrict__ fine, const uint8_t* __restrict__ bdist, uint8_t* __restrict__ fdist,
                                  uint8_t* __restrict__ ldist, uint32_t W, uint32_t H, uint32_t D, uint32_t Bx, uint32_t By) {
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
    if (x >= W) return;
    const uint32_t f = fine[(size_t)x + (size_t)y * W + (size_t)z * W * H];
    const size_t b = (size_t)(z >> 2) * By * Bx + (size_t)(y >> 2) * Bx + (x >> 2);
    const uint32_t d = bdist[b];
    uint32_t n = f;
    if (f != 0 && d != 0) n = max(f, (d - 1u) * 4u + 1u);
    n = min(n, VRT_SAFE_MAX);
    if (fdist) fdist[b * 64u + ((x & 3u) | ((y & 3u) << 2) | ((z & 3u) << 4))] = (uint8_t)n;
    if (ldist) ldist[(size_t)(x + 1u) + (size_t)(y + 1u) * (W + 2u) + (size_t)(z + 1u) * (W + 2u) * (H + 2u)] = (uint8_t)n;
}

cudaError_t launch_build_fdist(const uint8_t* vol, const uint8_t* bdist, uint32_t W, uint32_t H, uint32_t D, uint8_t* fdist, uint8_t* ldist,
                               uint8_t* a, uint8_t* t1, uint8_t* t2, cudaStream_t stream, int* launches) {
    const size_t n = (size_t)W * H * D;
    const uint32_t Bx = (W + 3) / 4, By = (H + 3) / 4, Bz = (D + 3) / 4;
    // either output may be null (only the field of the back-end asked for is built)
    cudaError_t e = fdist ? cudaMemsetAsync(fdist, 0, (size_t)Bx * By * Bz * 64u, stream) : cudaSuccess;   // padding voxels of edge bricks
    if (e != cudaSuccess) return e;
    e = ldist ? cudaMemsetAsync(ldist, (int)VRT_LIN_OOB, (size_t)(W + 2u) * (H + 2u) * (D + 2u), stream) : cudaSuccess;   // apron
    if (e != cudaSuccess) return e;
    fine_init_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(vol, a, n);
    int cnt = 1;
    const bool vec = (W % 4u) == 0;
    for (int r = 1; r < VRT_FINE_CAP; r++) {
        if (vec) {
            const int W4 = (int)(W / 4);
            const dim3 block(128), grid((W4 + 127) / 128, H, D);
            fine_pass4_kernel<0><<<grid, block, 0, stream>>>((const uint32_t*)a, (uint32_t*)t1, W4, (int)H, (int)D);
            fine_pass4_kernel<1><<<grid, block, 0, stream>>>((const uint32_t*)t1, (uint32_t*)t2, W4, (int)H, (int)D);
            fine_pass4_kernel<2><<<grid, block, 0, stream>>>((const uint32_t*)t2, (uint32_t*)a, W4, (int)H, (int)D);
        } else {
            const dim3 block(128), grid((W + 127) / 128, H, D);
            fine_pass_kernel<0><<<grid, block, 0, stream>>>(a, t1, (int)W, (int)H, (int)D);
            fine_pass_kernel<1><<<grid, block, 0, stream>>>(t1, t2, (int)W, (int)H, (int)D);
            fine_pass_kernel<2><<<grid, block, 0, stream>>>(t2, a, (int)W, (int)H, (int)D);
        }
        cnt += 3;
    }
    fine_merge_kernel<<<dim3((W + 127) / 128, H, D), 128, 0, stream>>>(a, bdist, fdist, ldist, W, H, D, Bx, By);
    cnt++;
    if (launches) *launches += cnt;
    return cudaGetLastError();
}

}  // namespace vrt
