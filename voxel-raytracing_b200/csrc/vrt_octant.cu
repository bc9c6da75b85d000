// vrt_octant.cu — builds the directional safe-step fields (SceneDev::odist, vrt_types.h) on the GPU.
//
// odist[o][v] = largest level n of VRT_OCT_LEVEL_LIST such that the cube of side n anchored at v and extending towards
// octant o (offsets 0..n-1 per axis, in the octant's sign direction) contains no solid and no out-of-bounds voxel; 0 for a
// solid voxel; apron = VRT_LIN_OOB. Computed exactly on bit-volumes (1 bit per voxel, 32 voxels along x per word):
//   E_1(v)      = voxel v is empty (out of bounds reads 0)
//   E_{m+d}(v)  = AND over (a,b,c) in {0,d}^3 of E_m(v + sign_o * (a,b,c))        for d <= m  (8 cubes of side m cover side m+d)
// one fused kernel per level (x: funnel shifts inside the row, y/z: row offsets), then one pass that counts the levels each
// voxel satisfies (bit-sliced adder) and writes the byte field. Load-time only: HBM-streaming, ~200 launches per scene.
#include <cuda_runtime.h>
#include "vrt_launch.h"

namespace vrt {

__constant__ uint8_t c_octLevels[VRT_OCT_NLEVELS + 1];

// level 0: emptiness bits
__global__ void oct_init_kernel(const uint8_t* __restrict__ vol, uint32_t* __restrict__ bits, uint32_t W, uint32_t H, uint32_t WW) {
    const uint32_t wx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y, z = blockIdx.z;
    if (wx >= WW || y >= H) return;
    const uint8_t* row = vol + (size_t)y * W + (size_t)z * W * H;
    uint32_t m = 0;
    const uint32_t x0 = wx * 32u;
    if (x0 + 32u <= W && (W & 15u) == 0) {
        const uint4 a = *reinterpret_cast<const uint4*>(row + x0), b = *reinterpret_cast<const uint4*>(row + x0 + 16);
        const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t v = w[k];
            m |= (((v & 0xffu) ? 0u : 1u) | ((v & 0xff00u) ? 0u : 2u) | ((v & 0xff0000u) ? 0u : 4u) | ((v & 0xff000000u) ? 0u : 8u)) << (4 * k);
        }
    } else {
        for (uint32_t b = 0; b < 32u && x0 + b < W; b++) if (row[x0 + b] == 0) m |= 1u << b;
    }
    bits[(size_t)wx + (size_t)y * WW + (size_t)z * WW * H] = m;
}

// row word wx of E shifted so that bit x holds E(x + sx*d): sx = +1 -> towards higher x
__device__ __forceinline__ uint32_t oct_shifted(const uint32_t* __restrict__ row, int wx, int WW, int d, int sx) {
    const int q = d >> 5, r = d & 31;
    if (sx > 0) {
        const int i0 = wx + q, i1 = wx + q + 1;
        const uint32_t lo = i0 < WW ? row[i0] : 0u, hi = (r && i1 < WW) ? row[i1] : 0u;
        return __funnelshift_r(lo, hi, r);
    }
    const int i1 = wx - q, i0 = wx - q - 1;
    const uint32_t hi = i1 >= 0 ? row[i1] : 0u, lo = (r && i0 >= 0) ? row[i0] : 0u;
    return __funnelshift_l(lo, hi, r);
}

// next = level m+d from prev = level m
__global__ void oct_level_kernel(const uint32_t* __restrict__ prev, uint32_t* __restrict__ next, int WW, int H, int D, int d, int sx, int sy, int sz) {
    const int wx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y, z = blockIdx.z;
    if (wx >= WW || y >= H) return;
    uint32_t acc = 0xffffffffu;
#pragma unroll
    for (int c = 0; c < 2; c++) {
        const int zz = z + (c ? sz * d : 0);
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const int yy = y + (b ? sy * d : 0);
            if (zz < 0 || zz >= D || yy < 0 || yy >= H) { acc = 0u; continue; }
            const uint32_t* row = prev + (size_t)yy * WW + (size_t)zz * WW * H;
            acc &= row[wx] & oct_shifted(row, wx, WW, d, sx);
        }
    }
    next[(size_t)wx + (size_t)y * WW + (size_t)z * WW * H] = acc;
}

// count the levels each voxel satisfies (levels are nested: E_n subset of E_m for n > m) and write the aproned byte field
__global__ void oct_final_kernel(const uint32_t* __restrict__ levels, size_t levelWords, uint8_t* __restrict__ out, uint32_t W, uint32_t H, uint32_t WW) {
    const uint32_t wx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y, z = blockIdx.z;
    if (wx >= WW || y >= H) return;
    const size_t i = (size_t)wx + (size_t)y * WW + (size_t)z * WW * H;
    uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0;
    for (int l = 0; l < VRT_OCT_NLEVELS; l++) {
        uint32_t w = levels[(size_t)l * levelWords + i], t;
        t = c0 & w; c0 ^= w; w = t;
        t = c1 & w; c1 ^= w; w = t;
        t = c2 & w; c2 ^= w; w = t;
        t = c3 & w; c3 ^= w; w = t;
        c4 ^= w;
    }
    uint8_t* dst = out + (size_t)(wx * 32u + 1u) + (size_t)(y + 1u) * (W + 2u) + (size_t)(z + 1u) * (W + 2u) * (H + 2u);
    for (uint32_t b = 0; b < 32u && wx * 32u + b < W; b++) {
        const uint32_t k = ((c0 >> b) & 1u) | (((c1 >> b) & 1u) << 1) | (((c2 >> b) & 1u) << 2) | (((c3 >> b) & 1u) << 3) | (((c4 >> b) & 1u) << 4);
        dst[b] = c_octLevels[k];
    }
}

size_t octant_scratch_bytes(uint32_t W, uint32_t H, uint32_t D) {
    const size_t WW = (W + 31u) / 32u;
    return (size_t)VRT_OCT_NLEVELS * WW * H * D * sizeof(uint32_t);
}

// odist[0..7]: (W+2)(H+2)(D+2) bytes each; scratch: octant_scratch_bytes()
cudaError_t launch_build_odist(const uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, uint8_t* const odist[8], uint32_t* scratch,
                               cudaStream_t stream, int* launches) {
    static const int lv[VRT_OCT_NLEVELS] = VRT_OCT_LEVEL_LIST;
    uint8_t table[VRT_OCT_NLEVELS + 1];
    table[0] = 0;
    for (int l = 0; l < VRT_OCT_NLEVELS; l++) table[l + 1] = (uint8_t)safe_encode((uint32_t)lv[l]);
    cudaError_t e = cudaMemcpyToSymbolAsync(c_octLevels, table, sizeof(table), 0, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;
    const uint32_t WW = (W + 31u) / 32u;
    const size_t levelWords = (size_t)WW * H * D;
    const dim3 block(32, 8), grid((WW + 31u) / 32u, (H + 7u) / 8u, D);
    oct_init_kernel<<<grid, block, 0, stream>>>(vol, scratch, W, H, WW);
    int cnt = 1;
    for (int o = 0; o < 8; o++) {
        const int sx = (o & 1) ? -1 : 1, sy = (o & 2) ? -1 : 1, sz = (o & 4) ? -1 : 1;
        e = cudaMemsetAsync(odist[o], (int)VRT_LIN_OOB, (size_t)(W + 2u) * (H + 2u) * (D + 2u), stream);   // apron
        if (e != cudaSuccess) return e;
        for (int l = 1; l < VRT_OCT_NLEVELS; l++) {
            oct_level_kernel<<<grid, block, 0, stream>>>(scratch + (size_t)(l - 1) * levelWords, scratch + (size_t)l * levelWords, (int)WW, (int)H, (int)D,
                                                         lv[l] - lv[l - 1], sx, sy, sz);
            cnt++;
        }
        oct_final_kernel<<<grid, block, 0, stream>>>(scratch, levelWords, odist[o], W, H, WW);
        cnt++;
    }
    if (launches) *launches += cnt;
    return cudaGetLastError();
}

}  // namespace vrt
