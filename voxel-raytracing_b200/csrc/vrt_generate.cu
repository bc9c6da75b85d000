// vrt_generate.cu — on-GPU procedural volumes for the synthetic configs (BASELINE.json configs 4 and 5). Pure
// functions of (x, y, z, seed) in 32-bit integer arithmetic, so every rank (and the numpy restatement the tests
// use, oracle/synth.py) produces identical bytes without broadcasting a 1-8 GiB volume. Layout is the shader's:
// index = x + y*W + z*W*H with +y up (what voxel_scene.cpp:99 produces from MagicaVoxel's z-up models).
#include <cuda_runtime.h>
#include "vrt_launch.h"
#include "vrt_generate.cuh"

namespace vrt {

__global__ void gen_terrain_kernel(uint32_t seed, uint8_t* __restrict__ vol, uint32_t W, uint32_t H, uint32_t D) {
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x, z = blockIdx.y;
    if (x >= W || z >= D) return;
    const uint32_t h = terrain_height(seed, x, z, H);
    // pillar of this 64x64 column cell (kept inside the cell)
    const uint32_t pc = hash3(seed ^ 0xA5A5u, x >> 6, z >> 6, 1u);
    uint32_t ptop = 0;
    if ((pc & 3u) == 0u) {
        const uint32_t px0 = (x & ~63u) + 8 + ((pc >> 4) % 40u), pz0 = (z & ~63u) + 8 + ((pc >> 12) % 40u);
        if (x >= px0 && x < px0 + 8 && z >= pz0 && z < pz0 + 8) ptop = terrain_height(seed, px0, pz0, H) + 16 + ((pc >> 20) % 64u);
    }
    uint8_t* col = vol + (size_t)x + (size_t)z * W * H;
    for (uint32_t y = 0; y < H; y++) {
        uint8_t v = 0;
        if (y < h) v = (uint8_t)(1 + (y * 8) / H);
        else if (y < ptop) v = 9;
        col[(size_t)y * W] = v;
    }
}

// ---- generator 2: sparse lattice shapes: one sphere or box in 1 of 8 cells of a 64^3 lattice ----
__global__ void gen_sparse_kernel(uint32_t seed, uint8_t* __restrict__ vol, uint32_t W, uint32_t H, uint32_t D) {
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
    if (x >= W) return;
    const uint32_t hc = mixh(hash3(seed, x >> 6, y >> 6, z >> 6), 0xC0FFEEu);
    uint8_t v = 0;
    if ((hc & 7u) == 0u) {
        const int r = 8 + (int)((hc >> 3) % 17u);
        const uint32_t span = 62u - 2u * (uint32_t)r;
        const int cx = (int)((x & ~63u) + r + 1 + ((hc >> 8) % span)), cy = (int)((y & ~63u) + r + 1 + ((hc >> 14) % span)),
                  cz = (int)((z & ~63u) + r + 1 + ((hc >> 20) % span));
        const int dx = (int)x - cx, dy = (int)y - cy, dz = (int)z - cz;
        bool in;
        if ((hc >> 26) & 1u) in = dx * dx + dy * dy + dz * dz <= r * r;
        else { const int q = r * 3 / 4; in = abs(dx) <= q && abs(dy) <= q && abs(dz) <= q; }
        if (in) v = (uint8_t)(1 + ((hc >> 27) % 12u));
    }
    vol[(size_t)x + (size_t)y * W + (size_t)z * W * H] = v;
}

cudaError_t launch_generate(int generator, uint32_t seed, uint8_t* vol, uint32_t W, uint32_t H, uint32_t D, cudaStream_t stream) {
    if (generator == 1) {
        gen_terrain_kernel<<<dim3((W + 127) / 128, D), 128, 0, stream>>>(seed, vol, W, H, D);
    } else if (generator == 2) {
        gen_sparse_kernel<<<dim3((W + 127) / 128, H, D), 128, 0, stream>>>(seed, vol, W, H, D);
    } else {
        return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// Palette of a generated scene: hashed sRGB-free colours (no pow: exact in binary32 everywhere), metallic bands.
void generator_palette(int generator, uint32_t seed, vrt_material palette[256]) {
    for (uint32_t i = 0; i < 256; i++) {
        const uint32_t h = hash3(seed, i, 0x9A1Eu, 0u);
        vrt_material& m = palette[i];
        for (int k = 0; k < 3; k++) m.diffuse[k] = (float)(64u + (((h >> (8 * k)) & 255u) * 3u) / 4u) / 255.0f;
        m.diffuse[3] = 1.0f;
        const bool metal = generator == 1 ? (i == 4 || i == 9) : (i % 4u == 0u);
        m.metallic = (i != 0 && metal) ? 0.5f : 0.0f;
        m.pad1 = m.pad2 = m.pad3 = 0.0f;
    }
    for (int k = 0; k < 4; k++) palette[0].diffuse[k] = 0.0f;
}

}  // namespace vrt
