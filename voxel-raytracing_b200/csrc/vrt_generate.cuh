// vrt_generate.cuh — the procedural volumes as a per-voxel device function (shared by the dense generator kernels in
// vrt_generate.cu and the brick-wise sparse generator in vrt_octant.cu). Pure 32-bit integer functions of (x, y, z, seed).
#pragma once
#include <stdint.h>

namespace vrt {

__host__ __device__ inline uint32_t lowbias32(uint32_t x) {   // Chris Wellons' lowbias32 (public domain)
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x;
}
__host__ __device__ inline uint32_t mixh(uint32_t h, uint32_t v) { return lowbias32(h ^ (v + 0x9e3779b9u + (h << 6) + (h >> 2))); }
__host__ __device__ inline uint32_t hash3(uint32_t seed, uint32_t a, uint32_t b, uint32_t c) { return mixh(mixh(mixh(lowbias32(seed), a), b), c); }

// ---- generator 1: noise terrain (5 octaves of bilinear value noise on hashed lattices, cells 256..16) + pillars ----
__device__ inline uint32_t terrain_height(uint32_t seed, uint32_t x, uint32_t z, uint32_t H) {
    uint32_t acc = 0;
    for (uint32_t o = 0; o < 5; o++) {
        const uint32_t sh = 8 - o, cell = 1u << sh;
        const uint32_t ix = x >> sh, fx = x & (cell - 1), iz = z >> sh, fz = z & (cell - 1);
        const uint32_t v00 = hash3(seed + o, ix, iz, 0x5EEDu) & 0xFFFFu, v10 = hash3(seed + o, ix + 1, iz, 0x5EEDu) & 0xFFFFu;
        const uint32_t v01 = hash3(seed + o, ix, iz + 1, 0x5EEDu) & 0xFFFFu, v11 = hash3(seed + o, ix + 1, iz + 1, 0x5EEDu) & 0xFFFFu;
        const uint32_t top = v00 * (cell - fx) + v10 * fx, bot = v01 * (cell - fx) + v11 * fx;   // < 2^24
        const uint32_t val = (top * (cell - fz) + bot * fz) >> (2 * sh);                          // < 2^32 before the shift
        acc += val >> o;
    }
    return H * 3 / 8 + ((acc * (H / 4)) >> 17);
}

// voxel (x, y, z) of generator 1 (terrain) / 2 (sparse lattice shapes); the value the dense kernels store
__device__ inline uint8_t gen_voxel(int generator, uint32_t seed, uint32_t x, uint32_t y, uint32_t z, uint32_t W, uint32_t H, uint32_t D) {
    (void)W; (void)D;
    if (generator == 1) {
        const uint32_t h = terrain_height(seed, x, z, H);
        if (y < h) return (uint8_t)(1 + (y * 8) / H);
        const uint32_t pc = hash3(seed ^ 0xA5A5u, x >> 6, z >> 6, 1u);
        if ((pc & 3u) == 0u) {
            const uint32_t px0 = (x & ~63u) + 8 + ((pc >> 4) % 40u), pz0 = (z & ~63u) + 8 + ((pc >> 12) % 40u);
            if (x >= px0 && x < px0 + 8 && z >= pz0 && z < pz0 + 8 && y < terrain_height(seed, px0, pz0, H) + 16 + ((pc >> 20) % 64u)) return 9;
        }
        return 0;
    }
    const uint32_t hc = mixh(hash3(seed, x >> 6, y >> 6, z >> 6), 0xC0FFEEu);
    if ((hc & 7u) != 0u) return 0;
    const int r = 8 + (int)((hc >> 3) % 17u);
    const uint32_t span = 62u - 2u * (uint32_t)r;
    const int cx = (int)((x & ~63u) + r + 1 + ((hc >> 8) % span)), cy = (int)((y & ~63u) + r + 1 + ((hc >> 14) % span)),
              cz = (int)((z & ~63u) + r + 1 + ((hc >> 20) % span));
    const int dx = (int)x - cx, dy = (int)y - cy, dz = (int)z - cz;
    bool in;
    if ((hc >> 26) & 1u) in = dx * dx + dy * dy + dz * dz <= r * r;
    else { const int q = r * 3 / 4; in = abs(dx) <= q && abs(dy) <= q && abs(dz) <= q; }
    return in ? (uint8_t)(1 + ((hc >> 27) % 12u)) : 0;
}

}  // namespace vrt
