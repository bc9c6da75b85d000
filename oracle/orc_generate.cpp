// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.h). CPU restatement of the synthetic-volume generators defined in
// DESIGN.md ("synthetic volumes"; product: voxel-raytracing_b200/csrc/vrt_generate.cu; numpy: oracle/synth.py) so
// the CPU baseline / reference arm can build BASELINE configs 4 and 5 without touching the GPU path. These are
// input-data generators (not part of the reference); tests check all three implementations byte for byte.
#include "vrt_oracle.h"
#include <omp.h>

namespace {
inline uint32_t lowbias32(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }
inline uint32_t mixh(uint32_t h, uint32_t v) { return lowbias32(h ^ (v + 0x9e3779b9u + (h << 6) + (h >> 2))); }
inline uint32_t hash3(uint32_t s, uint32_t a, uint32_t b, uint32_t c) { return mixh(mixh(mixh(lowbias32(s), a), b), c); }

uint32_t terrain_height(uint32_t seed, uint32_t x, uint32_t z, uint32_t H) {
    uint32_t acc = 0;
    for (uint32_t o = 0; o < 5; o++) {
        const uint32_t sh = 8 - o, cell = 1u << sh, ix = x >> sh, fx = x & (cell - 1), iz = z >> sh, fz = z & (cell - 1);
        const uint32_t v00 = hash3(seed + o, ix, iz, 0x5EEDu) & 0xFFFFu, v10 = hash3(seed + o, ix + 1, iz, 0x5EEDu) & 0xFFFFu;
        const uint32_t v01 = hash3(seed + o, ix, iz + 1, 0x5EEDu) & 0xFFFFu, v11 = hash3(seed + o, ix + 1, iz + 1, 0x5EEDu) & 0xFFFFu;
        const uint32_t top = v00 * (cell - fx) + v10 * fx, bot = v01 * (cell - fx) + v11 * fx;
        acc += ((top * (cell - fz) + bot * fz) >> (2 * sh)) >> o;
    }
    return H * 3 / 8 + ((acc * (H / 4)) >> 17);
}
}  // namespace

extern "C" int orc_generate(int generator, uint32_t seed, uint8_t* vol, uint32_t W, uint32_t H, uint32_t D) {
    if (!vol || (generator != 1 && generator != 2)) return -1;
#pragma omp parallel for schedule(static)
    for (int64_t z = 0; z < (int64_t)D; z++) {
        if (generator == 1) {
            for (uint32_t x = 0; x < W; x++) {
                const uint32_t h = terrain_height(seed, x, (uint32_t)z, H);
                const uint32_t pc = hash3(seed ^ 0xA5A5u, x >> 6, (uint32_t)z >> 6, 1u);
                uint32_t ptop = 0;
                if ((pc & 3u) == 0u) {
                    const uint32_t px0 = (x & ~63u) + 8 + ((pc >> 4) % 40u), pz0 = ((uint32_t)z & ~63u) + 8 + ((pc >> 12) % 40u);
                    if (x >= px0 && x < px0 + 8 && (uint32_t)z >= pz0 && (uint32_t)z < pz0 + 8) ptop = terrain_height(seed, px0, pz0, H) + 16 + ((pc >> 20) % 64u);
                }
                for (uint32_t y = 0; y < H; y++)
                    vol[(size_t)x + (size_t)y * W + (size_t)z * W * H] = y < h ? (uint8_t)(1 + (y * 8) / H) : (y < ptop ? 9 : 0);
            }
        } else {
            for (uint32_t y = 0; y < H; y++)
                for (uint32_t x = 0; x < W; x++) {
                    const uint32_t hc = mixh(hash3(seed, x >> 6, y >> 6, (uint32_t)z >> 6), 0xC0FFEEu);
                    uint8_t v = 0;
                    if ((hc & 7u) == 0u) {
                        const int r = 8 + (int)((hc >> 3) % 17u);
                        const uint32_t span = 62u - 2u * (uint32_t)r;
                        const int cx = (int)((x & ~63u) + r + 1 + ((hc >> 8) % span)), cy = (int)((y & ~63u) + r + 1 + ((hc >> 14) % span)),
                                  cz = (int)(((uint32_t)z & ~63u) + r + 1 + ((hc >> 20) % span));
                        const int dx = (int)x - cx, dy = (int)y - cy, dz = (int)z - cz;
                        bool in;
                        if ((hc >> 26) & 1u) in = dx * dx + dy * dy + dz * dz <= r * r;
                        else { const int q = r * 3 / 4; in = (dx < 0 ? -dx : dx) <= q && (dy < 0 ? -dy : dy) <= q && (dz < 0 ? -dz : dz) <= q; }
                        if (in) v = (uint8_t)(1 + ((hc >> 27) % 12u));
                    }
                    vol[(size_t)x + (size_t)y * W + (size_t)z * W * H] = v;
                }
        }
    }
    return 0;
}
