// vrt_types.h — POD parameter blocks shared by the host API and the kernels (device views, not ABI).
#pragma once
#include <stdint.h>
#include "../../include/vrt.h"

#if defined(__CUDACC__)
#define VRT_HD __host__ __device__ __forceinline__
#else
#define VRT_HD inline
#endif

namespace vrt {

// Brick acceleration (exact replay: it only elides voxel fetches, never DDA steps — SURVEY.md F11).
//   bmask[b] : 4x4x4 occupancy bits of brick b, bit = (x&3) | (y&3)<<2 | (z&3)<<4
//   bdist[b] : 0 if the brick holds a solid voxel or straddles the volume bounds; otherwise the Chebyshev
//              distance (in bricks, capped at VRT_DIST_CAP) to the nearest such brick or to outside the grid.
//              From any voxel of a brick with bdist = n >= 1 the next (n-1)*4 positions of ANY ray are empty and
//              in bounds, so (n-1)*4+1 DDA steps can run without looking at memory.
constexpr int VRT_DIST_CAP = 33;      // (33-1)*4+1 = 129 unchecked steps at most per lookup

struct SceneDev {
    const uint8_t* vol;               // dense u8[W*H*D], x fastest (voxel_scene.cpp:99)
    const unsigned long long* bmask;  // [Bz][By][Bx]
    const uint8_t* bdist;             // [Bz][By][Bx]
    uint32_t W, H, D;
    uint32_t Bx, By, Bz;
    const vrt_material* palette;      // 256 entries
    const uint8_t* noise;             // 512x512 RGBA8 or null
    const float* sky;                 // skyW*skyH RGBA32F or null
    uint32_t skyW, skyH;
};

struct PlanesDev {
    uint8_t* color;       // RGBA8
    float* depth;
    float* motion;
    uint8_t* mask;
    float* position;      // RGBA32F
    int8_t* normal;       // RGBA8S
    int16_t* hit;         // x,y,z,steps
    uint8_t* hit_palette;
    float* color_f32;     // RGB
    uint32_t* rays;
    uint32_t* fetches;
};

// Tile sharding: the image is cut into TILE x TILE pixel tiles, numbered row-major; rank r of `world` owns tiles
// t with t % world == r. With world == 1 planes are plain row-major frames; with world > 1 each rank stores its
// tiles compactly, tile-major (local tile j = t / world, TILE*TILE pixels each, row-major inside the tile).
constexpr int VRT_TILE = 32;

struct FrameParams {
    vrt_camera cam;
    uint32_t pass_mask, out_mask, traversal;
    uint32_t aoSamples;
    float ambientIntensity;
    float lightDir[3];
    float lightIntensity;
    float lightColor[3];
    int rank, world;
    int tilesX, tilesY;
    int localTiles;       // tiles owned by this rank
};

struct Counters { unsigned long long rays, fetches, pixels; };

}  // namespace vrt
