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
//   fdist[v] : per-voxel "safe step count", stored 4x4x4-bricked (64 B per brick). 0 = solid voxel. Otherwise a lower
//              bound >= 1 of the Chebyshev distance C(v) from v to the nearest solid or out-of-bounds voxel: one DDA
//              step moves at most 1 in Chebyshev distance, so the next fdist[v]-1 positions of ANY ray through v are
//              empty and in bounds and fdist[v] steps can run before memory is looked at again. Built as
//              max(min(C(v), VRT_FINE_CAP) [exact near surfaces], (bdist-1)*4+1 [far field]).
constexpr int VRT_FINE_CAP = 15;
//   ldist[v] : the same safe-step values stored LINEARLY with a one-voxel apron: ldist[(x+1) + (y+1)*LW + (z+1)*LW*LH],
//              LW = W+2, LH = H+2; apron voxels (the layer just outside the volume) hold VRT_LIN_OOB. A ray that does its
//              look-ups every fdist steps can be at most one voxel outside the volume when it looks again, so the
//              traversal needs no bounds test and no coordinates: one byte load at entry_offset + sum(count_axis * stride_axis).
//              Read as a SIGNED byte the field is: 0 solid, -1 outside, 1..VRT_SAFE_MAX safe steps (one compare ends a ray).
constexpr uint32_t VRT_LIN_OOB = 255;
constexpr uint32_t VRT_SAFE_MAX = 127;
//   odist[o][v] : DIRECTIONAL safe-step fields, one per direction octant o = (dir.x<0) | (dir.y<0)<<1 | (dir.z<0)<<2, same aproned
//              linear layout and value coding as ldist. A ray of octant o only ever moves by 0 or sign_o(axis) per axis and step,
//              so after k steps it is inside the cube [0,k]^3 (in octant direction) anchored at its voxel: odist[o][v] = n means the
//              cube of side n anchored at v (offsets 0..n-1 per axis towards o) holds no solid and no out-of-bounds voxel, hence n
//              steps can run before the field is looked at again. What lies behind or beside the ray no longer shortens its steps
//              (a ray LEAVING a surface gets its full budget at once). n is rounded down to a level of VRT_OCT_LEVELS.
//              Byte coding of every safe-step field: 0 solid, 255 outside, 1..127 that many steps, 128..254 -> 4*code - 384
//              steps (128, 132, .. 632): safe_decode(). Levels above 127 are multiples of 4.
constexpr int VRT_OCT_NLEVELS = 31;
#define VRT_OCT_LEVEL_LIST {1, 2, 3, 4, 5, 6, 7, 8, 10, 12, 14, 16, 20, 24, 28, 32, 40, 48, 56, 64, 80, 96, 112, 128, 160, 192, 256, 320, 384, 448, 512}
VRT_HD uint32_t safe_decode(uint32_t code) { return code < 128u ? code : code * 4u - 384u; }
VRT_HD uint32_t safe_encode(uint32_t n) { return n < 128u ? n : (n + 384u) / 4u; }

//   sdist[o]    : the SAME directional safe-step values in a two-level (sparse) layout, VRT_TRAVERSAL_SPARSE. The volume is cut into
//              8x8x8 bricks (coordinates shifted by one brick so that the one-voxel apron needs no special case); `top` holds one word
//              per brick: either the one byte code that stands for the whole brick (a solid brick: 0; a brick outside the volume:
//              255; a brick whose every voxel has at least VRT_SP_UNIFORM_MIN safe steps: the code of their MINIMUM, which is a valid
//              lower bound for each of them) or VRT_SP_FINE | index of a 512-byte block of per-voxel codes in `fine`. Only bricks near
//              a surface are stored per voxel: 8 x 8.6 GiB of linear fields for a 2048^3 scene become ~1 GiB, the top levels
//              (8 x 8.6 MB for 1024^3) stay resident in the 126 MB L2.
constexpr uint32_t VRT_SP_FINE = 0x80000000u;
#ifndef VRT_SP_UNIFORM_MIN_V
#define VRT_SP_UNIFORM_MIN_V 16
#endif
constexpr uint32_t VRT_SP_UNIFORM_MIN = VRT_SP_UNIFORM_MIN_V;
struct SparseField { const uint32_t* top; const uint8_t* fine; };

struct SceneDev {
    const uint8_t* vol;               // dense u8[W*H*D], x fastest (voxel_scene.cpp:99); null when the volume is kept two-level (svol)
    SparseField svol;                 // the palette volume in the two-level layout of sdist (bricks of one value + 512-byte bricks), scenes over the dense budget
    const unsigned long long* bmask;  // [Bz][By][Bx]
    const uint8_t* bdist;             // [Bz][By][Bx]
    const uint8_t* fdist;             // [Bz][By][Bx][4][4][4]
    const uint8_t* ldist;             // [D+2][H+2][W+2]
    const uint8_t* odist[8];          // per direction octant, [D+2][H+2][W+2] each
    SparseField sdist[8];             // per direction octant, two-level
    uint32_t TBx, TBy, TBz;           // bricks of the top level per axis: ((dim + 8) >> 3) + 1
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

// Fused exchange (vrt_render_exchange, multi-GPU): row-major W x H planes of the assembled frame on rank 0 (peer memory over NVLink for
// the other ranks). A kernel that stores a pixel of one of these planes stores it there as well: the pass's own kernels are the
// "collective", no push kernel runs afterwards. Null = the plane is not exchanged in-kernel.
// Strips (vrt_render_denoise_exchange, multi-GPU with the denoiser): the three planes the denoiser reads do not go to rank 0 but to the
// rank that DENOISES their rows. Rank q owns rows [q * stripRows, (q + 1) * stripRows) and holds, per slot, row-major buffers of the rows
// [lo_q, hi_q) = its own rows plus `halo` rows either side (clipped to the image): a pixel is stored into the buffer of every rank whose
// range holds its row (at most three). The table lives in device memory (one per slot), pointers as THIS rank sees them (own memory or
// CUDA-IPC peer mappings).
constexpr int VRT_MAX_WORLD = 16;
struct StripTable {
    uint8_t* color[VRT_MAX_WORLD]; int8_t* normal[VRT_MAX_WORLD]; float* position[VRT_MAX_WORLD];
    int world, stripRows, halo, W, H;
};
VRT_HD int strip_lo(int q, int stripRows, int halo) { const int v = q * stripRows - halo; return v < 0 ? 0 : v; }
VRT_HD int strip_hi(int q, int stripRows, int halo, int H) { const int v = (q + 1) * stripRows + halo; return v > H ? H : v; }
// f(q, row) for every rank q whose buffer holds image row py (row = its index in that buffer). With halo <= stripRows these can only be
// the owner of the row and its two neighbours (vrt_comm_alloc_strips refuses thinner strips); tests/test_host.py checks the window
// against all ranks.
template <class F>
VRT_HD void strip_for_each_holder(int py, int world, int stripRows, int halo, int H, F f) {
    const int o = py / stripRows;
    for (int d = -1; d <= 1; d++) {
        const int q = o + d;
        if (q < 0 || q >= world) continue;
        const int lo = strip_lo(q, stripRows, halo);
        if (py >= lo && py < strip_hi(q, stripRows, halo, H)) f(q, py - lo);
    }
}
struct XPlanes { uint8_t* color; float* depth; float* motion; uint8_t* mask; float* position; int8_t* normal; uint8_t* colorHost; /* the same pixel of the shared pinned HOST frame (zero-copy over this GPU's PCIe link) */
                 const StripTable* strips; /* or null */ };

struct FrameParams {
    vrt_camera cam;
    XPlanes x;
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

struct Counters { unsigned long long rays, fetches, pixels; unsigned long long stageRays[3], stageFetches[3]; unsigned long long dbg[24]; };   // dbg: -DVRT_ENGINE_STATS builds only

// ---- wavefront pipeline (vrt_wave.cuh / vrt_wave.cu) ----
// The shader's per-pixel call tree (primary -> <=5 reflection bounces -> per shaded hit: aoSamples AO rays + 1
// shadow ray) is cut into ray waves that run as separate kernels over COMPACT lists, so that every warp traces 32
// rays of one kind. Each ray is still traced exactly as the reference does; only the order between rays changes
// (rays are independent), so every output byte equals the per-pixel schedule's.
//   hit record (48 B = 3 x 16 B):  q0 = {pos.xyz, px | py << 16}
//                                  q1 = {normal.xyz, material | depth << 8}
//                                  q2 = {dir.xyz, link}    link: primary record -> deepest record of its reflection
//                                                          chain (lastIdx) or WAVE_NONE; bounce record -> parent record
//   records [0, pixCap) are primary hits (allocated compactly by the primary kernel), [pixCap, 6*pixCap) bounce hits.
struct alignas(16) RecQ { float x, y, z; uint32_t w; };
struct alignas(8) BounceEnt { uint32_t rec, root; };
constexpr uint32_t WAVE_NONE = 0xffffffffu;
struct WaveCtr {
    uint32_t n0;          // primary-hit records
    uint32_t nb;          // bounce records
    uint32_t nsq;         // bounce records queued for shading (AO + shadow rays)
    uint32_t chunk;       // any-hit work distribution
    uint32_t qn[6];       // bounce queue lengths per wave
    uint32_t pad[2];
    uint32_t bcur[8];     // bounce work distribution per wave (32-ray groups are handed out dynamically: ray lengths differ by 500x)
};
struct WaveBuf {
    RecQ* rec;            // 3 per record
    BounceEnt* bq[2];     // ping-pong bounce queues, pixCap entries each
    uint32_t* sq;         // bounce records to shade, 5*pixCap entries
    uint8_t* res;         // any-hit results: res[i * recCap + rec], i < aoSamples: AO ray i hit, i == aoSamples: shadowed
    WaveCtr* ctr;
    uint32_t pixCap, recCap;
};

}  // namespace vrt
