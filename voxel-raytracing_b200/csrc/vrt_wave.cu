// vrt_wave.cu — wavefront form of the voxel_volume pass for sm_100a (default pipeline). Kernels per frame:
//   1  wave_primary_kernel   one thread per pixel (warp = 8x4 patch): primary ray, G-buffer, compact hit records
//   2  wave_bounce_kernel x5 reflection bounce w of every chain still alive (compact queue, ping-pong)
//   3  wave_anyhit_kernel    all AO and shadow rays of all shaded hits: persistent warps pull 32-ray chunks of ONE
//                            ray kind (AO sample i of 32 neighbouring hits / 32 shadow rays) from a global counter
//   4  wave_resolve_kernel   shading + reflection sum per primary hit, colour store
// Versus the per-pixel megakernel (vrt_render.cu): ~40 instead of 117 registers per thread (the bounce stack lives in
// records, not registers) => 3-4x the resident warps to hide the DDA's dependent-issue latency, and warps stay full
// because sky pixels / non-metallic pixels / finished pixels do not occupy lanes in the secondary waves.
#include <cuda_runtime.h>
#include <stdlib.h>
#include <mutex>
#include "vrt_wave.cuh"
#include "vrt_engine.cuh"
#include "vrt_launch.h"

// blocks per SM the register allocation of the run-to-completion kernels aims at (48 / 64 registers, no spills; 12 / 10 spill and lose)
#ifndef WAVE_PRIMARY_MIN_BLOCKS
#define WAVE_PRIMARY_MIN_BLOCKS 10
#endif
#ifndef WAVE_BOUNCE_MIN_BLOCKS
#define WAVE_BOUNCE_MIN_BLOCKS 8
#endif

namespace vrt {

// Warp-aggregated slot allocation; must be called by all 32 lanes.
__device__ __forceinline__ uint32_t warp_alloc(uint32_t* ctr, bool want) {
    const unsigned m = __ballot_sync(0xffffffffu, want);
    if (m == 0) return 0;
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(ctr, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + __popc(m & ((1u << lane) - 1u));
}

__device__ __forceinline__ void add_totals(Counters* ctr, int stage, unsigned rays, unsigned fetches, unsigned pixels) {
    const unsigned r = __reduce_add_sync(0xffffffffu, rays);
    const unsigned f = __reduce_add_sync(0xffffffffu, fetches);
    const unsigned n = __reduce_add_sync(0xffffffffu, pixels);
    if ((threadIdx.x & 31) == 0) {
        if (r) { atomicAdd(&ctr->rays, (unsigned long long)r); atomicAdd(&ctr->stageRays[stage], (unsigned long long)r); }
        if (f) { atomicAdd(&ctr->fetches, (unsigned long long)f); atomicAdd(&ctr->stageFetches[stage], (unsigned long long)f); }
        if (n) atomicAdd(&ctr->pixels, (unsigned long long)n);
    }
}

// pixel (px, py) of a denoiser input plane into the strip buffer of every rank whose row range holds py (StripTable, vrt_types.h)
template <class T, class B>
__device__ __forceinline__ void strip_store(const StripTable* __restrict__ t, B* const* bases, int px, int py, T v) {
    const int W = t->W;
    strip_for_each_holder(py, t->world, t->stripRows, t->halo, t->H, [&](int q, int row) { reinterpret_cast<T*>(bases[q])[(size_t)row * W + px] = v; });
}

__device__ __forceinline__ void store_color(const FrameParams& fp, const PlanesDev& pl, size_t p, V3 c, int px, int py) {
    const uchar4 q = make_uchar4(to_unorm8(c.x), to_unorm8(c.y), to_unorm8(c.z), 0);
    reinterpret_cast<uchar4*>(pl.color)[p] = q;
    if (fp.x.strips) strip_store<uchar4>(fp.x.strips, fp.x.strips->color, px, py, q);
    if (fp.x.color) reinterpret_cast<uchar4*>(fp.x.color)[(size_t)py * fp.cam.screenSize[0] + px] = q;   // fused exchange: rank 0's frame
    if (fp.x.colorHost) reinterpret_cast<uchar4*>(fp.x.colorHost)[(size_t)py * fp.cam.screenSize[0] + px] = q;
    if (fp.out_mask & VRT_OUT_INSTRUMENT) { pl.color_f32[3 * p + 0] = c.x; pl.color_f32[3 * p + 1] = c.y; pl.color_f32[3 * p + 2] = c.z; }
}

template <int TRAVERSAL>
__global__ void __launch_bounds__(128, WAVE_PRIMARY_MIN_BLOCKS) wave_primary_kernel(const SceneDev sc, const FrameParams fp, const PlanesDev pl, const WaveBuf wb,
                                                           Counters* __restrict__ ctr) {
    const int W = fp.cam.screenSize[0], H = fp.cam.screenSize[1];
    const int j = blockIdx.x >> 3, sub = blockIdx.x & 7;           // local tile, 16x8 sub-block of the tile
    const int t = j * fp.world + fp.rank;                          // global tile (interleaved shard)
    const int tx = t % fp.tilesX, ty = t / fp.tilesX;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int lx = (sub & 1) * 16 + (warp & 1) * 8 + (lane & 7);
    const int ly = (sub >> 1) * 8 + (warp >> 1) * 4 + (lane >> 3);
    const int px = tx * VRT_TILE + lx, py = ty * VRT_TILE + ly;
    const bool active = px < W && py < H;

    int action = WAVE_PIXEL_FINAL;
    PixelOut o;
    V3 rayDir = mk3(0.0f, 0.0f, 0.0f);
    size_t p = 0;
    if (active) {
        action = wave_primary<TRAVERSAL>(sc, fp, px, py, o, rayDir);
        p = fp.world == 1 ? (size_t)py * W + px : (size_t)j * (VRT_TILE * VRT_TILE) + ly * VRT_TILE + lx;
        if (action == WAVE_PIXEL_FINAL) store_color(fp, pl, p, o.color, px, py);
        if (fp.out_mask & VRT_OUT_GBUFFER) {
            const float4 pos4 = make_float4(o.pos.x, o.pos.y, o.pos.z, 0.0f);
            const char4 n4 = make_char4(to_snorm8(o.normal.x), to_snorm8(o.normal.y), to_snorm8(o.normal.z), 0);
            pl.depth[p] = o.depth;
            reinterpret_cast<float2*>(pl.motion)[p] = make_float2(0.0f, 0.0f);
            pl.mask[p] = to_unorm8(o.mask);
            reinterpret_cast<float4*>(pl.position)[p] = pos4;
            reinterpret_cast<char4*>(pl.normal)[p] = n4;
            const size_t xp = (size_t)py * W + px;                  // fused exchange: the same pixel of rank 0's row-major planes
            if (fp.x.depth) fp.x.depth[xp] = o.depth;
            if (fp.x.motion) reinterpret_cast<float2*>(fp.x.motion)[xp] = make_float2(0.0f, 0.0f);
            if (fp.x.mask) fp.x.mask[xp] = to_unorm8(o.mask);
            if (fp.x.position) reinterpret_cast<float4*>(fp.x.position)[xp] = pos4;
            if (fp.x.normal) reinterpret_cast<char4*>(fp.x.normal)[xp] = n4;
            if (fp.x.strips && fp.x.strips->position[0]) { strip_store<float4>(fp.x.strips, fp.x.strips->position, px, py, pos4); strip_store<char4>(fp.x.strips, fp.x.strips->normal, px, py, n4); }
        }
        if (fp.out_mask & VRT_OUT_INSTRUMENT) {
            auto sat = [](int v) { return (short)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); };
            reinterpret_cast<short4*>(pl.hit)[p] = make_short4(sat(o.hx), sat(o.hy), sat(o.hz), (short)o.steps);
            pl.hit_palette[p] = (uint8_t)o.palette;
            pl.rays[p] = 1u;
            pl.fetches[p] = o.fetches;
        }
    }
    const bool shade = action != WAVE_PIXEL_FINAL, bounce = action == WAVE_PIXEL_SHADE_AND_BOUNCE;
    const uint32_t idx = warp_alloc(&wb.ctr->n0, shade);
    if (shade) {
        RecQ* r = wb.rec + 3 * (size_t)idx;
        r[0] = mkq(o.pos, (uint32_t)px | ((uint32_t)py << 16));
        r[1] = mkq(o.normal, o.palette);                           // depth 0
        r[2] = mkq(rayDir, WAVE_NONE);
    }
    const uint32_t qi = warp_alloc(&wb.ctr->qn[0], bounce);
    if (bounce) { BounceEnt e; e.rec = idx; e.root = idx; wb.bq[0][qi] = e; }
    add_totals(ctr, 0, active ? 1u : 0u, active ? o.fetches : 0u, active ? 1u : 0u);
}

// Reflection wave w: bounce w of every chain still alive. With `chain` set the wave also runs all LATER bounces of its rays in
// place (the lanes whose chain goes on keep going, the others idle): the last waves hold a few thousand rays each and cost a
// launch + a latency-bound tail apiece, which is more than their idle lanes cost.
template <int TRAVERSAL>
__global__ void __launch_bounds__(128, WAVE_BOUNCE_MIN_BLOCKS) wave_bounce_kernel(const SceneDev sc, const FrameParams fp, const PlanesDev pl, const WaveBuf wb,
                                                          Counters* __restrict__ ctr, const int w, const bool chain, const uint32_t roundCap) {
    const uint32_t n = wb.ctr->qn[w];
    const BounceEnt* __restrict__ qin = wb.bq[w & 1];
    BounceEnt* __restrict__ qout = wb.bq[(w + 1) & 1];
    const uint32_t lane = threadIdx.x & 31, ltMask = (1u << lane) - 1u;
    // Every queued ray yields exactly one record, so its index needs no allocation: records of wave w start after those of the
    // earlier waves. Only the in-place continuation bounces (chain) allocate, behind this wave's block.
    uint32_t recBase = wb.pixCap;
    for (int v = 0; v < w; v++) recBase += wb.ctr->qn[v];
    const uint32_t recBase2 = recBase + n;
    unsigned rays = 0, fetches = 0;
    // 32-ray groups are claimed one ahead: the counter's round trip overlaps the tracing of the current group
    // (only while plenty of groups are left: near the end a group held in reserve by a busy warp would lengthen the tail)
    const uint32_t reserve = 2u * 32u * (gridDim.x * (blockDim.x >> 5));
    bool have = false;
    uint32_t nextE0 = 0;
    for (;;) {
        if (!have && lane == 0) nextE0 = atomicAdd(&wb.ctr->bcur[w], 32u);
        const uint32_t e0 = __shfl_sync(0xffffffffu, nextE0, 0);
        if (e0 >= n) break;
        have = e0 + reserve < n;
        if (have && lane == 0) nextE0 = atomicAdd(&wb.ctr->bcur[w], 32u);
        const uint32_t e = e0 + lane;
        bool act = e < n;
        BounceEnt ent; ent.rec = 0; ent.root = 0;
        V3 pos = mk3(0.0f, 0.0f, 0.0f), normal = pos, dir = pos;
        uint32_t pxy = 0;
        if (act) {
            ent = qin[e];
            const RecQ* pr = wb.rec + 3 * (size_t)ent.rec;
            const RecQ p0 = pr[0], p1 = pr[1], p2 = pr[2];
            pxy = p0.w; pos = q3(p0); normal = q3(p1); dir = q3(p2);
        }
        for (uint32_t wl = (uint32_t)w;; wl++) {                    // warp-uniform depth of this bounce
            TraceRes r; r.material = 0; r.fetches = 0; r.pos = r.normal = mk3(0.0f, 0.0f, 0.0f);
            V3 rd = mk3(0.0f, 0.0f, 0.0f);
            bool cont = false, terminal = false;
            uint32_t cnt = 0;           // records of this chain that get AO + shadow rays now (the chain ends here)
            if (act) {
                r = wave_bounce<TRAVERSAL>(sc, pos, normal, dir, rd, roundCap);
                rays++; fetches += r.fetches;
                if (fp.out_mask & VRT_OUT_INSTRUMENT) {
                    const size_t p = plane_index(fp, (int)(pxy & 0xffffu), (int)(pxy >> 16));
                    atomicAdd(pl.rays + p, 1u); atomicAdd(pl.fetches + p, r.fetches);
                }
                terminal = r.material == 0 || sc.palette[r.material].metallic <= 0.0f;   // :289-292 lastIdx = w
                if (terminal) cnt = wl + (r.material != 0 ? 1u : 0u);   // every hit record below the primary hit
                else if (wl + 1 < MAX_REFLECTIONS) cont = true;          // else: 5 metallic bounces, lastIdx stays -1 (:298)
            }
            // all allocations of this bounce with ONE exposed round trip: lane 0 issues the (independent) atomics back to back
            const bool first = wl == (uint32_t)w;
            const unsigned mRec = first ? 0u : __ballot_sync(0xffffffffu, act), mQ = chain ? 0u : __ballot_sync(0xffffffffu, cont);
            uint32_t incl = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if ((int)lane >= d) incl += t; }
            const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
            uint32_t aRec = 0, aSq = 0, aQ = 0;
            if (lane == 0) {
                if (mRec) aRec = atomicAdd(&wb.ctr->nb, (uint32_t)__popc(mRec));
                if (total) aSq = atomicAdd(&wb.ctr->nsq, total);
                if (mQ) aQ = atomicAdd(&wb.ctr->qn[w + 1], (uint32_t)__popc(mQ));
            }
            aRec = __shfl_sync(0xffffffffu, aRec, 0); aSq = __shfl_sync(0xffffffffu, aSq, 0); aQ = __shfl_sync(0xffffffffu, aQ, 0);
            const uint32_t child = first ? recBase + e : recBase2 + aRec + __popc(mRec & ltMask);
            if (act) {
                RecQ* cr = wb.rec + 3 * (size_t)child;
                cr[0] = mkq(r.pos, pxy);
                cr[1] = mkq(r.normal, r.material | (wl << 8));
                cr[2] = mkq(rd, ent.rec);
                if (terminal) {
                    wb.rec[3 * (size_t)ent.root + 2].w = child;
                    uint32_t s = aSq + incl - cnt;
                    if (cnt) {
                        if (r.material != 0) wb.sq[s++] = child;
                        for (uint32_t k = ent.rec; k != ent.root; k = wb.rec[3 * (size_t)k + 2].w) wb.sq[s++] = k;
                    }
                }
            }
            if (!chain) {
                if (cont) { BounceEnt o; o.rec = child; o.root = ent.root; qout[aQ + __popc(mQ & ltMask)] = o; }
                break;
            }
            if (!__any_sync(0xffffffffu, cont)) break;
            act = cont; ent.rec = child; pos = r.pos; normal = r.normal; dir = rd;
        }
    }
    add_totals(ctr, 1, rays, fetches, 0u);
}

constexpr uint32_t ANYHIT_CHUNK = 8;   // 32-ray groups per work-counter atomic

template <int TRAVERSAL>
__global__ void __launch_bounds__(256) wave_anyhit_kernel(const SceneDev sc, const FrameParams fp, const PlanesDev pl, const WaveBuf wb,
                                                          Counters* __restrict__ ctr) {
    const uint32_t n0 = wb.ctr->n0, S = n0 + wb.ctr->nsq;
    const uint32_t aoN = (fp.pass_mask & VRT_PASS_AO) ? fp.aoSamples : 0u;
    const uint32_t R = aoN + ((fp.pass_mask & VRT_PASS_SHADOW) ? 1u : 0u);
    const unsigned long long total = (unsigned long long)S * R;
    const uint32_t lane = threadIdx.x & 31;
    unsigned rays = 0, fetches = 0;
    for (;;) {
        uint32_t c = 0;
        if (lane == 0) c = atomicAdd(&wb.ctr->chunk, 1u);
        c = __shfl_sync(0xffffffffu, c, 0);
        const unsigned long long t0 = (unsigned long long)c * (32u * ANYHIT_CHUNK);
        if (t0 >= total) break;
        const uint32_t i0 = (uint32_t)(t0 / S);                     // warp-uniform; lanes only add and wrap
        const uint32_t q0 = (uint32_t)(t0 - (unsigned long long)i0 * S);
        for (uint32_t g = 0; g < ANYHIT_CHUNK; g++) {
            if (t0 + g * 32u + lane >= total) break;
            uint32_t i = i0;
            unsigned long long qq = (unsigned long long)q0 + g * 32u + lane;
            while (qq >= S) { qq -= S; i++; }
            const uint32_t q = (uint32_t)qq;
            const uint32_t idx = q < n0 ? q : wb.sq[q - n0];
            const RecQ r0 = wb.rec[3 * (size_t)idx], r1 = wb.rec[3 * (size_t)idx + 1];
            const int px = (int)(r0.w & 0xffffu), py = (int)(r0.w >> 16);
            uint32_t f = 0;
            const bool hit = wave_anyhit<TRAVERSAL>(sc, fp, q3(r0), q3(r1), px, py, r1.w >> 8, i, aoN, f);
            wb.res[(size_t)i * wb.recCap + idx] = hit ? 1 : 0;
            rays++; fetches += f;
            if (fp.out_mask & VRT_OUT_INSTRUMENT) {
                const size_t p = plane_index(fp, px, py);
                atomicAdd(pl.rays + p, 1u); atomicAdd(pl.fetches + p, f);
            }
        }
    }
    add_totals(ctr, 2, rays, fetches, 0u);
}

// ---- any-hit wave on the ray engine (vrt_engine.cuh): LINEAR (isotropic field) or OCTANT (directional fields) ----
template <int TRAVERSAL, int NEAR_TOP, int FAR_TOP>
__global__ void __launch_bounds__(ENGINE_WARPS * 32, ENGINE_MIN_BLOCKS) wave_anyhit_engine_kernel(const SceneDev sc, const FrameParams fp, const PlanesDev pl, const WaveBuf wb,
                                                                                Counters* __restrict__ ctr, const uint32_t refillT, const uint32_t farShare, const int part, const uint32_t quota, const int topup) {
    __shared__ EngineStash stashAll[ENGINE_WARPS];
    __shared__ uint32_t blockChunk;
    __shared__ SparseField octFields[8];             // VRT_TRAVERSAL_SPARSE: the per-octant fields, indexed by a ray's octant
    constexpr bool SPARSE = TRAVERSAL == VRT_TRAVERSAL_SPARSE;
    if (SPARSE) { if (threadIdx.x < 8) octFields[threadIdx.x] = sc.sdist[threadIdx.x]; __syncthreads(); }
    EngineStash& st = stashAll[threadIdx.x >> 5];
    const uint32_t lane = threadIdx.x & 31, ltMask = (1u << lane) - 1u;
    // part 0: the rays of every shaded record, persistent grid, chunks from a global counter
    // part 1: the rays of the primary-hit records only, short-lived blocks: block b owns items [b, b+1) * ENGINE_BLOCK_QUOTA
    // part 2: the rays of the queued bounce records only, persistent grid
    const uint32_t n0 = wb.ctr->n0, qBase = part == 2 ? n0 : 0u, S = part == 1 ? n0 : (part == 2 ? wb.ctr->nsq : n0 + wb.ctr->nsq);
    const uint32_t aoN = (fp.pass_mask & VRT_PASS_AO) ? fp.aoSamples : 0u;
    const uint32_t R = aoN + ((fp.pass_mask & VRT_PASS_SHADOW) ? 1u : 0u);
    const unsigned long long total = (unsigned long long)S * R;
    if (part == 1) {
        if ((unsigned long long)blockIdx.x * quota >= total) return;     // the grid is sized for the worst case
        if (threadIdx.x == 0) blockChunk = 0;
        __syncthreads();
    }
    const int LW = (int)sc.W + 2, LWH = LW * ((int)sc.H + 2);
    const bool instrument = (fp.out_mask & VRT_OUT_INSTRUMENT) != 0;

    // warp-uniform work range [cur, end) and the (ray kind, record slot) of item `cur`
    unsigned long long cur = 0, end = 0;
    uint32_t ci = 0, cq = 0, stashN = 0;
    bool exhausted = total == 0;
    // lane state
    EngRay er; er.rem = 0; er.left = 0;
    er.l.sx = er.l.sy = er.l.sz = er.l.dx = er.l.dy = er.l.dz = er.l.sx0 = er.l.sy0 = er.l.sz0 = er.l.ax = er.l.ay = er.l.az = 0.0f;
    er.l.p0 = sc.vol; er.l.SX = er.l.SY = er.l.SZ = er.l.T0 = 0; er.l.maxSteps = 0;
    er.l.mx0 = er.l.my0 = er.l.mz0 = 0; er.l.stx = er.l.sty = er.l.stz = 1; er.l.i = 0; er.l.mask = 0;
    er.f.top = nullptr; er.f.fine = nullptr;
    bool alive = false, pending = false, hitRes = false;
    uint32_t idx = 0, kind = 0, fetchRes = 0;
    unsigned rays = 0, fetches = 0;
#ifdef VRT_ENGINE_STATS
    unsigned long long dbg[24] = {};
#define ENG_STAT(k, v) dbg[k] += (v)
#else
#define ENG_STAT(k, v)
#endif

    auto retire = [&](uint32_t rIdx, uint32_t rKind, bool hit, uint32_t f) {
        wb.res[(size_t)rKind * wb.recCap + rIdx] = hit ? 1 : 0;
        rays++; fetches += f;
        if (instrument) {
            const uint32_t pxy = wb.rec[3 * (size_t)rIdx].w;
            const size_t p = plane_index(fp, (int)(pxy & 0xffffu), (int)(pxy >> 16));
            atomicAdd(pl.rays + p, 1u); atomicAdd(pl.fetches + p, f);
        }
    };

    for (;;) {
        // ---- refill: idle lanes retire their finished ray and pop a prepared one
        const unsigned idleMask = __ballot_sync(0xffffffffu, !alive);
        const uint32_t nIdle = __popc(idleMask);
        const bool canRefill = stashN > 0 || !exhausted;
        if (nIdle == 32u && !canRefill) break;
        if (canRefill && (nIdle >= refillT || nIdle == 32u)) {
            ENG_STAT(7, lane == 0);
            if (!alive && pending) { retire(idx, kind, hitRes, fetchRes); pending = false; }
            // produce prepared rays, 32 at a time with all lanes, until the idle lanes can be served
            while (stashN < nIdle && !exhausted) {
                if (cur >= end) {
                    uint32_t c = 0;
                    if (lane == 0) c = part == 1 ? atomicAdd(&blockChunk, 1u) : atomicAdd(part == 2 ? &wb.ctr->pad[0] : &wb.ctr->chunk, 1u);
                    c = __shfl_sync(0xffffffffu, c, 0);
                    if (part == 1 && c >= quota / ENGINE_CHUNK) { exhausted = true; break; }
                    const unsigned long long t0 = (unsigned long long)c * ENGINE_CHUNK + (part == 1 ? (unsigned long long)blockIdx.x * quota : 0ull);
                    if (t0 >= total) { exhausted = true; break; }
                    cur = t0; end = t0 + ENGINE_CHUNK < total ? t0 + ENGINE_CHUNK : total;
                    ci = (uint32_t)(t0 / S); cq = (uint32_t)(t0 - (unsigned long long)ci * S);
                }
                const bool valid = cur + lane < end;
                bool push = false;
                LinRay nr = er.l;
                uint32_t nIdx = 0, i = ci, firstRem = 0;
                if (valid) {
                    unsigned long long qq = (unsigned long long)cq + lane;
                    while (qq >= S) { qq -= S; i++; }
                    const uint32_t q = qBase + (uint32_t)qq;
                    nIdx = q < n0 ? q : wb.sq[q - n0];
                    const RecQ r0 = wb.rec[3 * (size_t)nIdx], r1 = wb.rec[3 * (size_t)nIdx + 1];
                    V3 rs, rd, pos; uint32_t maxSteps; Dda s;
                    wave_anyhit_ray(sc, fp, q3(r0), q3(r1), (int)(r0.w & 0xffffu), (int)(r0.w >> 16), r1.w >> 8, i, aoN, rs, rd, maxSteps);
                    if (!lin_usable(dda_begin(sc, rs, rd, pos, s), s)) {
                        uint32_t f; const bool hit = engine_anyhit_literal(sc, s, maxSteps, f);
                        retire(nIdx, i, hit, f);
                    } else if ((SPARSE ? engine_begin_sparse(sc, s, rd, maxSteps, nr) : lin_begin(sc, TRAVERSAL == VRT_TRAVERSAL_OCTANT ? octant_field(sc, s) : sc.ldist, s, rd, maxSteps, nr)) == LIN_OOB)
                        retire(nIdx, i, false, 0u);
                    else {
                        // the first look-up, with the whole warp: a ray it ends (exhausted at once / starts in a solid voxel) is retired here
                        EngRay t; t.l = nr; t.rem = 0; t.left = maxSteps;
                        if (SPARSE) t.f = octFields[(s.stx < 0 ? 1 : 0) | (s.sty < 0 ? 2 : 0) | (s.stz < 0 ? 4 : 0)];
                        const int stt = engine_lookup<SPARSE>(t, sc);
                        ENG_STAT(17, 1); ENG_STAT(18, stt != ENG_GO);
                        if (stt != ENG_GO) retire(nIdx, i, stt == ENG_HIT, engine_steps_taken(t) + (stt == ENG_HIT ? 1u : 0u));
                        else { push = true; firstRem = (uint32_t)t.rem; }
                    }
                }
                const unsigned pm = __ballot_sync(0xffffffffu, push);
                if (push) engine_push(st, stashN + __popc(pm & ltMask), nr, nIdx, i, firstRem);
                stashN += __popc(pm);
                cur += 32u; cq += 32u;
                while (cq >= S) { cq -= S; ci++; }
            }
            __syncwarp();
            const uint32_t take = nIdle < stashN ? nIdle : stashN;
            const uint32_t r = __popc(idleMask & ltMask);
            if (!alive && r < take) { engine_pop<SPARSE>(st, stashN - 1u - r, er, LW, LWH, idx, kind, octFields); alive = true; }
            stashN -= take;
            __syncwarp();
        }
        // ---- look-up for lanes whose budget is spent or would run out inside the coming round (top-up, see engine_lookup)
        if (alive && er.rem < topup) {
            const int stt = engine_lookup<SPARSE>(er, sc);
            ENG_STAT(6, 1);
            ENG_STAT(8, stt == ENG_EXHAUSTED);
#ifdef VRT_ENGINE_STATS
            if (stt == ENG_GO) { const int n = er.rem; dbg[9 + (n <= 1 ? 0 : n <= 2 ? 1 : n <= 4 ? 2 : n <= 8 ? 3 : n <= 16 ? 4 : n <= 32 ? 5 : n <= 64 ? 6 : 7)]++; }
#endif
            if (stt != ENG_GO) {
                alive = false; pending = true; hitRes = stt == ENG_HIT; fetchRes = engine_steps_taken(er) + (hitRes ? 1u : 0u);
            }
        }
        // ---- lock-step DDA steps: a long round when at least farShare/8 of the live lanes can use all of it. (Choosing among 3, 7,
        // 15 and 31 steps by useful steps per issued instruction was measured and lost: 3.65 vs 3.44 ms.)
        const uint32_t nFar = __popc(__ballot_sync(0xffffffffu, alive && er.rem >= 2 * FAR_TOP - 1));
        if (nFar * 8u >= (32u - nIdle) * farShare && nFar) {
            ENG_STAT(1, lane == 0); ENG_STAT(3, alive); ENG_STAT(5, alive ? min(er.rem, 2 * FAR_TOP - 1) : 0);
            engine_round<FAR_TOP>(er.l, er.rem);
        } else {
            ENG_STAT(0, lane == 0); ENG_STAT(2, alive); ENG_STAT(4, alive ? min(er.rem, 2 * NEAR_TOP - 1) : 0);
            engine_round<NEAR_TOP>(er.l, er.rem);
        }
    }
    if (pending) retire(idx, kind, hitRes, fetchRes);
    add_totals(ctr, 2, rays, fetches, 0u);
#ifdef VRT_ENGINE_STATS
    for (int k = 0; k < 24; k++) if (dbg[k]) atomicAdd(&ctr->dbg[k], dbg[k]);
#endif
}

__global__ void __launch_bounds__(256) wave_resolve_kernel(const SceneDev sc, const FrameParams fp, const PlanesDev pl, const WaveBuf wb) {
    __shared__ float skyTab[64 * 3];          // skyColor of the 64 axis-code vectors (sky_of_normal)
    if (threadIdx.x < 64) {
        const V3 c = sky_color(sc, sky_axis_vector(threadIdx.x));
        skyTab[3 * threadIdx.x] = c.x; skyTab[3 * threadIdx.x + 1] = c.y; skyTab[3 * threadIdx.x + 2] = c.z;
    }
    __syncthreads();
    const uint32_t n0 = wb.ctr->n0;
    const uint32_t aoN = (fp.pass_mask & VRT_PASS_AO) ? fp.aoSamples : 0u;
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n0; idx += gridDim.x * blockDim.x) {
        const uint32_t pxy = wb.rec[3 * (size_t)idx].w;
        const V3 c = wave_resolve(sc, fp, wb, idx, aoN, skyTab);
        store_color(fp, pl, plane_index(fp, (int)(pxy & 0xffffu), (int)(pxy >> 16)), c, (int)(pxy & 0xffffu), (int)(pxy >> 16));
    }
}

// refill threshold of the ray engine (idle lanes that trigger a refill); VRT_ENGINE_T=0 selects the run-to-completion kernels
static int engineT() {
    static const int t = [] { const char* e = getenv("VRT_ENGINE_T"); const int v = e ? atoi(e) : 8; return v > 32 ? 32 : v; }();
    return t;
}

template <int T>
static cudaError_t launch_wave_t(const SceneDev& sc, const FrameParams& fp, const PlanesDev& pl, Counters* ctr, const WaveBuf& wb, int sms,
                                 cudaStream_t stream, const WaveSide* side, uint32_t* launches, cudaEvent_t* ev) {
    cudaError_t e = cudaMemsetAsync(wb.ctr, 0, sizeof(WaveCtr), stream);
    if (e != cudaSuccess) return e;
    static int occAny = 0, occBounce = 0, occEngine = 0;
    static std::once_flag occOnce;                        // contexts on different host threads may render at the same time
    std::call_once(occOnce, [] {
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occAny, wave_anyhit_kernel<T>, 256, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occBounce, wave_bounce_kernel<T>, 128, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occEngine, wave_anyhit_engine_kernel<T, ENGINE_NEAR_TOP, ENGINE_FAR_TOP>, ENGINE_WARPS * 32, 0);
        if (occAny < 1) occAny = 1;
        if (occBounce < 1) occBounce = 1;
        if (occEngine < 1) occEngine = 1;
    });
    static const uint32_t quota = getenv("VRT_ENGINE_QUOTA") ? (uint32_t)atoi(getenv("VRT_ENGINE_QUOTA")) / ENGINE_CHUNK * ENGINE_CHUNK : ENGINE_BLOCK_QUOTA;
    // the wave from which the rest of every reflection chain runs in place (MAX_REFLECTIONS - 1: one launch per bounce)
    static const int bounceChain = getenv("VRT_BOUNCE_CHAIN") ? min(max(atoi(getenv("VRT_BOUNCE_CHAIN")), 0), MAX_REFLECTIONS - 1) : 2;
    static const uint32_t bounceCapLate = getenv("VRT_BOUNCE_CAP_LATE") ? (uint32_t)atoi(getenv("VRT_BOUNCE_CAP_LATE")) : ROUND_CAP_SECONDARY;
    static const int farShare = getenv("VRT_ENGINE_FARSHARE") ? atoi(getenv("VRT_ENGINE_FARSHARE")) : 4;
    // measured (profiles/r1_overlap.log): beside-the-waves scheduling pays when the frame (or this rank's shard of it) is small and
    // the launches are tail-bound (1/8 of a 4K frame: -12 %), and costs ~1 % on a full 4K frame. VRT_OVERLAP=0/1 overrides.
    static const int overlapEnv = getenv("VRT_OVERLAP") ? atoi(getenv("VRT_OVERLAP")) : -1;
    const bool smallFrame = (size_t)fp.localTiles * (VRT_TILE * VRT_TILE) <= (size_t)1 << 21;
    wave_primary_kernel<T><<<(unsigned)fp.localTiles * 8u, 128, 0, stream>>>(sc, fp, pl, wb, ctr);
    uint32_t n = 1;
    if (ev) cudaEventRecord(ev[0], stream);
    const bool shade = (fp.pass_mask & VRT_PASS_SHADE) != 0;
    const bool reflect = shade && (fp.pass_mask & VRT_PASS_REFLECT);
    const uint32_t aoN = (fp.pass_mask & VRT_PASS_AO) ? fp.aoSamples : 0u;
    const uint32_t R = aoN + ((fp.pass_mask & VRT_PASS_SHADOW) ? 1u : 0u);
    const bool anyhit = shade && R;
    // the ray engine needs a linear safe-step field; the other back-ends keep the run-to-completion kernel (VRT_ENGINE_T=0 forces it)
    const bool engine = (T == VRT_TRAVERSAL_LINEAR || T == VRT_TRAVERSAL_OCTANT || T == VRT_TRAVERSAL_SPARSE) && engineT() > 0;
    const bool overlap = anyhit && reflect && engine && side && (overlapEnv < 0 ? smallFrame : overlapEnv != 0);
    auto launch_engine = [&](cudaStream_t s, int part, unsigned grid) {
        // part 1 shares the SMs with the reflection waves: padding its shared memory caps its blocks per SM so that the waves'
        // blocks find registers at once instead of waiting for a side block to retire
        static const int sidePad = getenv("VRT_SIDE_PAD_KB") ? atoi(getenv("VRT_SIDE_PAD_KB")) * 1024 : 0;
        static const int mainPad = getenv("VRT_ENGINE_PAD_KB") ? atoi(getenv("VRT_ENGINE_PAD_KB")) * 1024 : 0;   // measurement: caps the blocks per SM
        const size_t dyn = part == 1 ? (size_t)sidePad : (size_t)mainPad;
        if (dyn) cudaFuncSetAttribute(wave_anyhit_engine_kernel<T, ENGINE_NEAR_TOP, ENGINE_FAR_TOP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
        // look-ups join in when a lane's budget would run out inside a near round (engine_lookup); 1 = only when it is spent. Measured on
        // cfg4: 1 -> 3.48 ms, 3 -> 3.45, 7 -> 3.40, 12 -> 3.42, 31 -> 3.46 (profiles/r2_engine_experiments.txt)
        static const int topup = getenv("VRT_ENGINE_TOPUP") ? max(1, atoi(getenv("VRT_ENGINE_TOPUP"))) : 2 * ENGINE_NEAR_TOP - 1;
        wave_anyhit_engine_kernel<T, ENGINE_NEAR_TOP, ENGINE_FAR_TOP><<<grid, ENGINE_WARPS * 32, dyn, s>>>(sc, fp, pl, wb, ctr, (uint32_t)engineT(), (uint32_t)farShare, part, quota, topup);
        n++;
    };
    if (overlap) {
        const unsigned long long worst = (unsigned long long)fp.localTiles * (VRT_TILE * VRT_TILE) * R;       // every pixel a hit
        cudaEventRecord(side->fork, stream);
        cudaStreamWaitEvent(side->stream, side->fork, 0);
        launch_engine(side->stream, 1, (unsigned)((worst + quota - 1) / quota));
        cudaEventRecord(side->join, side->stream);
    }
    if (reflect)
        for (int w = 0; w <= bounceChain; w++) {
            wave_bounce_kernel<T><<<(unsigned)(sms * occBounce), 128, 0, stream>>>(sc, fp, pl, wb, ctr, w, w == bounceChain && w + 1 < MAX_REFLECTIONS, w == 0 ? ROUND_CAP_SECONDARY : bounceCapLate);
            n++;
        }
    if (ev) cudaEventRecord(ev[1], stream);
    if (anyhit) {
        if (overlap) { launch_engine(stream, 2, (unsigned)(sms * occEngine)); cudaStreamWaitEvent(stream, side->join, 0); }
        else if (engine) {
            // The persistent grid owns every block slot of the GPU until its work counter runs dry; a few slots are left free so
            // that the small kernels of OTHER frames in flight (tile pushes of the exchange, blits, the next frame's first blocks) are
            // not held back for the length of this kernel (VRT_ENGINE_GRID_SLACK, measured in profiles/r2_engine_experiments.txt)
            static const int slack = getenv("VRT_ENGINE_GRID_SLACK") ? max(0, atoi(getenv("VRT_ENGINE_GRID_SLACK"))) : 0;
            launch_engine(stream, 0, (unsigned)max(sms * occEngine - slack, sms));
        }
        else { wave_anyhit_kernel<T><<<(unsigned)(sms * occAny), 256, 0, stream>>>(sc, fp, pl, wb, ctr); n++; }
    }
    if (ev) cudaEventRecord(ev[2], stream);
    if (shade) { wave_resolve_kernel<<<(unsigned)(sms * 8), 256, 0, stream>>>(sc, fp, pl, wb); n++; }
    if (launches) *launches += n;
    return cudaGetLastError();
}

cudaError_t launch_wavefront(const SceneDev& sc, const FrameParams& fp, const PlanesDev& pl, Counters* ctr, const WaveBuf& wb, int sms,
                             cudaStream_t stream, const WaveSide* side, uint32_t* launches, cudaEvent_t* ev) {
    if (fp.localTiles <= 0) return cudaSuccess;
    if (fp.traversal == VRT_TRAVERSAL_DENSE_U8) return launch_wave_t<VRT_TRAVERSAL_DENSE_U8>(sc, fp, pl, ctr, wb, sms, stream, side, launches, ev);
    if (fp.traversal == VRT_TRAVERSAL_FINE) return launch_wave_t<VRT_TRAVERSAL_FINE>(sc, fp, pl, ctr, wb, sms, stream, side, launches, ev);
    if (fp.traversal == VRT_TRAVERSAL_LINEAR) return launch_wave_t<VRT_TRAVERSAL_LINEAR>(sc, fp, pl, ctr, wb, sms, stream, side, launches, ev);
    if (fp.traversal == VRT_TRAVERSAL_OCTANT) return launch_wave_t<VRT_TRAVERSAL_OCTANT>(sc, fp, pl, ctr, wb, sms, stream, side, launches, ev);
    if (fp.traversal == VRT_TRAVERSAL_SPARSE) return launch_wave_t<VRT_TRAVERSAL_SPARSE>(sc, fp, pl, ctr, wb, sms, stream, side, launches, ev);
    return launch_wave_t<VRT_TRAVERSAL_BRICK>(sc, fp, pl, ctr, wb, sms, stream, side, launches, ev);
}

}  // namespace vrt
