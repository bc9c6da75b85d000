// vrt_engine.cuh — persistent-warp ray engine (device only; sm_100a): lane refill + fixed-length lock-step rounds.
//
// Why (ncu r1c/r1d/r1e, profiles/): a warp that traces 32 rays to completion with the per-ray loop of trace_ray runs at 9-15 of
// 32 lanes although it is issue-bound, for two reasons: rays END after very different step counts (AO: 1..64, shadow /
// reflection: 1..512), and within one look-up round every lane runs a different number of steps (the unrolled step loop held 55 %
// of all issued instructions at 5-9 active lanes; it runs max-over-lanes iterations). The engine removes both:
//   * ray set-up (record load, direction, box entry, DDA set-up, IEEE divisions: ~270 instructions) always runs with all 32 lanes,
//     32 rays at a time, and pushes the resulting traversal states into a per-warp shared-memory stash;
//   * a refill event happens when >= refillT lanes are idle: those lanes first retire the result of their finished ray (so the
//     retire code also runs with >= refillT lanes), then pop a prepared state (LDS.128, conflict-free);
//   * a round = [lanes whose safe-step budget is spent look the field up] + a FIXED block sequence of DDA steps for the whole
//     warp (up to 7 steps, or up to 31 when enough live lanes have that much budget), fully unrolled; a lane runs the binary
//     blocks of its own budget (engine_round): 7 instructions per step.
// Every ray still performs exactly the float operations of traceRayInt in the same order (the select form of
// voxel_volume.frag:164-170) and looks the field up through the same lin_bits() offset recovery as trace_ray<LINEAR/OCTANT>, so
// results are bit-identical (tests/test_gpu.py runs every back-end against the oracle).
#pragma once
#include "vrt_wave.cuh"

namespace vrt {

constexpr int ENGINE_WARPS = 8;          // warps per block
constexpr uint32_t ENGINE_CHUNK = 64;    // rays per claim of the global work counter (small: the tail of a sharded frame is short)
constexpr uint32_t ENGINE_BLOCK_QUOTA = ENGINE_WARPS * 256u;   // rays per block when the kernel runs as many short-lived blocks
#ifndef ENGINE_NEAR_TOP
#define ENGINE_NEAR_TOP 4              // near round: up to 7 steps (blocks 4,2,1)
#endif
#ifndef ENGINE_FAR_TOP
#define ENGINE_FAR_TOP 16              // far round: up to 31 steps (blocks 16,8,4,2,1)
#endif
#ifndef ENGINE_MIN_BLOCKS
#define ENGINE_MIN_BLOCKS 4             // blocks per SM the register allocation aims at (64 registers)
#endif
constexpr int ENGINE_STASH = 64;         // prepared rays per warp (a refill asks for <= 32 more while < 32 are left)

// Traversal state of one ray: LinRay (vrt_core.cuh) + the step budget
struct EngRay {
    LinRay l;                            // l.i is not used here
    int rem;                             // steps that may still run before the field must be looked at again
    uint32_t left;                       // maxSteps - steps taken - rem
    SparseField f;                       // VRT_TRAVERSAL_SPARSE: the two-level field of the ray's octant (l.p0 / l.SX.. are unused then)
};
// VRT_TRAVERSAL_SPARSE keeps the entry voxel and the octant instead of a pointer into a linear field: packed into the stash's pointer slot
__device__ __forceinline__ const uint8_t* engine_pack_entry(int mx, int my, int mz, uint32_t oct) {
    return reinterpret_cast<const uint8_t*>((unsigned long long)(uint32_t)mx | ((unsigned long long)(uint32_t)my << 16) | ((unsigned long long)(uint32_t)mz << 32) | ((unsigned long long)oct << 48));
}
// entry state of a ray for the sparse back-end in LinRay form (SX/SY/SZ carry only the step signs); LIN_OOB when it starts outside
__device__ __forceinline__ int engine_begin_sparse(const SceneDev& sc, const Dda& s, V3 dir, uint32_t maxSteps, LinRay& r) {
    r.sx = r.sx0 = s.sx; r.sy = r.sy0 = s.sy; r.sz = r.sz0 = s.sz;
    r.dx = s.dx; r.dy = s.dy; r.dz = s.dz;
    r.ax = fabsf(dir.x); r.ay = fabsf(dir.y); r.az = fabsf(dir.z);
    r.mx0 = s.mx; r.my0 = s.my; r.mz0 = s.mz; r.stx = s.stx; r.sty = s.sty; r.stz = s.stz;
    r.SX = s.stx; r.SY = s.sty; r.SZ = s.stz; r.T0 = 0;
    r.i = 0; r.maxSteps = maxSteps; r.mask = 0;
    if (!in_bounds(sc, s)) return LIN_OOB;
    r.p0 = engine_pack_entry(s.mx, s.my, s.mz, (s.stx < 0 ? 1u : 0u) | (s.sty < 0 ? 2u : 0u) | (s.stz < 0 ? 4u : 0u));
    return LIN_CONTINUE;
}

// SoA of 16-byte quads: lanes popping consecutive entries read consecutive quads
struct alignas(16) EngineStash {
    float4 s[ENGINE_STASH];              // sideDist at entry (== sideDist0);  w: flags = first budget << 20 | maxSteps << 8 | sign bits (1: step -1) x,y,z
    float4 d[ENGINE_STASH];              // deltaDist;                         w: user word 0
    float4 a[ENGINE_STASH];              // |dir|;                             w: user word 1
    uint2 p[ENGINE_STASH];               // entry pointer into the ray's safe-step field
};

// rem: the budget the ray's FIRST look-up granted (the set-up code does that look-up with all 32 lanes: a ray that ends there —
// more than half of the AO rays exhaust their 64 steps at once — never enters the stash or occupies a lane of the step rounds)
__device__ __forceinline__ void engine_push(EngineStash& st, uint32_t slot, const LinRay& r, uint32_t u0, uint32_t u1, uint32_t rem) {
    const uint32_t flags = (rem << 20) | (r.maxSteps << 8) | (r.SX < 0 ? 1u : 0u) | (r.SY < 0 ? 2u : 0u) | (r.SZ < 0 ? 4u : 0u);
    st.s[slot] = make_float4(r.sx, r.sy, r.sz, __uint_as_float(flags));
    st.d[slot] = make_float4(r.dx, r.dy, r.dz, __uint_as_float(u0));
    st.a[slot] = make_float4(r.ax, r.ay, r.az, __uint_as_float(u1));
    const unsigned long long p = (unsigned long long)r.p0;
    st.p[slot] = make_uint2((uint32_t)p, (uint32_t)(p >> 32));
}

template <bool SPARSE>
__device__ __forceinline__ void engine_pop(const EngineStash& st, uint32_t slot, EngRay& e, int LW, int LWH, uint32_t& u0, uint32_t& u1, const SparseField* octFields) {
    const float4 s = st.s[slot], d = st.d[slot], a = st.a[slot];
    const uint2 p = st.p[slot];
    const uint32_t flags = __float_as_uint(s.w);
    LinRay& r = e.l;
    r.sx = r.sx0 = s.x; r.sy = r.sy0 = s.y; r.sz = r.sz0 = s.z;
    r.dx = d.x; r.dy = d.y; r.dz = d.z;
    r.ax = a.x; r.ay = a.y; r.az = a.z;
    u0 = __float_as_uint(d.w); u1 = __float_as_uint(a.w);
    r.maxSteps = (flags >> 8) & 0xfffu;
    if (SPARSE) {
        r.stx = (flags & 1u) ? -1 : 1; r.sty = (flags & 2u) ? -1 : 1; r.stz = (flags & 4u) ? -1 : 1;
        r.mx0 = (int)(p.x & 0xffffu); r.my0 = (int)(p.x >> 16); r.mz0 = (int)(p.y & 0xffffu);
        e.f = octFields[(p.y >> 16) & 7u];
    } else {
        r.SX = (flags & 1u) ? -1 : 1; r.SY = (flags & 2u) ? -LW : LW; r.SZ = (flags & 4u) ? -LWH : LWH;
        r.T0 = (int)(0u - (uint32_t)VRT_LIN_MAGIC_BITS * (uint32_t)(r.SX + r.SY));
        r.p0 = reinterpret_cast<const uint8_t*>((unsigned long long)p.x | ((unsigned long long)p.y << 32));
    }
    e.rem = (int)(flags >> 20); e.left = r.maxSteps - (flags >> 20);
}

enum { ENG_GO = 0, ENG_HIT = 1, ENG_OOB = 2, ENG_EXHAUSTED = 3 };

// Look-up for a lane whose budget is spent (rem == 0). Mirrors the loop head of traceRayInt (:152-162): i < maxSteps, bounds
// test, getVoxel != 0. On ENG_GO the budget is re-armed with min(safe steps, steps left).
// TOP-UP: the field may be looked at BEFORE the budget is spent (rem > 0): the ray's current voxel lies inside the region its last
// look-up cleared, so the value there is a plain budget (never solid / outside), valid from here, and the larger of the two budgets
// stands. The look-up code is issued for the whole warp anyway whenever one lane needs it; lanes whose budget would run out inside
// the coming round join it for free (one more byte load) instead of idling through the rest of that round.
template <bool SPARSE>
__device__ __forceinline__ int engine_lookup(EngRay& e, const SceneDev& sc) {
    if (e.left == 0u && e.rem == 0) return ENG_EXHAUSTED;
    const LinRay& r = e.l;
    uint32_t n;
    if (SPARSE) {
        const int x = r.mx0 + lin_count(r.sx, r.sx0, r.ax) * r.stx, y = r.my0 + lin_count(r.sy, r.sy0, r.ay) * r.sty, z = r.mz0 + lin_count(r.sz, r.sz0, r.az) * r.stz;
        n = sparse_fetch(sc, e.f, x, y, z);
    } else {
        const int bx = lin_bits(r.sx, r.sx0, r.ax), by = lin_bits(r.sy, r.sy0, r.ay), bz = lin_bits(r.sz, r.sz0, r.az);
        const int t = (int)((uint32_t)bx * (uint32_t)r.SX + (uint32_t)by * (uint32_t)r.SY + (uint32_t)r.T0);
        const long long off = (long long)(bz - VRT_LIN_MAGIC_BITS) * (long long)r.SZ + (long long)t;
        asm("ld.global.nc.u8 %0, [%1];" : "=r"(n) : "l"(r.p0 + off));
    }
    if (n - 1u >= 254u) return n == 0u ? ENG_HIT : ENG_OOB;
    n = safe_decode(n);
    // any-hit: a budget that covers all remaining steps ends the ray as a miss after maxSteps getVoxel calls (see lin_round)
    const uint32_t total = e.left + (uint32_t)e.rem;          // steps the ray may still take, its current budget included
    if (n >= total) { e.left = 0u; e.rem = 0; return ENG_EXHAUSTED; }
    if (n > (uint32_t)e.rem) { e.rem = (int)n; e.left = total - n; }
    return ENG_GO;
}

// A block of N DDA steps (:164-170, select form) for the lanes with `on` set; the other lanes keep their state. Written in PTX so
// that the guard rides on the three compares (FSETP.LE.AND) and the adds stay predicated: 7 instructions per step (FMNMX3,
// 3 FSETP, 3 FADD) + one ISETP per block. The step is bound by the ALU pipe (min + compares: 4 of the 7 run there at half rate), so
// a per-step budget compare would cost a fifth ALU slot: a round runs its lanes' budgets as binary blocks instead (16+8+4+2+1).
#define VRT_STEP1 "min.f32 m, %0, %1, %2;\n\tsetp.le.and.f32 px, %0, m, pa;\n\tsetp.le.and.f32 py, %1, m, pa;\n\tsetp.le.and.f32 pz, %2, m, pa;\n\t" \
                  "@px add.f32 %0, %0, %3;\n\t@py add.f32 %1, %1, %4;\n\t@pz add.f32 %2, %2, %5;\n\t"
#define VRT_STEP2 VRT_STEP1 VRT_STEP1
#define VRT_STEP4 VRT_STEP2 VRT_STEP2
#define VRT_STEP8 VRT_STEP4 VRT_STEP4
#define VRT_STEP16 VRT_STEP8 VRT_STEP8
#define VRT_BLOCK_ASM(STEPS) asm("{\n\t.reg .pred pa, px, py, pz;\n\t.reg .f32 m;\n\tsetp.ne.s32 pa, %6, 0;\n\t" STEPS "}" \
        : "+f"(r.sx), "+f"(r.sy), "+f"(r.sz) : "f"(r.dx), "f"(r.dy), "f"(r.dz), "r"(on))
template <int N>
__device__ __forceinline__ void engine_block(LinRay& r, int on) {
    static_assert(N == 1 || N == 2 || N == 4 || N == 8 || N == 16, "block sizes are powers of two");
    if (N == 1) VRT_BLOCK_ASM(VRT_STEP1);
    else if (N == 2) VRT_BLOCK_ASM(VRT_STEP2);
    else if (N == 4) VRT_BLOCK_ASM(VRT_STEP4);
    else if (N == 8) VRT_BLOCK_ASM(VRT_STEP8);
    else VRT_BLOCK_ASM(VRT_STEP16);
}
// A round of up to 2*TOP-1 steps: every lane runs min(rem, 2*TOP-1) of them, as the binary blocks TOP, TOP/2, .., 1
template <int TOP>
__device__ __forceinline__ void engine_round(LinRay& r, int& rem) {
    const int n = min(rem, 2 * TOP - 1);
    if (TOP >= 16) engine_block<16>(r, n & 16);
    if (TOP >= 8) engine_block<8>(r, n & 8);
    if (TOP >= 4) engine_block<4>(r, n & 4);
    if (TOP >= 2) engine_block<2>(r, n & 2);
    engine_block<1>(r, n & 1);
    rem -= n;
}
__device__ __forceinline__ uint32_t engine_steps_taken(const EngRay& e) { return e.l.maxSteps - e.left - (uint32_t)e.rem; }

// literal loop of traceRayInt for rays the select-form DDA cannot replay (zero / denormal / NaN components): any-hit form
__device__ __forceinline__ bool engine_anyhit_literal(const SceneDev& sc, Dda s, uint32_t maxSteps, uint32_t& fetches) {
    uint32_t i = 0; bool hit = false;
    for (; i < maxSteps; i++) {
        if (!in_bounds(sc, s)) break;
        if (fetch_voxel(sc, s) != 0) { hit = true; break; }
        dda_step_literal(s);
    }
    fetches = hit ? i + 1 : i;
    return hit;
}

}  // namespace vrt
