// vrt_core.cuh — per-ray / per-pixel arithmetic of the voxel_volume pass, written for the device and also
// compilable for the host (tests/ compiles it with g++ to check the kernel logic bit-for-bit without a GPU).
// Follows /root/reference/shader/voxel_volume.frag (line numbers cited per function). Exactness rules:
// compile with -fmad=false (nvcc) / -ffp-contract=off (g++); never -use_fast_math; IEEE div/sqrt.
#pragma once
#include <math.h>
#include <string.h>
#include "vrt_types.h"

namespace vrt {

struct V3 { float x, y, z; };
VRT_HD V3 mk3(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
VRT_HD V3 operator+(V3 a, V3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
VRT_HD V3 operator-(V3 a, V3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
VRT_HD V3 operator*(V3 a, V3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
VRT_HD V3 operator*(V3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
VRT_HD V3 operator*(float s, V3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
VRT_HD V3 operator/(V3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
VRT_HD V3 operator-(V3 a) { return mk3(-a.x, -a.y, -a.z); }
VRT_HD float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
VRT_HD float length3(V3 a) { return sqrtf(dot3(a, a)); }
VRT_HD V3 normalize3(V3 a) { float inv = 1.0f / sqrtf(dot3(a, a)); return a * inv; }   // glm: v * inversesqrt(dot)
VRT_HD V3 reflect3(V3 I, V3 N) { return I - N * dot3(N, I) * 2.0f; }                     // glm: I - N*dot(N,I)*2
VRT_HD float signf(float x) { return (float)((0.0f < x) - (x < 0.0f)); }

VRT_HD int f2i(float f) {  // cvt.rzi.s32.f32: truncate, saturate, NaN -> 0
#if defined(__CUDA_ARCH__)
    return __float2int_rz(f);
#else
    if (f != f) return 0;
    if (f >= 2147483648.0f) return 2147483647;
    if (f <= -2147483648.0f) return (-2147483647 - 1);
    return (int)f;
#endif
}
VRT_HD float min3f(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    float r; asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r;   // FMNMX3 on sm_100
#else
    return fminf(fminf(a, b), c);
#endif
}
VRT_HD bool finitef(float x) { return fabsf(x) <= 3.402823466e+38f; }
VRT_HD uint8_t to_unorm8(float f) {
    if (!(f == f)) return 0;
    float c = f < 0.0f ? 0.0f : (f > 1.0f ? 1.0f : f);
    return (uint8_t)rintf(c * 255.0f);
}
VRT_HD int8_t to_snorm8(float f) {
    if (!(f == f)) return 0;
    float c = f < -1.0f ? -1.0f : (f > 1.0f ? 1.0f : f);
    return (int8_t)rintf(c * 127.0f);
}

struct RayHit { uint32_t material; V3 pos, normal, dir; };   // voxel_volume.frag:43-49

struct TraceRes {
    uint32_t material;   // value of the last getVoxel (0 on a miss)
    V3 pos, normal;      // RayHit.pos / RayHit.normal (zero when material == 0: SURVEY F5)
    int mx, my, mz;      // mapPos at loop exit
    uint32_t steps;      // loop index i at exit
    uint32_t fetches;    // getVoxel calls of the reference loop
};

// ---- boxIntersection, voxel_volume.frag:109-125 ----
VRT_HD V3 box_intersection(const SceneDev& sc, V3 start, V3 dir) {
    V3 inv = mk3(1.0f / dir.x, 1.0f / dir.y, 1.0f / dir.z);
    V3 t1 = (-start) * inv;
    V3 t2 = (mk3((float)sc.W, (float)sc.H, (float)sc.D) - start) * inv;
    V3 tmn = mk3(fminf(t1.x, t2.x), fminf(t1.y, t2.y), fminf(t1.z, t2.z));
    V3 tmx = mk3(fmaxf(t1.x, t2.x), fmaxf(t1.y, t2.y), fmaxf(t1.z, t2.z));
    float tmin = fmaxf(tmn.x, fmaxf(tmn.y, tmn.z));
    float tmax = fminf(tmx.x, fminf(tmx.y, tmx.z));
    if (tmin >= 0.0f && tmax >= tmin) return start + (tmin + 0.1f) * dir;
    return start;
}

struct Dda {            // loop-carried state of traceRayInt (:127-174)
    float sx, sy, sz;   // sideDist
    float dx, dy, dz;   // deltaDist
    int mx, my, mz;     // mapPos
    int stx, sty, stz;  // rayStep
};

// One DDA step, reference lines :164-170. Select form == the literal float(mask)*deltaDist form whenever
// deltaDist is finite (0*finite = 0, sideDist is never -0); callers guarantee that (else: dda_step_literal).
VRT_HD void dda_step(Dda& s) {
    float m = min3f(s.sx, s.sy, s.sz);          // sx <= min(sy,sz)  <=>  sx <= min3 (also with NaNs: minNum)
    bool bx = s.sx <= m, by = s.sy <= m, bz = s.sz <= m;
    if (bx) { s.sx += s.dx; s.mx += s.stx; }
    if (by) { s.sy += s.dy; s.my += s.sty; }
    if (bz) { s.sz += s.dz; s.mz += s.stz; }
}
VRT_HD uint32_t dda_step_rec(Dda& s) {
    float m = min3f(s.sx, s.sy, s.sz);
    bool bx = s.sx <= m, by = s.sy <= m, bz = s.sz <= m;
    if (bx) { s.sx += s.dx; s.mx += s.stx; }
    if (by) { s.sy += s.dy; s.my += s.sty; }
    if (bz) { s.sz += s.dz; s.mz += s.stz; }
    return (bx ? 1u : 0u) | (by ? 2u : 0u) | (bz ? 4u : 0u);
}
// float-only steps for VRT_TRAVERSAL_FINE (mapPos is recovered from sideDist at lookups)
VRT_HD void dda_step_f(Dda& s) {
    float m = min3f(s.sx, s.sy, s.sz);
    if (s.sx <= m) s.sx += s.dx;
    if (s.sy <= m) s.sy += s.dy;
    if (s.sz <= m) s.sz += s.dz;
}
// k float-only steps. Unrolled by eight with a binary remainder: the plain `for` loop spent 2 of every 9 issued instructions on its
// counter and branch (ncu r2p1: 16 % of wave_primary_kernel), and the step is bound by instruction issue.
VRT_HD void dda_steps_f(Dda& s, uint32_t k) {
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (; k >= 8u; k -= 8u) { dda_step_f(s); dda_step_f(s); dda_step_f(s); dda_step_f(s); dda_step_f(s); dda_step_f(s); dda_step_f(s); dda_step_f(s); }
    if (k & 4u) { dda_step_f(s); dda_step_f(s); dda_step_f(s); dda_step_f(s); }
    if (k & 2u) { dda_step_f(s); dda_step_f(s); }
    if (k & 1u) dda_step_f(s);
}
VRT_HD uint32_t dda_step_f_rec(Dda& s) {
    float m = min3f(s.sx, s.sy, s.sz);
    bool bx = s.sx <= m, by = s.sy <= m, bz = s.sz <= m;
    if (bx) s.sx += s.dx;
    if (by) s.sy += s.dy;
    if (bz) s.sz += s.dz;
    return (bx ? 1u : 0u) | (by ? 2u : 0u) | (bz ? 4u : 0u);
}
VRT_HD uint32_t dda_step_literal(Dda& s) {   // exactly as written in the shader (handles inf/NaN deltaDist)
    bool bx = s.sx <= fminf(s.sy, s.sz), by = s.sy <= fminf(s.sz, s.sx), bz = s.sz <= fminf(s.sx, s.sy);
    float fx = bx ? 1.0f : 0.0f, fy = by ? 1.0f : 0.0f, fz = bz ? 1.0f : 0.0f;
    s.sx = s.sx + fx * s.dx; s.sy = s.sy + fy * s.dy; s.sz = s.sz + fz * s.dz;
    s.mx += f2i(fx) * s.stx; s.my += f2i(fy) * s.sty; s.mz += f2i(fz) * s.stz;
    return (bx ? 1u : 0u) | (by ? 2u : 0u) | (bz ? 4u : 0u);
}

VRT_HD bool in_bounds(const SceneDev& sc, const Dda& s) {
    return (unsigned)s.mx < sc.W && (unsigned)s.my < sc.H && (unsigned)s.mz < sc.D;
}
// one byte of a two-level field / volume (SceneDev::sdist, ::svol): the brick's word, then the per-voxel byte only for bricks that have one
VRT_HD uint32_t sparse_fetch(const SceneDev& sc, const SparseField& f, int x, int y, int z) {   // x in [-1, W], y in [-1, H], z in [-1, D]
    const uint32_t X = (uint32_t)(x + 8), Y = (uint32_t)(y + 8), Z = (uint32_t)(z + 8);
    const uint32_t e = f.top[(X >> 3) + ((Y >> 3) + (Z >> 3) * sc.TBy) * sc.TBx];
    if (!(e & VRT_SP_FINE)) return e;
    return f.fine[(size_t)(e & ~VRT_SP_FINE) * 512u + ((X & 7u) | ((Y & 7u) << 3) | ((Z & 7u) << 6))];
}
VRT_HD uint32_t fetch_voxel(const SceneDev& sc, const Dda& s) {   // getVoxel :73-77 as an integer fetch
    if (sc.vol) return sc.vol[(size_t)s.mx + (size_t)s.my * sc.W + (size_t)s.mz * sc.W * (size_t)sc.H];
    return sparse_fetch(sc, sc.svol, s.mx, s.my, s.mz);           // scenes over the dense budget: the volume is two-level too
}

// ---- VRT_TRAVERSAL_LINEAR: resumable ray over the aproned linear safe-step field (ldist) ----
// Look-up = 3 x (FADD, FFMA) + 3 integer multiply-adds + one byte load. The per-axis step count since entry is
// recovered as rint((sideDist - sideDist0) * |dir|) by adding 1.5*2^23 inside the FFMA: the sum's low mantissa bits are
// the integer (error of <= 512 accumulated adds < 0.02 step, so the rounding is exact). No coordinates, no bounds test
// (apron), no brick address arithmetic.
constexpr float VRT_LIN_MAGIC = 12582912.0f;
constexpr int VRT_LIN_MAGIC_BITS = 0x4B400000;
VRT_HD int lin_count(float s, float s0, float a) {
#if defined(__CUDA_ARCH__)
    return __float_as_int(__fmaf_rn(s - s0, a, VRT_LIN_MAGIC)) - VRT_LIN_MAGIC_BITS;
#else
    const float r = fmaf(s - s0, a, VRT_LIN_MAGIC);
    int b; memcpy(&b, &r, 4); return b - VRT_LIN_MAGIC_BITS;
#endif
}
VRT_HD int lin_bits(float s, float s0, float a) {   // VRT_LIN_MAGIC_BITS + count
#if defined(__CUDA_ARCH__)
    return __float_as_int(__fmaf_rn(s - s0, a, VRT_LIN_MAGIC));
#else
    const float r = fmaf(s - s0, a, VRT_LIN_MAGIC);
    int b; memcpy(&b, &r, 4); return b;
#endif
}
struct LinRay {
    float sx, sy, sz;          // sideDist
    float dx, dy, dz;          // deltaDist
    float sx0, sy0, sz0;       // sideDist at entry
    float ax, ay, az;          // |dir|
    const uint8_t* p0;         // &ldist[entry voxel]
    int SX, SY, SZ;            // signed ldist strides of one step per axis (|SZ| <= (W+2)(H+2) < 2^31)
    int T0;                    // -MAGIC_BITS * (SX + SY), 32-bit wrap-around
    uint32_t i, maxSteps;
    uint32_t mask;             // axes advanced by the last step
    int mx0, my0, mz0, stx, sty, stz;
};
enum { LIN_CONTINUE = 0, LIN_HIT = 1, LIN_OOB = 2, LIN_EXHAUSTED = 3 };

// the directional field of the ray's octant (rayStep components are +-1 here: zero components take the literal loop)
VRT_HD const uint8_t* octant_field(const SceneDev& sc, const Dda& s) {
    return sc.odist[(s.stx < 0 ? 1 : 0) | (s.sty < 0 ? 2 : 0) | (s.stz < 0 ? 4 : 0)];
}
// Entry state from the DDA set-up of traceRayInt (:131-150). Returns LIN_OOB when mapPos starts outside the volume.
VRT_HD int lin_begin(const SceneDev& sc, const uint8_t* field, const Dda& s, V3 dir, uint32_t maxSteps, LinRay& r) {
    r.sx = r.sx0 = s.sx; r.sy = r.sy0 = s.sy; r.sz = r.sz0 = s.sz;
    r.dx = s.dx; r.dy = s.dy; r.dz = s.dz;
    r.ax = fabsf(dir.x); r.ay = fabsf(dir.y); r.az = fabsf(dir.z);
    r.mx0 = s.mx; r.my0 = s.my; r.mz0 = s.mz; r.stx = s.stx; r.sty = s.sty; r.stz = s.stz;
    r.i = 0; r.maxSteps = maxSteps; r.mask = 0;
    if (!in_bounds(sc, s)) return LIN_OOB;
    const int LW = (int)sc.W + 2, LWH = LW * ((int)sc.H + 2);
    r.p0 = field + ((size_t)(s.mx + 1) + (size_t)(s.my + 1) * (size_t)LW + (size_t)(s.mz + 1) * (size_t)LWH);
    r.SX = s.stx; r.SY = s.sty * LW; r.SZ = s.stz * LWH;
    r.T0 = (int)(0u - (uint32_t)VRT_LIN_MAGIC_BITS * (uint32_t)(r.SX + r.SY));
#if defined(__CUDA_ARCH__)
    // keep the strides in registers: without this ptxas re-derives them from W and H (64-bit multiplies) at every look-up
    asm volatile("" : "+r"(r.SX), "+r"(r.SY), "+r"(r.SZ), "+r"(r.T0), "+l"(r.p0));
#endif
    return LIN_CONTINUE;
}
// One look-up and the steps it allows. STEP_REC: keep the mask of the last step (needed for the hit normal).
// ANYHIT: the caller does not use the DDA state of a ray that misses (traceRayHit :198-202 returns a bool; traceRay :180-195 fills
// RayHit only on a hit). Once the field grants at least as many steps as the ray has left, the loop of traceRayInt provably runs
// out of steps over empty in-bounds voxels (exactly maxSteps getVoxel calls, no hit): the remaining steps need not be executed
// because their only product, the DDA state, is discarded. Primary rays never use this (their exit state is instrumented).
template <bool STEP_REC>
VRT_HD int lin_round(const SceneDev& sc, LinRay& r, const bool ANYHIT = false, const uint32_t roundCap = 512u) {
    const int bx = lin_bits(r.sx, r.sx0, r.ax), by = lin_bits(r.sy, r.sy0, r.ay), bz = lin_bits(r.sz, r.sz0, r.az);
    const int t = (int)((uint32_t)bx * (uint32_t)r.SX + (uint32_t)by * (uint32_t)r.SY + (uint32_t)r.T0);   // cx*SX + cy*SY
    const long long off = (long long)(bz - VRT_LIN_MAGIC_BITS) * (long long)r.SZ + (long long)t;
#if defined(__CUDA_ARCH__)
    uint32_t n;
    asm("ld.global.nc.u8 %0, [%1];" : "=r"(n) : "l"(r.p0 + off));   // p0 went through an asm barrier: say "global" again
#else
    uint32_t n = r.p0[off];
#endif
    if (n == 0) return LIN_HIT;
    if (n == VRT_LIN_OOB) return LIN_OOB;
    n = safe_decode(n);
    const uint32_t left = r.maxSteps - r.i;
    if (ANYHIT && n >= left) { r.i = r.maxSteps; return LIN_EXHAUSTED; }
    n = n < left ? n : left;
    n = n < roundCap ? n : roundCap;
    r.i += n;
    Dda d; d.sx = r.sx; d.sy = r.sy; d.sz = r.sz; d.dx = r.dx; d.dy = r.dy; d.dz = r.dz;
    dda_steps_f(d, n - 1u);
    if (STEP_REC) r.mask = dda_step_f_rec(d); else dda_step_f(d);
    r.sx = d.sx; r.sy = d.sy; r.sz = d.sz;
    return r.i >= r.maxSteps ? LIN_EXHAUSTED : LIN_CONTINUE;
}
// mapPos of the current state
VRT_HD void lin_map_pos(const LinRay& r, int& mx, int& my, int& mz) {
    mx = r.mx0 + lin_count(r.sx, r.sx0, r.ax) * r.stx;
    my = r.my0 + lin_count(r.sy, r.sy0, r.ay) * r.sty;
    mz = r.mz0 + lin_count(r.sz, r.sz0, r.az) * r.stz;
}

// ---- VRT_TRAVERSAL_SPARSE: the same resumable ray over the two-level directional fields (SceneDev::sdist) ----
// The look-up needs the voxel itself (brick index + position inside the brick): per axis the entry coordinate plus the recovered step
// count. One 4-byte load of the top level (L2-resident), and a second, dependent byte load only for bricks near a surface.
VRT_HD SparseField sparse_octant_field(const SceneDev& sc, const Dda& s) {
    return sc.sdist[(s.stx < 0 ? 1 : 0) | (s.sty < 0 ? 2 : 0) | (s.stz < 0 ? 4 : 0)];
}
struct SpRay {
    float sx, sy, sz, dx, dy, dz, sx0, sy0, sz0, ax, ay, az;   // as LinRay
    SparseField f;
    int mx0, my0, mz0, stx, sty, stz;
    uint32_t i, maxSteps, mask;
};
VRT_HD int sp_begin(const SceneDev& sc, const Dda& s, V3 dir, uint32_t maxSteps, SpRay& r) {
    r.sx = r.sx0 = s.sx; r.sy = r.sy0 = s.sy; r.sz = r.sz0 = s.sz;
    r.dx = s.dx; r.dy = s.dy; r.dz = s.dz;
    r.ax = fabsf(dir.x); r.ay = fabsf(dir.y); r.az = fabsf(dir.z);
    r.mx0 = s.mx; r.my0 = s.my; r.mz0 = s.mz; r.stx = s.stx; r.sty = s.sty; r.stz = s.stz;
    r.i = 0; r.maxSteps = maxSteps; r.mask = 0;
    if (!in_bounds(sc, s)) return LIN_OOB;
    r.f = sparse_octant_field(sc, s);
    return LIN_CONTINUE;
}
VRT_HD void sp_map_pos(const SpRay& r, int& mx, int& my, int& mz) {
    mx = r.mx0 + lin_count(r.sx, r.sx0, r.ax) * r.stx;
    my = r.my0 + lin_count(r.sy, r.sy0, r.ay) * r.sty;
    mz = r.mz0 + lin_count(r.sz, r.sz0, r.az) * r.stz;
}
template <bool STEP_REC>
VRT_HD int sp_round(const SceneDev& sc, SpRay& r, const bool ANYHIT = false, const uint32_t roundCap = 512u) {
    int x, y, z; sp_map_pos(r, x, y, z);
    uint32_t n = sparse_fetch(sc, r.f, x, y, z);
    if (n == 0) return LIN_HIT;
    if (n == VRT_LIN_OOB) return LIN_OOB;
    n = safe_decode(n);
    const uint32_t left = r.maxSteps - r.i;
    if (ANYHIT && n >= left) { r.i = r.maxSteps; return LIN_EXHAUSTED; }
    n = n < left ? n : left;
    n = n < roundCap ? n : roundCap;
    r.i += n;
    Dda d; d.sx = r.sx; d.sy = r.sy; d.sz = r.sz; d.dx = r.dx; d.dy = r.dy; d.dz = r.dz;
    dda_steps_f(d, n - 1u);
    if (STEP_REC) r.mask = dda_step_f_rec(d); else dda_step_f(d);
    r.sx = d.sx; r.sy = d.sy; r.sz = d.sz;
    return r.i >= r.maxSteps ? LIN_EXHAUSTED : LIN_CONTINUE;
}

// traceRayInt + traceRay (:127-196). TRAVERSAL selects how voxel emptiness is learned; the float/int DDA state
// evolves identically in every variant (one dda_step per reference iteration).
// DDA set-up of traceRayInt (:131-150). Returns whether the fast select-form step may be used (finite deltaDist).
VRT_HD bool dda_begin(const SceneDev& sc, V3 start, V3 dir, V3& pos, Dda& s) {
    pos = box_intersection(sc, start, dir);
    s.mx = f2i(floorf(pos.x)); s.my = f2i(floorf(pos.y)); s.mz = f2i(floorf(pos.z));
    s.dx = fabsf(1.0f / dir.x); s.dy = fabsf(1.0f / dir.y); s.dz = fabsf(1.0f / dir.z);
    const float gx = signf(dir.x), gy = signf(dir.y), gz = signf(dir.z);
    s.stx = f2i(gx); s.sty = f2i(gy); s.stz = f2i(gz);
    s.sx = ((gx * ((float)s.mx - pos.x) + (gx * 0.5f)) + 0.5f) * s.dx;
    s.sy = ((gy * ((float)s.my - pos.y) + (gy * 0.5f)) + 0.5f) * s.dy;
    s.sz = ((gz * ((float)s.mz - pos.z) + (gz * 0.5f)) + 0.5f) * s.dz;
    return finitef(s.dx) && finitef(s.dy) && finitef(s.dz);
}
// LINEAR additionally needs a finite start (see trace_ray)
VRT_HD bool lin_usable(bool fast, const Dda& s) { return fast && finitef(s.sx) && finitef(s.sy) && finitef(s.sz); }

constexpr uint32_t ROUND_CAP_PRIMARY = 128u, ROUND_CAP_SECONDARY = 32u;   // profiles/r1_round_cap.log
template <int TRAVERSAL>
VRT_HD TraceRes trace_ray(const SceneDev& sc, V3 start, V3 dir, uint32_t maxSteps, const bool anyhit = false, const uint32_t roundCap = ROUND_CAP_PRIMARY) {
    TraceRes r;
    V3 pos;
    Dda s;
    bool fast = dda_begin(sc, start, dir, pos, s);

    uint32_t i = 0, mask = 0, material = 0;
    bool hit = false, oob = false;
    // a NaN / infinite start (origin built from a NaN normal) never moves: the literal loop replays that too
    if (TRAVERSAL == VRT_TRAVERSAL_LINEAR || TRAVERSAL == VRT_TRAVERSAL_OCTANT || TRAVERSAL == VRT_TRAVERSAL_SPARSE) fast = lin_usable(fast, s);
    if (!fast) {
        // literal loop: rays with a zero / denormal / NaN direction component
        for (; i < maxSteps; i++) {
            if (!in_bounds(sc, s)) { oob = true; break; }
            material = fetch_voxel(sc, s);
            if (material != 0) { hit = true; break; }
            mask = dda_step_literal(s);
        }
    } else if (TRAVERSAL == VRT_TRAVERSAL_DENSE_U8) {
        for (; i < maxSteps; i++) {
            if (!in_bounds(sc, s)) { oob = true; break; }
            material = fetch_voxel(sc, s);
            if (material != 0) { hit = true; break; }
            mask = dda_step_rec(s);
        }
    } else if (TRAVERSAL == VRT_TRAVERSAL_LINEAR || TRAVERSAL == VRT_TRAVERSAL_OCTANT) {
        LinRay lr;
        int st = lin_begin(sc, TRAVERSAL == VRT_TRAVERSAL_OCTANT ? octant_field(sc, s) : sc.ldist, s, dir, maxSteps, lr);
        while (st == LIN_CONTINUE) st = lin_round<true>(sc, lr, anyhit, roundCap);
        lin_map_pos(lr, s.mx, s.my, s.mz);
        s.sx = lr.sx; s.sy = lr.sy; s.sz = lr.sz;
        i = lr.i; mask = lr.mask;
        if (st == LIN_HIT) { material = fetch_voxel(sc, s); hit = true; }
        else if (st == LIN_OOB) oob = true;
    } else if (TRAVERSAL == VRT_TRAVERSAL_SPARSE) {
        SpRay sr;
        int st = sp_begin(sc, s, dir, maxSteps, sr);
        while (st == LIN_CONTINUE) st = sp_round<true>(sc, sr, anyhit, roundCap);
        sp_map_pos(sr, s.mx, s.my, s.mz);
        s.sx = sr.sx; s.sy = sr.sy; s.sz = sr.sz;
        i = sr.i; mask = sr.mask;
        if (st == LIN_HIT) { material = fetch_voxel(sc, s); hit = true; }
        else if (st == LIN_OOB) oob = true;
    } else if (TRAVERSAL == VRT_TRAVERSAL_FINE) {
        // Per-voxel safe-step field. mapPos is not carried through the steps: it is recovered at each lookup from
        // how far sideDist has advanced, c = rint((sideDist - sideDist0) / deltaDist) per axis. The accumulated
        // rounding error of <= 512 float adds is < 0.02 of a step, so the rounded count is exact.
        const float sx0 = s.sx, sy0 = s.sy, sz0 = s.sz;
        const float ax = fabsf(dir.x), ay = fabsf(dir.y), az = fabsf(dir.z);
        const int mx0 = s.mx, my0 = s.my, mz0 = s.mz;
        while (i < maxSteps) {
            s.mx = mx0 + f2i(rintf((s.sx - sx0) * ax)) * s.stx;
            s.my = my0 + f2i(rintf((s.sy - sy0) * ay)) * s.sty;
            s.mz = mz0 + f2i(rintf((s.sz - sz0) * az)) * s.stz;
            if (!in_bounds(sc, s)) { oob = true; break; }
            const uint32_t b = ((uint32_t)s.mz >> 2) * (sc.By * sc.Bx) + ((uint32_t)s.my >> 2) * sc.Bx + ((uint32_t)s.mx >> 2);
            const uint32_t in = ((uint32_t)s.mx & 3u) | (((uint32_t)s.my & 3u) << 2) | (((uint32_t)s.mz & 3u) << 4);
            uint32_t n = sc.fdist[(size_t)b * 64u + in];
            if (n == 0) { material = fetch_voxel(sc, s); hit = true; break; }
            const uint32_t left = maxSteps - i;
            n = n < left ? n : left;
            i += n;
            for (uint32_t k = n - 1; k > 0; k--) dda_step_f(s);
            mask = dda_step_f_rec(s);
        }
        if (!hit && !oob) {   // ran out of steps: mapPos of the exit state (instrumentation only)
            s.mx = mx0 + f2i(rintf((s.sx - sx0) * ax)) * s.stx;
            s.my = my0 + f2i(rintf((s.sy - sy0) * ay)) * s.sty;
            s.mz = mz0 + f2i(rintf((s.sz - sz0) * az)) * s.stz;
        }
    } else {
        while (i < maxSteps) {
            if (!in_bounds(sc, s)) { oob = true; break; }
            const uint32_t b = ((uint32_t)s.mz >> 2) * (sc.By * sc.Bx) + ((uint32_t)s.my >> 2) * sc.Bx + ((uint32_t)s.mx >> 2);
            const uint32_t dist = sc.bdist[b];
            uint32_t n;
            if (dist == 0) {
                const unsigned long long m64 = sc.bmask[b];
                const uint32_t bit = ((uint32_t)s.mx & 3u) | (((uint32_t)s.my & 3u) << 2) | (((uint32_t)s.mz & 3u) << 4);
                if ((m64 >> bit) & 1ull) { material = fetch_voxel(sc, s); hit = true; break; }
                n = 1;
            } else {
                n = (dist - 1) * 4 + 1;
            }
            const uint32_t left = maxSteps - i;
            n = n < left ? n : left;
            i += n;
            for (uint32_t k = n - 1; k > 0; k--) dda_step(s);
            mask = dda_step_rec(s);
        }
    }
    (void)oob;
    r.material = material;
    r.mx = s.mx; r.my = s.my; r.mz = s.mz;
    r.steps = i;
    r.fetches = hit ? i + 1 : i;
    if (material != 0) {
        // :186-192  normal = normalize(vec3(mask) * -vec3(rayStep)); d = length(vec3(mask)*(sideDist-deltaDist))
        const float fx = (mask & 1u) ? 1.0f : 0.0f, fy = (mask & 2u) ? 1.0f : 0.0f, fz = (mask & 4u) ? 1.0f : 0.0f;
        r.normal = normalize3(mk3(fx * -(float)s.stx, fy * -(float)s.sty, fz * -(float)s.stz));
        const float d = length3(mk3(fx * (s.sx - s.dx), fy * (s.sy - s.dy), fz * (s.sz - s.dz)));
        r.pos = pos + d * dir;
    } else {
        r.normal = mk3(0.0f, 0.0f, 0.0f);
        r.pos = mk3(0.0f, 0.0f, 0.0f);
    }
    return r;
}

}  // namespace vrt
