// vrt_core.cuh — per-ray / per-pixel arithmetic of the voxel_volume pass, written for the device and also
// compilable for the host (tests/ compiles it with g++ to check the kernel logic bit-for-bit without a GPU).
// Follows /root/reference/shader/voxel_volume.frag (line numbers cited per function). Exactness rules:
// compile with -fmad=false (nvcc) / -ffp-contract=off (g++); never -use_fast_math; IEEE div/sqrt.
#pragma once
#include <math.h>
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
VRT_HD uint32_t fetch_voxel(const SceneDev& sc, const Dda& s) {   // getVoxel :73-77 as an integer fetch
    return sc.vol[(size_t)s.mx + (size_t)s.my * sc.W + (size_t)s.mz * sc.W * (size_t)sc.H];
}

// traceRayInt + traceRay (:127-196). TRAVERSAL selects how voxel emptiness is learned; the float/int DDA state
// evolves identically in every variant (one dda_step per reference iteration).
template <int TRAVERSAL>
VRT_HD TraceRes trace_ray(const SceneDev& sc, V3 start, V3 dir, uint32_t maxSteps) {
    TraceRes r;
    V3 pos = box_intersection(sc, start, dir);
    Dda s;
    s.mx = f2i(floorf(pos.x)); s.my = f2i(floorf(pos.y)); s.mz = f2i(floorf(pos.z));
    s.dx = fabsf(1.0f / dir.x); s.dy = fabsf(1.0f / dir.y); s.dz = fabsf(1.0f / dir.z);
    const float gx = signf(dir.x), gy = signf(dir.y), gz = signf(dir.z);
    s.stx = f2i(gx); s.sty = f2i(gy); s.stz = f2i(gz);
    s.sx = ((gx * ((float)s.mx - pos.x) + (gx * 0.5f)) + 0.5f) * s.dx;
    s.sy = ((gy * ((float)s.my - pos.y) + (gy * 0.5f)) + 0.5f) * s.dy;
    s.sz = ((gz * ((float)s.mz - pos.z) + (gz * 0.5f)) + 0.5f) * s.dz;

    uint32_t i = 0, mask = 0, material = 0;
    bool hit = false, oob = false;
    const bool fast = finitef(s.dx) && finitef(s.dy) && finitef(s.dz);
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
