// ORACLE — TEST INFRASTRUCTURE ONLY. Not a product path. Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference leg may build, link or call anything in oracle/.
//
// orc_math.h: the GLSL / glm scalar semantics the restated shaders run on, pinned so that a
// g++ build (-O2 -ffp-contract=off, no fast-math) is a deterministic IEEE-754 binary32 program.
// PARITY UNPINNED by the reference: it has no tests or golden vectors for this path
// (/root/reference/test/test_test.cpp:1-12 is a factorial placeholder), so the decisions below
// ARE the definition; each is listed in DESIGN.md "pinned semantics".
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };
struct ivec3 { int x, y, z; };
struct bvec3 { bool x, y, z; };

// ---- componentwise arithmetic (GLSL operators) ----
static inline vec3 v3(float a) { return {a, a, a}; }
static inline vec3 v3(float x, float y, float z) { return {x, y, z}; }
static inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
static inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline vec3 operator*(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
static inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
static inline vec3 operator/(float s, vec3 a) { return {s / a.x, s / a.y, s / a.z}; }
static inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
static inline vec3 operator+(vec3 a, float s) { return {a.x + s, a.y + s, a.z + s}; }
static inline vec3 operator-(vec3 a, float s) { return {a.x - s, a.y - s, a.z - s}; }

static inline vec4 operator+(vec4 a, vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
static inline vec4 operator-(vec4 a, vec4 b) { return {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
static inline vec4 operator*(vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
static inline vec4 operator/(vec4 a, float s) { return {a.x / s, a.y / s, a.z / s, a.w / s}; }

// ---- GLSL built-ins, pinned ----
// min/max: IEEE minNum/maxNum (what FMNMX does on NVIDIA hardware; GLSL leaves NaN undefined). SURVEY F9.
static inline float gmin(float a, float b) { return fminf(a, b); }
static inline float gmax(float a, float b) { return fmaxf(a, b); }
static inline vec3 gmin(vec3 a, vec3 b) { return {gmin(a.x, b.x), gmin(a.y, b.y), gmin(a.z, b.z)}; }
static inline vec3 gmax(vec3 a, vec3 b) { return {gmax(a.x, b.x), gmax(a.y, b.y), gmax(a.z, b.z)}; }
static inline vec3 gabs(vec3 a) { return {fabsf(a.x), fabsf(a.y), fabsf(a.z)}; }
static inline vec3 gfloor(vec3 a) { return {floorf(a.x), floorf(a.y), floorf(a.z)}; }
// sign(): -1, 0, +1 (0 for +-0; NaN -> 0)
static inline float gsign(float x) { return (float)((0.0f < x) - (x < 0.0f)); }
static inline vec3 gsign(vec3 a) { return {gsign(a.x), gsign(a.y), gsign(a.z)}; }
// dot: products summed left to right (glm compute_dot<vec3>: tmp.x + tmp.y + tmp.z)
static inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline float dot(vec4 a, vec4 b) { return ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w; }
static inline float length(vec3 a) { return sqrtf(dot(a, a)); }
// normalize: glm semantics v * inversesqrt(dot(v,v)), inversesqrt(x) = 1/sqrt(x) (correctly rounded ops)
static inline vec3 normalize(vec3 a) { float inv = 1.0f / sqrtf(dot(a, a)); return a * inv; }
// reflect: glm func_geometric.inl  I - N * dot(N, I) * 2
static inline vec3 reflect(vec3 I, vec3 N) { return I - N * dot(N, I) * 2.0f; }
static inline vec3 cross(vec3 a, vec3 b) {
    return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
// mod(x, 1.0) = x - 1.0*floor(x/1.0)
static inline float gmod1(float x) { return x - floorf(x); }

// float -> int conversion (GLSL int(float) is undefined out of range / NaN). Pinned to the
// CUDA cvt.rzi.s32.f32 behaviour: truncate, saturate, NaN -> 0.
static inline int f2i(float f) {
    if (f != f) return 0;
    if (f >= 2147483648.0f) return 2147483647;
    if (f <= -2147483648.0f) return (-2147483647 - 1);
    return (int)f;
}

// atan(y,x) / asin(x): GLSL gives only an accuracy bound. Pinned to the correctly rounded
// binary32 value, obtained by evaluating in binary64 and narrowing (skybox lookup is NEAREST,
// so 1-ulp differences could flip a texel; this makes g++ and nvcc agree).
static inline float atan2_cr(float y, float x) { return (float)atan2((double)y, (double)x); }
static inline float asin_cr(float x) { return (float)asin((double)x); }

// Render-target store conversions (Vulkan spec "Conversion from floating point to normalized fixed-point":
// clamp, scale, round to nearest; NaN -> 0). Pinned to round-half-even (rintf).
static inline uint8_t to_unorm8(float f) {
    if (!(f == f)) return 0;
    float c = f < 0.0f ? 0.0f : (f > 1.0f ? 1.0f : f);
    return (uint8_t)rintf(c * 255.0f);
}
static inline int8_t to_snorm8(float f) {
    if (!(f == f)) return 0;
    float c = f < -1.0f ? -1.0f : (f > 1.0f ? 1.0f : f);
    return (int8_t)rintf(c * 127.0f);
}
static inline float from_unorm8(uint8_t v) { return (float)v / 255.0f; }
static inline float from_snorm8(int8_t v) { float f = (float)v / 127.0f; return f < -1.0f ? -1.0f : f; }

}  // namespace orc
