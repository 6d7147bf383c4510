// vec.h -- small float3 math + constants shared by every stage (host+device so the
// host simulator in tests/hostsim can run the same per-thread bodies on the CPU).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <float.h>

#define HD __host__ __device__ __forceinline__

namespace mtsb {

// include/mitsuba/core/math.h:36-38
constexpr float kEpsilon       = 5.9604644775390625e-8f;        // 2^-24
constexpr float kRayEpsilon    = kEpsilon * 1500.f;
constexpr float kShadowEpsilon = kRayEpsilon * 10.f;
constexpr float kPi       = 3.14159265358979323846f;
constexpr float kInvPi    = 0.31830988618379067154f;
constexpr float kInvTwoPi = 0.15915494309189533577f;
constexpr float kInvSqrtPi = 0.56418958354775628695f;

HD float3 mk3(float x, float y, float z) { return make_float3(x, y, z); }
HD float3 mk3(float s) { return make_float3(s, s, s); }
HD float3 operator+(float3 a, float3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
HD float3 operator-(float3 a, float3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
HD float3 operator*(float3 a, float3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
HD float3 operator/(float3 a, float3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
HD float3 operator*(float3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
HD float3 operator*(float s, float3 a) { return mk3(a.x * s, a.y * s, a.z * s); }
HD float3 operator/(float3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
HD float3 operator-(float3 a) { return mk3(-a.x, -a.y, -a.z); }
HD float3 &operator+=(float3 &a, float3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
HD float3 &operator*=(float3 &a, float3 b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; return a; }
HD float3 &operator*=(float3 &a, float s) { a.x *= s; a.y *= s; a.z *= s; return a; }
HD float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
HD float3 cross(float3 a, float3 b) {
    return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
HD float sqr(float x) { return x * x; }
HD float norm(float3 a) { return sqrtf(dot(a, a)); }
HD float3 normalize(float3 a) { return a * (1.f / norm(a)); }
HD float hmax(float3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
HD float3 fmin3(float3 a, float3 b) { return mk3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
HD float3 fmax3(float3 a, float3 b) { return mk3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
HD float3 abs3(float3 a) { return mk3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
HD float safe_sqrt(float x) { return sqrtf(fmaxf(x, 0.f)); }
HD float safe_acos(float x) { return acosf(fminf(1.f, fmaxf(-1.f, x))); }
HD float lerpf(float a, float b, float t) { return fmaf(b, t, fmaf(-a, t, a)); }   // enoki lerp: fmadd(b,t,fnmadd(a,t,a))
HD float3 xyz(float4 a) { return mk3(a.x, a.y, a.z); }
HD float4 mk4(float3 a, float w) { return make_float4(a.x, a.y, a.z, w); }

HD uint32_t f2u(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; return c.u;
#endif
}
HD float u2f(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
// enoki mulsign(a, b) = a with sign flipped if b negative; mulsign_neg flips if b positive
HD float mulsign(float a, float b) { return u2f(f2u(a) ^ (f2u(b) & 0x80000000u)); }
HD float mulsign_neg(float a, float b) { return u2f(f2u(a) ^ ((~f2u(b)) & 0x80000000u)); }

// include/mitsuba/core/vector.h:116-136 (Duff et al.)
HD void coordinate_system(float3 n, float3 &s, float3 &t) {
    float sign = copysignf(1.f, n.z), a = -1.f / (sign + n.z), b = n.x * n.y * a;
    s = mk3(mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    t = mk3(b, sign + sqr(n.y) * a, -n.y);
}

struct Frame {
    float3 s, t, n;
    HD float3 to_local(float3 v) const { return mk3(dot(v, s), dot(v, t), dot(v, n)); }
    HD float3 to_world(float3 v) const { return s * v.x + t * v.y + n * v.z; }
};

// row-major 4x4 helpers
HD float3 xform_point_affine(const float *m, float3 p) {
    return mk3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3],
               m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
               m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
HD float3 xform_point_persp(const float *m, float3 p) {
    float w = m[12] * p.x + m[13] * p.y + m[14] * p.z + m[15];
    return xform_point_affine(m, p) * (1.f / w);
}
HD float3 xform_vec(const float *m, float3 v) {
    return mk3(m[0] * v.x + m[1] * v.y + m[2] * v.z,
               m[4] * v.x + m[5] * v.y + m[6] * v.z,
               m[8] * v.x + m[9] * v.y + m[10] * v.z);
}

// ---- RNG: PCG32 (ext/enoki/include/enoki/random.h:62-101), TEA (include/mitsuba/core/random.h:94-113)
struct Pcg32 {
    uint64_t state, inc;
    HD void seed(uint64_t initstate, uint64_t initseq) {
        state = 0; inc = (initseq << 1) | 1u;
        next_u32(); state += initstate; next_u32();
    }
    HD uint32_t next_u32() {
        uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        uint32_t xs = (uint32_t) (((old >> 18) ^ old) >> 27);
        uint32_t rot = (uint32_t) (old >> 59);
        return (xs >> rot) | (xs << ((32 - rot) & 31));
    }
    HD float next_f32() { return u2f((next_u32() >> 9) | 0x3f800000u) - 1.f; }
};
HD uint64_t tea64(uint32_t v0, uint32_t v1) {
    uint32_t sum = 0;
    for (int i = 0; i < 4; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return (uint64_t) v0 + ((uint64_t) v1 << 32);
}
// IndependentSampler::seed, wavefront branch (src/samplers/independent.cpp:61-70): stream idx of a wavefront
// seeded with scalar `seed`: seed_value = seed + idx, rng.seed(tea64(seed_value, idx), tea64(idx, seed_value)).
HD Pcg32 sampler_seed(uint64_t seed, uint32_t idx) {
    uint32_t sv = (uint32_t) (seed + idx);
    Pcg32 r; r.seed(tea64(sv, idx), tea64(idx, sv));
    return r;
}

} // namespace mtsb
