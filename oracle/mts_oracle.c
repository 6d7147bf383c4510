/* mts_oracle.c -- CPU oracle (TEST INFRASTRUCTURE ONLY; see mts_oracle.h). Plain C99 + OpenMP. */
#include "mts_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <float.h>
#include "../mitsuba2_b200/csrc/cie_tables.h"   /* published CIE 1931 / D65 data tables (data only) */

typedef struct { float x, y, z; } v3;
static inline v3 V(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vmul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 vscale(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline float vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 vcross(v3 a, v3 b) { return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline float vnorm(v3 a) { return sqrtf(vdot(a, a)); }
static inline v3 vnormalize(v3 a) { return vscale(a, 1.f / vnorm(a)); }
static inline float vhmax(v3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
static inline v3 vabs(v3 a) { return V(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
static inline float sqrf(float x) { return x * x; }
static inline float safe_sqrtf(float x) { return sqrtf(fmaxf(x, 0.f)); }
static inline float safe_acosf(float x) { return acosf(fminf(1.f, fmaxf(-1.f, x))); }
static inline float lerp_(float a, float b, float t) { return fmaf(b, t, fmaf(-a, t, a)); }
static inline v3 ld3(const float *a, size_t i) { return V(a[3 * i], a[3 * i + 1], a[3 * i + 2]); }
static inline uint32_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float bitsf(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline float mulsign(float a, float b) { return bitsf(fbits(a) ^ (fbits(b) & 0x80000000u)); }
static inline float mulsign_neg(float a, float b) { return bitsf(fbits(a) ^ ((~fbits(b)) & 0x80000000u)); }
/* Spectrum: Color<float,3> in rgb mode, 4 wavelengths in spectral mode (include/mitsuba/core/spectrum.h); the 4th
 * channel is carried along unused in rgb mode, reductions take the channel count */
typedef struct { float c[4]; } col;
static inline col C1(float v) { col r = { { v, v, v, v } }; return r; }
static inline col C3(v3 a) { col r = { { a.x, a.y, a.z, 0.f } }; return r; }
static inline col cmul(col a, col b) { col r; for (int i = 0; i < 4; ++i) r.c[i] = a.c[i] * b.c[i]; return r; }
static inline col cadd(col a, col b) { col r; for (int i = 0; i < 4; ++i) r.c[i] = a.c[i] + b.c[i]; return r; }
static inline col cscale(col a, float s) { col r; for (int i = 0; i < 4; ++i) r.c[i] = a.c[i] * s; return r; }
static inline float chmax(col a, int n) { float m = a.c[0]; for (int i = 1; i < n; ++i) m = fmaxf(m, a.c[i]); return m; }
static inline int cblack(col a, int n) { for (int i = 0; i < n; ++i) if (a.c[i] != 0.f) return 0; return 1; }

/* include/mitsuba/core/math.h:36-38 */
#define EPS_F (FLT_EPSILON / 2.f)
#define RAY_EPS (EPS_F * 1500.f)
#define SHADOW_EPS (RAY_EPS * 10.f)
#define PI_F 3.14159265358979323846f
#define INV_PI_F 0.31830988618379067154f
#define INV_TWOPI_F 0.15915494309189533577f
#define INV_SQRT_PI_F 0.56418958354775628695f

/* ---------------- PCG32: ext/enoki/include/enoki/random.h:62-101 ------------------------------- */
typedef struct { uint64_t state, inc; } pcg32;
static inline uint32_t pcg_next(pcg32 *r) {
    uint64_t old = r->state;
    r->state = old * 0x5851f42d4c957f2dULL + r->inc;
    uint32_t xs = (uint32_t) (((old >> 18u) ^ old) >> 27u), rot = (uint32_t) (old >> 59u);
    return (xs >> rot) | (xs << ((~rot + 1u) & 31));
}
static inline void pcg_seed(pcg32 *r, uint64_t initstate, uint64_t initseq) {
    r->state = 0; r->inc = (initseq << 1u) | 1u; pcg_next(r); r->state += initstate; pcg_next(r);
}
static inline float pcg_float(pcg32 *r) { return bitsf((pcg_next(r) >> 9) | 0x3f800000u) - 1.f; }

void orc_pcg32_floats(uint64_t n, const uint64_t *initstate, const uint64_t *initseq, uint32_t n_draws, float *out) {
    for (uint64_t i = 0; i < n; ++i) {
        pcg32 r; pcg_seed(&r, initstate[i], initseq[i]);
        for (uint32_t k = 0; k < n_draws; ++k) out[i * n_draws + k] = pcg_float(&r);
    }
}
/* include/mitsuba/core/random.h:73-113 */
uint32_t orc_tea32(uint32_t v0, uint32_t v1, int rounds) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return v1;
}
uint64_t orc_tea64(uint32_t v0, uint32_t v1, int rounds) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return (uint64_t) v0 + ((uint64_t) v1 << 32);
}
/* src/samplers/independent.cpp:61-70, wavefront branch (32-bit TEA inputs) */
static inline void sampler_seed(pcg32 *r, uint64_t seed, uint32_t idx) {
    uint32_t sv = (uint32_t) (seed + idx);
    pcg_seed(r, orc_tea64(sv, idx, 4), orc_tea64(idx, sv, 4));
}
void orc_sampler_seed(uint64_t n, uint64_t seed, uint64_t *state, uint64_t *inc) {
    for (uint64_t i = 0; i < n; ++i) { pcg32 r; sampler_seed(&r, seed, (uint32_t) i); state[i] = r.state; inc[i] = r.inc; }
}

/* ---------------- warps: include/mitsuba/core/warp.h -------------------------------------------- */
static void disk_concentric(float sx, float sy, float *ox, float *oy) {      /* :54-90 */
    float x = fmaf(2.f, sx, -1.f), y = fmaf(2.f, sy, -1.f);
    int is_zero = (x == 0.f && y == 0.f), q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = .25f * PI_F * rp / r;
    if (q13) phi = .5f * PI_F - phi;
    if (is_zero) phi = 0.f;
    *ox = r * cosf(phi); *oy = r * sinf(phi);
}
static v3 cosine_hemisphere(float sx, float sy) {                             /* :332-340 */
    float px, py; disk_concentric(sx, sy, &px, &py);
    return V(px, py, safe_sqrtf(1.f - (px * px + py * py)));
}
void orc_square_to_uniform_disk_concentric(float sx, float sy, float *o) { disk_concentric(sx, sy, o, o + 1); }
void orc_square_to_cosine_hemisphere(float sx, float sy, float *o) { v3 v = cosine_hemisphere(sx, sy); o[0] = v.x; o[1] = v.y; o[2] = v.z; }
void orc_square_to_uniform_triangle(float sx, float sy, float *o) { float t = safe_sqrtf(1.f - sx); o[0] = 1.f - t; o[1] = t * sy; } /* :153-156 */
static float interval_to_linear(float v0, float v1, float s) {                /* :367-373 */
    return fabsf(v0 - v1) > 1e-4f * (v0 + v1) ? (v0 - safe_sqrtf(lerp_(sqrf(v0), sqrf(v1), s))) / (v0 - v1) : s;
}
/* include/mitsuba/core/vector.h:116-136 */
static void coordinate_system(v3 n, v3 *s, v3 *t) {
    float sign = copysignf(1.f, n.z), a = -1.f / (sign + n.z), b = n.x * n.y * a;
    *s = V(mulsign(sqrf(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    *t = V(b, sign + sqrf(n.y) * a, -n.y);
}
void orc_coordinate_system(const float *n, float *s, float *t) {
    v3 a, b; coordinate_system(V(n[0], n[1], n[2]), &a, &b);
    s[0] = a.x; s[1] = a.y; s[2] = a.z; t[0] = b.x; t[1] = b.y; t[2] = b.z;
}

/* ---------------- Fresnel: include/mitsuba/render/fresnel.h:34-116 ------------------------------- */
static void fresnel(float cos_i, float eta, float *r, float *cos_t, float *eta_it, float *eta_ti) {
    int outside = cos_i >= 0.f;
    float rcp_eta = 1.f / eta;
    *eta_it = outside ? eta : rcp_eta; *eta_ti = outside ? rcp_eta : eta;
    float cos_t_sqr = fmaf(-fmaf(-cos_i, cos_i, 1.f), *eta_ti * *eta_ti, 1.f);
    float ci = fabsf(cos_i), ct = safe_sqrtf(cos_t_sqr);
    int index_matched = eta == 1.f, special = index_matched || ci == 0.f;
    float a_s = fmaf(-*eta_it, ct, ci) / fmaf(*eta_it, ct, ci);
    float a_p = fmaf(-*eta_it, ci, ct) / fmaf(*eta_it, ci, ct);
    *r = .5f * (a_s * a_s + a_p * a_p);
    if (special) *r = index_matched ? 0.f : 1.f;
    *cos_t = mulsign_neg(ct, cos_i);
}
void orc_fresnel(float c, float eta, float *o) { fresnel(c, eta, o, o + 1, o + 2, o + 3); }
float orc_fresnel_conductor(float cos_i, float eta_r, float eta_i) {
    float c2 = cos_i * cos_i, s2 = 1.f - c2, s4 = s2 * s2;
    float temp_1 = eta_r * eta_r - eta_i * eta_i - s2;
    float a_2_pb_2 = safe_sqrtf(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r);
    float a = safe_sqrtf(.5f * (a_2_pb_2 + temp_1));
    float term_1 = a_2_pb_2 + c2, term_2 = 2.f * cos_i * a;
    float r_s = (term_1 - term_2) / (term_1 + term_2);
    float term_3 = a_2_pb_2 * c2 + s4, term_4 = term_2 * s2;
    float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return .5f * (r_s + r_p);
}

/* ---------------- Microfacet: include/mitsuba/render/microfacet.h:184-413 ------------------------ */
float orc_erfinv(float x) {   /* ext/enoki/include/enoki/special.h:222-247 (Giles) */
    float w = -logf((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = sqrtf(w) - 3.f;
    static const float c1[9] = { 1.50140941f, 0.246640727f, -0.00417768164f, -0.00125372503f, 0.00021858087f, -4.39150654e-06f, -3.5233877e-06f, 3.43273939e-07f, 2.81022636e-08f };
    static const float c2[9] = { 2.83297682f, 1.00167406f, 0.00943887047f, -0.0076224613f, 0.00573950773f, -0.00367342844f, 0.00134934322f, 0.000100950558f, -0.000200214257f };
    float p1 = c1[8], p2 = c2[8];
    for (int i = 7; i >= 0; --i) { p1 = fmaf(p1, w1, c1[i]); p2 = fmaf(p2, w2, c2[i]); }
    return (w < 5.f ? p1 : p2) * x;
}
typedef struct { int type; float au, av; int visible; } mf_t;
static mf_t mf_make(int type, float au, float av, int visible) { mf_t m = { type, fmaxf(au, 1e-4f), fmaxf(av, 1e-4f), visible }; return m; }
static float mf_eval(const mf_t *d, v3 m) {
    float alpha_uv = d->au * d->av, cos_theta = m.z, cos_theta_2 = cos_theta * cos_theta, result;
    if (d->type == 0)
        result = expf(-(sqrf(m.x / d->au) + sqrf(m.y / d->av)) / cos_theta_2) / (PI_F * alpha_uv * sqrf(cos_theta_2));
    else
        result = 1.f / (PI_F * alpha_uv * sqrf(sqrf(m.x / d->au) + sqrf(m.y / d->av) + sqrf(m.z)));
    return result * cos_theta > 1e-20f ? result : 0.f;
}
static float mf_g1(const mf_t *d, v3 v, v3 m) {
    float xy_alpha_2 = sqrf(d->au * v.x) + sqrf(d->av * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqrf(v.z), result;
    if (d->type == 0) {
        float a = 1.f / sqrtf(tan_theta_alpha_2), a_sqr = a * a;
        result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
    } else
        result = 2.f / (1.f + sqrtf(1.f + tan_theta_alpha_2));
    if (xy_alpha_2 == 0.f) result = 1.f;
    if (vdot(v, m) * v.z <= 0.f) result = 0.f;
    return result;
}
static float mf_pdf(const mf_t *d, v3 wi, v3 m) {
    float result = mf_eval(d, m);
    if (d->visible) result *= mf_g1(d, wi, m) * fabsf(vdot(wi, m)) / wi.z; else result *= m.z;
    return result;
}
static void mf_sample_visible_11(const mf_t *d, float cos_theta_i, float sx, float sy, float *ox, float *oy) {
    if (d->type == 0) {
        float tan_theta_i = safe_sqrtf(fmaf(-cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i, cot_theta_i = 1.f / tan_theta_i;
        float maxval = erff(cot_theta_i);
        sx = fmaxf(fminf(sx, 1.f - 1e-6f), 1e-6f); sy = fmaxf(fminf(sy, 1.f - 1e-6f), 1e-6f);
        float x = maxval - (maxval + 1.f) * erff(sqrtf(-logf(sx)));
        sx *= 1.f + maxval + INV_SQRT_PI_F * tan_theta_i * expf(-sqrf(cot_theta_i));
        for (int i = 0; i < 3; ++i) {
            float slope = orc_erfinv(x), value = 1.f + x + INV_SQRT_PI_F * tan_theta_i * expf(-sqrf(slope)) - sx,
                  derivative = 1.f - slope * tan_theta_i;
            x -= value / derivative;
        }
        *ox = orc_erfinv(x); *oy = orc_erfinv(fmaf(2.f, sy, -1.f));
    } else {
        float px, py; disk_concentric(sx, sy, &px, &py);
        float s = .5f * (1.f + cos_theta_i);
        py = lerp_(safe_sqrtf(1.f - sqrf(px)), py, s);
        float x = px, y = py, z = safe_sqrtf(1.f - (px * px + py * py));
        float sin_theta_i = safe_sqrtf(1.f - sqrf(cos_theta_i));
        float nrm = 1.f / fmaf(sin_theta_i, y, cos_theta_i * z);
        *ox = fmaf(cos_theta_i, y, -(sin_theta_i * z)) * nrm; *oy = x * nrm;
    }
}
static v3 mf_sample(const mf_t *d, v3 wi, float sx, float sy, float *pdf) {
    if (!d->visible) {
        float sin_phi, cos_phi, cos_theta, cos_theta_2, alpha_2;
        if (d->au == d->av) { sin_phi = sinf(2.f * PI_F * sy); cos_phi = cosf(2.f * PI_F * sy); alpha_2 = d->au * d->au; }
        else {
            float ratio = d->av / d->au, tmp = ratio * tanf(2.f * PI_F * sy);
            cos_phi = 1.f / sqrtf(fmaf(tmp, tmp, 1.f));
            cos_phi = mulsign(cos_phi, fabsf(sy - .5f) - .25f);
            sin_phi = cos_phi * tmp;
            alpha_2 = 1.f / (sqrf(cos_phi / d->au) + sqrf(sin_phi / d->av));
        }
        if (d->type == 0) {
            cos_theta = 1.f / sqrtf(fmaf(-alpha_2, logf(1.f - sx), 1.f)); cos_theta_2 = sqrf(cos_theta);
            float cos_theta_3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
            *pdf = (1.f - sx) / (PI_F * d->au * d->av * cos_theta_3);
        } else {
            float tan_theta_m_2 = alpha_2 * sx / (1.f - sx);
            cos_theta = 1.f / sqrtf(1.f + tan_theta_m_2); cos_theta_2 = sqrf(cos_theta);
            float temp = 1.f + tan_theta_m_2 / alpha_2, cos_theta_3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
            *pdf = 1.f / (PI_F * d->au * d->av * cos_theta_3 * sqrf(temp));
        }
        float sin_theta = sqrtf(1.f - cos_theta_2);
        return V(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
    }
    v3 wi_p = vnormalize(V(d->au * wi.x, d->av * wi.y, wi.z));
    /* Frame::sincos_phi, include/mitsuba/core/frame.h:107-118 */
    float sin_theta_2 = fmaf(-wi_p.z, wi_p.z, 1.f), inv_sin_theta = 1.f / sqrtf(sin_theta_2);
    float cos_phi = fminf(fmaxf(wi_p.x * inv_sin_theta, -1.f), 1.f), sin_phi = fminf(fmaxf(wi_p.y * inv_sin_theta, -1.f), 1.f);
    if (fabsf(sin_theta_2) <= 4.f * EPS_F) { cos_phi = 1.f; sin_phi = 0.f; }
    float slx, sly; mf_sample_visible_11(d, wi_p.z, sx, sy, &slx, &sly);
    float rx = fmaf(cos_phi, slx, -(sin_phi * sly)) * d->au, ry = fmaf(sin_phi, slx, cos_phi * sly) * d->av;
    v3 m = vnormalize(V(-rx, -ry, 1.f));
    *pdf = mf_eval(d, m) * mf_g1(d, wi, m) * fabsf(vdot(wi, m)) / wi.z;
    return m;
}
float orc_microfacet_eval(int type, float au, float av, const float *m) { mf_t d = mf_make(type, au, av, 1); return mf_eval(&d, V(m[0], m[1], m[2])); }
float orc_microfacet_pdf(int type, float au, float av, int vis, const float *wi, const float *m) { mf_t d = mf_make(type, au, av, vis); return mf_pdf(&d, V(wi[0], wi[1], wi[2]), V(m[0], m[1], m[2])); }
float orc_microfacet_smith_g1(int type, float au, float av, const float *v, const float *m) { mf_t d = mf_make(type, au, av, 1); return mf_g1(&d, V(v[0], v[1], v[2]), V(m[0], m[1], m[2])); }
void orc_microfacet_sample(int type, float au, float av, int vis, const float *wi, float sx, float sy, float *o) {
    mf_t d = mf_make(type, au, av, vis); float pdf; v3 m = mf_sample(&d, V(wi[0], wi[1], wi[2]), sx, sy, &pdf);
    o[0] = m.x; o[1] = m.y; o[2] = m.z; o[3] = pdf;
}

/* ---------------- reconstruction filters: src/rfilters/{box,tent,gaussian,mitchell,catmullrom,lanczos}.cpp ---- */
static float cubic_bc(float x, float B, float C) {                                 /* mitchell.cpp:40-53 = catmullrom.cpp:29-43 */
    x = fabsf(x); float x2 = x * x, x3 = x2 * x;
    float r = (1.f / 6.f) * (x < 1.f ? (12.f - 9.f * B - 6.f * C) * x3 + (-18.f + 12.f * B + 6.f * C) * x2 + (6.f - 2.f * B)
                                     : (-B - 6.f * C) * x3 + (6.f * B + 30.f * C) * x2 + (-12.f * B - 48.f * C) * x + (8.f * B + 24.f * C));
    return x < 2.f ? r : 0.f;
}
float orc_rfilter_eval2(int type, float radius, float stddev, float c, float x) {
    if (type == MTSB_RFILTER_MITCHELL) return cubic_bc(x, stddev, c);
    if (type == MTSB_RFILTER_CATMULLROM) return cubic_bc(x, 0.f, 0.5f);
    if (type == MTSB_RFILTER_LANCZOS) {                                            /* lanczos.cpp:44-54 */
        x = fabsf(x); float x1 = PI_F * x, x2 = x1 / radius;
        return x < EPS_F ? 1.f : (x > radius ? 0.f : (sinf(x1) * sinf(x2)) / (x1 * x2));
    }
    return orc_rfilter_eval(type, radius, stddev, x);
}
float orc_rfilter_eval(int type, float radius, float stddev, float x) {
    if (type == MTSB_RFILTER_BOX) return fabsf(x) <= radius ? 1.f : 0.f;           /* radius includes +RayEpsilon, box.cpp:30 */
    if (type == MTSB_RFILTER_TENT) return fmaxf(0.f, 1.f - fabsf(x / radius));
    float alpha = -1.f / (2.f * stddev * stddev);
    return fmaxf(0.f, expf(alpha * x * x) - expf(alpha * radius * radius));
}

/* ---------------- PerspectiveCamera::update_camera_transforms: src/sensors/perspective.cpp:106-133 (no crop) */
static void mat4_mul(const double *a, const double *b, double *c) {
    double r[16];
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { double s = 0; for (int k = 0; k < 4; ++k) s += a[4 * i + k] * b[4 * k + j]; r[4 * i + j] = s; }
    memcpy(c, r, sizeof(r));
}
void orc_perspective_sample_to_camera(float fov, float near_, float far_, float aspect, float *out16) {
    /* inverse of: scale(-0.5, -0.5*aspect, 1) * translate(-1, -1/aspect, 0) * perspective(fov, near, far);
       Transform::perspective inverse from include/mitsuba/core/transform.h:203-220 */
    float tanv = tanf(fov * .5f * (PI_F / 180.f));
    double inv_persp[16] = { tanv, 0, 0, 0,  0, tanv, 0, 0,  0, 0, 0, 1,  0, 0, (near_ - far_) / (far_ * near_), 1.0 / near_ };
    double inv_trans[16] = { 1, 0, 0, 1,  0, 1, 0, 1.0 / aspect,  0, 0, 1, 0,  0, 0, 0, 1 };
    double inv_scale[16] = { -2, 0, 0, 0,  0, -2.0 / aspect, 0, 0,  0, 0, 1, 0,  0, 0, 0, 1 };
    double t[16]; mat4_mul(inv_persp, inv_trans, t); mat4_mul(t, inv_scale, t);
    for (int i = 0; i < 16; ++i) out16[i] = (float) t[i];
}

/* =============================================================================================== */
/* Scene                                                                                            */
/* =============================================================================================== */
typedef struct { uint32_t face_begin, face_count, vert_begin, vert_count; int bsdf, emitter, has_normals, has_uv; int shape; } omesh;
/* analytic shape after the plugin's update(): sphere.cpp:112-135, cylinder.cpp:103-127, disk.cpp:81-98, rectangle.cpp:85-95 */
typedef struct { int type, flip; float to_world[16], to_object[16]; v3 center; float radius, length; v3 n, dp_du, dp_dv; float inv_area; } oshape;
typedef struct { float lo[3], hi[3]; uint32_t left, count; /* count>0: leaf [left, left+count) in order[] ; else children left,left+1 */ } onode;
typedef struct { int n_levels; float **levels; uint32_t *widths; } oh2d;
typedef struct {
    int type; v3 radiance, position; int radiance_tex; int mesh; float *cdf; float cdf_sum, normalization; uint32_t valid_lo, valid_hi;
    uint32_t w, h; float *rgba; float to_world[16], to_local[16]; float scale; oh2d warp; float bsphere_radius;
} oemitter;
typedef struct { int type; v3 value; uint32_t w, h, ch; float *data; float to_uv[9]; float lambda_min, lambda_max, d65_scale; } otex;
struct orc_scene {
    uint32_t n_meshes, n_bsdfs, n_tex, n_emitters, n_sensors, n_faces, n_verts;
    omesh *meshes; mtsb_bsdf_desc *bsdfs; otex *tex; oemitter *emitters; mtsb_sensor_desc *sensors;
    float *pos, *nrm, *uv; uint32_t *faces /* 3 per face, global vertex ids */; uint32_t *face_mesh;
    onode *nodes; uint32_t n_nodes; uint32_t *order;
    uint32_t n_shapes; oshape *shapes;   /* primitive n_faces + k, meshes[n_meshes + k] */
    int env_emitter; int max_depth, rr_depth;
    int spectral, nch;        /* MTSB_MODE_SPECTRAL: 4 wavelengths per path */
    v3 bbox_lo, bbox_hi;
    float *rough_tab;         /* per BSDF: 64 external transmittances + mean internal reflectance (roughplastic.cpp) */
};
static void rough_plastic_tables(const mtsb_bsdf_desc *b, float *tab);
static float rough_lerp(const float *data, float x);

/* ---- a plain median-split BVH so the oracle finishes in seconds on 1M triangles; hit semantics are those of
 *      ShapeKDTree::ray_intersect_scalar (kdtree.h:2080-2175): closest t in [mint,maxt], test `t <= maxt`. */
static void tri_bounds(const orc_scene *s, uint32_t f, float *lo, float *hi) {
    for (int a = 0; a < 3; ++a) { lo[a] = INFINITY; hi[a] = -INFINITY; }
    for (int k = 0; k < 3; ++k) { const float *p = s->pos + 3 * (size_t) s->faces[3 * f + k];
        for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], p[a]); hi[a] = fmaxf(hi[a], p[a]); } }
}
static void shape_bounds(const oshape *sh, float *lo, float *hi);
static float *g_centroid; static int g_axis;
static int cmp_centroid(const void *a, const void *b) {
    float ca = g_centroid[3 * (size_t) *(const uint32_t *) a + g_axis], cb = g_centroid[3 * (size_t) *(const uint32_t *) b + g_axis];
    return ca < cb ? -1 : (ca > cb ? 1 : 0);
}
static void build_rec(orc_scene *s, uint32_t node, uint32_t begin, uint32_t end, const float *blo, const float *bhi) {
    onode *n = &s->nodes[node];
    for (int a = 0; a < 3; ++a) { n->lo[a] = INFINITY; n->hi[a] = -INFINITY; }
    for (uint32_t i = begin; i < end; ++i) for (int a = 0; a < 3; ++a) {
        n->lo[a] = fminf(n->lo[a], blo[3 * (size_t) s->order[i] + a]); n->hi[a] = fmaxf(n->hi[a], bhi[3 * (size_t) s->order[i] + a]); }
    if (end - begin <= 4) { n->left = begin; n->count = end - begin; return; }
    int axis = 0; float ext = -1.f;
    for (int a = 0; a < 3; ++a) if (n->hi[a] - n->lo[a] > ext) { ext = n->hi[a] - n->lo[a]; axis = a; }
    g_axis = axis;
    qsort(s->order + begin, end - begin, sizeof(uint32_t), cmp_centroid);
    uint32_t mid = (begin + end) / 2;
    uint32_t l = s->n_nodes; s->n_nodes += 2;
    n = &s->nodes[node]; n->left = l; n->count = 0;
    build_rec(s, l, begin, mid, blo, bhi); build_rec(s, l + 1, mid, end, blo, bhi);
}
static void build_accel(orc_scene *s) {
    free(s->nodes); free(s->order); s->nodes = NULL; s->order = NULL; s->n_nodes = 0;
    s->bbox_lo = V(INFINITY, INFINITY, INFINITY); s->bbox_hi = V(-INFINITY, -INFINITY, -INFINITY);
    for (uint32_t k = 0; k < s->n_shapes; ++k) { float lo[3], hi[3]; shape_bounds(&s->shapes[k], lo, hi);
        s->bbox_lo = V(fminf(s->bbox_lo.x, lo[0]), fminf(s->bbox_lo.y, lo[1]), fminf(s->bbox_lo.z, lo[2]));
        s->bbox_hi = V(fmaxf(s->bbox_hi.x, hi[0]), fmaxf(s->bbox_hi.y, hi[1]), fmaxf(s->bbox_hi.z, hi[2])); }
    if (s->n_faces == 0) return;
    float *blo = malloc(sizeof(float) * 3 * (size_t) s->n_faces), *bhi = malloc(sizeof(float) * 3 * (size_t) s->n_faces);
    g_centroid = malloc(sizeof(float) * 3 * (size_t) s->n_faces);
    s->order = malloc(sizeof(uint32_t) * s->n_faces);
    for (uint32_t f = 0; f < s->n_faces; ++f) {
        tri_bounds(s, f, blo + 3 * (size_t) f, bhi + 3 * (size_t) f);
        for (int a = 0; a < 3; ++a) g_centroid[3 * (size_t) f + a] = .5f * (blo[3 * (size_t) f + a] + bhi[3 * (size_t) f + a]);
        s->order[f] = f;
    }
    s->nodes = malloc(sizeof(onode) * (2 * (size_t) s->n_faces + 1)); s->n_nodes = 1;
    build_rec(s, 0, 0, s->n_faces, blo, bhi);
    s->bbox_lo = V(fminf(s->bbox_lo.x, s->nodes[0].lo[0]), fminf(s->bbox_lo.y, s->nodes[0].lo[1]), fminf(s->bbox_lo.z, s->nodes[0].lo[2]));
    s->bbox_hi = V(fmaxf(s->bbox_hi.x, s->nodes[0].hi[0]), fmaxf(s->bbox_hi.y, s->nodes[0].hi[1]), fmaxf(s->bbox_hi.z, s->nodes[0].hi[2]));
    free(blo); free(bhi); free(g_centroid); g_centroid = NULL;
}

/* ---- analytic shapes --------------------------------------------------------------------------------------------------- */
static int invert4(const float *m, float *inv);
static v3 xf_point(const float *m, v3 p) { return V(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7], m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]); }
static v3 xf_dir(const float *m, v3 v) { return V(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z, m[8] * v.x + m[9] * v.y + m[10] * v.z); }
static void m3_inv_transpose(double a[3][3], double r[3][3]) {
    double det = a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) - a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) + a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) {
        int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        r[i][j] = (a[i1][j1] * a[i2][j2] - a[i1][j2] * a[i2][j1]) / det;      /* cofactor / det = (A^-1)^T */
    }
}
/* enoki transform_decompose: polar decomposition A = Q P (scaled Newton, 10 steps), det(Q) > 0 */
static void polar(double a[3][3], double q[3][3], double p[3][3]) {
    memcpy(q, a, sizeof(double) * 9);
    for (int it = 0; it < 10; ++it) {
        double qi[3][3]; m3_inv_transpose(q, qi);
        double nq = 0, ni = 0;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { nq += q[i][j] * q[i][j]; ni += qi[i][j] * qi[i][j]; }
        double g = sqrt(sqrt(ni) / sqrt(nq));
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) q[i][j] = .5 * g * q[i][j] + .5 / g * qi[i][j];
    }
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { p[i][j] = 0; for (int k = 0; k < 3; ++k) p[i][j] += q[k][i] * a[k][j]; }
    double det = q[0][0] * (q[1][1] * q[2][2] - q[1][2] * q[2][1]) - q[0][1] * (q[1][0] * q[2][2] - q[1][2] * q[2][0]) + q[0][2] * (q[1][0] * q[2][1] - q[1][1] * q[2][0]);
    if (det < 0) for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { q[i][j] = -q[i][j]; p[i][j] = -p[i][j]; }
}
static int shape_setup(const mtsb_shape_desc *d, oshape *sh) {
    memset(sh, 0, sizeof(*sh));
    sh->type = d->type; sh->flip = d->flip_normals != 0;
    float w[16]; memcpy(w, d->to_world, sizeof(w));
    if (d->type == MTSB_SHAPE_SPHERE || d->type == MTSB_SHAPE_CYLINDER) {
        double a[3][3], q[3][3], p[3][3];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) a[i][j] = w[4 * i + j];
        polar(a, q, p);
        double radius = p[0][0];
        if (radius <= 0) { radius = -radius; sh->flip = !sh->flip; }
        sh->radius = (float) radius; sh->length = (float) p[2][2];
        sh->center = V(w[3], w[7], w[11]);
        double k = d->type == MTSB_SHAPE_SPHERE ? (double) sh->radius : 1.0;      /* transform_compose(S, Q, T) */
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) w[4 * i + j] = (float) (q[i][j] * k);
        w[12] = w[13] = w[14] = 0.f; w[15] = 1.f;
        sh->inv_area = d->type == MTSB_SHAPE_SPHERE ? 1.f / (4.f * PI_F * sh->radius * sh->radius) : 1.f / (2.f * PI_F * sh->radius * sh->length);
    } else if (d->type == MTSB_SHAPE_DISK || d->type == MTSB_SHAPE_RECTANGLE) {
        if (d->flip_normals) for (int i = 0; i < 4; ++i) w[4 * i + 2] = -w[4 * i + 2];
        sh->flip = 0;
    } else return 0;
    memcpy(sh->to_world, w, sizeof(w));
    if (!invert4(sh->to_world, sh->to_object)) return 0;
    if (d->type == MTSB_SHAPE_DISK || d->type == MTSB_SHAPE_RECTANGLE) {
        float k = d->type == MTSB_SHAPE_RECTANGLE ? 2.f : 1.f;
        sh->dp_du = xf_dir(sh->to_world, V(k, 0.f, 0.f)); sh->dp_dv = xf_dir(sh->to_world, V(0.f, k, 0.f));
        sh->n = vnormalize(V(sh->to_object[8], sh->to_object[9], sh->to_object[10]));     /* normals transform by the inverse transpose */
        if (d->type == MTSB_SHAPE_RECTANGLE) sh->inv_area = 1.f / vnorm(vcross(sh->dp_du, sh->dp_dv));
        else {
            float du = vnorm(sh->dp_du), dv = vnorm(sh->dp_dv);
            v3 fs = vscale(sh->dp_du, 1.f / du), ft = vscale(sh->dp_dv, 1.f / dv);
            float h = sqrtf(sqrf(dv) - sqrf(vdot(vscale(ft, dv), fs)));             /* disk.cpp:112-116 */
            sh->inv_area = 1.f / (PI_F * du * h);
        }
    }
    return 1;
}
static void shape_bounds(const oshape *sh, float *lo, float *hi) {
    v3 a = V(INFINITY, INFINITY, INFINITY), b = V(-INFINITY, -INFINITY, -INFINITY);
    if (sh->type == MTSB_SHAPE_SPHERE) { a = vsub(sh->center, V(sh->radius, sh->radius, sh->radius)); b = vadd(sh->center, V(sh->radius, sh->radius, sh->radius)); }
    else if (sh->type == MTSB_SHAPE_CYLINDER) {
        v3 x1 = xf_dir(sh->to_world, V(sh->radius, 0, 0)), x2 = xf_dir(sh->to_world, V(0, sh->radius, 0));
        v3 x = V(sqrtf(x1.x * x1.x + x2.x * x2.x), sqrtf(x1.y * x1.y + x2.y * x2.y), sqrtf(x1.z * x1.z + x2.z * x2.z));
        v3 p0 = xf_point(sh->to_world, V(0, 0, 0)), p1 = xf_point(sh->to_world, V(0, 0, sh->length));
        a = V(fminf(p0.x - x.x, p1.x - x.x), fminf(p0.y - x.y, p1.y - x.y), fminf(p0.z - x.z, p1.z - x.z));
        b = V(fmaxf(p0.x + x.x, p1.x + x.x), fmaxf(p0.y + x.y, p1.y + x.y), fmaxf(p0.z + x.z, p1.z + x.z));
    } else for (int k = 0; k < 4; ++k) {
        v3 p = xf_point(sh->to_world, V(k & 1 ? 1.f : -1.f, k & 2 ? 1.f : -1.f, 0.f));
        a = V(fminf(a.x, p.x), fminf(a.y, p.y), fminf(a.z, p.z)); b = V(fmaxf(b.x, p.x), fmaxf(b.y, p.y), fmaxf(b.z, p.z));
    }
    lo[0] = a.x; lo[1] = a.y; lo[2] = a.z; hi[0] = b.x; hi[1] = b.y; hi[2] = b.z;
}
/* math::solve_quadratic, core/math.h:371-413, on doubles */
static int solve_quadratic(double a, double b, double c, double *x0, double *x1) {
    int linear = a == 0.0, ok = !linear || b > 0.0;
    *x0 = *x1 = -c / b;
    double disc = b * b - 4.0 * a * c;
    ok = ok && (linear || disc >= 0.0);
    if (ok && !linear) {
        double root = sqrt(disc), temp = -0.5 * (b + copysign(root, b));
        double r0 = temp / a, r1 = c / temp;
        *x0 = fmin(r0, r1); *x1 = fmax(r0, r1);
    }
    return ok;
}
/* ray_intersect of the four plugins; (cu, cv) = the cache floats (disk / rectangle: object-space x, y) */
static int shape_hit(const oshape *sh, v3 o, v3 d, float mint, float maxt, float *t, float *cu, float *cv) {
    *cu = *cv = 0.f;
    if (sh->type == MTSB_SHAPE_SPHERE) {                                        /* sphere.cpp:276-300 */
        double ox = (double) o.x - sh->center.x, oy = (double) o.y - sh->center.y, oz = (double) o.z - sh->center.z, dx = d.x, dy = d.y, dz = d.z;
        double near_t, far_t;
        int found = solve_quadratic(dx * dx + dy * dy + dz * dz, 2.0 * (ox * dx + oy * dy + oz * dz), ox * ox + oy * oy + oz * oz - (double) sh->radius * sh->radius, &near_t, &far_t);
        int out_bounds = !(near_t <= maxt && far_t >= mint), in_bounds = near_t < mint && far_t > maxt;
        if (!found || out_bounds || in_bounds) return 0;
        *t = (float) (near_t < mint ? far_t : near_t);
        return 1;
    }
    v3 lo = xf_point(sh->to_object, o), ld = xf_dir(sh->to_object, d);
    if (sh->type == MTSB_SHAPE_CYLINDER) {                                      /* cylinder.cpp:248-287 */
        double ox = lo.x, oy = lo.y, oz = lo.z, dx = ld.x, dy = ld.y, dz = ld.z, near_t, far_t, len = sh->length;
        int found = solve_quadratic(dx * dx + dy * dy, 2.0 * (dx * ox + dy * oy), ox * ox + oy * oy - (double) sh->radius * sh->radius, &near_t, &far_t);
        int out_bounds = !(near_t <= maxt && far_t >= mint), in_bounds = near_t < mint && far_t > maxt;
        double zn = oz + dz * near_t, zf = oz + dz * far_t;
        int use_near = zn >= 0 && zn <= len && near_t >= mint, use_far = zf >= 0 && zf <= len && far_t <= maxt;
        if (!found || out_bounds || in_bounds || !(use_near || use_far)) return 0;
        *t = (float) (use_near ? near_t : far_t);
        return 1;
    }
    float tt = -lo.z * (1.f / ld.z);                                            /* disk.cpp:144-165, rectangle.cpp:142-164 */
    float x = fmaf(ld.x, tt, lo.x), y = fmaf(ld.y, tt, lo.y);
    int inside = sh->type == MTSB_SHAPE_DISK ? x * x + y * y <= 1.f : (fabsf(x) <= 1.f && fabsf(y) <= 1.f);
    if (!(tt >= mint && tt <= maxt && inside)) return 0;
    *t = tt; *cu = x; *cv = y;
    return 1;
}
/* Mesh::ray_intersect_triangle, include/mitsuba/render/mesh.h:195-221 (Moeller-Trumbore) */
static int mt_intersect(const orc_scene *s, uint32_t f, v3 o, v3 d, float mint, float maxt, float *u, float *v, float *t) {
    v3 p0 = ld3(s->pos, s->faces[3 * f]), p1 = ld3(s->pos, s->faces[3 * f + 1]), p2 = ld3(s->pos, s->faces[3 * f + 2]);
    v3 e1 = vsub(p1, p0), e2 = vsub(p2, p0);
    v3 pvec = vcross(d, e2);
    float inv_det = 1.f / vdot(e1, pvec);
    v3 tvec = vsub(o, p0);
    float uu = vdot(tvec, pvec) * inv_det;
    if (!(uu >= 0.f && uu <= 1.f)) return 0;
    v3 qvec = vcross(tvec, e1);
    float vv = vdot(d, qvec) * inv_det;
    if (!(vv >= 0.f && uu + vv <= 1.f)) return 0;
    float tt = vdot(e2, qvec) * inv_det;
    if (!(tt >= mint && tt <= maxt)) return 0;
    *u = uu; *v = vv; *t = tt; return 1;
}
static int box_hit(const onode *n, v3 o, v3 id, float mint, float maxt) {
    float t0 = mint, t1 = maxt;
    float a, b;
    a = (n->lo[0] - o.x) * id.x; b = (n->hi[0] - o.x) * id.x; t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
    a = (n->lo[1] - o.y) * id.y; b = (n->hi[1] - o.y) * id.y; t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
    a = (n->lo[2] - o.z) * id.z; b = (n->hi[2] - o.z) * id.z; t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
    return t0 <= t1 * 1.00001f + 1e-30f;
}
typedef struct { float t, u, v; uint32_t prim; } ohit;
/* analytic shapes are few: tested one by one, outside the triangle BVH (primitive id n_faces + k) */
static int trace_shapes(const orc_scene *s, v3 o, v3 d, float mint, float maxt, int any, ohit *h) {
    int found = 0;
    for (uint32_t k = 0; k < s->n_shapes; ++k) { float t, cu, cv;
        if (shape_hit(&s->shapes[k], o, d, mint, maxt, &t, &cu, &cv)) {
            if (any) { h->t = t; return 1; }
            maxt = t; found = 1; h->t = t; h->u = cu; h->v = cv; h->prim = s->n_faces + k;
        } }
    return found;
}
static int trace(const orc_scene *s, v3 o, v3 d, float mint, float maxt, int any, ohit *h) {
    h->t = INFINITY; h->u = h->v = 0.f; h->prim = 0xffffffffu;
    int found = 0;
    if (s->n_shapes) { found = trace_shapes(s, o, d, mint, maxt, any, h); if (found) { if (any) return 1; maxt = h->t; } }
    if (s->n_faces == 0) return found;
    v3 id = V(1.f / d.x, 1.f / d.y, 1.f / d.z);
    uint32_t stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        const onode *n = &s->nodes[stack[--sp]];
        if (!box_hit(n, o, id, mint, maxt)) continue;
        if (n->count) {
            for (uint32_t i = n->left; i < n->left + n->count; ++i) {
                float u, v, t;
                if (mt_intersect(s, s->order[i], o, d, mint, maxt, &u, &v, &t)) {
                    if (any) { h->t = t; return 1; }
                    maxt = t; found = 1; h->t = t; h->u = u; h->v = v; h->prim = s->order[i];
                }
            }
        } else { stack[sp++] = n->left; stack[sp++] = n->left + 1; }
    }
    return found;
}
static int trace_naive(const orc_scene *s, v3 o, v3 d, float mint, float maxt, ohit *h) {   /* kdtree.h:2304-2329 */
    h->t = INFINITY; h->u = h->v = 0.f; h->prim = 0xffffffffu; int found = 0;
    for (uint32_t f = 0; f < s->n_faces; ++f) { float u, v, t;
        if (mt_intersect(s, f, o, d, mint, maxt, &u, &v, &t)) { maxt = t; found = 1; h->t = t; h->u = u; h->v = v; h->prim = f; } }
    if (s->n_shapes) found |= trace_shapes(s, o, d, mint, maxt, 0, h);
    return found;
}
void orc_intersect_closest(const orc_scene *s, uint64_t n, const float *o, const float *d, float *hits, int threads) {
    #pragma omp parallel for schedule(dynamic, 1024) num_threads(threads > 0 ? threads : 1)
    for (int64_t i = 0; i < (int64_t) n; ++i) { ohit h; trace(s, V(o[4 * i], o[4 * i + 1], o[4 * i + 2]), V(d[4 * i], d[4 * i + 1], d[4 * i + 2]), o[4 * i + 3], d[4 * i + 3], 0, &h);
        hits[4 * i] = h.t; hits[4 * i + 1] = h.u; hits[4 * i + 2] = h.v; hits[4 * i + 3] = bitsf(h.prim); }
}
void orc_intersect_any(const orc_scene *s, uint64_t n, const float *o, const float *d, uint32_t *occ, int threads) {
    #pragma omp parallel for schedule(dynamic, 1024) num_threads(threads > 0 ? threads : 1)
    for (int64_t i = 0; i < (int64_t) n; ++i) { ohit h; occ[i] = (uint32_t) trace(s, V(o[4 * i], o[4 * i + 1], o[4 * i + 2]), V(d[4 * i], d[4 * i + 1], d[4 * i + 2]), o[4 * i + 3], d[4 * i + 3], 1, &h); }
}
void orc_intersect_naive(const orc_scene *s, uint64_t n, const float *o, const float *d, float *hits, int threads) {
    #pragma omp parallel for schedule(dynamic, 64) num_threads(threads > 0 ? threads : 1)
    for (int64_t i = 0; i < (int64_t) n; ++i) { ohit h; trace_naive(s, V(o[4 * i], o[4 * i + 1], o[4 * i + 2]), V(d[4 * i], d[4 * i + 1], d[4 * i + 2]), o[4 * i + 3], d[4 * i + 3], &h);
        hits[4 * i] = h.t; hits[4 * i + 1] = h.u; hits[4 * i + 2] = h.v; hits[4 * i + 3] = bitsf(h.prim); }
}

/* closest hit + the surface interaction built from it, 16 floats per ray: t, uv, primitive id bits, p, n, sh_frame.n, wi
 * (the record oracle/ref_query.cpp writes for Scene::ray_intersect) */
void orc_surface_interaction(const orc_scene *s, uint64_t n, const float *o, const float *d, float *out16, int threads);

/* ---- tables ------------------------------------------------------------------------------------ */
static void build_area_cdf(orc_scene *s, oemitter *e) {    /* distr_1d.h:49-88 over mesh.h:108-116 */
    const omesh *m = &s->meshes[e->mesh];
    free(e->cdf); e->cdf = malloc(sizeof(float) * (m->face_count ? m->face_count : 1));
    double sum = 0.0; e->valid_lo = e->valid_hi = (uint32_t) -1;
    for (uint32_t i = 0; i < m->face_count; ++i) {
        uint32_t f = m->face_begin + i;
        v3 p0 = ld3(s->pos, s->faces[3 * f]), p1 = ld3(s->pos, s->faces[3 * f + 1]), p2 = ld3(s->pos, s->faces[3 * f + 2]);
        float area = .5f * vnorm(vcross(vsub(p1, p0), vsub(p2, p0)));
        sum += (double) area; e->cdf[i] = (float) sum;
        if (area > 0.f) { if (e->valid_lo == (uint32_t) -1) e->valid_lo = i; e->valid_hi = i; }
    }
    e->cdf_sum = (float) sum; e->normalization = (float) (1.0 / sum);
}
static uint32_t log2i_ceil(uint32_t v) { uint32_t r = 0; while ((1u << r) < v) ++r; return r; }
static uint32_t h2d_index(uint32_t x, uint32_t y, uint32_t width) { return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + ((y & ~1u) * width); }
static void build_h2d(oh2d *t, const float *data, uint32_t w, uint32_t h) {    /* distr_2d.h:206-310 */
    uint32_t npx = w - 1, npy = h - 1, max_level = log2i_ceil(npx > npy ? npx : npy);
    t->n_levels = (int) max_level + 2;
    t->levels = calloc((size_t) t->n_levels, sizeof(float *)); t->widths = calloc((size_t) t->n_levels, sizeof(uint32_t));
    size_t *sizes = calloc((size_t) t->n_levels, sizeof(size_t));
    t->levels[0] = calloc((size_t) w * h, sizeof(float)); t->widths[0] = w; sizes[0] = (size_t) w * h;
    uint32_t lx = npx, ly = npy; int li = 1;
    for (int level = (int) max_level; level >= 0; --level) {
        lx += lx & 1u; ly += ly & 1u;
        t->levels[li] = calloc((size_t) lx * ly, sizeof(float)); t->widths[li] = lx; sizes[li] = (size_t) lx * ly; li++;
        lx >>= 1; ly >>= 1;
    }
    double sum = 0.0; const float *in = data;
    for (uint32_t y = 0; y < npy; ++y) { for (uint32_t x = 0; x < npx; ++x) {
            float avg = (in[0] + in[1] + in[w] + in[w + 1]) * .25f; sum += (double) avg;
            t->levels[1][h2d_index(x, y, t->widths[1])] = avg; ++in; } ++in; }
    float scale = (float) ((double) (npx * npy) / sum);
    for (size_t i = 0; i < sizes[0]; ++i) t->levels[0][i] = data[i] * scale;
    for (size_t i = 0; i < sizes[1]; ++i) t->levels[1][i] *= scale;
    lx = npx; ly = npy;
    for (uint32_t level = 2; level <= max_level + 1; ++level) {
        lx = (lx + 1) >> 1; ly = (ly + 1) >> 1;
        for (uint32_t y = 0; y < ly; ++y) for (uint32_t x = 0; x < lx; ++x) {
            const float *d0 = t->levels[level - 1] + h2d_index(x * 2, y * 2, t->widths[level - 1]);
            t->levels[level][h2d_index(x, y, t->widths[level])] = d0[0] + d0[1] + d0[2] + d0[3]; }
    }
    free(sizes);
}
static int invert4(const float *m, float *inv) {
    double a[4][8];
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { a[i][j] = m[4 * i + j]; a[i][4 + j] = i == j; }
    for (int c = 0; c < 4; ++c) {
        int p = c; for (int r = c + 1; r < 4; ++r) if (fabs(a[r][c]) > fabs(a[p][c])) p = r;
        if (a[p][c] == 0.0) return 0;
        if (p != c) for (int j = 0; j < 8; ++j) { double t = a[c][j]; a[c][j] = a[p][j]; a[p][j] = t; }
        double iv = 1.0 / a[c][c]; for (int j = 0; j < 8; ++j) a[c][j] *= iv;
        for (int r = 0; r < 4; ++r) if (r != c) { double f = a[r][c]; for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j]; }
    }
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) inv[4 * i + j] = (float) a[i][4 + j];
    return 1;
}
static void update_bsphere(orc_scene *s) {   /* envmap.cpp:126-130 */
    v3 c = vscale(vadd(s->bbox_lo, s->bbox_hi), .5f);
    float r = (s->n_faces + s->n_shapes) ? vnorm(vsub(s->bbox_hi, c)) : 0.f;
    r = fmaxf(RAY_EPS, r * (1.f + RAY_EPS));
    for (uint32_t i = 0; i < s->n_emitters; ++i) if (s->emitters[i].type == MTSB_EMITTER_ENVMAP || s->emitters[i].type == MTSB_EMITTER_CONSTANT) s->emitters[i].bsphere_radius = r;
}

orc_scene *orc_scene_create(const mtsb_scene_desc *d) {
    orc_scene *s = calloc(1, sizeof(orc_scene));
    s->n_meshes = d->n_meshes; s->n_bsdfs = d->n_bsdfs; s->n_tex = d->n_textures; s->n_emitters = d->n_emitters; s->n_sensors = d->n_sensors;
    s->meshes = calloc(d->n_meshes + d->n_shapes + 1, sizeof(omesh));
    s->n_shapes = d->n_shapes; s->shapes = calloc(d->n_shapes + 1, sizeof(oshape));
    for (uint32_t i = 0; i < d->n_meshes; ++i) { s->n_faces += d->meshes[i].face_count; s->n_verts += d->meshes[i].vertex_count; }
    s->pos = calloc(3 * (size_t) s->n_verts + 1, sizeof(float)); s->nrm = calloc(3 * (size_t) s->n_verts + 1, sizeof(float)); s->uv = calloc(2 * (size_t) s->n_verts + 1, sizeof(float));
    s->faces = calloc(3 * (size_t) s->n_faces + 1, sizeof(uint32_t)); s->face_mesh = calloc((size_t) s->n_faces + 1, sizeof(uint32_t));
    uint32_t vb = 0, fb = 0;
    for (uint32_t i = 0; i < d->n_meshes; ++i) {
        const mtsb_mesh_desc *m = &d->meshes[i]; omesh *om = &s->meshes[i];
        om->face_begin = fb; om->face_count = m->face_count; om->vert_begin = vb; om->vert_count = m->vertex_count;
        om->bsdf = m->bsdf_id; om->emitter = m->emitter_id; om->has_normals = m->normals != NULL; om->has_uv = m->texcoords != NULL; om->shape = -1;
        memcpy(s->pos + 3 * (size_t) vb, m->positions, sizeof(float) * 3 * (size_t) m->vertex_count);
        if (m->normals) memcpy(s->nrm + 3 * (size_t) vb, m->normals, sizeof(float) * 3 * (size_t) m->vertex_count);
        if (m->texcoords) memcpy(s->uv + 2 * (size_t) vb, m->texcoords, sizeof(float) * 2 * (size_t) m->vertex_count);
        for (uint32_t f = 0; f < m->face_count; ++f) { for (int k = 0; k < 3; ++k) s->faces[3 * (size_t) (fb + f) + k] = m->faces[3 * (size_t) f + k] + vb; s->face_mesh[fb + f] = i; }
        vb += m->vertex_count; fb += m->face_count;
    }
    for (uint32_t k = 0; k < d->n_shapes; ++k) {
        omesh *om = &s->meshes[d->n_meshes + k];
        if (!shape_setup(&d->shapes[k], &s->shapes[k])) { fprintf(stderr, "oracle: bad shape %u\n", k); }
        om->face_begin = fb + k; om->face_count = 1; om->bsdf = d->shapes[k].bsdf_id; om->emitter = d->shapes[k].emitter_id; om->has_uv = 1; om->shape = (int) k;
    }
    s->bsdfs = calloc(d->n_bsdfs ? d->n_bsdfs : 1, sizeof(mtsb_bsdf_desc)); memcpy(s->bsdfs, d->bsdfs, sizeof(mtsb_bsdf_desc) * d->n_bsdfs);
    s->rough_tab = calloc((d->n_bsdfs ? d->n_bsdfs : 1) * 65, sizeof(float));
    for (uint32_t i = 0; i < d->n_bsdfs; ++i) if (d->bsdfs[i].type == MTSB_BSDF_ROUGHPLASTIC) rough_plastic_tables(&d->bsdfs[i], s->rough_tab + 65 * (size_t) i);
    s->sensors = calloc(d->n_sensors ? d->n_sensors : 1, sizeof(mtsb_sensor_desc)); memcpy(s->sensors, d->sensors, sizeof(mtsb_sensor_desc) * d->n_sensors);
    s->spectral = d->mode == MTSB_MODE_SPECTRAL; s->nch = s->spectral ? 4 : 3;
    s->tex = calloc(d->n_textures ? d->n_textures : 1, sizeof(otex));
    for (uint32_t i = 0; i < d->n_textures; ++i) {
        const mtsb_texture_desc *t = &d->textures[i]; otex *ot = &s->tex[i];
        ot->type = t->type; ot->value = V(t->value[0], t->value[1], t->value[2]);
        ot->lambda_min = t->lambda_min; ot->lambda_max = t->lambda_max; ot->d65_scale = t->d65_scale; ot->w = t->width;
        if (t->type == MTSB_TEX_REGULAR || t->type == MTSB_TEX_IRREGULAR) { size_t n = (size_t) t->width * (t->type == MTSB_TEX_IRREGULAR ? 2 : 1);
            ot->data = malloc(sizeof(float) * n); memcpy(ot->data, t->data, sizeof(float) * n); }
        if (t->type == MTSB_TEX_MESH_ATTRIBUTE) { ot->h = t->height; ot->ch = t->channels; size_t n = (size_t) t->width * t->channels;
            ot->data = malloc(sizeof(float) * n); memcpy(ot->data, t->data, sizeof(float) * n); }
        if (t->type == MTSB_TEX_BITMAP) { ot->w = t->width; ot->h = t->height; ot->ch = t->channels; memcpy(ot->to_uv, t->to_uv, sizeof(ot->to_uv));
            size_t n = (size_t) t->width * t->height * t->channels; ot->data = malloc(sizeof(float) * n); memcpy(ot->data, t->data, sizeof(float) * n); }
    }
    build_accel(s);
    s->emitters = calloc(d->n_emitters ? d->n_emitters : 1, sizeof(oemitter)); s->env_emitter = -1;
    for (uint32_t i = 0; i < d->n_emitters; ++i) {
        const mtsb_emitter_desc *e = &d->emitters[i]; oemitter *oe = &s->emitters[i];
        oe->type = e->type; oe->radiance_tex = e->radiance_tex;
        if (e->type == MTSB_EMITTER_AREA) { oe->radiance = V(e->radiance[0], e->radiance[1], e->radiance[2]); oe->mesh = e->mesh_id;
            if (s->meshes[e->mesh_id].shape >= 0) oe->normalization = s->shapes[s->meshes[e->mesh_id].shape].inv_area; else build_area_cdf(s, oe); }
        else if (e->type == MTSB_EMITTER_CONSTANT) { oe->radiance = V(e->radiance[0], e->radiance[1], e->radiance[2]); oe->mesh = -1; s->env_emitter = (int) i; }  /* constant.cpp:33-49 */
        else if (e->type == MTSB_EMITTER_POINT) { oe->radiance = V(e->radiance[0], e->radiance[1], e->radiance[2]); oe->mesh = -1;
            oe->position = V(e->position[0], e->position[1], e->position[2]); }                                    /* point.cpp:38-58 */
        else {
            s->env_emitter = (int) i; oe->w = e->width; oe->h = e->height; oe->scale = e->scale;
            memcpy(oe->to_world, e->to_world, sizeof(oe->to_world)); invert4(e->to_world, oe->to_local);
            size_t n = (size_t) e->width * e->height; oe->rgba = malloc(sizeof(float) * 4 * n); memcpy(oe->rgba, s->spectral ? e->coeff_scale : e->rgba, sizeof(float) * 4 * n);
            float *lum = malloc(sizeof(float) * n);
            for (uint32_t y = 0; y < e->height; ++y) {     /* envmap.cpp:87-111 */
                float sin_theta = sinf((float) y / (float) (e->height - 1) * PI_F);
                for (uint32_t x = 0; x < e->width; ++x) { const float *p = e->rgba + 4 * ((size_t) y * e->width + x);
                    lum[(size_t) y * e->width + x] = (p[0] * 0.212671f + p[1] * 0.715160f + p[2] * 0.072169f) * sin_theta; }
            }
            build_h2d(&oe->warp, lum, e->width, e->height); free(lum);
        }
    }
    update_bsphere(s);
    s->max_depth = d->integrator.max_depth < 0 ? (1 << 30) : d->integrator.max_depth; s->rr_depth = d->integrator.rr_depth;
    return s;
}
void orc_scene_destroy(orc_scene *s) {
    if (!s) return;
    for (uint32_t i = 0; i < s->n_tex; ++i) free(s->tex[i].data);
    for (uint32_t i = 0; i < s->n_emitters; ++i) { oemitter *e = &s->emitters[i]; free(e->cdf); free(e->rgba);
        for (int l = 0; l < e->warp.n_levels; ++l) free(e->warp.levels[l]); free(e->warp.levels); free(e->warp.widths); }
    free(s->rough_tab); free(s->meshes); free(s->bsdfs); free(s->tex); free(s->emitters); free(s->sensors); free(s->pos); free(s->nrm); free(s->uv);
    free(s->faces); free(s->face_mesh); free(s->nodes); free(s->order); free(s->shapes); free(s);
}
void orc_scene_update_texture(orc_scene *s, uint32_t id, const float *data) {
    otex *t = &s->tex[id];
    if (t->type == MTSB_TEX_CONSTANT) t->value = V(data[0], data[1], data[2]);
    else memcpy(t->data, data, sizeof(float) * (size_t) t->w * t->h * t->ch);
}

/* =============================================================================================== */
/* Surface interaction, textures, BSDFs, emitters                                                   */
/* =============================================================================================== */
typedef struct { v3 p, n, sh_s, sh_t, sh_n, wi; float u, v, t; uint32_t prim; int mesh, valid; float b1, b2; } osi;
static v3 to_local(const osi *si, v3 w) { return V(vdot(w, si->sh_s), vdot(w, si->sh_t), vdot(w, si->sh_n)); }
static v3 to_world(const osi *si, v3 w) { return vadd(vadd(vscale(si->sh_s, w.x), vscale(si->sh_t, w.y)), vscale(si->sh_n, w.z)); }

/* fill_surface_interaction of the analytic shapes (sphere.cpp:329-373, disk.cpp:167-205, cylinder.cpp:330-361,
 * rectangle.cpp:180-205) followed by the frame completion of kdtree.h:2356-2362 */
static void shape_si(const orc_scene *s, uint32_t k, v3 o, v3 d, const ohit *h, osi *si) {
    const oshape *sh = &s->shapes[k];
    si->mesh = (int) (s->n_meshes + k); si->b1 = h->u; si->b2 = h->v;
    v3 at = V(fmaf(d.x, h->t, o.x), fmaf(d.y, h->t, o.y), fmaf(d.z, h->t, o.z)), dp_du;
    if (sh->type == MTSB_SHAPE_SPHERE) {
        v3 n = vnormalize(vsub(at, sh->center));
        si->p = V(fmaf(n.x, sh->radius, sh->center.x), fmaf(n.y, sh->radius, sh->center.y), fmaf(n.z, sh->radius, sh->center.z));
        v3 l = xf_point(sh->to_object, si->p);
        float half = asinf(.5f * vnorm(V(l.x, l.y, l.z - mulsign(1.f, l.z)))) * 2.f;      /* unit_angle_z */
        float theta = l.z >= 0.f ? half : PI_F - half, phi = atan2f(l.y, l.x);
        if (phi < 0.f) phi += 2.f * PI_F;
        si->u = phi * (.5f / PI_F); si->v = theta * (1.f / PI_F);
        dp_du = vscale(xf_dir(sh->to_world, V(-l.y, l.x, 0.f)), 2.f * PI_F);
        si->n = sh->flip ? vneg(n) : n;
    } else if (sh->type == MTSB_SHAPE_DISK) {
        float r = sqrtf(h->u * h->u + h->v * h->v), inv_r = 1.f / r, ang = atan2f(h->v, h->u) * (.5f / PI_F);
        if (ang < 0.f) ang += 1.f;
        float c = r != 0.f ? h->u * inv_r : 1.f, sn = r != 0.f ? h->v * inv_r : 0.f;
        dp_du = xf_dir(sh->to_world, V(c, sn, 0.f));
        si->n = sh->n; si->u = r; si->v = ang; si->p = at;
    } else if (sh->type == MTSB_SHAPE_CYLINDER) {
        v3 l = xf_point(sh->to_object, at);
        float phi = atan2f(l.y, l.x);
        if (phi < 0.f) phi += 2.f * PI_F;
        si->u = phi * (.5f / PI_F); si->v = l.z / sh->length;
        dp_du = xf_dir(sh->to_world, vscale(V(-l.y, l.x, 0.f), 2.f * PI_F));
        v3 dp_dv = xf_dir(sh->to_world, V(0.f, 0.f, sh->length));
        v3 n = vnormalize(vcross(dp_du, dp_dv));
        si->p = vadd(at, vscale(n, sh->radius - sqrtf(l.x * l.x + l.y * l.y)));
        si->n = sh->flip ? vneg(n) : n;
    } else {
        si->n = sh->n; dp_du = sh->dp_du; si->p = at;
        si->u = fmaf(h->u, .5f, .5f); si->v = fmaf(h->v, .5f, .5f);
    }
    si->sh_n = si->n;
    si->sh_s = vnormalize(vsub(dp_du, vscale(si->n, vdot(si->n, dp_du))));
    si->sh_t = vcross(si->n, si->sh_s);
    si->wi = to_local(si, vneg(d));
}

/* kdtree.h:2334-2368 + mesh.cpp:399-462; miss: scene_native.inl:29-31 */
static void make_si(const orc_scene *s, v3 o, v3 d, const ohit *h, osi *si) {
    memset(si, 0, sizeof(*si));
    si->t = h->t; si->prim = h->prim; si->valid = h->t != INFINITY; si->mesh = -1;
    if (!si->valid) { si->wi = vneg(d); si->p = o; return; }
    if (h->prim >= s->n_faces) { shape_si(s, h->prim - s->n_faces, o, d, h, si); return; }
    uint32_t f = h->prim; const omesh *m = &s->meshes[s->face_mesh[f]]; si->mesh = (int) s->face_mesh[f];
    uint32_t i0 = s->faces[3 * f], i1 = s->faces[3 * f + 1], i2 = s->faces[3 * f + 2];
    float b1 = h->u, b2 = h->v, b0 = 1.f - b1 - b2; si->b1 = b1; si->b2 = b2;
    v3 p0 = ld3(s->pos, i0), p1 = ld3(s->pos, i1), p2 = ld3(s->pos, i2);
    v3 dp0 = vsub(p1, p0), dp1 = vsub(p2, p0);
    si->p = vadd(vadd(vscale(p0, b0), vscale(p1, b1)), vscale(p2, b2));
    v3 n = vnormalize(vcross(dp0, dp1)); si->n = n;
    v3 dp_du, dp_dv; coordinate_system(n, &dp_du, &dp_dv);
    si->u = b1; si->v = b2;
    if (m->has_uv) {
        float u0 = s->uv[2 * i0], v0 = s->uv[2 * i0 + 1], u1 = s->uv[2 * i1], v1 = s->uv[2 * i1 + 1], u2 = s->uv[2 * i2], v2 = s->uv[2 * i2 + 1];
        si->u = u0 * b0 + u1 * b1 + u2 * b2; si->v = v0 * b0 + v1 * b1 + v2 * b2;
        float duv0x = u1 - u0, duv0y = v1 - v0, duv1x = u2 - u0, duv1y = v2 - v0;
        float det = fmaf(duv0x, duv1y, -(duv0y * duv1x));
        if (det != 0.f) { float inv_det = 1.f / det;
            dp_du = vscale(vsub(vscale(dp0, duv1y), vscale(dp1, duv0y)), inv_det);
            dp_dv = vscale(vsub(vscale(dp1, duv0x), vscale(dp0, duv1x)), inv_det); }
    }
    v3 ns = n;
    if (m->has_normals) ns = vnormalize(vadd(vadd(vscale(ld3(s->nrm, i0), b0), vscale(ld3(s->nrm, i1), b1)), vscale(ld3(s->nrm, i2), b2)));
    si->sh_n = ns;
    si->sh_s = vnormalize(vsub(dp_du, vscale(ns, vdot(ns, dp_du))));
    si->sh_t = vcross(ns, si->sh_s);
    si->wi = to_local(si, vneg(d));
}
void orc_surface_interaction(const orc_scene *s, uint64_t n, const float *o, const float *d, float *out, int threads) {
    #pragma omp parallel for schedule(dynamic, 1024) num_threads(threads > 0 ? threads : 1)
    for (int64_t i = 0; i < (int64_t) n; ++i) {
        v3 ro = V(o[4 * i], o[4 * i + 1], o[4 * i + 2]), rd = V(d[4 * i], d[4 * i + 1], d[4 * i + 2]);
        ohit h; trace(s, ro, rd, o[4 * i + 3], d[4 * i + 3], 0, &h);
        osi si; make_si(s, ro, rd, &h, &si);
        float *q = out + 16 * i;
        memset(q, 0, 16 * sizeof(float));
        q[0] = h.t; q[3] = bitsf(h.prim);
        if (!si.valid) continue;
        q[1] = si.u; q[2] = si.v;
        q[4] = si.p.x; q[5] = si.p.y; q[6] = si.p.z; q[7] = si.n.x; q[8] = si.n.y; q[9] = si.n.z;
        q[10] = si.sh_n.x; q[11] = si.sh_n.y; q[12] = si.sh_n.z; q[13] = si.wi.x; q[14] = si.wi.y; q[15] = si.wi.z;
    }
}
/* src/textures/bitmap.cpp:250-293, src/spectra/srgb.cpp */
/* ---- spectral mode arithmetic -------------------------------------------------------------------------------- */
/* sample_wavelength (core/spectrum.h:304-316) = sample_rgb_spectrum (:270-285) of math::sample_shifted (core/math.h:418-442) */
static void sample_wavelengths(float u, float *wav, float *weight) {
    for (int i = 0; i < 4; ++i) {
        float v = u + (float) i / 4.f; if (v > 1.f) v -= 1.f;
        wav[i] = 538.f - atanhf(0.8569106254698279f - 1.8275019724092267f * v) * 138.88888888888889f;
        float t = coshf(0.0072f * (wav[i] - 538.f));
        weight[i] = 253.82f * t * t;
    }
}
void orc_sample_wavelength(float u, float *wav4, float *weight4) { sample_wavelengths(u, wav4, weight4); }
/* srgb_model_eval, include/mitsuba/render/srgb.h:8-24 */
static float srgb_model_eval(const float *c, float l) {
    float v = fmaf(fmaf(c[0], l, c[1]), l, c[2]);
    if (isinf(c[2])) return fmaf(copysignf(1.f, c[2]), .5f, .5f);
    return fmaxf(0.f, fmaf(.5f * v, 1.f / sqrtf(fmaf(v, v, 1.f)), .5f));
}
float orc_srgb_model_eval(const float *coeff3, float lambda) { return srgb_model_eval(coeff3, lambda); }
/* ContinuousDistribution::eval_pdf, core/distr_1d.h:378-393 (`regular` spectrum, src/spectra/regular.cpp:77-84) */
static float regular_eval(const float *values, uint32_t n, float lo, float hi, float x) {
    if (!(x >= lo && x <= hi)) return 0.f;
    float inv_interval = (float) (n - 1) / (hi - lo);       /* m_inv_interval_size, distr_1d.h:316-318 */
    x = (x - lo) * inv_interval;
    int32_t idx = (int32_t) x; if (idx < 0) idx = 0; if (idx > (int32_t) n - 2) idx = (int32_t) n - 2;
    float w1 = x - (float) idx, w0 = 1.f - w1;
    return fmaf(w0, values[idx], w1 * values[idx + 1]);
}
float orc_regular_eval(const float *values, uint32_t n, float lo, float hi, float x) { return regular_eval(values, n, lo, hi, x); }
/* IrregularContinuousDistribution::eval_pdf, core/distr_1d.h:655-677 (`irregular` spectrum) */
static float irregular_eval(const float *nodes, const float *values, uint32_t n, float x) {
    if (!(x >= nodes[0] && x <= nodes[n - 1])) return 0.f;
    uint32_t start = 0, end = n, iterations = 0; { uint32_t v = n; while (v) { ++iterations; v >>= 1; } }   /* log2i(n) + 1 */
    for (uint32_t i = 0; i < iterations; ++i) {             /* enoki::binary_search, array_utils.h:153-171 */
        uint32_t middle = (start + end) >> 1;
        int pred = middle < n && nodes[middle] < x;
        if (pred) start = middle + 1 < end ? middle + 1 : end; else end = middle;
    }
    uint32_t idx = start < n - 1 ? start : n - 1; if (idx < 1) idx = 1; idx -= 1;
    float x0 = nodes[idx], x1 = nodes[idx + 1], y0 = values[idx], y1 = values[idx + 1];
    x = (x - x0) / (x1 - x0);
    return fmaf(x, y1 - y0, y0);
}
float orc_irregular_eval(const float *nodes, const float *values, uint32_t n, float x) { return irregular_eval(nodes, values, n, x); }
/* D65 whitepoint: `d65` expands to `regular` over data * scale / 10568 (src/spectra/d65.cpp:56-73) */
static float d65_eval(float l) {
    static float tbl[95]; static int init = 0;
    if (!init) {
        #pragma omp critical
        { for (int i = 0; i < 95; ++i) tbl[i] = kCieD65[i] * (1.f / 10568.f); init = 1; }
    }
    return regular_eval(tbl, 95, 360.f, 830.f, l);
}
float orc_d65_eval(float lambda) { return d65_eval(lambda); }
/* cie1931_xyz, core/spectrum.h:147-176 */
static v3 cie1931_xyz(float l) {
    float t = (l - 360.f) * (94.f / (830.f - 360.f));
    if (!(l >= 360.f && l <= 830.f)) return V(0, 0, 0);
    int32_t i0 = (int32_t) t; if (i0 < 0) i0 = 0; if (i0 > 93) i0 = 93;
    float w1 = t - (float) i0, w0 = 1.f - w1;
    return V(fmaf(w0, kCie1931[i0], w1 * kCie1931[i0 + 1]), fmaf(w0, kCie1931[95 + i0], w1 * kCie1931[95 + i0 + 1]),
             fmaf(w0, kCie1931[190 + i0], w1 * kCie1931[190 + i0 + 1]));
}
/* spectrum_to_xyz, core/spectrum.h:209-217: hmean over the 4 wavelengths (pairwise horizontal sum) */
static v3 spectrum_to_xyz(const float *value, const float *wav) {
    v3 m[4]; for (int i = 0; i < 4; ++i) m[i] = cie1931_xyz(wav[i]);
    return V(((m[0].x * value[0] + m[1].x * value[1]) + (m[2].x * value[2] + m[3].x * value[3])) * .25f,
             ((m[0].y * value[0] + m[1].y * value[1]) + (m[2].y * value[2] + m[3].y * value[3])) * .25f,
             ((m[0].z * value[0] + m[1].z * value[1]) + (m[2].z * value[2] + m[3].z * value[3])) * .25f);
}
void orc_spectrum_to_xyz(const float *value4, const float *wav4, float *xyz) { v3 r = spectrum_to_xyz(value4, wav4); xyz[0] = r.x; xyz[1] = r.y; xyz[2] = r.z; }

/* Texture::eval.  rgb mode: src/spectra/srgb.cpp:44-50 (constant), src/textures/bitmap.cpp:250-293.  Spectral mode (wav != NULL):
 * constants and 3-channel texels are sRGB-model coefficients (srgb.cpp:44-48, bitmap.cpp:262-275), tabulated spectra
 * (regular.cpp, irregular.cpp), uniform.cpp; srgb_d65.cpp:60-62 multiplies by the D65 spectrum. */
static col tex_eval(const orc_scene *s, int id, float u, float v, const float *wav) {
    const otex *t = &s->tex[id]; col r = C1(0.f);
    if (t->type != MTSB_TEX_BITMAP) {
        if (!s->spectral) return C3(t->value);
        float cf[3] = { t->value.x, t->value.y, t->value.z };
        for (int i = 0; i < 4; ++i)
            r.c[i] = t->type == MTSB_TEX_CONSTANT ? srgb_model_eval(cf, wav[i])
                   : t->type == MTSB_TEX_REGULAR ? regular_eval(t->data, t->w, t->lambda_min, t->lambda_max, wav[i])
                   : t->type == MTSB_TEX_IRREGULAR ? irregular_eval(t->data, t->data + t->w, t->w, wav[i]) : t->value.x;
    } else {
        float tu = t->to_uv[0] * u + t->to_uv[1] * v + t->to_uv[2], tv = t->to_uv[3] * u + t->to_uv[4] * v + t->to_uv[5];
        tu -= floorf(tu); tv -= floorf(tv);
        tu *= (float) (t->w - 1); tv *= (float) (t->h - 1);
        uint32_t px = (uint32_t) tu, py = (uint32_t) tv; if (px > t->w - 2) px = t->w - 2; if (py > t->h - 2) py = t->h - 2;
        float w1x = tu - (float) px, w1y = tv - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
        size_t idx = px + (size_t) py * t->w;
        if (s->spectral && t->ch == 3) {
            const float *c00 = t->data + 3 * idx, *c10 = c00 + 3, *c01 = c00 + 3 * (size_t) t->w, *c11 = c01 + 3;
            for (int i = 0; i < 4; ++i) {
                float v0 = fmaf(w0x, srgb_model_eval(c00, wav[i]), w1x * srgb_model_eval(c10, wav[i])),
                      v1 = fmaf(w0x, srgb_model_eval(c01, wav[i]), w1x * srgb_model_eval(c11, wav[i]));
                r.c[i] = fmaf(w0y, v0, w1y * v1);
            }
        } else {
            for (uint32_t c = 0; c < (uint32_t) s->nch; ++c) {
                uint32_t cc = t->ch == 1 ? 0 : c;
                float v00 = t->data[idx * t->ch + cc], v10 = t->data[(idx + 1) * t->ch + cc], v01 = t->data[(idx + t->w) * t->ch + cc], v11 = t->data[(idx + t->w + 1) * t->ch + cc];
                float v0 = fmaf(w0x, v00, w1x * v10), v1 = fmaf(w0x, v01, w1x * v11);
                r.c[c] = fmaf(w0y, v0, w1y * v1);
            }
        }
    }
    if (s->spectral && t->d65_scale != 0.f) for (int i = 0; i < 4; ++i) r.c[i] = d65_eval(wav[i]) * t->d65_scale * r.c[i];
    return r;
}
/* `mesh_attribute` (src/textures/mesh_attribute.cpp -> Mesh::eval_attribute / interpolate_attribute, mesh.h:278-318) */
static col tex_eval_si(const orc_scene *s, int id, const osi *si, const float *wav) {
    const otex *t = &s->tex[id];
    if (t->type != MTSB_TEX_MESH_ATTRIBUTE) return tex_eval(s, id, si->u, si->v, wav);
    col r = C1(0.f);
    if (!si->valid) return r;
    const omesh *m = &s->meshes[si->mesh];
    col v[3]; int nv = t->h == 1 ? 1 : 3;
    for (int k = 0; k < nv; ++k) {
        uint32_t i = t->h == 1 ? si->prim - m->face_begin : s->faces[3 * si->prim + k] - m->vert_begin;
        if (t->ch == 1) v[k] = C1(t->data[i]);
        else if (s->spectral) { for (int c = 0; c < 4; ++c) v[k].c[c] = srgb_model_eval(t->data + 3 * (size_t) i, wav[c]); }
        else v[k] = C3(V(t->data[3 * (size_t) i], t->data[3 * (size_t) i + 1], t->data[3 * (size_t) i + 2]));
    }
    if (nv == 1) r = v[0];
    else { float b0 = 1.f - si->b1 - si->b2; for (int c = 0; c < 4; ++c) r.c[c] = v[0].c[c] * b0 + (v[1].c[c] * si->b1 + v[2].c[c] * si->b2); }
    return cscale(r, t->value.x);
}
typedef struct { v3 wo; col weight; float pdf, eta; int delta; } obs;
/* fresnel_diffuse_reflectance (include/mitsuba/render/fresnel.h:324-358) */
static float fresnel_diffuse_reflectance(float eta) {
    if (eta < 1.f) return -1.4399f * (eta * eta) + 0.7099f * eta + 0.6681f + 0.0636f / eta;
    float i1 = 1.f / eta, i2 = i1 * i1, i3 = i2 * i1, i4 = i3 * i1, i5 = i4 * i1;
    return 0.919317f - 3.4793f * i1 + 6.75335f * i2 - 7.80989f * i3 + 4.98554f * i4 - 1.36881f * i5;
}
static col fresnel_conductor_col(float c, col eta, col k) { col r; for (int i = 0; i < 4; ++i) r.c[i] = orc_fresnel_conductor(c, eta.c[i], k.c[i]); return r; }
static void bsdf_eval_pdf_nested(const orc_scene *s, const mtsb_bsdf_desc *b, const osi *si, v3 wo, col *val, float *pdf, const float *wav);
static obs bsdf_sample_nested(const orc_scene *s, const mtsb_bsdf_desc *b, const osi *si, float s1, float sx, float sy, const float *wav) {
    obs bs; memset(&bs, 0, sizeof(bs)); bs.eta = 1.f;
    float cos_theta_i = si->wi.z;
    if (b->type == MTSB_BSDF_DIFFUSE) {                /* diffuse.cpp:78-103 */
        if (!(cos_theta_i > 0.f)) return bs;
        bs.wo = cosine_hemisphere(sx, sy); bs.pdf = INV_PI_F * bs.wo.z;
        if (bs.pdf > 0.f) bs.weight = tex_eval_si(s, b->reflectance_tex, si, wav);
        return bs;
    }
    if (b->type == MTSB_BSDF_ROUGHCONDUCTOR) {         /* roughconductor.cpp:196-276 */
        if (!(cos_theta_i > 0.f)) return bs;
        mf_t distr = mf_make(b->distribution, b->alpha_u, b->alpha_v, b->sample_visible);
        v3 m = mf_sample(&distr, si->wi, sx, sy, &bs.pdf);
        bs.wo = vsub(vscale(m, 2.f * vdot(si->wi, m)), si->wi);
        int active = bs.pdf != 0.f && bs.wo.z > 0.f;
        float weight = b->sample_visible ? mf_g1(&distr, bs.wo, m)
                                         : mf_g1(&distr, si->wi, m) * mf_g1(&distr, bs.wo, m) * vdot(si->wi, m) / (cos_theta_i * m.z);
        bs.pdf /= 4.f * vdot(bs.wo, m);
        col eta = tex_eval_si(s, b->eta_tex, si, wav), k = tex_eval_si(s, b->k_tex, si, wav);
        col F = fresnel_conductor_col(vdot(si->wi, m), eta, k);
        if (b->specular_reflectance_tex >= 0) F = cmul(F, tex_eval_si(s, b->specular_reflectance_tex, si, wav));
        bs.weight = active ? cscale(F, weight) : C1(0.f);
        return bs;
    }
    if (b->type == MTSB_BSDF_CONDUCTOR) {              /* conductor.cpp: sample (delta reflection, weight = R * F(cos_i)) */
        if (!(cos_theta_i > 0.f)) return bs;
        bs.wo = V(-si->wi.x, -si->wi.y, si->wi.z); bs.pdf = 1.f; bs.delta = 1;
        col eta = tex_eval_si(s, b->eta_tex, si, wav), k = tex_eval_si(s, b->k_tex, si, wav);
        col F = fresnel_conductor_col(cos_theta_i, eta, k);
        if (b->specular_reflectance_tex >= 0) F = cmul(tex_eval_si(s, b->specular_reflectance_tex, si, wav), F);
        bs.weight = F; return bs;
    }
    if (b->type == MTSB_BSDF_PLASTIC) {                /* plastic.cpp:178-245 */
        if (!(cos_theta_i > 0.f)) return bs;
        float f_i, ct, e1, e2; fresnel(cos_theta_i, b->eta, &f_i, &ct, &e1, &e2);
        float ssw = b->specular_sampling_weight;
        float prob_specular = f_i * ssw, prob_diffuse = (1.f - f_i) * (1.f - ssw);
        prob_specular = prob_specular / (prob_specular + prob_diffuse); prob_diffuse = 1.f - prob_specular;
        if (s1 < prob_specular) {
            bs.wo = V(-si->wi.x, -si->wi.y, si->wi.z); bs.pdf = prob_specular; bs.delta = 1;
            col value = C1(f_i / bs.pdf);
            if (b->specular_reflectance_tex >= 0) value = cmul(value, tex_eval_si(s, b->specular_reflectance_tex, si, wav));
            bs.weight = value;
        } else {
            bs.wo = cosine_hemisphere(sx, sy); bs.pdf = prob_diffuse * (INV_PI_F * bs.wo.z);
            float f_o; fresnel(bs.wo.z, b->eta, &f_o, &ct, &e1, &e2);
            float fdr_int = fresnel_diffuse_reflectance(1.f / b->eta), inv_eta_2 = 1.f / (b->eta * b->eta);
            col value = tex_eval_si(s, b->reflectance_tex, si, wav);
            for (int i = 0; i < 4; ++i) value.c[i] /= 1.f - (b->nonlinear ? value.c[i] * fdr_int : fdr_int);
            bs.weight = cscale(value, inv_eta_2 * (1.f - f_i) * (1.f - f_o) / prob_diffuse);
        }
        return bs;
    }
    if (b->type == MTSB_BSDF_ROUGHPLASTIC) {           /* roughplastic.cpp: sample (choose a lobe; weight = eval / pdf over both) */
        if (!(cos_theta_i > 0.f)) return bs;
        const float *tab = s->rough_tab + 65 * (size_t) (b - s->bsdfs);
        float t_i = rough_lerp(tab, cos_theta_i), ssw = b->specular_sampling_weight;
        float prob_specular = (1.f - t_i) * ssw, prob_diffuse = t_i * (1.f - ssw);
        prob_specular = prob_specular / (prob_specular + prob_diffuse);
        if (s1 < prob_specular) {
            mf_t distr = mf_make(b->distribution, b->alpha_u, b->alpha_u, b->sample_visible);
            float unused; v3 m = mf_sample(&distr, si->wi, sx, sy, &unused);
            bs.wo = vsub(vscale(m, 2.f * vdot(si->wi, m)), si->wi);
        } else
            bs.wo = cosine_hemisphere(sx, sy);
        col val; bsdf_eval_pdf_nested(s, b, si, bs.wo, &val, &bs.pdf, wav);
        bs.weight = bs.pdf > 0.f ? cscale(val, 1.f / bs.pdf) : C1(0.f);
        if (!(bs.pdf > 0.f)) bs.pdf = 0.f;
        return bs;
    }
    if (b->type == MTSB_BSDF_ROUGHDIELECTRIC) {        /* roughdielectric.cpp: sample */
        if (cos_theta_i == 0.f) return bs;
        mf_t distr = mf_make(b->distribution, b->alpha_u, b->alpha_v, b->sample_visible), sample_distr = distr;
        if (!b->sample_visible) { float sc = 1.2f - .2f * sqrtf(fabsf(cos_theta_i)); sample_distr.au *= sc; sample_distr.av *= sc; }
        v3 wi_up = V(mulsign(si->wi.x, cos_theta_i), mulsign(si->wi.y, cos_theta_i), mulsign(si->wi.z, cos_theta_i));
        v3 m = mf_sample(&sample_distr, wi_up, sx, sy, &bs.pdf);
        if (bs.pdf == 0.f) return bs;
        float F, cos_theta_t, eta_it, eta_ti, wim = vdot(si->wi, m);
        fresnel(wim, b->eta, &F, &cos_theta_t, &eta_it, &eta_ti);
        int selected_r = s1 <= F;
        col weight = C1(1.f);
        bs.pdf *= selected_r ? F : 1.f - F;
        bs.eta = selected_r ? 1.f : eta_it;
        float dwh_dwo;
        if (selected_r) {
            bs.wo = vsub(vscale(m, 2.f * wim), si->wi);
            if (b->specular_reflectance_tex >= 0) weight = cmul(weight, tex_eval_si(s, b->specular_reflectance_tex, si, wav));
            dwh_dwo = 1.f / (4.f * vdot(bs.wo, m));
        } else {
            bs.wo = vsub(vscale(m, fmaf(wim, eta_ti, cos_theta_t)), vscale(si->wi, eta_ti));
            col factor = C1(sqrf(eta_ti));
            if (b->specular_transmittance_tex >= 0) factor = cmul(factor, tex_eval_si(s, b->specular_transmittance_tex, si, wav));
            weight = cmul(weight, factor);
            dwh_dwo = (sqrf(bs.eta) * vdot(bs.wo, m)) / sqrf(wim + bs.eta * vdot(bs.wo, m));
        }
        if (b->sample_visible) weight = cscale(weight, mf_g1(&distr, bs.wo, m));
        else weight = cscale(weight, mf_g1(&distr, si->wi, m) * mf_g1(&distr, bs.wo, m) * wim / (cos_theta_i * m.z));
        bs.pdf *= fabsf(dwh_dwo);
        bs.weight = weight; return bs;
    }
    /* dielectric.cpp:201-316 */
    float r_i, cos_theta_t, eta_it, eta_ti; fresnel(cos_theta_i, b->eta, &r_i, &cos_theta_t, &eta_it, &eta_ti);
    float t_i = 1.f - r_i; int selected_r = s1 <= r_i;
    bs.pdf = selected_r ? r_i : t_i; bs.delta = 1;
    bs.wo = selected_r ? V(-si->wi.x, -si->wi.y, si->wi.z) : V(-eta_ti * si->wi.x, -eta_ti * si->wi.y, cos_theta_t);
    bs.eta = selected_r ? 1.f : eta_it;
    col w = C1(1.f);
    if (selected_r) { if (b->specular_reflectance_tex >= 0) w = cmul(w, tex_eval_si(s, b->specular_reflectance_tex, si, wav)); }
    else { if (b->specular_transmittance_tex >= 0) w = cmul(w, tex_eval_si(s, b->specular_transmittance_tex, si, wav)); w = cscale(w, sqrf(eta_ti)); }
    bs.weight = w; return bs;
}
static void bsdf_eval_pdf_nested(const orc_scene *s, const mtsb_bsdf_desc *b, const osi *si, v3 wo, col *val, float *pdf, const float *wav) {
    *val = C1(0.f); *pdf = 0.f;
    float cos_theta_i = si->wi.z, cos_theta_o = wo.z;
    if (b->type == MTSB_BSDF_DIFFUSE) {                /* diffuse.cpp:105-135 */
        if (cos_theta_i > 0.f && cos_theta_o > 0.f) { *val = cscale(cscale(tex_eval_si(s, b->reflectance_tex, si, wav), INV_PI_F), cos_theta_o); *pdf = INV_PI_F * cos_theta_o; }
        return;
    }
    if (b->type == MTSB_BSDF_ROUGHCONDUCTOR) {         /* roughconductor.cpp:278-394 */
        if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return;
        v3 H = vnormalize(vadd(wo, si->wi));
        mf_t distr = mf_make(b->distribution, b->alpha_u, b->alpha_v, b->sample_visible);
        float D = mf_eval(&distr, H);
        if (D != 0.f) {
            float G = mf_g1(&distr, si->wi, H) * mf_g1(&distr, wo, H), result = D * G / (4.f * cos_theta_i);
            col eta = tex_eval_si(s, b->eta_tex, si, wav), k = tex_eval_si(s, b->k_tex, si, wav);
            col F = fresnel_conductor_col(vdot(si->wi, H), eta, k);
            if (b->specular_reflectance_tex >= 0) F = cmul(F, tex_eval_si(s, b->specular_reflectance_tex, si, wav));
            *val = cscale(F, result);
        }
        if (vdot(si->wi, H) > 0.f && vdot(wo, H) > 0.f)
            *pdf = b->sample_visible ? D * mf_g1(&distr, si->wi, H) / (4.f * cos_theta_i) : mf_pdf(&distr, si->wi, H) / (4.f * vdot(wo, H));
        return;
    }
    if (b->type == MTSB_BSDF_ROUGHDIELECTRIC) {        /* roughdielectric.cpp: eval, pdf */
        if (cos_theta_i == 0.f) return;
        int reflect = cos_theta_i * cos_theta_o > 0.f;
        float eta = cos_theta_i > 0.f ? b->eta : 1.f / b->eta, inv_eta = cos_theta_i > 0.f ? 1.f / b->eta : b->eta;
        v3 m = vnormalize(vadd(si->wi, vscale(wo, reflect ? 1.f : eta)));
        m = V(mulsign(m.x, m.z), mulsign(m.y, m.z), mulsign(m.z, m.z));
        mf_t distr = mf_make(b->distribution, b->alpha_u, b->alpha_v, b->sample_visible), sample_distr = distr;
        float wim = vdot(si->wi, m), wom = vdot(wo, m);
        float D = mf_eval(&distr, m), F, ct, e1, e2; fresnel(wim, b->eta, &F, &ct, &e1, &e2);
        float G = mf_g1(&distr, si->wi, m) * mf_g1(&distr, wo, m);
        if (reflect) {
            *val = C1(F * D * G / (4.f * fabsf(cos_theta_i)));
            if (b->specular_reflectance_tex >= 0) *val = cmul(*val, tex_eval_si(s, b->specular_reflectance_tex, si, wav));
        } else {
            *val = C1(fabsf((sqrf(inv_eta) * (1.f - F) * D * G * eta * eta * wim * wom) / (cos_theta_i * sqrf(wim + eta * wom))));
            if (b->specular_transmittance_tex >= 0) *val = cmul(*val, tex_eval_si(s, b->specular_transmittance_tex, si, wav));
        }
        if (!(wim * cos_theta_i > 0.f && wom * cos_theta_o > 0.f)) return;
        float dwh_dwo = reflect ? 1.f / (4.f * wom) : (eta * eta * wom) / sqrf(wim + eta * wom);
        if (!b->sample_visible) { float sc = 1.2f - .2f * sqrtf(fabsf(cos_theta_i)); sample_distr.au *= sc; sample_distr.av *= sc; }
        float prob = mf_pdf(&sample_distr, V(mulsign(si->wi.x, cos_theta_i), mulsign(si->wi.y, cos_theta_i), mulsign(si->wi.z, cos_theta_i)), m);
        prob *= reflect ? F : 1.f - F;
        *pdf = prob * fabsf(dwh_dwo);
        return;
    }
    if (b->type == MTSB_BSDF_ROUGHPLASTIC) {           /* roughplastic.cpp: eval, pdf */
        if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return;
        const float *tab = s->rough_tab + 65 * (size_t) (b - s->bsdfs);
        mf_t distr = mf_make(b->distribution, b->alpha_u, b->alpha_u, b->sample_visible);
        v3 H = vnormalize(vadd(wo, si->wi));
        float D = mf_eval(&distr, H), F, ct, e1, e2; fresnel(vdot(si->wi, H), b->eta, &F, &ct, &e1, &e2);
        col value = C1(F * D * (mf_g1(&distr, si->wi, H) * mf_g1(&distr, wo, H)) / (4.f * cos_theta_i));
        if (b->specular_reflectance_tex >= 0) value = cmul(value, tex_eval_si(s, b->specular_reflectance_tex, si, wav));
        float t_i = rough_lerp(tab, cos_theta_i), t_o = rough_lerp(tab, cos_theta_o);
        col diff = tex_eval_si(s, b->reflectance_tex, si, wav);
        for (int i = 0; i < 4; ++i) diff.c[i] /= 1.f - (b->nonlinear ? diff.c[i] * tab[64] : tab[64]);
        diff = cscale(diff, INV_PI_F * (1.f / (b->eta * b->eta)) * cos_theta_o * t_i * t_o);
        for (int i = 0; i < 4; ++i) val->c[i] = value.c[i] + diff.c[i];
        float ssw = b->specular_sampling_weight, prob_specular = (1.f - t_i) * ssw, prob_diffuse = t_i * (1.f - ssw);
        prob_specular = prob_specular / (prob_specular + prob_diffuse); prob_diffuse = 1.f - prob_specular;
        float r = b->sample_visible ? D * mf_g1(&distr, si->wi, H) / (4.f * cos_theta_i) : mf_pdf(&distr, si->wi, H) / (4.f * vdot(wo, H));
        *pdf = r * prob_specular + prob_diffuse * (INV_PI_F * cos_theta_o);
        return;
    }
    if (b->type == MTSB_BSDF_PLASTIC) {                /* plastic.cpp:247-296 (the diffuse lobe; the specular one is a delta) */
        if (!(cos_theta_i > 0.f && cos_theta_o > 0.f)) return;
        float f_i, f_o, ct, e1, e2; fresnel(cos_theta_i, b->eta, &f_i, &ct, &e1, &e2); fresnel(cos_theta_o, b->eta, &f_o, &ct, &e1, &e2);
        float fdr_int = fresnel_diffuse_reflectance(1.f / b->eta), inv_eta_2 = 1.f / (b->eta * b->eta);
        col diff = tex_eval_si(s, b->reflectance_tex, si, wav);
        for (int i = 0; i < 4; ++i) diff.c[i] /= 1.f - (b->nonlinear ? diff.c[i] * fdr_int : fdr_int);
        *val = cscale(diff, (INV_PI_F * cos_theta_o) * inv_eta_2 * (1.f - f_i) * (1.f - f_o));
        float ssw = b->specular_sampling_weight, prob_specular = f_i * ssw, prob_diffuse = (1.f - f_i) * (1.f - ssw);
        *pdf = (INV_PI_F * cos_theta_o) * (prob_diffuse / (prob_specular + prob_diffuse));
    }
}
/* TwoSidedBRDF::sample / eval / pdf (twosided.cpp:98-200) with the same nested BRDF on both sides: on the back side flip
 * wi (and wo) about the surface; cos(wi) == 0 is neither front nor back */
/* ---- roughplastic.cpp: precomputed boundary integrals (parameters_changed; microfacet.h:454-552) and lerp_gather ---- */
static void gauss_legendre(int n, double *x, double *w) {      /* quad::gauss_legendre: Newton on P_n from the Chebyshev guess */
    for (int i = 0; i < (n + 1) / 2; ++i) {
        double z = -cos((2 * i + 1) / (double) (2 * n) * 3.14159265358979323846), dp = 1.0;
        for (int it = 0; it < 100; ++it) {
            double p0 = 1.0, p1 = z;
            for (int k = 2; k <= n; ++k) { double p2 = ((2 * k - 1) * z * p1 - (k - 1) * p0) / k; p0 = p1; p1 = p2; }
            dp = n * (z * p1 - p0) / (z * z - 1.0);
            double step = p1 / dp; z -= step;
            if (fabs(step) <= 4e-16 * fabs(z)) break;
        }
        double p0 = 1.0, p1 = z;
        for (int k = 2; k <= n; ++k) { double p2 = ((2 * k - 1) * z * p1 - (k - 1) * p0) / k; p0 = p1; p1 = p2; }
        dp = n * (z * p1 - p0) / (z * z - 1.0);
        x[i] = z; x[n - 1 - i] = -z; w[i] = w[n - 1 - i] = 2.0 / ((1.0 - z * z) * dp * dp);
    }
}
static float rough_boundary_integral(int type, float alpha, float eta, float mu, int transmit) {
    int res = eta > 1.f ? 32 : 128;
    double xs[128], ws[128]; gauss_legendre(res, xs, ws);
    mf_t distr = mf_make(type, alpha, alpha, 1);
    v3 wi = V(sqrtf(1.f - mu * mu), 0.f, mu);
    float accum = 0.f;
    for (int a = 0; a < res; ++a) for (int b = 0; b < res; ++b) {
        float pdf; v3 m = mf_sample(&distr, wi, fmaf((float) xs[b], .5f, .5f), fmaf((float) xs[a], .5f, .5f), &pdf);
        float f, ct, eit, eti, wim = vdot(wi, m); fresnel(wim, eta, &f, &ct, &eit, &eti);
        float smith;
        if (transmit) { v3 wo = vsub(vscale(m, fmaf(wim, eti, ct)), vscale(wi, eti)); smith = wo.z * wi.z >= 0.f ? 0.f : mf_g1(&distr, wo, m) * (1.f - f); }
        else { v3 wo = vsub(vscale(m, 2.f * wim), wi); smith = (wo.z <= 0.f || wi.z <= 0.f) ? 0.f : mf_g1(&distr, wo, m) * f; }
        accum += smith * ((float) ws[a] * (float) ws[b]);
    }
    return accum * .25f;
}
static void rough_plastic_tables(const mtsb_bsdf_desc *b, float *tab /* 64 transmittances, then the internal reflectance */) {
    double acc = 0;
    for (int i = 0; i < 64; ++i) {
        float mu = fmaxf(1e-6f, (float) i / 63.f);
        tab[i] = rough_boundary_integral(b->distribution, b->alpha_u, b->eta, mu, 1);
        acc += (double) rough_boundary_integral(b->distribution, b->alpha_u, 1.f / b->eta, mu, 0) * mu;
    }
    tab[64] = (float) (acc / 64.0) * 2.f;
}
static float rough_lerp(const float *data, float x) {
    x *= 63.f; uint32_t index = (uint32_t) x; if (index > 62u) index = 62u;
    return lerp_(data[index], data[index + 1], x - (float) index);
}
/* MaskBSDF::eval_opacity (mask.cpp): clamp(opacity->eval_1(si), 0, 1); the texture is single-channel */
static float mask_opacity(const orc_scene *s, const mtsb_bsdf_desc *b, const osi *si, const float *wav) {
    return fminf(fmaxf(tex_eval_si(s, b->opacity_tex_plus1 - 1, si, wav).c[0], 0.f), 1.f);
}
static obs bsdf_sample(const orc_scene *s, const mtsb_bsdf_desc *b, const osi *si, float s1, float sx, float sy, const float *wav) {
    if (b->opacity_tex_plus1 > 0) {                          /* mask.cpp:95-128 */
        float opacity = mask_opacity(s, b, si, wav);
        if (!(s1 < opacity)) {
            obs bs; memset(&bs, 0, sizeof(bs)); bs.wo = V(-si->wi.x, -si->wi.y, -si->wi.z); bs.eta = 1.f; bs.pdf = 1.f - opacity; bs.delta = 1; bs.weight = C1(1.f);
            return bs;
        }
        s1 /= opacity;
    }
    if (!b->two_sided || si->wi.z > 0.f) return bsdf_sample_nested(s, b, si, s1, sx, sy, wav);
    obs bs; memset(&bs, 0, sizeof(bs)); bs.eta = 1.f;
    if (!(si->wi.z < 0.f)) return bs;
    osi flipped = *si; flipped.wi.z *= -1.f;
    bs = bsdf_sample_nested(s, b, &flipped, s1, sx, sy, wav);
    bs.wo.z *= -1.f;
    return bs;
}
static void bsdf_eval_pdf_unmasked(const orc_scene *s, const mtsb_bsdf_desc *b, const osi *si, v3 wo, col *val, float *pdf, const float *wav);
static void bsdf_eval_pdf(const orc_scene *s, const mtsb_bsdf_desc *b, const osi *si, v3 wo, col *val, float *pdf, const float *wav) {
    bsdf_eval_pdf_unmasked(s, b, si, wo, val, pdf, wav);
    if (b->opacity_tex_plus1 > 0) { float opacity = mask_opacity(s, b, si, wav); *val = cscale(*val, opacity); *pdf *= opacity; }    /* mask.cpp:130-155 */
}
static void bsdf_eval_pdf_unmasked(const orc_scene *s, const mtsb_bsdf_desc *b, const osi *si, v3 wo, col *val, float *pdf, const float *wav) {
    if (!b->two_sided || si->wi.z > 0.f) { bsdf_eval_pdf_nested(s, b, si, wo, val, pdf, wav); return; }
    *val = C1(0.f); *pdf = 0.f;
    if (!(si->wi.z < 0.f)) return;
    osi flipped = *si; flipped.wi.z *= -1.f; wo.z *= -1.f;
    bsdf_eval_pdf_nested(s, b, &flipped, wo, val, pdf, wav);
}
/* ---- emitters ---- */
static v3 xf_vec(const float *m, v3 v) { return V(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z, m[8] * v.x + m[9] * v.y + m[10] * v.z); }
static col env_lookup(const orc_scene *s, const oemitter *e, float u, float v, const float *wav) {     /* envmap.cpp:267-317 */
    u *= (float) (e->w - 1); v *= (float) (e->h - 1);
    uint32_t px = (uint32_t) u, py = (uint32_t) v; if (px > e->w - 2) px = e->w - 2; if (py > e->h - 2) py = e->h - 2;
    float w1x = u - (float) px, w1y = v - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    size_t idx = px + (size_t) py * e->w; col out = C1(0.f);
    const float *t00 = e->rgba + 4 * idx, *t10 = t00 + 4, *t01 = t00 + 4 * (size_t) e->w, *t11 = t01 + 4;
    if (s->spectral) {       /* texels = model coefficients + scale; result = s * wp * f * m_scale (envmap.cpp:285-306) */
        float f = fmaf(w0y, fmaf(w0x, t00[3], w1x * t10[3]), w1y * fmaf(w0x, t01[3], w1x * t11[3]));
        for (int i = 0; i < 4; ++i) {
            float s0 = fmaf(w0x, srgb_model_eval(t00, wav[i]), w1x * srgb_model_eval(t10, wav[i])),
                  s1 = fmaf(w0x, srgb_model_eval(t01, wav[i]), w1x * srgb_model_eval(t11, wav[i]));
            out.c[i] = fmaf(w0y, s0, w1y * s1) * d65_eval(wav[i]) * f * e->scale;
        }
        return out;
    }
    for (int c = 0; c < 3; ++c) {
        float v0 = fmaf(w0x, t00[c], w1x * t10[c]), v1 = fmaf(w0x, t01[c], w1x * t11[c]);
        out.c[c] = fmaf(w0y, v0, w1y * v1) * e->scale;
    }
    return out;
}
/* radiance of area / constant emitters, intensity of point lights: rgb triple, or (spectral) a spectrum texture */
static col emitter_radiance(const orc_scene *s, const oemitter *e, const float *wav) {
    return s->spectral ? tex_eval(s, e->radiance_tex, 0.f, 0.f, wav) : C3(e->radiance);
}
static void env_uv(const oemitter *e, v3 dw, float *u, float *v, v3 *dl) {   /* envmap.cpp:132-145 */
    *dl = xf_vec(e->to_local, dw);
    *u = atan2f(dl->x, -dl->z) * INV_TWOPI_F; *v = safe_acosf(dl->y) * INV_PI_F;
    *u -= floorf(*u); *v -= floorf(*v);
}
static col env_eval(const orc_scene *s, const oemitter *e, v3 dw, const float *wav) { float u, v; v3 dl; env_uv(e, dw, &u, &v, &dl); return env_lookup(s, e, u, v, wav); }
static float h2d_eval(const oemitter *e, float u, float v) {      /* distr_2d.h:495-530 */
    u = fminf(fmaxf(u, 0.f), 1.f) * (float) (e->w - 1); v = fminf(fmaxf(v, 0.f), 1.f) * (float) (e->h - 1);
    uint32_t ox = (uint32_t) (int) u, oy = (uint32_t) (int) v; if (ox > e->w - 2) ox = e->w - 2; if (oy > e->h - 2) oy = e->h - 2;
    u -= (float) (int) ox; v -= (float) (int) oy;
    const float *d0 = e->warp.levels[0]; uint32_t w0 = e->warp.widths[0]; size_t i = ox + (size_t) oy * w0;
    return lerp_(lerp_(d0[i], d0[i + 1], u), lerp_(d0[i + w0], d0[i + w0 + 1], u), v);
}
static void h2d_sample(const oemitter *e, float sx, float sy, float *u, float *v, float *pdf) {   /* distr_2d.h:317-403 */
    sx = fminf(fmaxf(sx, 0.f), 1.f); sy = fminf(fmaxf(sy, 0.f), 1.f);
    uint32_t ox = 0, oy = 0;
    for (int l = e->warp.n_levels - 2; l > 0; --l) {
        const float *d = e->warp.levels[l]; ox <<= 1; oy <<= 1;
        size_t i = h2d_index(ox, oy, e->warp.widths[l]);
        float v00 = d[i], v10 = d[i + 1], v01 = d[i + 2], v11 = d[i + 3];
        sx = fminf(fmaxf(sx, 0.f), 1.f); sy = fminf(fmaxf(sy, 0.f), 1.f);
        float r0 = v00 + v10, r1 = v01 + v11;
        sy *= r0 + r1; int mask = sy > r0;
        if (mask) { oy += 1; sy -= r0; }
        sy /= mask ? r1 : r0;
        float c0 = mask ? v01 : v00, c1 = mask ? v11 : v10;
        sx *= c0 + c1; int mx = sx > c0;
        if (mx) sx -= c0;
        sx /= mx ? c1 : c0;
        if (mx) ox += 1;
    }
    const float *d0 = e->warp.levels[0]; uint32_t w0 = e->warp.widths[0]; size_t i = ox + (size_t) oy * w0;
    float v00 = d0[i], v10 = d0[i + 1], v01 = d0[i + w0], v11 = d0[i + w0 + 1];
    float r0 = v00 + v10, r1 = v01 + v11;                      /* warp.h:398-414 */
    sy = interval_to_linear(r0, r1, sy);
    float c0 = lerp_(v00, v01, sy), c1 = lerp_(v10, v11, sy);
    sx = interval_to_linear(c0, c1, sx);
    *pdf = lerp_(c0, c1, sx);
    *u = ((float) (int) ox + sx) * (1.f / (float) (e->w - 1)); *v = ((float) (int) oy + sy) * (1.f / (float) (e->h - 1));
}
static float env_pdf_direction(const oemitter *e, v3 dw) {        /* envmap.cpp:189-206 */
    float u, v; v3 d; env_uv(e, dw, &u, &v, &d);
    float inv_sin_theta = 1.f / sqrtf(fmaxf(sqrf(d.x) + sqrf(d.z), sqrf(EPS_F)));
    return h2d_eval(e, u, v) * inv_sin_theta * (1.f / (2.f * sqrf(PI_F)));
}
typedef struct { v3 p, n, d; col value; float dist, pdf; int delta; } ods;
static ods env_sample_direction(const orc_scene *s, const oemitter *e, v3 ref_p, float sx, float sy, const float *wav) {   /* envmap.cpp:154-187 */
    ods ds; float u, v, pdf; h2d_sample(e, sx, sy, &u, &v, &pdf);
    float theta = v * PI_F, phi = u * (2.f * PI_F);
    v3 d = V(cosf(phi) * sinf(theta), sinf(phi) * sinf(theta), cosf(theta));
    d = V(d.y, d.z, -d.x);
    float dist = 2.f * e->bsphere_radius;
    float inv_sin_theta = 1.f / sqrtf(fmaxf(sqrf(d.x) + sqrf(d.z), sqrf(EPS_F)));
    d = xf_vec(e->to_world, d);
    ds.p = vadd(ref_p, vscale(d, dist)); ds.n = vneg(d); ds.d = d; ds.dist = dist;
    ds.pdf = pdf > 0.f ? pdf * inv_sin_theta * (1.f / (2.f * sqrf(PI_F))) : 0.f;
    ds.value = cscale(env_lookup(s, e, u, v, wav), 1.f / ds.pdf);
    return ds;
}
static uint32_t ek_binary_search(uint32_t start, uint32_t end, const float *cdf, float value) {   /* array_utils.h:153-171 */
    uint32_t iterations = 0;
    if (start < end) { uint32_t x = end - start, l = 0; while (x >>= 1) l++; iterations = l + 1; }
    for (uint32_t i = 0; i < iterations; ++i) {
        uint32_t middle = (start + end) >> 1;
        if (cdf[middle] < value) { start = middle + 1 < end ? middle + 1 : end; } else end = middle;
    }
    return start;
}
static v3 square_to_uniform_sphere(float sx, float sy) {    /* warp.h:262-267 */
    float z = 1.f - 2.f * sy, r = safe_sqrtf(1.f - z * z), phi = 2.f * PI_F * sx;
    return V(r * cosf(phi), r * sinf(phi), z);
}
/* Shape::sample_direction of an analytic shape -> (p, n, d, dist, pdf).  Sphere: sphere.cpp:182-255; the others go through
 * Shape::sample_direction (shape.cpp:252-270) over their sample_position (disk.cpp:127-141, cylinder.cpp:214-236,
 * rectangle.cpp:114-127). */
static void shape_sample_direction(const oshape *sh, v3 ref_p, float sx, float sy, ods *ds) {
    if (sh->type == MTSB_SHAPE_SPHERE) {
        v3 dc = vsub(sh->center, ref_p), dir;
        float dc2 = vdot(dc, dc), radius_adj = sh->radius * (sh->flip ? 1.f + RAY_EPS : 1.f - RAY_EPS);
        if (dc2 > sqrf(radius_adj)) {            /* reference point outside: uniform over the cone of directions that see the sphere */
            float inv_dc = 1.f / sqrtf(dc2), smax = sh->radius * inv_dc, smax2 = sqrf(smax), inv_smax = 1.f / smax, cmax = safe_sqrtf(1.f - smax2);
            float st2 = smax2 > 0.00068523f ? 1.f - sqrf(fmaf(cmax - 1.f, sx, 1.f)) : smax2 * sx, ct = safe_sqrtf(1.f - st2);
            float cos_alpha = st2 * inv_smax + ct * safe_sqrtf(fmaf(-st2, sqrf(inv_smax), 1.f)), sin_alpha = safe_sqrtf(fmaf(-cos_alpha, cos_alpha, 1.f));
            float sphi = sinf(sy * (2.f * PI_F)), cphi = cosf(sy * (2.f * PI_F));
            v3 fn = vscale(dc, -inv_dc), fs, ft; coordinate_system(fn, &fs, &ft);
            v3 l = V(cphi * sin_alpha, sphi * sin_alpha, cos_alpha);
            dir = vadd(vadd(vscale(fs, l.x), vscale(ft, l.y)), vscale(fn, l.z));
            ds->p = V(fmaf(dir.x, sh->radius, sh->center.x), fmaf(dir.y, sh->radius, sh->center.y), fmaf(dir.z, sh->radius, sh->center.z));
            ds->d = vsub(ds->p, ref_p);
            float dist2 = vdot(ds->d, ds->d); ds->dist = sqrtf(dist2); ds->d = vscale(ds->d, 1.f / ds->dist);
            ds->pdf = ds->dist == 0.f ? 0.f : (.5f / PI_F) / (1.f - cmax);
        } else {                                  /* inside: uniform over the surface */
            dir = square_to_uniform_sphere(sx, sy);
            ds->p = V(fmaf(dir.x, sh->radius, sh->center.x), fmaf(dir.y, sh->radius, sh->center.y), fmaf(dir.z, sh->radius, sh->center.z));
            ds->d = vsub(ds->p, ref_p);
            float dist2 = vdot(ds->d, ds->d); ds->dist = sqrtf(dist2); ds->d = vscale(ds->d, 1.f / ds->dist);
            ds->pdf = sh->inv_area * dist2 / fabsf(vdot(ds->d, dir));
        }
        ds->n = sh->flip ? vneg(dir) : dir;
        return;
    }
    if (sh->type == MTSB_SHAPE_DISK) { float x, y; disk_concentric(sx, sy, &x, &y); ds->p = xf_point(sh->to_world, V(x, y, 0.f)); ds->n = sh->n; }
    else if (sh->type == MTSB_SHAPE_CYLINDER) {
        float st = sinf(2.f * PI_F * sy), ct = cosf(2.f * PI_F * sy);
        ds->p = xf_point(sh->to_world, V(ct * sh->radius, st * sh->radius, sx * sh->length));
        v3 n = V(ct, st, 0.f); if (sh->flip) n = vneg(n);
        ds->n = n;      /* cylinder.cpp:222-232 returns the OBJECT-space normal (only the position goes through to_world); kept as is */
    } else { ds->p = xf_point(sh->to_world, V(fmaf(sx, 2.f, -1.f), fmaf(sy, 2.f, -1.f), 0.f)); ds->n = sh->n; }
    ds->d = vsub(ds->p, ref_p);
    float dist2 = vdot(ds->d, ds->d); ds->dist = sqrtf(dist2); ds->d = vscale(ds->d, 1.f / ds->dist);
    float dp = fabsf(vdot(ds->d, ds->n));
    ds->pdf = sh->inv_area * (dp != 0.f ? dist2 / dp : 0.f);
}
/* Sphere::pdf_direction (sphere.cpp:257-270); others: Shape::pdf_direction (shape.cpp:272-283) */
static float shape_pdf_direction(const oshape *sh, v3 ref_p, v3 d, v3 n, float dist) {
    float dp = fabsf(vdot(d, n));
    if (sh->type != MTSB_SHAPE_SPHERE) return sh->inv_area * (dp != 0.f ? (dist * dist) / dp : 0.f);
    float sin_alpha = sh->radius * (1.f / vnorm(vsub(sh->center, ref_p))), cos_alpha = safe_sqrtf(1.f - sin_alpha * sin_alpha);
    return sin_alpha < 1.f - EPS_F ? (.5f / PI_F) / (1.f - cos_alpha) : sh->inv_area * sqrf(dist) / dp;
}
static ods area_sample_direction(const orc_scene *s, const oemitter *e, v3 ref_p, float sx, float sy, const float *wav) {
    /* area.cpp:103-117 -> shape.cpp:252-270 -> mesh.cpp:320-365 -> distr_1d.h:144-203 */
    ods ds; const omesh *m = &s->meshes[e->mesh];
    if (m->shape >= 0) {
        memset(&ds, 0, sizeof(ds));
        shape_sample_direction(&s->shapes[m->shape], ref_p, sx, sy, &ds);
        int on = vdot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
        ds.value = on ? cscale(emitter_radiance(s, e, wav), 1.f / ds.pdf) : C1(0.f);
        return ds;
    }
    uint32_t idx = ek_binary_search(e->valid_lo, e->valid_hi, e->cdf, sy * e->cdf_sum);
    uint32_t f = m->face_begin + idx, i0 = s->faces[3 * f], i1 = s->faces[3 * f + 1], i2 = s->faces[3 * f + 2];
    v3 p0 = ld3(s->pos, i0), p1 = ld3(s->pos, i1), p2 = ld3(s->pos, i2), e0 = vsub(p1, p0), e1 = vsub(p2, p0);
    float pmf = .5f * vnorm(vcross(e0, e1)) * e->normalization, cdf_prev = idx > 0 ? e->cdf[idx - 1] * e->normalization : 0.f;
    sy = (sy - cdf_prev) / pmf;
    float t = safe_sqrtf(1.f - sx), bx = 1.f - t, by = t * sy;
    ds.p = vadd(vadd(p0, vscale(e0, bx)), vscale(e1, by));
    if (m->has_normals) ds.n = vnormalize(vadd(vadd(vscale(ld3(s->nrm, i0), 1.f - bx - by), vscale(ld3(s->nrm, i1), bx)), vscale(ld3(s->nrm, i2), by)));
    else ds.n = vnormalize(vcross(e0, e1));
    ds.d = vsub(ds.p, ref_p);
    float dist_squared = vdot(ds.d, ds.d); ds.dist = sqrtf(dist_squared); ds.d = vscale(ds.d, 1.f / ds.dist);
    float dp = fabsf(vdot(ds.d, ds.n));
    ds.pdf = e->normalization * (dp != 0.f ? dist_squared / dp : 0.f);
    int active = vdot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
    ds.value = active ? cscale(emitter_radiance(s, e, wav), 1.f / ds.pdf) : C1(0.f);
    return ds;
}
static float area_pdf_direction(const orc_scene *s, const oemitter *e, v3 ref_p, v3 d, v3 n, float dist) {   /* area.cpp:119-125, shape.cpp:272-283 */
    int shape = s->meshes[e->mesh].shape;
    if (shape >= 0) return vdot(d, n) < 0.f ? shape_pdf_direction(&s->shapes[shape], ref_p, d, n, dist) : 0.f;
    float dp = fabsf(vdot(d, n)), pdf = e->normalization * (dp != 0.f ? (dist * dist) / dp : 0.f);
    return vdot(d, n) < 0.f ? pdf : 0.f;
}
static ods constant_sample_direction(const orc_scene *s, const oemitter *e, v3 ref_p, float sx, float sy, const float *wav) {   /* constant.cpp:77-103 */
    ods ds; memset(&ds, 0, sizeof(ds));
    v3 d = square_to_uniform_sphere(sx, sy);
    ds.dist = 2.f * e->bsphere_radius; ds.p = vadd(ref_p, vscale(d, ds.dist)); ds.n = vneg(d); ds.d = d;
    ds.pdf = .25f * INV_PI_F; ds.value = cscale(emitter_radiance(s, e, wav), 1.f / ds.pdf);
    return ds;
}
static ods point_sample_direction(const orc_scene *s, const oemitter *e, v3 ref_p, const float *wav) {    /* point.cpp:74-98 */
    ods ds; memset(&ds, 0, sizeof(ds));
    ds.p = e->position; ds.pdf = 1.f; ds.delta = 1;
    ds.d = vsub(ds.p, ref_p); ds.dist = vnorm(ds.d);
    float inv_dist = 1.f / ds.dist; ds.d = vscale(ds.d, inv_dist);
    ds.value = cscale(emitter_radiance(s, e, wav), inv_dist * inv_dist);
    return ds;
}
static ods sample_emitter_direction(const orc_scene *s, v3 ref_p, float sx, float sy, const float *wav) {   /* scene.cpp:141-176 */
    ods ds; memset(&ds, 0, sizeof(ds));
    if (s->n_emitters == 0) return ds;
    uint32_t index = 0; float emitter_pdf = 1.f;
    if (s->n_emitters > 1) {
        emitter_pdf = 1.f / (float) s->n_emitters;
        index = (uint32_t) (sx * (float) s->n_emitters); if (index > s->n_emitters - 1) index = s->n_emitters - 1;
        sx = (sx - (float) index * emitter_pdf) * (float) s->n_emitters;
    }
    const oemitter *e = &s->emitters[index];
    if (e->type == MTSB_EMITTER_AREA) ds = area_sample_direction(s, e, ref_p, sx, sy, wav);
    else if (e->type == MTSB_EMITTER_ENVMAP) ds = env_sample_direction(s, e, ref_p, sx, sy, wav);
    else if (e->type == MTSB_EMITTER_CONSTANT) ds = constant_sample_direction(s, e, ref_p, sx, sy, wav);
    else ds = point_sample_direction(s, e, ref_p, wav);
    ds.delta = e->type == MTSB_EMITTER_POINT;
    if (s->n_emitters > 1) { ds.pdf *= emitter_pdf; ds.value = cscale(ds.value, 1.f / emitter_pdf); }
    return ds;
}

/* =============================================================================================== */
/* Sensor, path integrator, film                                                                    */
/* =============================================================================================== */
static void sensor_ray(const mtsb_sensor_desc *sn, float px, float py, v3 *o, v3 *d, float *mint, float *maxt) {
    /* perspective.cpp:151-188 */
    const float *m = sn->sample_to_camera;
    float w = m[12] * px + m[13] * py + m[15];
    v3 near_p = vscale(V(m[0] * px + m[1] * py + m[3], m[4] * px + m[5] * py + m[7], m[8] * px + m[9] * py + m[11]), 1.f / w);
    v3 dl = vnormalize(near_p);
    float inv_z = 1.f / dl.z;
    *mint = sn->near_clip * inv_z; *maxt = sn->far_clip * inv_z;
    *o = V(sn->to_world[3], sn->to_world[7], sn->to_world[11]); *d = xf_vec(sn->to_world, dl);
}
void orc_sensor_sample_ray(const orc_scene *s, uint32_t id, uint64_t n, const float *pos2, float *o_mint, float *d_maxt) {
    for (uint64_t i = 0; i < n; ++i) { v3 o, d; float a, b; sensor_ray(&s->sensors[id], pos2[2 * i], pos2[2 * i + 1], &o, &d, &a, &b);
        o_mint[4 * i] = o.x; o_mint[4 * i + 1] = o.y; o_mint[4 * i + 2] = o.z; o_mint[4 * i + 3] = a;
        d_maxt[4 * i] = d.x; d_maxt[4 * i + 1] = d.y; d_maxt[4 * i + 2] = d.z; d_maxt[4 * i + 3] = b; }
}
static float mis_weight(float a, float b) { a *= a; b *= b; return a > 0.f ? a / (a + b) : 0.f; }   /* path.cpp:223-227 */

/* PathIntegrator::sample, src/integrators/path.cpp:100-211 */
static void path_sample(const orc_scene *s, pcg32 *rng, v3 o, v3 d, float mint, float maxt, const float *wav, float *out5, uint64_t *cnt) {
    float eta = 1.f, emission_weight = 1.f; const int nch = s->nch;
    col throughput = C1(1.f), result = C1(0.f);
    ohit h; trace(s, o, d, mint, maxt, 0, &h); cnt[1]++;
    osi si; make_si(s, o, d, &h, &si);
    int valid_ray = si.valid;
    int emitter = si.valid ? s->meshes[si.mesh].emitter : s->env_emitter;
    v3 ray_d = d;
    for (int depth = 1;; ++depth) {
        cnt[3]++;
        if (emitter >= 0) {
            const oemitter *e = &s->emitters[emitter];
            col Le = e->type == MTSB_EMITTER_AREA ? (si.wi.z > 0.f ? emitter_radiance(s, e, wav) : C1(0.f))       /* area.cpp:71-79 */
                   : e->type == MTSB_EMITTER_ENVMAP ? env_eval(s, e, ray_d, wav) : emitter_radiance(s, e, wav);   /* envmap.cpp:132-145, constant.cpp:105-108 */
            result = cadd(result, cscale(cmul(throughput, Le), emission_weight));
        }
        if (!si.valid) break;
        if (depth > s->rr_depth) {
            float q = fminf(chmax(throughput, nch) * sqrf(eta), .95f);
            int survive = pcg_float(rng) < q;
            throughput = cscale(throughput, 1.f / q);
            if (!survive) break;
        }
        if ((uint32_t) depth >= (uint32_t) s->max_depth) break;
        const mtsb_bsdf_desc *bsdf = &s->bsdfs[s->meshes[si.mesh].bsdf];
        if (bsdf->type != MTSB_BSDF_DIELECTRIC && bsdf->type != MTSB_BSDF_CONDUCTOR) {     /* has a Smooth lobe */
            float sx = pcg_float(rng), sy = pcg_float(rng);
            if (s->n_emitters > 0) {
                ods ds = sample_emitter_direction(s, si.p, sx, sy, wav);
                if (ds.pdf != 0.f) {
                    col emitter_val = ds.value;
                    if (!cblack(emitter_val, nch)) {
                        /* visibility: scene.cpp:178-182 */
                        ohit sh; cnt[2]++;
                        if (trace(s, si.p, ds.d, RAY_EPS * (1.f + vhmax(vabs(si.p))), ds.dist * (1.f - SHADOW_EPS), 1, &sh)) emitter_val = C1(0.f);
                    }
                    v3 wo = to_local(&si, ds.d); col bsdf_val; float bsdf_pdf;
                    bsdf_eval_pdf(s, bsdf, &si, wo, &bsdf_val, &bsdf_pdf, wav);
                    float mis = ds.delta ? 1.f : mis_weight(ds.pdf, bsdf_pdf);      /* path.cpp:169 */
                    result = cadd(result, cscale(cmul(cmul(throughput, bsdf_val), emitter_val), mis));
                }
            }
        }
        float s1 = pcg_float(rng), s2x = pcg_float(rng), s2y = pcg_float(rng);
        obs bs = bsdf_sample(s, bsdf, &si, s1, s2x, s2y, wav);
        throughput = cmul(throughput, bs.weight);
        if (cblack(throughput, nch)) break;
        eta *= bs.eta;
        v3 no = si.p, nd = to_world(&si, bs.wo);
        ohit h2; trace(s, no, nd, (1.f + vhmax(vabs(si.p))) * RAY_EPS, INFINITY, 0, &h2); cnt[1]++;
        osi si_bsdf; make_si(s, no, nd, &h2, &si_bsdf);
        emitter = si_bsdf.valid ? s->meshes[si_bsdf.mesh].emitter : s->env_emitter;
        if (emitter >= 0) {
            const oemitter *e = &s->emitters[emitter];
            float emitter_pdf = 0.f;
            if (!bs.delta) {
                if (e->type == MTSB_EMITTER_AREA) {
                    v3 dd = vsub(si_bsdf.p, si.p); float dist = vnorm(dd); dd = vscale(dd, 1.f / dist);
                    emitter_pdf = area_pdf_direction(s, e, si.p, dd, si_bsdf.sh_n, dist);
                } else if (e->type == MTSB_EMITTER_ENVMAP) emitter_pdf = env_pdf_direction(e, nd);   /* ds.d = -si_bsdf.wi = ray.d on a miss (records.h:173) */
                else emitter_pdf = .25f * INV_PI_F;                                                  /* constant.cpp:98-103 */
                if (s->n_emitters > 1) emitter_pdf *= 1.f / (float) s->n_emitters;
            }
            emission_weight = mis_weight(bs.pdf, emitter_pdf);
        }
        si = si_bsdf; ray_d = nd;
    }
    for (int i = 0; i < 4; ++i) out5[i] = result.c[i];
    out5[4] = valid_ray ? 1.f : 0.f;
}

/* ImageBlock::put, src/librender/imageblock.cpp:79-172, block = full film, border 0, analytic filter eval */
static void film_put(const mtsb_sensor_desc *sn, float px, float py, const float *value5, float *xyzaw) {
    float radius = sn->rfilter == MTSB_RFILTER_BOX ? sn->rfilter_radius + RAY_EPS : sn->rfilter_radius;
    float x = px - .5f, y = py - .5f; int W = (int) sn->film_width, H = (int) sn->film_height;
    for (int k = 0; k < 5; ++k) if (!(value5[k] >= -1e-5f) || !isfinite(value5[k])) return;     /* imageblock.cpp:84-108: invalid samples are dropped */
    if (radius > 1.f) {
        int lox = (int) ceilf(x - radius), loy = (int) ceilf(y - radius); if (lox < 0) lox = 0; if (loy < 0) loy = 0;
        int hix = (int) floorf(x + radius), hiy = (int) floorf(y + radius); if (hix > W - 1) hix = W - 1; if (hiy > H - 1) hiy = H - 1;
        int n = (int) ceilf((radius - 2.f * RAY_EPS) * 2.f);
        for (int yr = 0; yr < n; ++yr) { int yy = loy + yr; if (yy > hiy) break;
            float wy = orc_rfilter_eval2(sn->rfilter, radius, sn->rfilter_stddev, sn->rfilter_c, (float) loy - y + (float) yr);
            for (int xr = 0; xr < n; ++xr) { int xx = lox + xr; if (xx > hix) break;
                float w = wy * orc_rfilter_eval2(sn->rfilter, radius, sn->rfilter_stddev, sn->rfilter_c, (float) lox - x + (float) xr);
                float *p = xyzaw + 5 * ((size_t) yy * W + xx);
                for (int k = 0; k < 5; ++k) {
                    #pragma omp atomic
                    p[k] += value5[k] * w;
                } } }
    } else {
        int lx = (int) ceilf(x - .5f), ly = (int) ceilf(y - .5f);
        if (lx >= 0 && ly >= 0 && lx < W && ly < H) { float *p = xyzaw + 5 * ((size_t) ly * W + lx);
            for (int k = 0; k < 5; ++k) {
                #pragma omp atomic
                p[k] += value5[k];
            } }
    }
}
void orc_film_put(const orc_scene *s, uint32_t id, uint64_t n, const float *pos2, const float *value5, float *xyzaw) {
    for (uint64_t i = 0; i < n; ++i) film_put(&s->sensors[id], pos2[2 * i], pos2[2 * i + 1], value5 + 5 * i, xyzaw);
}
void orc_film_develop(const orc_scene *s, uint32_t id, const float *xyzaw, float *rgb) {   /* hdrfilm.cpp:249-320, spectrum.h:230-236 */
    size_t n = (size_t) s->sensors[id].film_width * s->sensors[id].film_height;
    for (size_t i = 0; i < n; ++i) {
        float w = xyzaw[5 * i + 4], inv = w != 0.f ? 1.f / w : 0.f, X = xyzaw[5 * i] * inv, Y = xyzaw[5 * i + 1] * inv, Z = xyzaw[5 * i + 2] * inv;
        rgb[3 * i] = 3.240479f * X - 1.537150f * Y - 0.498535f * Z;
        rgb[3 * i + 1] = -0.969256f * X + 1.875991f * Y + 0.041556f * Z;
        rgb[3 * i + 2] = 0.055648f * X - 0.204043f * Y + 1.057311f * Z;
    }
}
/* SamplingIntegrator::render_sample (integrator.cpp:224-271) for stream `stream` of pixel `pixel` */
/* rad4: rgb mode = radiance rgb + alpha; spectral mode = XYZ + alpha (integrator.cpp:249-262: ray_weight * result -> spectrum_to_xyz) */
static void render_one(const orc_scene *s, const mtsb_sensor_desc *sn, uint32_t pixel, uint32_t stream, uint64_t seed, float *pos2, float *rad4, uint64_t *cnt) {
    pcg32 rng; sampler_seed(&rng, seed, stream);
    float jx = pcg_float(&rng), jy = pcg_float(&rng);
    float wavelength_sample = pcg_float(&rng);       /* integrator.cpp:237 */
    float wav[4] = { 0, 0, 0, 0 }, wav_weight[4] = { 1, 1, 1, 1 };
    if (s->spectral) sample_wavelengths(wavelength_sample, wav, wav_weight);     /* perspective.cpp:196 */
    float fx = (float) (pixel % sn->film_width) + jx, fy = (float) (pixel / sn->film_width) + jy;
    v3 o, d; float mint, maxt;
    sensor_ray(sn, fx / (float) sn->film_width, fy / (float) sn->film_height, &o, &d, &mint, &maxt);
    cnt[0]++;
    float r5[5];
    path_sample(s, &rng, o, d, mint, maxt, wav, r5, cnt);
    if (s->spectral) {
        for (int i = 0; i < 4; ++i) r5[i] = wav_weight[i] * r5[i];
        v3 xyz = spectrum_to_xyz(r5, wav); rad4[0] = xyz.x; rad4[1] = xyz.y; rad4[2] = xyz.z;
    } else { rad4[0] = r5[0]; rad4[1] = r5[1]; rad4[2] = r5[2]; }
    rad4[3] = r5[4];
    pos2[0] = fx; pos2[1] = fy;
}
void orc_trace_sample(const orc_scene *s, uint32_t id, uint32_t pixel, uint32_t stream, uint64_t seed, float *out4) {
    float pos[2]; uint64_t cnt[4] = { 0, 0, 0, 0 }; render_one(s, &s->sensors[id], pixel, stream, seed, pos, out4, cnt);
}
void orc_render(const orc_scene *s, uint32_t id, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total, uint64_t seed,
                float *xyzaw, int threads, uint64_t *counts) {
    const mtsb_sensor_desc *sn = &s->sensors[id];
    int64_t npix = (int64_t) sn->film_width * sn->film_height;
    uint64_t c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    #pragma omp parallel for schedule(dynamic, 64) num_threads(threads > 0 ? threads : 1) reduction(+ : c0, c1, c2, c3)
    for (int64_t p = 0; p < npix; ++p) {
        for (uint32_t k = 0; k < spp_count; ++k) {
            float pos[2], r[4]; uint64_t cnt[4] = { 0, 0, 0, 0 };
            uint64_t idx = (uint64_t) p * spp_total + spp_begin + k;      /* >= 2^32 samples: the high half goes into the seed */
            render_one(s, sn, (uint32_t) p, (uint32_t) idx, seed + (idx >> 32), pos, r, cnt);
            float v5[5] = { 0.412453f * r[0] + 0.357580f * r[1] + 0.180423f * r[2], 0.212671f * r[0] + 0.715160f * r[1] + 0.072169f * r[2],
                            0.019334f * r[0] + 0.119193f * r[1] + 0.950227f * r[2], r[3], 1.f };     /* spectrum.h:220-227 */
            if (s->spectral) { v5[0] = r[0]; v5[1] = r[1]; v5[2] = r[2]; }
            film_put(sn, pos[0], pos[1], v5, xyzaw);
            c0 += cnt[0]; c1 += cnt[1]; c2 += cnt[2]; c3 += cnt[3];
        }
    }
    if (counts) { counts[0] = c0; counts[1] = c1; counts[2] = c2; counts[3] = c3; }
}

/* ---- batched stage hooks ---- */
static void local_si(osi *si, const float *wi, const float *uv) {
    memset(si, 0, sizeof(*si)); si->n = V(0, 0, 1); si->sh_s = V(1, 0, 0); si->sh_t = V(0, 1, 0); si->sh_n = V(0, 0, 1);
    si->wi = V(wi[0], wi[1], wi[2]); si->u = uv[0]; si->v = uv[1]; si->t = 1.f; si->valid = 1;
}
void orc_bsdf_sample(const orc_scene *s, uint32_t id, uint64_t n, const float *wi, const float *uv, const float *s1, const float *s2, float *out9) {
    for (uint64_t i = 0; i < n; ++i) { osi si; local_si(&si, wi + 3 * i, uv + 2 * i);
        obs bs = bsdf_sample(s, &s->bsdfs[id], &si, s1[i], s2[2 * i], s2[2 * i + 1], NULL); float *o = out9 + 9 * i;
        o[0] = bs.wo.x; o[1] = bs.wo.y; o[2] = bs.wo.z; o[3] = bs.pdf; o[4] = bs.eta; o[5] = bs.weight.c[0]; o[6] = bs.weight.c[1]; o[7] = bs.weight.c[2]; o[8] = bs.delta ? 1.f : 0.f; }
}
void orc_bsdf_eval_pdf(const orc_scene *s, uint32_t id, uint64_t n, const float *wi, const float *uv, const float *wo, float *eval3, float *pdf) {
    for (uint64_t i = 0; i < n; ++i) { osi si; local_si(&si, wi + 3 * i, uv + 2 * i); col val;
        bsdf_eval_pdf(s, &s->bsdfs[id], &si, V(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), &val, &pdf[i], NULL);
        eval3[3 * i] = val.c[0]; eval3[3 * i + 1] = val.c[1]; eval3[3 * i + 2] = val.c[2]; }
}
void orc_sample_emitter_direction(const orc_scene *s, uint64_t n, const float *ref_p, const float *s2, float *out11) {
    for (uint64_t i = 0; i < n; ++i) { ods ds = sample_emitter_direction(s, V(ref_p[3 * i], ref_p[3 * i + 1], ref_p[3 * i + 2]), s2[2 * i], s2[2 * i + 1], NULL);
        float *o = out11 + 11 * i; o[0] = ds.d.x; o[1] = ds.d.y; o[2] = ds.d.z; o[3] = ds.dist; o[4] = ds.pdf; o[5] = ds.value.c[0]; o[6] = ds.value.c[1]; o[7] = ds.value.c[2];
        o[8] = ds.n.x; o[9] = ds.n.y; o[10] = ds.n.z; }
}

/* Mesh::parameters_changed (mesh.cpp:794-804) + recompute_vertex_normals (mesh.cpp:199-276) */
void orc_scene_update_vertices(orc_scene *s, uint32_t mesh_id, const float *positions) {
    omesh *m = &s->meshes[mesh_id];
    memcpy(s->pos + 3 * (size_t) m->vert_begin, positions, sizeof(float) * 3 * (size_t) m->vert_count);
    if (m->has_normals) {
        double *acc = calloc(3 * (size_t) m->vert_count + 1, sizeof(double));
        for (uint32_t f = m->face_begin; f < m->face_begin + m->face_count; ++f) {
            uint32_t idx[3] = { s->faces[3 * f], s->faces[3 * f + 1], s->faces[3 * f + 2] };
            v3 v[3] = { ld3(s->pos, idx[0]), ld3(s->pos, idx[1]), ld3(s->pos, idx[2]) };
            v3 n = vcross(vsub(v[1], v[0]), vsub(v[2], v[0])); float l2 = vdot(n, n);
            if (!(l2 > 0.f)) continue;
            n = vscale(n, 1.f / sqrtf(l2));
            for (int k = 0; k < 3; ++k) {
                v3 d0 = vnormalize(vsub(v[(k + 1) % 3], v[k])), d1 = vnormalize(vsub(v[(k + 2) % 3], v[k]));
                float dp = vdot(d0, d1);
                v3 df = vsub(d1, dp < 0.f ? vneg(d0) : d0);
                float temp = 2.f * asinf(fminf(1.f, .5f * vnorm(df))), ang = dp >= 0.f ? temp : PI_F - temp;   /* enoki unit_angle */
                size_t li = idx[k] - m->vert_begin;
                acc[3 * li] += n.x * ang; acc[3 * li + 1] += n.y * ang; acc[3 * li + 2] += n.z * ang;
            }
        }
        for (uint32_t i = 0; i < m->vert_count; ++i) {
            double x = acc[3 * i], y = acc[3 * i + 1], z = acc[3 * i + 2], l = sqrt(x * x + y * y + z * z);
            float *o = s->nrm + 3 * (size_t) (m->vert_begin + i);
            if (l != 0.0) { o[0] = (float) (x / l); o[1] = (float) (y / l); o[2] = (float) (z / l); } else { o[0] = 1.f; o[1] = 0.f; o[2] = 0.f; }
        }
        free(acc);
    }
    build_accel(s);
    for (uint32_t i = 0; i < s->n_emitters; ++i) if (s->emitters[i].type == MTSB_EMITTER_AREA) build_area_cdf(s, &s->emitters[i]);
    update_bsphere(s);
}

/* =============================================================================================== */
/* Distribution hooks (unit parity against the reference's Hierarchical2D / DiscreteDistribution)   */
/* =============================================================================================== */
static void free_h2d(oh2d *t) { for (int i = 0; i < t->n_levels; ++i) free(t->levels[i]); free(t->levels); free(t->widths); }
void orc_h2d_sample(const float *data, uint32_t w, uint32_t h, uint64_t n, const float *s2, float *out3) {
    oemitter e; memset(&e, 0, sizeof(e)); e.w = w; e.h = h; build_h2d(&e.warp, data, w, h);
    for (uint64_t i = 0; i < n; ++i) h2d_sample(&e, s2[2 * i], s2[2 * i + 1], &out3[3 * i], &out3[3 * i + 1], &out3[3 * i + 2]);
    free_h2d(&e.warp);
}
void orc_h2d_eval(const float *data, uint32_t w, uint32_t h, uint64_t n, const float *uv, float *out) {
    oemitter e; memset(&e, 0, sizeof(e)); e.w = w; e.h = h; build_h2d(&e.warp, data, w, h);
    for (uint64_t i = 0; i < n; ++i) out[i] = h2d_eval(&e, uv[2 * i], uv[2 * i + 1]);
    free_h2d(&e.warp);
}
/* DiscreteDistribution::sample_reuse (distr_1d.h:49-88 update, :144-203 sample): the same arithmetic the area-light
 * triangle pick uses (area_sample_direction above), exposed on a bare pmf */
void orc_discrete_sample_reuse(const float *pmf, uint32_t n_pmf, uint64_t n, const float *u, uint32_t *idx, float *u_out) {
    float *cdf = malloc(sizeof(float) * (n_pmf ? n_pmf : 1)); double sum = 0.0; uint32_t lo = (uint32_t) -1, hi = (uint32_t) -1;
    for (uint32_t i = 0; i < n_pmf; ++i) { sum += (double) pmf[i]; cdf[i] = (float) sum; if (pmf[i] > 0.f) { if (lo == (uint32_t) -1) lo = i; hi = i; } }
    float fsum = (float) sum, norm = (float) (1.0 / sum);
    for (uint64_t k = 0; k < n; ++k) {
        uint32_t j = ek_binary_search(lo, hi, cdf, u[k] * fsum);
        float p = pmf[j] * norm, prev = j > 0 ? cdf[j - 1] * norm : 0.f;
        idx[k] = j; u_out[k] = (u[k] - prev) / p;
    }
    free(cdf);
}
