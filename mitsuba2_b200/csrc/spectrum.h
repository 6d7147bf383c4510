// spectrum.h -- colour representation of the two render modes and the spectral-mode arithmetic (HD bodies).
//   Rgb      : Spectrum = Color<float,3>  (scalar_rgb / cuda_b200_rgb)
//   Spectral : Spectrum = 4 wavelengths   (scalar_spectral / cuda_b200_spectral), include/mitsuba/core/spectrum.h
#pragma once
#include "scene.h"

namespace mtsb {

// ---- 4-wide arithmetic (a Spectrum<float,4> or its 4 wavelengths) ----------------------------------------------
HD float4 mk4(float s) { return make_float4(s, s, s, s); }
HD float4 operator+(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
HD float4 operator*(float4 a, float4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
HD float4 operator*(float4 a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }
HD float4 operator/(float4 a, float s) { return make_float4(a.x / s, a.y / s, a.z / s, a.w / s); }
HD float4 &operator+=(float4 &a, float4 b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; return a; }
HD float4 &operator*=(float4 &a, float4 b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; a.w *= b.w; return a; }
HD float4 &operator*=(float4 &a, float s) { a.x *= s; a.y *= s; a.z *= s; a.w *= s; return a; }
HD float hmax(float4 a) { return fmaxf(fmaxf(a.x, a.y), fmaxf(a.z, a.w)); }
HD bool is_black(float3 a) { return a.x == 0.f && a.y == 0.f && a.z == 0.f; }
HD bool is_black(float4 a) { return a.x == 0.f && a.y == 0.f && a.z == 0.f && a.w == 0.f; }
template <typename F> HD float3 map_spec(float3 a, float3 b, F f) { return mk3(f(a.x, b.x), f(a.y, b.y), f(a.z, b.z)); }
template <typename F> HD float4 map_spec(float4 a, float4 b, F f) { return make_float4(f(a.x, b.x), f(a.y, b.y), f(a.z, b.z), f(a.w, b.w)); }
template <typename F> HD float4 map_spec(float4 a, F f) { return make_float4(f(a.x), f(a.y), f(a.z), f(a.w)); }

// `ext`: the stage bodies also handle the SURVEY-8f-N4 plugins (smooth conductor, plastic, rough dielectric, mask).  Scenes
// that use none of them run the lean instances, whose shade kernel keeps its register budget (48 registers / 10 blocks per SM
// in rgb); the host picks the instance per scene.
struct Rgb      { typedef float3 Spec; static constexpr bool spectral = false, ext = false; static HD float3 splat(float v) { return mk3(v); } };
struct Spectral { typedef float4 Spec; static constexpr bool spectral = true,  ext = false; static HD float4 splat(float v) { return mk4(v); } };
struct RgbX : Rgb           { static constexpr bool ext = true; };
struct SpectralX : Spectral { static constexpr bool ext = true; };

// ---- wavelength sampling: sample_wavelength = sample_rgb_spectrum(sample_shifted(u)) (core/spectrum.h:270-285, 304-316;
//      core/math.h:418-442) --------------------------------------------------------------------------------------
HD void sample_wavelengths(float u, float4 &wav, float4 &weight) {
    float4 s = make_float4(u, u + 0.25f, u + 0.5f, u + 0.75f);
    s = map_spec(s, [](float v) { return v > 1.f ? v - 1.f : v; });
    wav = map_spec(s, [](float v) { return 538.f - atanhf(0.8569106254698279f - 1.8275019724092267f * v) * 138.88888888888889f; });
    weight = map_spec(wav, [](float l) { float t = coshf(0.0072f * (l - 538.f)); return 253.82f * t * t; });
}

// ---- srgb_model_eval (include/mitsuba/render/srgb.h:8-24): sigmoid of a quadratic in the wavelength ---------------
HD float srgb_model_eval1(float3 c, float l) {
    float v = fmaf(fmaf(c.x, l, c.y), l, c.z);
    if (isinf(c.z)) return fmaf(copysignf(1.f, c.z), 0.5f, 0.5f);
    return fmaxf(0.f, fmaf(0.5f * v, 1.f / sqrtf(fmaf(v, v, 1.f)), 0.5f));
}
HD float4 srgb_model_eval(float3 c, float4 wav) {
    return make_float4(srgb_model_eval1(c, wav.x), srgb_model_eval1(c, wav.y), srgb_model_eval1(c, wav.z), srgb_model_eval1(c, wav.w));
}

// ---- tabulated spectra: ContinuousDistribution::eval_pdf (core/distr_1d.h:378-393, `regular` plugin) and
//      IrregularContinuousDistribution::eval_pdf (:655-677, `irregular` plugin) ----------------------------------
HD float regular_eval1(const float *values, uint32_t n, float lo, float hi, float inv_interval, float x) {
    if (!(x >= lo && x <= hi)) return 0.f;
    x = (x - lo) * inv_interval;
    uint32_t idx = min((uint32_t) max((int) x, 0), n - 2);
    float w1 = x - (float) idx, w0 = 1.f - w1;
    return fmaf(w0, values[idx], w1 * values[idx + 1]);
}
// buckets (optional, host_tables.h): nb start indices + scale after the values; the interval found is the one the reference's
// binary search finds (first node >= x, minus one), so the value is identical
HD float irregular_eval1(const float *nodes, const float *values, uint32_t n, float x, const float *buckets = nullptr, uint32_t nb = 0) {
    if (!(x >= nodes[0] && x <= nodes[n - 1])) return 0.f;
    uint32_t idx;
    if (nb != 0) {
        uint32_t j = min((uint32_t) ((x - nodes[0]) * buckets[nb]), nb - 1u);
        idx = (uint32_t) buckets[j];
        while (idx + 2 < n && nodes[idx + 1] < x) ++idx;
    } else {
        uint32_t start = 0, end = n;                               // enoki::binary_search (array_utils.h:153-171)
        uint32_t it = (uint32_t) (31 - clz32(n)) + 1u;
        for (uint32_t i = 0; i < it; ++i) {
            uint32_t mid = (start + end) >> 1;
            if (nodes[min(mid, n - 1)] < x && mid < n) start = min(mid + 1, end); else end = mid;
        }
        idx = max(min(start, n - 1), 1u) - 1u;
    }
    float x0 = nodes[idx], x1 = nodes[idx + 1], y0 = values[idx], y1 = values[idx + 1];
    x = (x - x0) / (x1 - x0);
    return fmaf(x, y1 - y0, y0);
}
// D65 whitepoint spectrum, Texture::D65(1) = `regular` over kCieD65 / 10568 (src/spectra/d65.cpp:56-73)
HD float4 d65_eval(const SceneDev &sc, float4 wav) {
    const float *t = sc.cie + 3 * 95;
    return map_spec(wav, [&](float l) { return regular_eval1(t, 95, 360.f, 830.f, 94.f / 470.f, l); });
}

// ---- cie1931_xyz + spectrum_to_xyz (core/spectrum.h:147-176, 209-217) ---------------------------------------------
HD float3 cie1931_xyz1(const float *tbl, float l) {
    float t = (l - 360.f) * (94.f / (830.f - 360.f));
    if (!(l >= 360.f && l <= 830.f)) return mk3(0.f);
    int i0 = min(max((int) t, 0), 93);
    float w1 = t - (float) i0, w0 = 1.f - w1;
    return mk3(fmaf(w0, tbl[i0], w1 * tbl[i0 + 1]), fmaf(w0, tbl[95 + i0], w1 * tbl[95 + i0 + 1]),
               fmaf(w0, tbl[190 + i0], w1 * tbl[190 + i0 + 1]));
}
HD float3 spectrum_to_xyz(const SceneDev &sc, float4 value, float4 wav) {
    float3 a = cie1931_xyz1(sc.cie, wav.x), b = cie1931_xyz1(sc.cie, wav.y), c = cie1931_xyz1(sc.cie, wav.z), d = cie1931_xyz1(sc.cie, wav.w);
    // hmean = hsum * (1/4), hsum of a 4-array = (v0+v1)+(v2+v3)
    return mk3(((a.x * value.x + b.x * value.y) + (c.x * value.z + d.x * value.w)) * 0.25f,
               ((a.y * value.x + b.y * value.y) + (c.y * value.z + d.y * value.w)) * 0.25f,
               ((a.z * value.x + b.z * value.y) + (c.z * value.z + d.z * value.w)) * 0.25f);
}

} // namespace mtsb
