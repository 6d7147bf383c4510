// ref_harness.cpp -- TEST INFRASTRUCTURE.  C wrappers around the REFERENCE's own header-only arithmetic for the hot
// path, compiled from the headers where they lie under /root/reference (see oracle/Makefile, target `ref`) into
// oracle/_ref/libmts_ref.so.  It is used to (a) validate the plain-C restatement in mts_oracle.c and (b) generate the
// golden vectors in tests/golden/ref_vectors.npz (tests/golden/make_ref_vectors.py).  The full reference (plugins,
// scene loading, kd-tree, integrator) needs its cmake build + generated sources and is treated as unbuildable here.
//
// Each ref_* function has the signature of the orc_* function it pins.
#include <mitsuba/core/warp.h>
#include <mitsuba/core/random.h>
#include <mitsuba/core/vector.h>
#include <mitsuba/core/transform.h>
#include <mitsuba/core/distr_1d.h>
#include <mitsuba/core/distr_2d.h>
#include <mitsuba/core/bbox.h>
#include <mitsuba/core/ray.h>
#include <mitsuba/core/spectrum.h>
#include <mitsuba/render/fresnel.h>
#include <mitsuba/render/microfacet.h>
#include <mitsuba/render/srgb.h>
#include <enoki/random.h>
#include <enoki/special.h>
#include <cstdint>
#include <cstring>

#include <stdexcept>
// The two libmitsuba-core symbols the headers reference (error reporting); the harness turns errors into exceptions.
#undef Throw
namespace mitsuba {
const Class *m_class = nullptr;
namespace detail {
[[noreturn]] void Throw(LogLevel, const Class *, const char *file, int line, const std::string &msg) {
    throw std::runtime_error(std::string(file) + ":" + std::to_string(line) + ": " + msg);
}
}  // namespace detail
}  // namespace mitsuba

// The CIE 1931 table pointers live in libmitsuba-core (src/libcore/spectrum.cpp:187-189); the harness points them at the
// same published data (mitsuba2_b200/csrc/cie_tables.h is generated from that file, tools/gen_cie_tables.py).
#include "../mitsuba2_b200/csrc/cie_tables.h"
namespace mitsuba {
const float *cie1931_x_data = kCie1931, *cie1931_y_data = kCie1931 + 95, *cie1931_z_data = kCie1931 + 190;
}  // namespace mitsuba

using namespace mitsuba;
using Float = float;
using Spec = Color<float, 3>;
using P2 = Point<float, 2>;
using V3 = Vector<float, 3>;
using N3 = Normal<float, 3>;
using MF = MicrofacetDistribution<Float, Spec>;

extern "C" {

// enoki::PCG32 (ext/enoki/include/enoki/random.h:62-101)
void ref_pcg32_floats(uint64_t n, const uint64_t *initstate, const uint64_t *initseq, uint32_t n_draws, float *out) {
    for (uint64_t i = 0; i < n; ++i) {
        enoki::PCG32<float> rng(initstate[i], initseq[i]);
        for (uint32_t k = 0; k < n_draws; ++k) out[i * n_draws + k] = rng.next_float32();
    }
}
void ref_pcg32_default_floats(uint32_t n_draws, float *out) {
    enoki::PCG32<float> rng;
    for (uint32_t k = 0; k < n_draws; ++k) out[k] = rng.next_float32();
}
// sample_tea_32 / sample_tea_64 (include/mitsuba/core/random.h:67-107)
uint32_t ref_tea32(uint32_t v0, uint32_t v1, int rounds) { return sample_tea_32<uint32_t>(v0, v1, rounds); }
uint64_t ref_tea64(uint32_t v0, uint32_t v1, int rounds) { return sample_tea_64<uint32_t>(v0, v1, rounds); }

// warps (include/mitsuba/core/warp.h)
void ref_square_to_uniform_disk_concentric(float sx, float sy, float *o) { auto p = warp::square_to_uniform_disk_concentric(P2(sx, sy)); o[0] = p.x(); o[1] = p.y(); }
void ref_square_to_cosine_hemisphere(float sx, float sy, float *o) { auto p = warp::square_to_cosine_hemisphere(P2(sx, sy)); o[0] = p.x(); o[1] = p.y(); o[2] = p.z(); }
void ref_square_to_uniform_triangle(float sx, float sy, float *o) { auto p = warp::square_to_uniform_triangle(P2(sx, sy)); o[0] = p.x(); o[1] = p.y(); }
void ref_square_to_uniform_sphere(float sx, float sy, float *o) { auto p = warp::square_to_uniform_sphere(P2(sx, sy)); o[0] = p.x(); o[1] = p.y(); o[2] = p.z(); }
// coordinate_system (include/mitsuba/core/vector.h:116-136)
void ref_coordinate_system(const float *n, float *s, float *t) {
    auto [a, b] = coordinate_system(V3(n[0], n[1], n[2]));
    s[0] = a.x(); s[1] = a.y(); s[2] = a.z(); t[0] = b.x(); t[1] = b.y(); t[2] = b.z();
}
// fresnel / fresnel_conductor (include/mitsuba/render/fresnel.h:34-116)
void ref_fresnel(float cos_theta_i, float eta, float *o) {
    auto [r, ct, eit, eti] = fresnel(cos_theta_i, eta);
    o[0] = r; o[1] = ct; o[2] = eit; o[3] = eti;
}
float ref_fresnel_conductor(float cos_theta_i, float eta_r, float eta_i) {
    return fresnel_conductor(cos_theta_i, enoki::Complex<float>(eta_r, eta_i));
}
// MicrofacetDistribution (include/mitsuba/render/microfacet.h:184-413); type: 0 = Beckmann, 1 = GGX
static MF make_mf(int type, float au, float av, int visible) { return MF(type == 0 ? MicrofacetType::Beckmann : MicrofacetType::GGX, au, av, visible != 0); }
float ref_microfacet_eval(int type, float au, float av, const float *m) { return make_mf(type, au, av, 1).eval(V3(m[0], m[1], m[2])); }
float ref_microfacet_pdf(int type, float au, float av, int visible, const float *wi, const float *m) {
    return make_mf(type, au, av, visible).pdf(V3(wi[0], wi[1], wi[2]), V3(m[0], m[1], m[2]));
}
float ref_microfacet_smith_g1(int type, float au, float av, const float *v, const float *m) {
    return make_mf(type, au, av, 1).smith_g1(V3(v[0], v[1], v[2]), V3(m[0], m[1], m[2]));
}
void ref_microfacet_sample(int type, float au, float av, int visible, const float *wi, float sx, float sy, float *o) {
    auto [m, pdf] = make_mf(type, au, av, visible).sample(V3(wi[0], wi[1], wi[2]), P2(sx, sy));
    o[0] = m.x(); o[1] = m.y(); o[2] = m.z(); o[3] = pdf;
}
float ref_erfinv(float x) { return enoki::erfinv(x); }
float ref_erf(float x) { return enoki::erf(x); }

// Transform::perspective + the sensor's sample_to_camera chain (include/mitsuba/render/sensor.h perspective_projection)
void ref_perspective_sample_to_camera(float x_fov_deg, float near_clip, float far_clip, float aspect, float *out16) {
    using T4 = Transform<Point<float, 4>>;
    T4 t = T4::scale(V3(-0.5f, -0.5f * aspect, 1.f)) * T4::translate(V3(-1.f, -1.f / aspect, 0.f)) * T4::perspective(x_fov_deg, near_clip, far_clip);
    T4 inv = t.inverse();
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out16[4 * r + c] = inv.matrix(r, c);
}

// Hierarchical2D<float, 0> (include/mitsuba/core/distr_2d.h:179-530)
void ref_h2d_sample(const float *data, uint32_t w, uint32_t h, uint64_t n, const float *s2, float *out3) {
    Hierarchical2D<float, 0> d(data, Vector<uint32_t, 2>(w, h));
    for (uint64_t i = 0; i < n; ++i) {
        auto [p, pdf] = d.sample(P2(s2[2 * i], s2[2 * i + 1]));
        out3[3 * i] = p.x(); out3[3 * i + 1] = p.y(); out3[3 * i + 2] = pdf;
    }
}
void ref_h2d_eval(const float *data, uint32_t w, uint32_t h, uint64_t n, const float *uv, float *out) {
    Hierarchical2D<float, 0> d(data, Vector<uint32_t, 2>(w, h));
    for (uint64_t i = 0; i < n; ++i) out[i] = d.eval(P2(uv[2 * i], uv[2 * i + 1]));
}
// DiscreteDistribution<float>::sample_reuse (include/mitsuba/core/distr_1d.h:144-203)
void ref_discrete_sample_reuse(const float *pmf, uint32_t n_pmf, uint64_t n, const float *u, uint32_t *idx, float *u_out) {
    DiscreteDistribution<float> d(pmf, n_pmf);
    for (uint64_t k = 0; k < n; ++k) { auto [i, r] = d.sample_reuse(u[k]); idx[k] = i; u_out[k] = r; }
}
// BoundingBox::ray_intersect (include/mitsuba/core/bbox.h:302-325): out = {hit, mint, maxt}
void ref_bbox_ray_intersect(const float *lo, const float *hi, const float *o, const float *d, float *out3) {
    BoundingBox<Point<float, 3>> bb(Point<float, 3>(lo[0], lo[1], lo[2]), Point<float, 3>(hi[0], hi[1], hi[2]));
    Ray<Point<float, 3>, Spec> ray; ray.o = Point<float, 3>(o[0], o[1], o[2]); ray.d = V3(d[0], d[1], d[2]); ray.update();
    auto [hit, mint, maxt] = bb.ray_intersect(ray);
    out3[0] = hit ? 1.f : 0.f; out3[1] = mint; out3[2] = maxt;
}
// srgb_to_xyz / xyz_to_srgb (include/mitsuba/core/spectrum.h:220-245)
void ref_srgb_to_xyz(const float *rgb, float *xyz) { auto r = srgb_to_xyz(Color<float, 3>(rgb[0], rgb[1], rgb[2])); xyz[0] = r.x(); xyz[1] = r.y(); xyz[2] = r.z(); }
void ref_xyz_to_srgb(const float *xyz, float *rgb) { auto r = xyz_to_srgb(Color<float, 3>(xyz[0], xyz[1], xyz[2])); rgb[0] = r.x(); rgb[1] = r.y(); rgb[2] = r.z(); }

// ---- spectral variants (Spectrum = mitsuba::Spectrum<float, 4>) ----
using Spec4 = mitsuba::Spectrum<float, 4>;
// sample_wavelength (include/mitsuba/core/spectrum.h:304-316)
void ref_sample_wavelength(float u, float *wav4, float *weight4) {
    auto [w, wt] = sample_wavelength<float, Spec4>(u);
    for (int i = 0; i < 4; ++i) { wav4[i] = w[i]; weight4[i] = wt[i]; }
}
// srgb_model_eval (include/mitsuba/render/srgb.h:8-24), one wavelength replicated over the 4 lanes
float ref_srgb_model_eval(const float *c, float lambda) {
    return srgb_model_eval<Spec4>(Array<float, 3>(c[0], c[1], c[2]), Spec4(lambda))[0];
}
// ContinuousDistribution / IrregularContinuousDistribution::eval_pdf (include/mitsuba/core/distr_1d.h:378-393, 655-677)
float ref_regular_eval(const float *values, uint32_t n, float lo, float hi, float x) {
    ContinuousDistribution<float> d(Vector<float, 2>(lo, hi), values, n);
    return d.eval_pdf(x);
}
float ref_irregular_eval(const float *nodes, const float *values, uint32_t n, float x) {
    IrregularContinuousDistribution<float> d(nodes, values, n);
    return d.eval_pdf(x);
}
// spectrum_to_xyz (include/mitsuba/core/spectrum.h:209-217)
void ref_spectrum_to_xyz(const float *value4, const float *wav4, float *xyz) {
    Spec4 v(value4[0], value4[1], value4[2], value4[3]), w(wav4[0], wav4[1], wav4[2], wav4[3]);
    auto r = spectrum_to_xyz(v, w);
    xyz[0] = r.x(); xyz[1] = r.y(); xyz[2] = r.z();
}

}  // extern "C"
