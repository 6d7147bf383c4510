/*
 * mts_oracle.h -- CPU ORACLE for the mtsb200 hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C, scalar restatement of the reference's algorithm for the path (Mitsuba 2, `scalar_rgb`
 * arithmetic: float32 throughout, double only where the reference uses double).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library;
 * the product (mitsuba2_b200/) never does.  Every function cites the reference file:line it follows.
 *
 * Pinning status (see DESIGN.md "Oracle"): RNG, TEA, warps, Fresnel, microfacet, filters, area-light and
 * intersection KATs from the reference's own tests are checked in tests/test_oracle_kat.py; the header-only
 * reference math (enoki PCG32, mitsuba warp/fresnel/microfacet templates) is additionally compiled from
 * /root/reference into oracle/_ref/ and compared in tests/test_oracle_vs_ref.py + tests/golden/.
 * End-to-end images: the reference binary cannot be built under this tier's rules (needs cmake + generated
 * config.h), so image-level parity is pinned only by the reference's stored golden means where a scene is
 * expressible; gradients: "parity unpinned" (no reference implementation exists).
 */
#ifndef MTS_ORACLE_H
#define MTS_ORACLE_H
#include "../include/mtsb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_scene orc_scene;

/* PCG32 / TEA / sampler seeding */
void orc_pcg32_floats(uint64_t n, const uint64_t *initstate, const uint64_t *initseq, uint32_t n_draws, float *out);
uint32_t orc_tea32(uint32_t v0, uint32_t v1, int rounds);
uint64_t orc_tea64(uint32_t v0, uint32_t v1, int rounds);
void orc_sampler_seed(uint64_t n, uint64_t seed, uint64_t *state, uint64_t *inc);

/* warps + small math (unit KATs) */
void orc_square_to_uniform_disk_concentric(float sx, float sy, float *out2);
void orc_square_to_cosine_hemisphere(float sx, float sy, float *out3);
void orc_square_to_uniform_triangle(float sx, float sy, float *out2);
void orc_coordinate_system(const float *n, float *s, float *t);
void orc_fresnel(float cos_theta_i, float eta, float *out4 /* r, cos_theta_t, eta_it, eta_ti */);
float orc_fresnel_conductor(float cos_theta_i, float eta_r, float eta_i);
float orc_microfacet_eval(int type, float au, float av, const float *m);
float orc_microfacet_pdf(int type, float au, float av, int visible, const float *wi, const float *m);
float orc_microfacet_smith_g1(int type, float au, float av, const float *v, const float *m);
void orc_microfacet_sample(int type, float au, float av, int visible, const float *wi, float sx, float sy, float *out4 /* m, pdf */);
float orc_erfinv(float x);
float orc_rfilter_eval(int type, float radius, float stddev, float x);
float orc_rfilter_eval2(int type, float radius, float stddev, float c, float x);   /* + mitchell (B = stddev, C = c), catmullrom, lanczos */
void orc_perspective_sample_to_camera(float x_fov_deg, float near_clip, float far_clip, float aspect, float *out16);

/* distribution hooks: Hierarchical2D (distr_2d.h) over a w x h float image, DiscreteDistribution::sample_reuse (distr_1d.h) */
void orc_h2d_sample(const float *data, uint32_t w, uint32_t h, uint64_t n, const float *s2, float *out3 /* u, v, pdf */);
void orc_h2d_eval(const float *data, uint32_t w, uint32_t h, uint64_t n, const float *uv, float *out);
void orc_discrete_sample_reuse(const float *pmf, uint32_t n_pmf, uint64_t n, const float *u, uint32_t *idx, float *u_out);

/* scene */
orc_scene *orc_scene_create(const mtsb_scene_desc *desc);
void orc_scene_destroy(orc_scene *s);
void orc_scene_update_vertices(orc_scene *s, uint32_t mesh_id, const float *positions);
void orc_scene_update_texture(orc_scene *s, uint32_t tex_id, const float *data);

/* Scene::ray_intersect / ray_test / ray_intersect_naive with the reference's Moeller-Trumbore test */
void orc_intersect_closest(const orc_scene *s, uint64_t n, const float *o_mint, const float *d_maxt, float *hits, int threads);
void orc_intersect_any(const orc_scene *s, uint64_t n, const float *o_mint, const float *d_maxt, uint32_t *occluded, int threads);
void orc_surface_interaction(const orc_scene *s, uint64_t n, const float *o_mint, const float *d_maxt, float *out16 /* t, uv, prim bits, p, n, sh_n, wi */, int threads);
void orc_intersect_naive(const orc_scene *s, uint64_t n, const float *o_mint, const float *d_maxt, float *hits, int threads);

/* stage hooks (same signatures as the mtsb_* parity hooks) */
void orc_sensor_sample_ray(const orc_scene *s, uint32_t sensor_id, uint64_t n, const float *pos2, float *o_mint, float *d_maxt);
void orc_bsdf_sample(const orc_scene *s, uint32_t bsdf_id, uint64_t n, const float *wi, const float *uv, const float *s1, const float *s2, float *out9);
void orc_bsdf_eval_pdf(const orc_scene *s, uint32_t bsdf_id, uint64_t n, const float *wi, const float *uv, const float *wo, float *eval3, float *pdf);
void orc_sample_emitter_direction(const orc_scene *s, uint64_t n, const float *ref_p, const float *s2, float *out11);
void orc_film_put(const orc_scene *s, uint32_t sensor_id, uint64_t n, const float *pos2, const float *value5, float *xyzaw);
void orc_film_develop(const orc_scene *s, uint32_t sensor_id, const float *xyzaw, float *rgb);

/* SamplingIntegrator::render with per-sample PCG32 streams (the reference's wavefront seeding); accumulates
 * into xyzaw (width*height*5).  counts: [samples, closest rays, shadow rays, segments] (may be NULL) */
void orc_render(const orc_scene *s, uint32_t sensor_id, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total,
                uint64_t seed, float *xyzaw, int threads, uint64_t *counts);
/* one path, for debugging parity: returns radiance rgb + alpha of stream `stream` at pixel `pixel` */
void orc_trace_sample(const orc_scene *s, uint32_t sensor_id, uint32_t pixel, uint32_t stream, uint64_t seed, float *out4);

/* spectral mode arithmetic (scalar_spectral): core/spectrum.h:147-217, 270-316; render/srgb.h:8-24; core/distr_1d.h:378-393, 655-677;
 * src/spectra/d65.cpp */
void orc_sample_wavelength(float u, float *wav4, float *weight4);
float orc_srgb_model_eval(const float *coeff3, float lambda);
float orc_regular_eval(const float *values, uint32_t n, float lo, float hi, float x);
float orc_irregular_eval(const float *nodes, const float *values, uint32_t n, float x);
float orc_d65_eval(float lambda);
void orc_spectrum_to_xyz(const float *value4, const float *wav4, float *xyz);

#ifdef __cplusplus
}
#endif
#endif
