/*
 * mtsb200.h -- C ABI of the B200-native rendering backend for Mitsuba 2's hot path.
 *
 * Drop-in boundary.  Every entry point cites the reference
 * interface it replaces (paths relative to the Mitsuba 2 tree).  Plain pointers and
 * sizes only: no C++ types, no torch types.  All functions return 0 on success and a
 * negative MTSB_E_* code on failure; mtsb_last_error() gives the message of the last
 * failure on the calling thread (the reference throws std::runtime_error via Throw(),
 * include/mitsuba/core/logger.h; the adapter re-throws).
 *
 * Memory spaces: every bulk pointer argument is accompanied (per call) by a `mem`
 * selector: MTSB_MEM_HOST (library stages through pinned memory and copies) or
 * MTSB_MEM_DEVICE (pointer is a CUDA device pointer of the context's device).
 */
#ifndef MTSB200_H
#define MTSB200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MTSB_ABI_VERSION 7

enum { MTSB_OK = 0, MTSB_E_INVALID = -1, MTSB_E_CUDA = -2, MTSB_E_NOMEM = -3, MTSB_E_CANCELLED = -4 };
enum { MTSB_MEM_HOST = 0, MTSB_MEM_DEVICE = 1 };

/* ---- scene description (flattened `Scene`, src/librender/scene.cpp:22-99) ---------- */

/* Texture: `Texture::eval` (include/mitsuba/render/texture.h:40-143).
 * CONSTANT = src/spectra/srgb.cpp / srgb_d65.cpp / uniform.cpp in the rgb variant,
 * BITMAP   = src/textures/bitmap.cpp:250-293 (bilinear, repeat wrap, 3x3 uv transform). */
enum { MTSB_TEX_CONSTANT = 0, MTSB_TEX_BITMAP = 1,
       /* spectral mode only (MTSB_MODE_SPECTRAL): tabulated spectra, wavelengths in nm */
       MTSB_TEX_REGULAR = 2,    /* src/spectra/regular.cpp: `width` samples evenly spaced over [lambda_min, lambda_max] */
       MTSB_TEX_IRREGULAR = 3,  /* src/spectra/irregular.cpp: data = `width` wavelengths followed by `width` values */
       MTSB_TEX_UNIFORM = 4,    /* src/spectra/uniform.cpp: value[0] at every wavelength */
       /* src/textures/mesh_attribute.cpp (-> Mesh::eval_attribute, include/mitsuba/render/mesh.h:278-318): attribute buffer of the
        * mesh(es) whose BSDF uses the texture: `width` elements of `channels` (1 | 3) floats in `data`, `height` = 1: one
        * element per face, 0: per vertex (barycentric interpolation); value[0] = `scale`.  Any colour mode (since ABI 6). */
       MTSB_TEX_MESH_ATTRIBUTE = 5 };
/* In spectral mode the host hands over what the reference's texture plugins store in that mode: CONSTANT `value` and
   3-channel BITMAP texels are sRGB-model coefficients (srgb_model_fetch, src/librender/srgb.cpp:15-42; srgb.cpp:44-48,
   bitmap.cpp:262-275), not colours.  d65_scale != 0 multiplies the value by the D65 whitepoint spectrum times d65_scale:
   that is `srgb_d65` (src/spectra/srgb_d65.cpp:36-62), the spectrum of an emitter given as an rgb colour. */
typedef struct mtsb_texture_desc {
    int32_t  type;
    float    value[3];          /* CONSTANT: rgb (spectral mode: model coefficients); UNIFORM: value[0] */
    uint32_t width, height;     /* BITMAP; REGULAR / IRREGULAR: width = number of samples */
    uint32_t channels;          /* BITMAP: 1 or 3 */
    const float *data;          /* BITMAP: row-major, channels interleaved (host pointer); REGULAR / IRREGULAR: samples */
    float    to_uv[9];          /* BITMAP: row-major 3x3 affine uv transform */
    float    lambda_min, lambda_max;   /* REGULAR */
    float    d65_scale;         /* spectral mode: 0 = plain; else value *= D65(lambda) * d65_scale */
} mtsb_texture_desc;

/* BSDF: `BSDF::sample/eval/pdf` (include/mitsuba/render/bsdf.h).
 * DIFFUSE src/bsdfs/diffuse.cpp:78-135, ROUGHCONDUCTOR src/bsdfs/roughconductor.cpp:196-394,
 * DIELECTRIC src/bsdfs/dielectric.cpp:201-326, CONDUCTOR src/bsdfs/conductor.cpp (smooth: delta reflection),
 * PLASTIC src/bsdfs/plastic.cpp (smooth dielectric coating over a diffuse base).
 * Texture ids index mtsb_scene_desc.textures; -1 = absent. */
enum { MTSB_BSDF_DIFFUSE = 0, MTSB_BSDF_ROUGHCONDUCTOR = 1, MTSB_BSDF_DIELECTRIC = 2, MTSB_BSDF_CONDUCTOR = 3, MTSB_BSDF_PLASTIC = 4,
       MTSB_BSDF_ROUGHDIELECTRIC = 5,     /* src/bsdfs/roughdielectric.cpp: microfacet reflection + transmission */
       MTSB_BSDF_ROUGHPLASTIC = 6 };      /* src/bsdfs/roughplastic.cpp: rough dielectric coating over a diffuse base (alpha_u,
                                             distribution, sample_visible, eta, reflectance_tex = diffuse_reflectance, nonlinear,
                                             specular_sampling_weight); the transmittance tables are computed at scene creation */
enum { MTSB_MF_BECKMANN = 0, MTSB_MF_GGX = 1 };
typedef struct mtsb_bsdf_desc {
    int32_t type;
    int32_t reflectance_tex;        /* DIFFUSE: reflectance; PLASTIC: diffuse_reflectance */
    int32_t eta_tex, k_tex;         /* ROUGHCONDUCTOR / CONDUCTOR: complex IOR (rgb) */
    int32_t specular_reflectance_tex;   /* ROUGHCONDUCTOR / CONDUCTOR / DIELECTRIC / PLASTIC, -1 = 1.0 */
    int32_t specular_transmittance_tex; /* DIELECTRIC, -1 = 1.0 */
    float   alpha_u, alpha_v;       /* ROUGHCONDUCTOR / ROUGHDIELECTRIC */
    int32_t distribution;           /* MTSB_MF_* */
    int32_t sample_visible;         /* bool */
    float   eta;                    /* DIELECTRIC / PLASTIC / ROUGHDIELECTRIC: int_ior / ext_ior */
    int32_t two_sided;              /* bool: wrapped in `twosided` with the same BRDF on both sides
                                       (src/bsdfs/twosided.cpp:66-200); not allowed for DIELECTRIC */
    int32_t nonlinear;              /* PLASTIC: `nonlinear` (plastic.cpp:154)                          (since ABI 5) */
    float   specular_sampling_weight;   /* PLASTIC: s_mean / (d_mean + s_mean) of the texture means (plastic.cpp:165-176) */
    int32_t opacity_tex_plus1;      /* 0 = plain BSDF (so that zero-initialised descriptors stay valid); otherwise 1 + the id of
                                       the single-channel opacity texture of a `mask` adapter around it (src/bsdfs/mask.cpp) */
} mtsb_bsdf_desc;

/* Emitter: `Emitter::sample_direction/pdf_direction/eval`.
 * AREA src/emitters/area.cpp:71-125 (constant radiance), ENVMAP src/emitters/envmap.cpp:132-317.
 * The order of the array is the order of `Scene::m_emitters` (scene.cpp:23-48). */
enum { MTSB_EMITTER_AREA = 0, MTSB_EMITTER_ENVMAP = 1,
       MTSB_EMITTER_CONSTANT = 2,  /* src/emitters/constant.cpp:77-108: uniform background, `radiance` */
       MTSB_EMITTER_POINT = 3 };   /* src/emitters/point.cpp:80-105: isotropic point light, `radiance` = intensity */
typedef struct mtsb_emitter_desc {
    int32_t  type;
    float    radiance[3];       /* AREA / CONSTANT: radiance; POINT: intensity */
    int32_t  mesh_id;           /* AREA: index into meshes[] */
    uint32_t width, height;     /* ENVMAP */
    const float *rgba;          /* ENVMAP: width*height*4 float32 (alpha ignored), host pointer */
    float    to_world[16];      /* ENVMAP: row-major 4x4 */
    float    scale;             /* ENVMAP */
    float    position[3];       /* POINT: world-space position */
    const float *coeff_scale;   /* ENVMAP, spectral mode: width*height*4, per texel srgb_model_fetch(rgb / max(1e-8, 2 hmax(rgb)))
                                   and the scale 2 hmax(rgb) (src/emitters/envmap.cpp:98-110); `rgba` still drives the importance map */
    int32_t  radiance_tex;      /* spectral mode: texture holding the radiance / intensity spectrum (AREA, CONSTANT, POINT);
                                   ENVMAP texels are then model coefficients + scale (src/emitters/envmap.cpp:98-110) */
} mtsb_emitter_desc;

/* Mesh: `Mesh` buffers (include/mitsuba/render/mesh.h:328-332); world-space positions. */
typedef struct mtsb_mesh_desc {
    const uint32_t *faces;      /* 3*face_count */
    const float *positions;     /* 3*vertex_count */
    const float *normals;       /* 3*vertex_count or NULL */
    const float *texcoords;     /* 2*vertex_count or NULL */
    uint32_t face_count, vertex_count;
    int32_t  bsdf_id;           /* index into bsdfs[] */
    int32_t  emitter_id;        /* index into emitters[] or -1 */
    int32_t  mem;               /* MTSB_MEM_HOST (0, default) or MTSB_MEM_DEVICE: where faces / positions / normals / texcoords
                                   live.  Either way the buffers are copied straight into the scene's device arrays (no host
                                   staging copy); device buffers make scene creation a set of device-to-device copies (since ABI 6) */
} mtsb_mesh_desc;

/* Analytic shapes as BVH leaf primitives (since ABI 7): src/shapes/sphere.cpp:276-300, disk.cpp:144-165, cylinder.cpp:248-287,
 * rectangle.cpp:142-164 (`ray_intersect` / `ray_test`), their `fill_surface_interaction`, `sample_position` /
 * `sample_direction` / `pdf_direction`.  `to_world` is the plugin's object-to-world transform with `center` / `radius`
 * (sphere.cpp:104-106) or `p0` / `p1` / `radius` (cylinder.cpp:88-97) already folded in by the host; the library applies
 * the plugin's update() (sphere, cylinder: decomposition into scale / rotation / translation, no shear; disk, rectangle:
 * flip_normals = to_world * scale(1, 1, -1)).
 * Global primitive index: after the faces of all meshes, one primitive per shape in array order (`m_primitive_map`,
 * kdtree.h:2382-2400, with the meshes listed first).  Closest-hit records of a disk / rectangle carry the object-space
 * hit position (x, y) in (u, v) (the two floats the plugins cache); sphere / cylinder: 0.
 * An AREA emitter refers to shape k through mesh_id = n_meshes + k.  Gradient calls reject scenes with shapes. */
enum { MTSB_SHAPE_SPHERE = 1, MTSB_SHAPE_DISK = 2, MTSB_SHAPE_CYLINDER = 3, MTSB_SHAPE_RECTANGLE = 4 };
typedef struct mtsb_shape_desc {
    int32_t  type;              /* MTSB_SHAPE_* */
    int32_t  flip_normals;
    float    to_world[16];      /* row-major 4x4 */
    int32_t  bsdf_id;           /* index into bsdfs[] */
    int32_t  emitter_id;        /* index into emitters[] or -1 */
} mtsb_shape_desc;

/* Sensor + film: src/sensors/perspective.cpp:106-222, src/films/hdrfilm.cpp, rfilters. */
/* src/rfilters/{box,tent,gaussian,mitchell,catmullrom,lanczos}.cpp */
enum { MTSB_RFILTER_BOX = 0, MTSB_RFILTER_TENT = 1, MTSB_RFILTER_GAUSSIAN = 2, MTSB_RFILTER_MITCHELL = 3, MTSB_RFILTER_CATMULLROM = 4,
       MTSB_RFILTER_LANCZOS = 5 };
typedef struct mtsb_sensor_desc {
    float    to_world[16];          /* row-major camera-to-world */
    float    sample_to_camera[16];  /* row-major, perspective.cpp:125-133 */
    float    near_clip, far_clip;
    uint32_t film_width, film_height;   /* size of the rendered block = the film's crop window; the window's placement in the
                                         * full film is folded into sample_to_camera (perspective.cpp:106-133) by the host */
    int32_t  rfilter;               /* MTSB_RFILTER_* */
    float    rfilter_radius;        /* box: 0.5 (+RayEpsilon added internally), tent: 1, gaussian: 4*stddev, mitchell and
                                     * catmullrom: 2, lanczos: number of lobes (default 3) */
    float    rfilter_stddev;        /* gaussian: stddev; mitchell: B (default 1/3) */
    float    rfilter_c;             /* mitchell: C (default 1/3)                      (since ABI 4) */
} mtsb_sensor_desc;

/* `MonteCarloIntegrator` props (src/librender/integrator.cpp:285-298). */
typedef struct mtsb_integrator_desc {
    int32_t max_depth;  /* -1 = unbounded (clamped internally to 2^30) */
    int32_t rr_depth;   /* default 5 */
} mtsb_integrator_desc;

enum { MTSB_MODE_RGB = 0, MTSB_MODE_SPECTRAL = 1 };
typedef struct mtsb_scene_desc {
    uint32_t n_meshes, n_bsdfs, n_textures, n_emitters, n_sensors;
    const mtsb_mesh_desc    *meshes;
    const mtsb_bsdf_desc    *bsdfs;
    const mtsb_texture_desc *textures;
    const mtsb_emitter_desc *emitters;
    const mtsb_sensor_desc  *sensors;
    mtsb_integrator_desc integrator;
    int32_t mode;               /* MTSB_MODE_RGB (scalar_rgb semantics) or MTSB_MODE_SPECTRAL (scalar_spectral: 4 wavelengths
                                   per path, include/mitsuba/core/spectrum.h:304-316, film in XYZ via cie1931_xyz :147-217) */
    uint32_t n_shapes;          /* analytic shapes (since ABI 7) */
    const mtsb_shape_desc *shapes;
} mtsb_scene_desc;

/* ---- SoA ray / hit records (128-bit loads) ------------------------------------------ */
/* ray i = { o_mint[4i..4i+3] = (o.x,o.y,o.z,mint), d_maxt[4i..4i+3] = (d.x,d.y,d.z,maxt) }
 * closest hit i = hits[4i..4i+3] = (t, u, v, bitcast<float>(global_prim_index)); miss: t = +inf, prim = 0xffffffff.
 * any hit i = occluded[i] (uint32 0/1).  global prim index = prefix sum of face counts over meshes
 * (`m_primitive_map`, include/mitsuba/render/kdtree.h:2382-2400). */

typedef struct mtsb_ctx mtsb_ctx;
typedef struct mtsb_scene mtsb_scene;

/* Per-call statistics (what bench.py turns into Mrays/s, Msamples/s and roofline bytes). */
typedef struct mtsb_stats {
    uint64_t samples;           /* camera samples started */
    uint64_t closest_rays;      /* closest-hit rays traced */
    uint64_t shadow_rays;       /* any-hit rays traced */
    uint64_t segments;          /* shade-kernel invocations (path vertices incl. misses) */
    uint64_t kernel_launches;   /* kernels launched by this call */
    float    ms_total;          /* device time of the call (CUDA events) */
    float    ms_closest, ms_any, ms_shade, ms_raygen, ms_film; /* filled when profiling flag set */
    float    ms_build;          /* BVH build (scene_create / update) */
    uint32_t bvh_nodes;         /* wide nodes */
    uint32_t bvh_depth;
} mtsb_stats;

const char *mtsb_last_error(void);
int mtsb_abi_version(void);

/* context = one GPU / one process rank */
int mtsb_ctx_create(int device, mtsb_ctx **out);
int mtsb_ctx_destroy(mtsb_ctx *ctx);
/* `Integrator::cancel()` (include/mitsuba/render/integrator.h:51): async-signal-safe flag */
int mtsb_cancel(mtsb_ctx *ctx);
/* `timeout` property of SamplingIntegrator (src/librender/integrator.cpp:36, should_stop() integrator.h:143-146): a render
 * that has run longer than `seconds` stops starting new batches and returns MTSB_OK with what was rendered
 * (mtsb_stats.samples < requested); <= 0 disables.  Like mtsb_cancel it acts within two wavefront batches. */
int mtsb_set_timeout(mtsb_ctx *ctx, float seconds);
/* per-stage CUDA-event timing (the reference's SIGPROF phase profiler, src/libcore/profiler.cpp, re-stated as
 * event pairs around every launch on the launching stream); fills mtsb_stats.ms_{raygen,closest,shade,any,film} */
int mtsb_set_profiling(mtsb_ctx *ctx, int on);

/* `Scene::Scene` + `accel_init_*` (scene.cpp:22-99, scene_native.inl:3-10): copies everything, builds the BVH on the GPU */
int mtsb_scene_create(mtsb_ctx *ctx, const mtsb_scene_desc *desc, mtsb_scene **out);
int mtsb_scene_destroy(mtsb_scene *scene);
int mtsb_scene_stats(mtsb_scene *scene, mtsb_stats *out);
/* `Mesh::parameters_changed` (src/librender/mesh.cpp:794-804): new positions, normals recomputed if the
 * mesh had computed normals, area CDF rebuilt, BVH rebuilt */
int mtsb_scene_update_vertices(mtsb_scene *scene, uint32_t mesh_id, const float *positions, int mem);
/* How mtsb_scene_update_vertices brings the acceleration structure up to date (the reference rebuilds its kd-tree,
 * src/librender/scene.cpp:84-99 `parameters_changed` -> accel_release / accel_init): refit = 0 (default) full GPU rebuild;
 * refit = 1 keeps the topology of the last build and refits leaf triangles and node boxes to the new positions (an
 * optimisation loop moves vertices a little per step), with a full rebuild on every `rebuild_every`-th update (0 = never). */
int mtsb_scene_update_config(mtsb_scene *scene, int refit, uint32_t rebuild_every);
/* `BitmapTextureImpl::parameters_changed` (src/textures/bitmap.cpp:301-331) */
int mtsb_scene_update_texture(mtsb_scene *scene, uint32_t tex_id, const float *data, int mem);

/* `Scene::ray_intersect` (scene.cpp:108-116) / `Scene::ray_test` (scene.cpp:131-139), batched */
int mtsb_intersect_closest(mtsb_scene *scene, uint64_t n, const float *o_mint, const float *d_maxt,
                           float *hits, int mem, mtsb_stats *stats);
int mtsb_intersect_any(mtsb_scene *scene, uint64_t n, const float *o_mint, const float *d_maxt,
                       uint32_t *occluded, int mem, mtsb_stats *stats);
/* brute-force variant of the above = `Scene::ray_intersect_naive` (scene.cpp:118-129); test hook */
int mtsb_intersect_naive(mtsb_scene *scene, uint64_t n, const float *o_mint, const float *d_maxt,
                         float *hits, int mem);
/* Allocation-free variant of mtsb_intersect_closest.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same kernel, same results (bit-identical hits), but the call performs no cudaMalloc / cudaFree /
 * allocating library call: the traversal's work counter and -- for MTSB_MEM_HOST callers -- the staging copies of rays
 * and hits live in one fixed file-scope device array (12 MiB + 256 B per GPU, held from library load until exit).
 * Limit: n <= mtsb_noalloc_max_rays() = 262144 rays per call; larger n fails with MTSB_E_INVALID and a message (no
 * fallback to the allocating path).  The array is shared by every context and scene on the device and keeps the last
 * call's contents (the work counter is cleared at the start of every call): calls that overlap on one GPU race.
 * With MTSB_MEM_DEVICE the caller's buffers are used directly and only the 256-byte counter block is touched. */
int mtsb_intersect_closest_noalloc(mtsb_scene *scene, uint64_t n, const float *o_mint, const float *d_maxt,
                                   float *hits, int mem, mtsb_stats *stats);
uint64_t mtsb_noalloc_max_rays(void);

/* `SamplingIntegrator::render` (src/librender/integrator.cpp:52-176), GPU branch semantics (:144-169):
 * renders samples [spp_begin, spp_begin+spp_count) of every pixel of sensor `sensor_id`, sample s of pixel p
 * uses PCG32 stream (p*spp_total + s) seeded through TEA with `seed` (independent.cpp:61-70), and ACCUMULATES
 * into the scene's device film (XYZAW sums).  mtsb_film_* read / clear / reduce it. */
int mtsb_render(mtsb_scene *scene, uint32_t sensor_id, uint32_t spp_begin, uint32_t spp_count,
                uint32_t spp_total, uint64_t seed, mtsb_stats *stats);
int mtsb_film_clear(mtsb_scene *scene, uint32_t sensor_id);
/* `moment` integrator (src/integrators/moment.cpp:55-90): when enabled, every splat also accumulates the squares of the
 * sample's X, Y, Z with the same filter weight -- the `m2_<nested>.X/Y/Z` AOV channels.  mtsb_film_read_moments returns
 * width*height*3 float32 SUMS (divide by the film's W channel like every other channel); mtsb_film_clear clears them. */
int mtsb_film_enable_moments(mtsb_scene *scene, uint32_t sensor_id, int on);
int mtsb_film_read_moments(mtsb_scene *scene, uint32_t sensor_id, float *m2_xyz, int mem);
/* `Film::put(ImageBlock*)`-side export: packed float32 XYZAW, width*height*5 (integrator.cpp:72-74) */
int mtsb_film_read(mtsb_scene *scene, uint32_t sensor_id, float *xyzaw, int mem);
/* add an external XYZAW buffer into the device film (used after the NCCL reduce and for resume) */
int mtsb_film_accumulate(mtsb_scene *scene, uint32_t sensor_id, const float *xyzaw, int mem);
/* `HDRFilm::bitmap()/develop()` (src/films/hdrfilm.cpp:249-320): weight-normalise + XYZ->linear sRGB, float32 RGB */
int mtsb_film_develop(mtsb_scene *scene, uint32_t sensor_id, float *rgb, int mem);
/* device pointer of the internal packed XYZAW staging buffer (for NCCL all-reduce by the host layer) */
int mtsb_film_device_ptr(mtsb_scene *scene, uint32_t sensor_id, void **ptr, uint64_t *n_floats);

/* Gradient pass (the adjoint of mtsb_render; replaces autodiff through path.cpp, src/python/python/autodiff.py:6-194).
 * dL_dimage: width*height*3 float32 = dL/d(developed linear RGB).  Outputs are ACCUMULATED:
 * grad_texture (size of texture `tex_id`), grad_vertices (3*vertex_count of mesh `mesh_id`, the parameter the reference
 * exposes as `<mesh>.vertex_positions_buf`, src/librender/mesh.cpp:781-804); either may be NULL.
 * Vertex gradients = interior term (path-space adjoint) + boundary terms (edge sampling on the film and from x1), see
 * mitsuba2_b200/csrc/adjoint.h.  Any reconstruction filter: under a filter wider than a pixel a sample's adjoint weight is the
 * gather of dL_dimage over its footprint divided by the film's weight channel of all spp_total samples (hdrfilm.cpp develop).
 * MTSB_MODE_RGB is required; `mesh_id` may carry an area light (its emitter samples move with the mesh) unless it has vertex
 * normals; analytic shapes are not differentiated. */
int mtsb_render_grad(mtsb_scene *scene, uint32_t sensor_id, uint32_t spp_begin, uint32_t spp_count,
                     uint32_t spp_total, uint64_t seed, const float *dL_dimage,
                     int32_t tex_id, float *grad_texture, int32_t mesh_id, float *grad_vertices,
                     int mem, mtsb_stats *stats);
/* Options of the vertex-gradient pass: boundary = 0 interior term only, 1 = + edge sampling on the film (primary
 * visibility), 2 (default) = + edge sampling seen from the first path vertex (shadow / occlusion boundaries), 3 = + the same
 * term at every later path vertex (a shadow seen through a bounce; each depth takes half the samples of the one before);
 * edge_fraction = edge samples of each kind per path sample (default 0.25). */
int mtsb_grad_config(mtsb_scene *scene, int boundary, float edge_fraction);

/* ---- batched parity hooks (device math called exactly as the kernels call it) ------- */
/* `PCG32::seed/next_float32` (ext/enoki/include/enoki/random.h:62-101): out[i*n_draws + k] */
int mtsb_pcg32_floats(mtsb_ctx *ctx, uint64_t n, const uint64_t *initstate, const uint64_t *initseq,
                      uint32_t n_draws, float *out);
/* sampler seeding of the wavefront: state/inc of stream idx (independent.cpp:67-70, core/random.h:94-113) */
int mtsb_sampler_seed(mtsb_ctx *ctx, uint64_t n, uint64_t seed, uint64_t *state, uint64_t *inc);
/* `Sensor::sample_ray` (perspective.cpp:151-188): pos = position_sample in [0,1]^2 */
int mtsb_sensor_sample_ray(mtsb_scene *scene, uint32_t sensor_id, uint64_t n, const float *pos2,
                           float *o_mint, float *d_maxt);
/* `BSDF::sample` / `eval` / `pdf` in local frame. wi: 3n, uv: 2n, sample1: n, sample2: 2n.
 * out_sample: 8n = (wo.xyz, pdf, eta, weight.rgb)... see DESIGN.md; eval: 3n; pdf: n */
int mtsb_bsdf_sample(mtsb_scene *scene, uint32_t bsdf_id, uint64_t n, const float *wi, const float *uv,
                     const float *sample1, const float *sample2, float *out9);
int mtsb_bsdf_eval_pdf(mtsb_scene *scene, uint32_t bsdf_id, uint64_t n, const float *wi, const float *uv,
                       const float *wo, float *eval3, float *pdf);
/* `Scene::sample_emitter_direction` without the visibility test (scene.cpp:141-176):
 * ref_p: 3n, sample: 2n -> out: 11n = (d.xyz, dist, pdf, value.rgb, n.xyz) */
int mtsb_sample_emitter_direction(mtsb_scene *scene, uint64_t n, const float *ref_p, const float *sample2,
                                  float *out11);
/* `ImageBlock::put(pos, value)` (src/librender/imageblock.cpp:79-172) into the device film: pos 2n, value 5n */
int mtsb_film_put(mtsb_scene *scene, uint32_t sensor_id, uint64_t n, const float *pos2, const float *value5);

#ifdef __cplusplus
}
#endif
#endif /* MTSB200_H */
