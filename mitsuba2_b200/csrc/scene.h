// scene.h -- device-side scene snapshot (POD structs of pointers) shared by all stages.
#pragma once
#include "bvh.h"

namespace mtsb {

// type 0: constant (rgb colour; spectral mode: sRGB-model coefficients, src/spectra/srgb.cpp), 1: bitmap (spectral mode,
// 3 channels: coefficient texels, src/textures/bitmap.cpp:262-275), 2: `regular` tabulated spectrum (w samples over
// [lambda_min, lambda_max]), 3: `irregular` (w nodes then w values in data), 4: `uniform` spectrum (value.x).
// d65_scale != 0: the value is multiplied by the D65 whitepoint spectrum times d65_scale (src/spectra/srgb_d65.cpp).
struct TexDev {
    int type; float3 value;
    uint32_t w, h, ch; const float *data; float to_uv[9];
    float lambda_min, lambda_max, d65_scale;
};
struct BsdfDev {
    int type; int refl, eta, k, spec_r, spec_t;
    float alpha_u, alpha_v; int distr; int sample_visible; float eta_f;
    int two_sided;          // `twosided` adapter, same BRDF on both sides (src/bsdfs/twosided.cpp)
    int nonlinear; float spec_weight;   // plastic (type 4): `nonlinear`, specular sampling weight; type 3 = smooth conductor
    int opacity;            // >= 0: `mask` adapter (src/bsdfs/mask.cpp) with this single-channel opacity texture
    const float *rough_t;   // rough plastic (type 6): external transmittance at 64 cosines (roughplastic.cpp parameters_changed)
    float internal_refl;    // rough plastic: mean internal reflectance
};
constexpr int kMaxEnvLevels = 20;
struct EmitterDev {
    int type;               // 0 area, 1 envmap, 2 constant background, 3 point
    float3 position;        // point
    // area (radiance doubles as the constant emitter's radiance and the point light's intensity)
    float3 radiance; int radiance_tex;   // spectral mode: radiance / intensity spectrum = texture radiance_tex
    int mesh; const float *cdf; float cdf_sum, normalization; uint32_t valid_lo, valid_hi;
    // envmap
    uint32_t w, h; const float4 *rgba;   // rgb + 1; spectral mode: sRGB-model coefficients + scale (envmap.cpp:98-110)
    float to_world[16], to_local[16]; float scale;
    int n_levels; const float *level_data[kMaxEnvLevels]; uint32_t level_width[kMaxEnvLevels];
    float bsphere_radius;
};
struct MeshDev {
    uint32_t face_begin, face_count, vert_begin, vert_count;
    int bsdf, emitter; int16_t has_normals, has_uv;
    int shape;              // >= 0: an analytic shape (one primitive), index into SceneDev::shapes; -1: triangle mesh
};
static_assert(sizeof(MeshDev) == 32, "MeshDev is fetched as two 128-bit loads");
// reconstruction filter + film size, the part of SensorDev the gradient pass's film gather needs (path.h film_gather)
struct FilmFilter { int rfilter; float radius, stddev, filter_c; uint32_t width, height; };
struct SensorDev {
    float to_world[16], sample_to_camera[16];
    float near_clip, far_clip;
    uint32_t width, height;
    int rfilter; float radius, stddev;   // radius already includes box RayEpsilon; stddev doubles as Mitchell's B
    float filter_c;                      // Mitchell's C
    float4 *film_xyza; float *film_w;    // device film accumulators
    float4 *film_m2;                     // optional second-moment plane (sum of w X^2, w Y^2, w Z^2): `moment` integrator AOVs
};
struct SceneDev {
    Bvh bvh;
    const float *pos, *nrm, *uv;
    const uint4 *faces;
    const ShapeDev *shapes;
    const MeshDev *meshes; const TexDev *tex; const BsdfDev *bsdfs; const EmitterDev *emitters;
    int n_emitters, env_emitter;
    int max_depth, rr_depth;
    const float *cie;       // spectral mode: x-bar, y-bar, z-bar (3 x 95) then D65 / 10568 (95), 360..830 nm
};

} // namespace mtsb
