// emit.h -- area + envmap emitters and the scene-level emitter sampling (HD bodies).
#pragma once
#include "shade.h"

namespace mtsb {

template <typename M> struct DirSampleT { float3 p, n, d; float dist, pdf; typename M::Spec value; bool delta; int etype; /* EmitterDev::type */ };
typedef DirSampleT<Rgb> DirSample;

// radiance (area, constant) / intensity (point) of an emitter: an rgb triple, or in spectral mode a texture
// (srgb_d65 for rgb input, or any tabulated spectrum)
template <typename M> HD typename M::Spec emitter_radiance(const SceneDev &sc, const EmitterDev &e, float4 wav) {
    if constexpr (M::spectral) return tex_eval_spectral(sc, e.radiance_tex, 0.f, 0.f, wav); else return e.radiance;
}

// enoki::binary_search (ext/enoki/include/enoki/array_utils.h:153-171)
template <typename Pred> HD uint32_t binary_search(uint32_t start, uint32_t end, Pred pred) {
    uint32_t iterations = (start < end) ? (uint32_t) (31 - clz32(end - start)) + 1u : 0u;
    for (uint32_t i = 0; i < iterations; ++i) {
        uint32_t middle = (start + end) >> 1;
        if (pred(middle)) start = min(middle + 1, end); else end = middle;
    }
    return start;
}

// ---- envmap (src/emitters/envmap.cpp:132-206, 267-317) -------------------------------------------
HD float3 env_lookup(const EmitterDev &e, float u, float v) {
    u *= (float) (e.w - 1); v *= (float) (e.h - 1);
    uint32_t px = min((uint32_t) u, e.w - 2), py = min((uint32_t) v, e.h - 2);
    float w1x = u - (float) px, w1y = v - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    uint32_t idx = px + py * e.w;
    float4 v00 = e.rgba[idx], v10 = e.rgba[idx + 1], v01 = e.rgba[idx + e.w], v11 = e.rgba[idx + e.w + 1];
    float3 a = mk3(fmaf(w0x, v00.x, w1x * v10.x), fmaf(w0x, v00.y, w1x * v10.y), fmaf(w0x, v00.z, w1x * v10.z));
    float3 b = mk3(fmaf(w0x, v01.x, w1x * v11.x), fmaf(w0x, v01.y, w1x * v11.y), fmaf(w0x, v01.z, w1x * v11.z));
    return mk3(fmaf(w0y, a.x, w1y * b.x), fmaf(w0y, a.y, w1y * b.y), fmaf(w0y, a.z, w1y * b.z)) * e.scale;
}
// spectral branch of EnvironmentMapEmitter::eval_spectrum (envmap.cpp:285-306): texels = model coefficients + scale
HD float4 env_lookup_spectral(const SceneDev &sc, const EmitterDev &e, float u, float v, float4 wav) {
    u *= (float) (e.w - 1); v *= (float) (e.h - 1);
    uint32_t px = min((uint32_t) u, e.w - 2), py = min((uint32_t) v, e.h - 2);
    float w1x = u - (float) px, w1y = v - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    uint32_t idx = px + py * e.w;
    float4 v00 = e.rgba[idx], v10 = e.rgba[idx + 1], v01 = e.rgba[idx + e.w], v11 = e.rgba[idx + e.w + 1];
    auto bx = [&](float a, float b) { return fmaf(w0x, a, w1x * b); };
    auto by = [&](float a, float b) { return fmaf(w0y, a, w1y * b); };
    float4 s0 = map_spec(srgb_model_eval(xyz(v00), wav), srgb_model_eval(xyz(v10), wav), bx),
           s1 = map_spec(srgb_model_eval(xyz(v01), wav), srgb_model_eval(xyz(v11), wav), bx);
    float f = by(bx(v00.w, v10.w), bx(v01.w, v11.w));
    return map_spec(s0, s1, by) * d65_eval(sc, wav) * f * e.scale;
}
template <typename M> HD typename M::Spec env_lookup_m(const SceneDev &sc, const EmitterDev &e, float u, float v, float4 wav) {
    if constexpr (M::spectral) return env_lookup_spectral(sc, e, u, v, wav); else return env_lookup(e, u, v);
}
HD void env_dir_to_uv(const EmitterDev &e, float3 dworld, float &u, float &v, float3 &dl) {
    dl = xform_vec(e.to_local, dworld);
    u = atan2f(dl.x, -dl.z) * kInvTwoPi; v = safe_acos(dl.y) * kInvPi;
    u -= floorf(u); v -= floorf(v);
}
template <typename M = Rgb>
HD typename M::Spec env_eval(const SceneDev &sc, const EmitterDev &e, float3 dworld, float4 wav = float4()) {     // direction pointing away from the scene
    float u, v; float3 dl; env_dir_to_uv(e, dworld, u, v, dl);
    return env_lookup_m<M>(sc, e, u, v, wav);
}
// Hierarchical2D<Float,0> (include/mitsuba/core/distr_2d.h:317-403, 495-530); level 0 = bilinear data,
// levels 1.. = MIP pyramid in 2x2-blocked layout (Level::index, :583-586)
HD uint32_t h2d_index(uint32_t x, uint32_t y, uint32_t width) {
    return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + ((y & ~1u) * width);
}
HD float interval_to_linear(float v0, float v1, float s) {
    return fabsf(v0 - v1) > 1e-4f * (v0 + v1) ? (v0 - safe_sqrt(lerpf(sqr(v0), sqr(v1), s))) / (v0 - v1) : s;
}
HD void h2d_sample(const EmitterDev &e, float sx, float sy, float &u, float &v, float &pdf) {
    sx = fminf(fmaxf(sx, 0.f), 1.f); sy = fminf(fmaxf(sy, 0.f), 1.f);
    uint32_t ox = 0, oy = 0;
    for (int l = e.n_levels - 2; l > 0; --l) {
        const float *d = e.level_data[l];
        ox <<= 1; oy <<= 1;
        uint32_t i = h2d_index(ox, oy, e.level_width[l]);
        float v00 = d[i], v10 = d[i + 1], v01 = d[i + 2], v11 = d[i + 3];
        sx = fminf(fmaxf(sx, 0.f), 1.f); sy = fminf(fmaxf(sy, 0.f), 1.f);
        float r0 = v00 + v10, r1 = v01 + v11;
        sy *= r0 + r1;
        bool mask = sy > r0;
        if (mask) { oy += 1; sy -= r0; }
        sy /= mask ? r1 : r0;
        float c0 = mask ? v01 : v00, c1 = mask ? v11 : v10;
        sx *= c0 + c1;
        bool mx = sx > c0;
        if (mx) sx -= c0;
        sx /= mx ? c1 : c0;
        if (mx) ox += 1;
    }
    const float *d0 = e.level_data[0]; uint32_t w0 = e.level_width[0];
    uint32_t i = ox + oy * w0;
    float v00 = d0[i], v10 = d0[i + 1], v01 = d0[i + w0], v11 = d0[i + w0 + 1];
    // warp::square_to_bilinear (include/mitsuba/core/warp.h:398-414)
    float r0 = v00 + v10, r1 = v01 + v11;
    sy = interval_to_linear(r0, r1, sy);
    float c0 = lerpf(v00, v01, sy), c1 = lerpf(v10, v11, sy);
    sx = interval_to_linear(c0, c1, sx);
    pdf = lerpf(c0, c1, sx);
    u = ((float) (int) ox + sx) * (1.f / (float) (e.w - 1));
    v = ((float) (int) oy + sy) * (1.f / (float) (e.h - 1));
}
HD float h2d_eval(const EmitterDev &e, float u, float v) {
    u = fminf(fmaxf(u, 0.f), 1.f) * (float) (e.w - 1); v = fminf(fmaxf(v, 0.f), 1.f) * (float) (e.h - 1);
    uint32_t ox = min((uint32_t) (int) u, e.w - 2), oy = min((uint32_t) (int) v, e.h - 2);
    u -= (float) (int) ox; v -= (float) (int) oy;
    const float *d0 = e.level_data[0]; uint32_t w0 = e.level_width[0];
    uint32_t i = ox + oy * w0;
    float v00 = d0[i], v10 = d0[i + 1], v01 = d0[i + w0], v11 = d0[i + w0 + 1];
    return lerpf(lerpf(v00, v10, u), lerpf(v01, v11, u), v);
}
HD float env_pdf_direction(const EmitterDev &e, float3 dworld) {
    float u, v; float3 d; env_dir_to_uv(e, dworld, u, v, d);
    float inv_sin = 1.f / sqrtf(fmaxf(sqr(d.x) + sqr(d.z), sqr(kEpsilon)));
    return h2d_eval(e, u, v) * inv_sin * (1.f / (2.f * sqr(kPi)));
}
template <typename M>
HD DirSampleT<M> env_sample_direction(const SceneDev &sc, const EmitterDev &e, float3 ref_p, float sx, float sy, float4 wav) {
    DirSampleT<M> ds;
    float u, v, pdf; h2d_sample(e, sx, sy, u, v, pdf);
    float st, ct, sp, cp; sincospif(v, &st, &ct); sincospif(2.f * u, &sp, &cp);      // theta = v pi, phi = 2 u pi
    float3 d = mk3(cp * st, sp * st, ct);         // math::sphdir
    d = mk3(d.y, d.z, -d.x);
    float dist = 2.f * e.bsphere_radius;
    float inv_sin = 1.f / sqrtf(fmaxf(sqr(d.x) + sqr(d.z), sqr(kEpsilon)));
    d = xform_vec(e.to_world, d);
    ds.p = ref_p + d * dist; ds.n = -d; ds.d = d; ds.dist = dist;
    ds.pdf = pdf > 0.f ? pdf * inv_sin * (1.f / (2.f * sqr(kPi))) : 0.f;
    ds.value = env_lookup_m<M>(sc, e, u, v, wav) / ds.pdf;
    return ds;
}

// ---- area light on a mesh (src/emitters/area.cpp:103-125, src/librender/shape.cpp:252-283,
//      src/librender/mesh.cpp:320-365, include/mitsuba/core/distr_1d.h:144-203) ---------------------
// warp::square_to_uniform_sphere (include/mitsuba/core/warp.h:262-267)
HD float3 square_to_uniform_sphere(float sx, float sy) {
    float z = fmaf(-2.f, sy, 1.f), r = safe_sqrt(fmaf(-z, z, 1.f));
    float sn, cs; sincospif(2.f * sx, &sn, &cs);
    return mk3(r * cs, r * sn, z);
}
// Shape::sample_direction of an analytic shape: sphere.cpp:182-255 (cone of directions subtended by the sphere, uniform
// inside), the others Shape::sample_direction (shape.cpp:252-270) over sample_position (disk.cpp:127-141,
// cylinder.cpp:214-236, rectangle.cpp:114-127).  Fills p, n, d, dist, pdf.
static __host__ __device__ __noinline__ void shape_sample_direction(const ShapeDev &sh, float3 ref_p, float sx, float sy,
                                                                   float3 &p, float3 &n, float3 &d, float &dist, float &pdf) {
    if (sh.type == kShapeSphere) {
        float3 dc_v = sh.center - ref_p;
        float dc_2 = dot(dc_v, dc_v);
        float radius_adj = sh.radius * (sh.flip ? (1.f + kRayEpsilon) : (1.f - kRayEpsilon));
        float3 dir;
        if (dc_2 > sqr(radius_adj)) {
            float inv_dc = 1.f / sqrtf(dc_2), sin_theta_max = sh.radius * inv_dc, sin_theta_max_2 = sqr(sin_theta_max),
                  inv_sin_theta_max = 1.f / sin_theta_max, cos_theta_max = safe_sqrt(1.f - sin_theta_max_2);
            float sin_theta_2 = sin_theta_max_2 > 0.00068523f ? 1.f - sqr(fmaf(cos_theta_max - 1.f, sx, 1.f)) : sin_theta_max_2 * sx,
                  cos_theta = safe_sqrt(1.f - sin_theta_2);
            float cos_alpha = sin_theta_2 * inv_sin_theta_max + cos_theta * safe_sqrt(fmaf(-sin_theta_2, sqr(inv_sin_theta_max), 1.f)),
                  sin_alpha = safe_sqrt(fmaf(-cos_alpha, cos_alpha, 1.f));
            float sin_phi, cos_phi; sincosf(sy * (2.f * kPi), &sin_phi, &cos_phi);
            Frame fr; fr.n = dc_v * -inv_dc; coordinate_system(fr.n, fr.s, fr.t);
            dir = fr.to_world(mk3(cos_phi * sin_alpha, sin_phi * sin_alpha, cos_alpha));
            p = mk3(fmaf(dir.x, sh.radius, sh.center.x), fmaf(dir.y, sh.radius, sh.center.y), fmaf(dir.z, sh.radius, sh.center.z));
            d = p - ref_p;
            float dist2 = dot(d, d);
            dist = sqrtf(dist2); d = d / dist;
            pdf = kInvTwoPi / (1.f - cos_theta_max);
            if (dist == 0.f) pdf = 0.f;
        } else {
            dir = square_to_uniform_sphere(sx, sy);
            p = mk3(fmaf(dir.x, sh.radius, sh.center.x), fmaf(dir.y, sh.radius, sh.center.y), fmaf(dir.z, sh.radius, sh.center.z));
            d = p - ref_p;
            float dist2 = dot(d, d);
            dist = sqrtf(dist2); d = d / dist;
            pdf = sh.inv_area * dist2 / fabsf(dot(d, dir));
        }
        n = sh.flip ? -dir : dir;
        return;
    }
    if (sh.type == kShapeDisk) {
        float px, py; square_to_uniform_disk_concentric(sx, sy, px, py);
        p = xform_point_affine(sh.to_world, mk3(px, py, 0.f)); n = sh.n;
    } else if (sh.type == kShapeCylinder) {
        float st, ct; sincosf(2.f * kPi * sy, &st, &ct);
        p = xform_point_affine(sh.to_world, mk3(ct * sh.radius, st * sh.radius, sx * sh.length));
        n = sh.flip ? mk3(-ct, -st, 0.f) : mk3(ct, st, 0.f);      // cylinder.cpp:222-232 returns the object-space normal as is; kept for parity
    } else {
        p = xform_point_affine(sh.to_world, mk3(fmaf(sx, 2.f, -1.f), fmaf(sy, 2.f, -1.f), 0.f)); n = sh.n;
    }
    d = p - ref_p;
    float dist2 = dot(d, d);
    dist = sqrtf(dist2); d = d / dist;
    float dp = fabsf(dot(d, n));
    pdf = sh.inv_area * (dp != 0.f ? dist2 / dp : 0.f);
}
// Sphere::pdf_direction (sphere.cpp:257-270); the other shapes use Shape::pdf_direction (shape.cpp:272-283)
static __host__ __device__ __noinline__ float shape_pdf_direction(const ShapeDev &sh, float3 ref_p, float3 d, float3 n, float dist) {
    float dp = fabsf(dot(d, n));
    float area_pdf = sh.inv_area * (dp != 0.f ? (dist * dist) / dp : 0.f);
    if (sh.type != kShapeSphere) return area_pdf;
    float sin_alpha = sh.radius * (1.f / norm(sh.center - ref_p)), cos_alpha = safe_sqrt(1.f - sin_alpha * sin_alpha);
    return sin_alpha < 1.f - kEpsilon ? kInvTwoPi / (1.f - cos_alpha) : sh.inv_area * sqr(dist) / dp;
}

template <typename M>
HD DirSampleT<M> area_sample_direction(const SceneDev &sc, const EmitterDev &e, float3 ref_p, float sx, float sy, float4 wav) {
    DirSampleT<M> ds;
    const MeshDev m = sc.meshes[e.mesh];
    if (M::ext && m.shape >= 0) {
        shape_sample_direction(sc.shapes[m.shape], ref_p, sx, sy, ds.p, ds.n, ds.d, ds.dist, ds.pdf);
        bool active = dot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
        ds.value = active ? emitter_radiance<M>(sc, e, wav) / ds.pdf : M::splat(0.f);
        return ds;
    }
    float value = sy * e.cdf_sum;
    const float *cdf = e.cdf;
    uint32_t idx = binary_search(e.valid_lo, e.valid_hi, [&](uint32_t i) { return cdf[i] < value; });
    uint4 f = sc.faces[m.face_begin + idx];
    float3 p0 = ld3(sc.pos, f.x), p1 = ld3(sc.pos, f.y), p2 = ld3(sc.pos, f.z);
    float3 e0 = p1 - p0, e1 = p2 - p0;
    float pmf = 0.5f * norm(cross(e0, e1)) * e.normalization;
    float cdf_prev = idx > 0 ? cdf[idx - 1] * e.normalization : 0.f;
    sy = (sy - cdf_prev) / pmf;
    float t = safe_sqrt(1.f - sx), bx = 1.f - t, by = t * sy;
    ds.p = p0 + e0 * bx + e1 * by;
    if (m.has_normals) {
        float3 n0 = ld3(sc.nrm, f.x), n1 = ld3(sc.nrm, f.y), n2 = ld3(sc.nrm, f.z);
        ds.n = normalize(n0 * (1.f - bx - by) + n1 * bx + n2 * by);
    } else
        ds.n = normalize(cross(e0, e1));
    ds.d = ds.p - ref_p;
    float dist2 = dot(ds.d, ds.d);
    ds.dist = sqrtf(dist2);
    ds.d = ds.d / ds.dist;
    float dp = fabsf(dot(ds.d, ds.n));
    ds.pdf = e.normalization * (dp != 0.f ? dist2 / dp : 0.f);
    bool active = dot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
    ds.value = active ? emitter_radiance<M>(sc, e, wav) / ds.pdf : M::splat(0.f);
    return ds;
}
template <bool Shapes = true>
HD float area_pdf_direction(const SceneDev &sc, const EmitterDev &e, float3 ref_p, float3 d, float3 n, float dist) {
    if (Shapes) {
        int shape = sc.meshes[e.mesh].shape;
        if (shape >= 0) return dot(d, n) < 0.f ? shape_pdf_direction(sc.shapes[shape], ref_p, d, n, dist) : 0.f;
    }
    float dp = fabsf(dot(d, n));
    float pdf = e.normalization * (dp != 0.f ? (dist * dist) / dp : 0.f);
    return dot(d, n) < 0.f ? pdf : 0.f;
}

// ---- Scene::sample_emitter_direction (src/librender/scene.cpp:141-176), without the shadow ray ---
// ConstantBackgroundEmitter::sample_direction (src/emitters/constant.cpp:77-103); pdf_direction = 1/(4 pi)
template <typename M>
HD DirSampleT<M> constant_sample_direction(const SceneDev &sc, const EmitterDev &e, float3 ref_p, float sx, float sy, float4 wav) {
    DirSampleT<M> ds; float3 d = square_to_uniform_sphere(sx, sy);
    ds.dist = 2.f * e.bsphere_radius; ds.p = ref_p + d * ds.dist; ds.n = -d; ds.d = d;
    ds.pdf = 0.25f * kInvPi; ds.value = emitter_radiance<M>(sc, e, wav) * (1.f / ds.pdf); ds.delta = false;
    return ds;
}
// PointLight::sample_direction (src/emitters/point.cpp:74-98): delta position, pdf 1, intensity / dist^2
template <typename M>
HD DirSampleT<M> point_sample_direction(const SceneDev &sc, const EmitterDev &e, float3 ref_p, float4 wav) {
    DirSampleT<M> ds; ds.p = e.position; ds.n = mk3(0.f); ds.pdf = 1.f; ds.delta = true;
    ds.d = ds.p - ref_p; ds.dist = norm(ds.d);
    float inv = 1.f / ds.dist; ds.d = ds.d * inv;
    ds.value = emitter_radiance<M>(sc, e, wav) * (inv * inv);
    return ds;
}

template <typename M = Rgb>
HD DirSampleT<M> sample_emitter_direction(const SceneDev &sc, float3 ref_p, float sx, float sy, float4 wav = float4()) {
    DirSampleT<M> ds; ds.delta = false; ds.etype = 0; ds.pdf = 0.f; ds.value = M::splat(0.f); ds.d = mk3(0.f); ds.dist = 0.f; ds.p = ds.n = mk3(0.f);
    if (sc.n_emitters == 0) return ds;
    uint32_t index = 0; float epdf = 1.f;
    if (sc.n_emitters > 1) {
        float ne = (float) sc.n_emitters;
        epdf = 1.f / ne;
        index = min((uint32_t) (sx * ne), (uint32_t) sc.n_emitters - 1);
        sx = (sx - (float) index * epdf) * ne;
    }
    const EmitterDev &e = sc.emitters[index];
    if (e.type == 0) ds = area_sample_direction<M>(sc, e, ref_p, sx, sy, wav);
    else if (e.type == 1) ds = env_sample_direction<M>(sc, e, ref_p, sx, sy, wav);
    else if (e.type == 2) ds = constant_sample_direction<M>(sc, e, ref_p, sx, sy, wav);
    else ds = point_sample_direction<M>(sc, e, ref_p, wav);
    ds.delta = e.type == 3; ds.etype = e.type;
    if (sc.n_emitters > 1) { ds.pdf *= epdf; ds.value *= 1.f / epdf; }
    return ds;
}

} // namespace mtsb
