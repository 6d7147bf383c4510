// shade.h -- surface interaction reconstruction, textures and the three in-scope BSDFs (HD bodies).
#pragma once
#include "spectrum.h"

namespace mtsb {

struct SI {
    float3 p, n; Frame sh; float u, v; float3 wi; float t; uint32_t prim; int mesh; bool valid;
    float b1, b2;      // barycentrics of the hit (mesh_attribute textures interpolate with them)
};

HD float3 ld3(const float *a, uint32_t i) { return mk3(a[3 * i], a[3 * i + 1], a[3 * i + 2]); }

// ShapeKDTree::create_surface_interaction (include/mitsuba/render/kdtree.h:2334-2368) +
// Mesh::fill_surface_interaction (src/librender/mesh.cpp:399-462); miss: scene_native.inl:29-31.
// Shape::fill_surface_interaction of the analytic shapes (sphere.cpp:329-373, disk.cpp:167-205, cylinder.cpp:330-361,
// rectangle.cpp:180-205): position, normal, uv and dp_du; (cx, cy) = the floats ray_intersect cached (local x, y).
// Out of line: triangle scenes never pay for it in registers.
static __host__ __device__ __noinline__ void shape_fill_si(const ShapeDev &sh, float3 o, float3 d, float t, float cx, float cy,
                                                           float3 &p, float3 &n, float &u, float &v, float3 &dp_du) {
    float3 hitp = mk3(fmaf(d.x, t, o.x), fmaf(d.y, t, o.y), fmaf(d.z, t, o.z));
    if (sh.type == kShapeSphere) {
        n = normalize(hitp - sh.center);
        p = mk3(fmaf(n.x, sh.radius, sh.center.x), fmaf(n.y, sh.radius, sh.center.y), fmaf(n.z, sh.radius, sh.center.z));
        float3 l = xform_point_affine(sh.to_object, p);
        // unit_angle_z (include/mitsuba/core/vector.h)
        float temp = 2.f * asinf(0.5f * norm(mk3(l.x, l.y, l.z - mulsign(1.f, l.z))));
        float theta = l.z >= 0.f ? temp : kPi - temp;
        float phi = atan2f(l.y, l.x);
        if (phi < 0.f) phi += 2.f * kPi;
        u = phi * kInvTwoPi; v = theta * kInvPi;
        dp_du = xform_vec(sh.to_world, mk3(-l.y, l.x, 0.f)) * (2.f * kPi);
        if (sh.flip) n = -n;
    } else if (sh.type == kShapeDisk) {
        float r = sqrtf(cx * cx + cy * cy), inv_r = 1.f / r;
        float vv = atan2f(cy, cx) * kInvTwoPi;
        if (vv < 0.f) vv += 1.f;
        float cp = r != 0.f ? cx * inv_r : 1.f, sp = r != 0.f ? cy * inv_r : 0.f;
        dp_du = xform_vec(sh.to_world, mk3(cp, sp, 0.f));
        n = sh.n; u = r; v = vv; p = hitp;
    } else if (sh.type == kShapeCylinder) {
        float3 l = xform_point_affine(sh.to_object, hitp);
        float phi = atan2f(l.y, l.x);
        if (phi < 0.f) phi += 2.f * kPi;
        u = phi * kInvTwoPi; v = l.z / sh.length;
        dp_du = xform_vec(sh.to_world, mk3(-l.y, l.x, 0.f) * (2.f * kPi));
        float3 dp_dv = xform_vec(sh.to_world, mk3(0.f, 0.f, sh.length));
        n = normalize(cross(dp_du, dp_dv));
        p = hitp + n * (sh.radius - sqrtf(l.x * l.x + l.y * l.y));
        if (sh.flip) n = -n;
    } else {
        n = sh.n; dp_du = sh.dp_du; p = hitp;
        u = fmaf(cx, 0.5f, 0.5f); v = fmaf(cy, 0.5f, 0.5f);
    }
}

constexpr uint32_t kPrimMask = 0x3fffffffu;     // primitive ids use 30 bits; the top two carry the shade class in the queues
// Shapes = false: the scene has no analytic primitive (the lean Rgb / Spectral shade instances; scenes with shapes run the
// RgbX / SpectralX ones)
template <bool Shapes = true>
HD SI make_si(const SceneDev &sc, float3 o, float3 d, float4 hit) {
    SI si;
    si.t = hit.x; si.prim = f2u(hit.w) & kPrimMask; si.valid = hit.x != INFINITY;
    si.b1 = hit.y; si.b2 = hit.z;
    if (!si.valid) { si.wi = -d; si.mesh = -1; si.p = o; si.n = mk3(0.f); si.u = si.v = 0.f; si.sh.s = si.sh.t = si.sh.n = mk3(0.f); return si; }
    uint4 f = sc.faces[si.prim];
    si.mesh = (int) f.w;
    const MeshDev m = sc.meshes[f.w];
    if (Shapes && f.y == kAnalyticTag) {
        float3 dp_du;
        shape_fill_si(sc.shapes[f.x], o, d, hit.x, hit.y, hit.z, si.p, si.n, si.u, si.v, dp_du);
        si.sh.n = si.n;
        si.sh.s = normalize(dp_du - si.n * dot(si.n, dp_du));
        si.sh.t = cross(si.n, si.sh.s);
        si.wi = si.sh.to_local(-d);
        return si;
    }
    float b1 = hit.y, b2 = hit.z, b0 = 1.f - b1 - b2;
    float3 p0 = ld3(sc.pos, f.x), p1 = ld3(sc.pos, f.y), p2 = ld3(sc.pos, f.z);
    float3 dp0 = p1 - p0, dp1 = p2 - p0;
    si.p = p0 * b0 + p1 * b1 + p2 * b2;
    float3 n = normalize(cross(dp0, dp1));
    si.n = n;
    float3 dp_du, dp_dv;
    coordinate_system(n, dp_du, dp_dv);
    si.u = b1; si.v = b2;
    if (m.has_uv) {
        float u0 = sc.uv[2 * f.x], v0 = sc.uv[2 * f.x + 1], u1 = sc.uv[2 * f.y], v1 = sc.uv[2 * f.y + 1],
              u2 = sc.uv[2 * f.z], v2 = sc.uv[2 * f.z + 1];
        si.u = u0 * b0 + u1 * b1 + u2 * b2; si.v = v0 * b0 + v1 * b1 + v2 * b2;
        float du0 = u1 - u0, dv0 = v1 - v0, du1 = u2 - u0, dv1 = v2 - v0;
        float det = fmaf(du0, dv1, -(dv0 * du1));
        if (det != 0.f) {
            float inv = 1.f / det;
            dp_du = (dp0 * dv1 - dp1 * dv0) * inv;
            dp_dv = (dp1 * du0 - dp0 * du1) * inv;
        }
    }
    float3 ns = n;
    if (m.has_normals) {
        float3 n0 = ld3(sc.nrm, f.x), n1 = ld3(sc.nrm, f.y), n2 = ld3(sc.nrm, f.z);
        ns = normalize(n0 * b0 + n1 * b1 + n2 * b2);
    }
    si.sh.n = ns;
    si.sh.s = normalize(dp_du - ns * dot(ns, dp_du));
    si.sh.t = cross(ns, si.sh.s);
    si.wi = si.sh.to_local(-d);
    return si;
}

// src/textures/bitmap.cpp:250-293 (bilinear, repeat) and constant rgb (src/spectra/srgb.cpp)
HD float3 tex_eval_rgb(const SceneDev &sc, int id, float u, float v) {
    const TexDev &t = sc.tex[id];
    if (t.type != 1) return t.value;
    float tu = t.to_uv[0] * u + t.to_uv[1] * v + t.to_uv[2], tv = t.to_uv[3] * u + t.to_uv[4] * v + t.to_uv[5];
    tu -= floorf(tu); tv -= floorf(tv);
    tu *= (float) (t.w - 1); tv *= (float) (t.h - 1);
    uint32_t px = min((uint32_t) tu, t.w - 2), py = min((uint32_t) tv, t.h - 2);
    float w1x = tu - (float) px, w1y = tv - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    uint32_t idx = px + py * t.w;
    if (t.ch == 1) {
        float v00 = t.data[idx], v10 = t.data[idx + 1], v01 = t.data[idx + t.w], v11 = t.data[idx + t.w + 1];
        float v0 = fmaf(w0x, v00, w1x * v10), v1 = fmaf(w0x, v01, w1x * v11);
        return mk3(fmaf(w0y, v0, w1y * v1));
    }
    float3 v00 = ld3(t.data, idx), v10 = ld3(t.data, idx + 1), v01 = ld3(t.data, idx + t.w), v11 = ld3(t.data, idx + t.w + 1);
    float3 v0 = mk3(fmaf(w0x, v00.x, w1x * v10.x), fmaf(w0x, v00.y, w1x * v10.y), fmaf(w0x, v00.z, w1x * v10.z));
    float3 v1 = mk3(fmaf(w0x, v01.x, w1x * v11.x), fmaf(w0x, v01.y, w1x * v11.y), fmaf(w0x, v01.z, w1x * v11.z));
    return mk3(fmaf(w0y, v0.x, w1y * v1.x), fmaf(w0y, v0.y, w1y * v1.y), fmaf(w0y, v0.z, w1y * v1.z));
}
// Spectral mode: the same textures hold sRGB-model coefficients (srgb.cpp:44-48; bitmap.cpp:262-275 evaluates the model at
// the four texels and blends the spectra), or a tabulated / uniform spectrum; emitters multiply by D65 (srgb_d65.cpp:60-62).
// one out-of-line copy: the five texture kinds times the call sites of a shade launch otherwise inline to ~39 k SASS lines
#ifndef MTSB_TEXSPEC_INLINE
__host__ __device__ __noinline__
#else
HD
#endif
float4 tex_eval_spectral(const SceneDev &sc, int id, float u, float v, float4 wav) {
    const TexDev &t = sc.tex[id];
    float4 r;
    if (t.type == 0) r = srgb_model_eval(t.value, wav);
    else if (t.type == 2) { float inv = (float) (t.w - 1) / (t.lambda_max - t.lambda_min);
        r = map_spec(wav, [&](float l) { return regular_eval1(t.data, t.w, t.lambda_min, t.lambda_max, inv, l); }); }
    else if (t.type == 3) r = map_spec(wav, [&](float l) { return irregular_eval1(t.data, t.data + t.w, t.w, l, t.data + 2 * t.w, t.h); });
    else if (t.type == 4) r = mk4(t.value.x);
    else {
        float tu = t.to_uv[0] * u + t.to_uv[1] * v + t.to_uv[2], tv = t.to_uv[3] * u + t.to_uv[4] * v + t.to_uv[5];
        tu -= floorf(tu); tv -= floorf(tv);
        tu *= (float) (t.w - 1); tv *= (float) (t.h - 1);
        uint32_t px = min((uint32_t) tu, t.w - 2), py = min((uint32_t) tv, t.h - 2);
        float w1x = tu - (float) px, w1y = tv - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
        uint32_t idx = px + py * t.w;
        auto blend = [&](float a, float b) { return fmaf(w0x, a, w1x * b); };
        if (t.ch == 1) {
            float v0 = blend(t.data[idx], t.data[idx + 1]), v1 = blend(t.data[idx + t.w], t.data[idx + t.w + 1]);
            r = mk4(fmaf(w0y, v0, w1y * v1));
        } else {
            float4 c00 = srgb_model_eval(ld3(t.data, idx), wav), c10 = srgb_model_eval(ld3(t.data, idx + 1), wav),
                   c01 = srgb_model_eval(ld3(t.data, idx + t.w), wav), c11 = srgb_model_eval(ld3(t.data, idx + t.w + 1), wav);
            float4 c0 = map_spec(c00, c10, blend), c1 = map_spec(c01, c11, blend);
            r = map_spec(c0, c1, [&](float a, float b) { return fmaf(w0y, a, w1y * b); });
        }
    }
    if (t.d65_scale != 0.f) r = d65_eval(sc, wav) * t.d65_scale * r;
    return r;
}
template <typename M> HD typename M::Spec tex_eval_m(const SceneDev &sc, int id, float u, float v, float4 wav) {
    if constexpr (M::spectral) return tex_eval_spectral(sc, id, u, v, wav); else return tex_eval_rgb(sc, id, u, v);
}
HD float3 tex_eval(const SceneDev &sc, int id, float u, float v) { return tex_eval_rgb(sc, id, u, v); }
// `mesh_attribute` texture (src/textures/mesh_attribute.cpp -> Mesh::eval_attribute, include/mitsuba/render/mesh.h:278-318):
// a per-vertex (barycentric interpolation, fmadd(v0, b0, fmadd(v1, b1, v2 b2))) or per-face attribute of the mesh that was
// hit, 1 or 3 floats, times `scale`.  TexDev: type 5, data = the attribute buffer of that mesh, h = 1 per face / 0 per
// vertex, ch = 1 | 3, value.x = scale.  Spectral mode: 3-float attributes are sRGB-model coefficients (mesh.h:297-306).
template <typename M>
HD typename M::Spec mesh_attribute_eval(const SceneDev &sc, const TexDev &t, const SI &si, float4 wav) {
    typedef typename M::Spec Spec;
    if (!si.valid) return M::splat(0.f);
    const MeshDev m = sc.meshes[si.mesh];
    auto fetch = [&](uint32_t i) -> Spec {
        if (t.ch == 1) return M::splat(t.data[i]);
        float3 c = ld3(t.data, i);
        if constexpr (M::spectral) return srgb_model_eval(c, wav); else return c;
    };
    Spec r;
    if (t.h == 1) r = fetch(si.prim - m.face_begin);
    else {
        uint4 f = sc.faces[si.prim];
        float b0 = 1.f - si.b1 - si.b2;
        Spec v0 = fetch(f.x - m.vert_begin), v1 = fetch(f.y - m.vert_begin), v2 = fetch(f.z - m.vert_begin);
        r = v0 * b0 + (v1 * si.b1 + v2 * si.b2);
    }
    return r * t.value.x;
}
template <typename M> HD typename M::Spec tex_fetch(const SceneDev &sc, int id, const SI &si, float4 wav) {
    if constexpr (M::ext) { const TexDev &t = sc.tex[id]; if (t.type == 5) return mesh_attribute_eval<M>(sc, t, si, wav); }
    return tex_eval_m<M>(sc, id, si.u, si.v, wav);
}
// texel footprint of a bitmap lookup (for the adjoint scatter): indices of the 4 texels + weights
HD bool tex_footprint(const TexDev &t, float u, float v, uint32_t idx[4], float w[4]) {
    if (t.type != 1) return false;
    float tu = t.to_uv[0] * u + t.to_uv[1] * v + t.to_uv[2], tv = t.to_uv[3] * u + t.to_uv[4] * v + t.to_uv[5];
    tu -= floorf(tu); tv -= floorf(tv);
    tu *= (float) (t.w - 1); tv *= (float) (t.h - 1);
    uint32_t px = min((uint32_t) tu, t.w - 2), py = min((uint32_t) tv, t.h - 2);
    float w1x = tu - (float) px, w1y = tv - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    uint32_t i = px + py * t.w;
    idx[0] = i; idx[1] = i + 1; idx[2] = i + t.w; idx[3] = i + t.w + 1;
    w[0] = w0x * w0y; w[1] = w1x * w0y; w[2] = w0x * w1y; w[3] = w1x * w1y;
    return true;
}

// ---- warps (include/mitsuba/core/warp.h:54-90, 153-156, 332-356) ---------------------------------
HD void square_to_disk_concentric(float sx, float sy, float &ox, float &oy) {
    float x = fmaf(2.f, sx, -1.f), y = fmaf(2.f, sy, -1.f);
    bool is_zero = x == 0.f && y == 0.f, q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    // angle in units of pi: sincospif needs no large-argument reduction path (smaller kernel, no local-memory table)
    float phi = 0.25f * rp / r;
    if (q13) phi = 0.5f - phi;
    if (is_zero) phi = 0.f;
    float s, c; sincospif(phi, &s, &c);
    ox = r * c; oy = r * s;
}
HD float3 square_to_cosine_hemisphere(float sx, float sy) {
    float px, py; square_to_disk_concentric(sx, sy, px, py);
    return mk3(px, py, safe_sqrt(1.f - (px * px + py * py)));
}

template <typename M> struct BsdfSampleT { float3 wo; float pdf, eta; bool delta; typename M::Spec weight; };
typedef BsdfSampleT<RgbX> BsdfSample;       // stand-alone hooks handle every plugin

// ---- fresnel (include/mitsuba/render/fresnel.h:34-76, 92-116) ------------------------------------
HD void fresnel_dielectric(float cos_i, float eta, float &r, float &cos_t, float &eta_it, float &eta_ti) {
    bool outside = cos_i >= 0.f;
    float rcp_eta = 1.f / eta;
    eta_it = outside ? eta : rcp_eta; eta_ti = outside ? rcp_eta : eta;
    float cos_t_sqr = fmaf(-fmaf(-cos_i, cos_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = fabsf(cos_i), ct = safe_sqrt(cos_t_sqr);
    bool index_matched = eta == 1.f, special = index_matched || ci == 0.f;
    float r_sc = index_matched ? 0.f : 1.f;
    float a_s = fmaf(-eta_it, ct, ci) / fmaf(eta_it, ct, ci);
    float a_p = fmaf(-eta_it, ci, ct) / fmaf(eta_it, ci, ct);
    r = 0.5f * (a_s * a_s + a_p * a_p);
    if (special) r = r_sc;
    cos_t = mulsign_neg(ct, cos_i);
}
HD float fresnel_conductor(float cos_i, float eta_r, float eta_i) {
    float c2 = cos_i * cos_i, s2 = 1.f - c2, s4 = s2 * s2;
    float temp_1 = eta_r * eta_r - eta_i * eta_i - s2;
    float a2pb2 = safe_sqrt(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r);
    float a = safe_sqrt(0.5f * (a2pb2 + temp_1));
    float term_1 = a2pb2 + c2, term_2 = 2.f * cos_i * a;
    float r_s = (term_1 - term_2) / (term_1 + term_2);
    float term_3 = a2pb2 * c2 + s4, term_4 = term_2 * s2;
    float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return 0.5f * (r_s + r_p);
}

HD float3 fresnel_conductor(float c, float3 eta, float3 k) { return mk3(fresnel_conductor(c, eta.x, k.x), fresnel_conductor(c, eta.y, k.y), fresnel_conductor(c, eta.z, k.z)); }
HD float4 fresnel_conductor(float c, float4 eta, float4 k) { return make_float4(fresnel_conductor(c, eta.x, k.x), fresnel_conductor(c, eta.y, k.y), fresnel_conductor(c, eta.z, k.z), fresnel_conductor(c, eta.w, k.w)); }

// ---- microfacet (include/mitsuba/render/microfacet.h:184-413) ------------------------------------
HD float erfinv_giles(float x) {   // ext/enoki/include/enoki/special.h:222-247
    float w = -logf((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = sqrtf(w) - 3.f;
    float p1 = 2.81022636e-08f;
    p1 = fmaf(p1, w1, 3.43273939e-07f); p1 = fmaf(p1, w1, -3.5233877e-06f); p1 = fmaf(p1, w1, -4.39150654e-06f);
    p1 = fmaf(p1, w1, 0.00021858087f); p1 = fmaf(p1, w1, -0.00125372503f); p1 = fmaf(p1, w1, -0.00417768164f);
    p1 = fmaf(p1, w1, 0.246640727f); p1 = fmaf(p1, w1, 1.50140941f);
    float p2 = -0.000200214257f;
    p2 = fmaf(p2, w2, 0.000100950558f); p2 = fmaf(p2, w2, 0.00134934322f); p2 = fmaf(p2, w2, -0.00367342844f);
    p2 = fmaf(p2, w2, 0.00573950773f); p2 = fmaf(p2, w2, -0.0076224613f); p2 = fmaf(p2, w2, 0.00943887047f);
    p2 = fmaf(p2, w2, 1.00167406f); p2 = fmaf(p2, w2, 2.83297682f);
    return (w < 5.f ? p1 : p2) * x;
}
struct Microfacet {
    int type; float au, av; bool visible;
    HD Microfacet(int t, float a_u, float a_v, bool vis) : type(t), au(fmaxf(a_u, 1e-4f)), av(fmaxf(a_v, 1e-4f)), visible(vis) {}
    HD float eval(float3 m) const {
        float auv = au * av, ct = m.z, ct2 = ct * ct, result;
        if (type == 0)
            result = expf(-(sqr(m.x / au) + sqr(m.y / av)) / ct2) / (kPi * auv * sqr(ct2));
        else
            result = 1.f / (kPi * auv * sqr(sqr(m.x / au) + sqr(m.y / av) + sqr(m.z)));
        return result * ct > 1e-20f ? result : 0.f;
    }
    HD float smith_g1(float3 v, float3 m) const {
        float xy2 = sqr(au * v.x) + sqr(av * v.y), tan2 = xy2 / sqr(v.z), result;
        if (type == 0) {
            float a = 1.f / sqrtf(tan2), a2 = a * a;
            result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a2) / (1.f + 2.276f * a + 2.577f * a2);
        } else
            result = 2.f / (1.f + sqrtf(1.f + tan2));
        if (xy2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    HD float G(float3 wi, float3 wo, float3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }
    HD float pdf(float3 wi, float3 m) const {
        float r = eval(m);
        if (visible) r *= smith_g1(wi, m) * fabsf(dot(wi, m)) / wi.z; else r *= m.z;
        return r;
    }
    HD void sample_visible_11(float cos_i, float sx, float sy, float &ox, float &oy) const {
        if (type == 0) {
            float tan_i = safe_sqrt(fmaf(-cos_i, cos_i, 1.f)) / cos_i, cot_i = 1.f / tan_i;
            float maxval = erff(cot_i);
            sx = fmaxf(fminf(sx, 1.f - 1e-6f), 1e-6f); sy = fmaxf(fminf(sy, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * erff(sqrtf(-logf(sx)));
            sx *= 1.f + maxval + kInvSqrtPi * tan_i * expf(-sqr(cot_i));
            for (int i = 0; i < 3; ++i) {
                float slope = erfinv_giles(x);
                float value = 1.f + x + kInvSqrtPi * tan_i * expf(-sqr(slope)) - sx;
                float deriv = 1.f - slope * tan_i;
                x -= value / deriv;
            }
            ox = erfinv_giles(x); oy = erfinv_giles(fmaf(2.f, sy, -1.f));
        } else {
            float px, py; square_to_disk_concentric(sx, sy, px, py);
            float s = 0.5f * (1.f + cos_i);
            py = lerpf(safe_sqrt(1.f - px * px), py, s);
            float x = px, y = py, z = safe_sqrt(1.f - (px * px + py * py));
            float sin_i = safe_sqrt(1.f - cos_i * cos_i);
            float nrm = 1.f / fmaf(sin_i, y, cos_i * z);
            ox = fmaf(cos_i, y, -(sin_i * z)) * nrm; oy = x * nrm;
        }
    }
    HD float3 sample(float3 wi, float sx, float sy, float &pdf_out) const {
        if (!visible) {
            float sin_phi, cos_phi, cos_theta, cos_theta_2, alpha_2;
            if (au == av) { sincospif(2.f * sy, &sin_phi, &cos_phi); alpha_2 = au * au; }
            else {
                float ratio = av / au, ts, tc; sincospif(2.f * sy, &ts, &tc); float tmp = ratio * (ts / tc);
                cos_phi = 1.f / sqrtf(fmaf(tmp, tmp, 1.f));
                cos_phi = mulsign(cos_phi, fabsf(sy - 0.5f) - 0.25f);
                sin_phi = cos_phi * tmp;
                alpha_2 = 1.f / (sqr(cos_phi / au) + sqr(sin_phi / av));
            }
            if (type == 0) {
                cos_theta = 1.f / sqrtf(fmaf(-alpha_2, logf(1.f - sx), 1.f));
                cos_theta_2 = cos_theta * cos_theta;
                float c3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdf_out = (1.f - sx) / (kPi * au * av * c3);
            } else {
                float t2 = alpha_2 * sx / (1.f - sx);
                cos_theta = 1.f / sqrtf(1.f + t2); cos_theta_2 = cos_theta * cos_theta;
                float temp = 1.f + t2 / alpha_2, c3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdf_out = 1.f / (kPi * au * av * c3 * sqr(temp));
            }
            float sin_theta = sqrtf(1.f - cos_theta_2);
            return mk3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
        }
        float3 wi_p = normalize(mk3(au * wi.x, av * wi.y, wi.z));
        // Frame::sincos_phi (include/mitsuba/core/frame.h): degenerate -> (0, 1)
        float sin_theta_2 = fmaf(-wi_p.z, wi_p.z, 1.f), inv = 1.f / sqrtf(sin_theta_2);
        float sin_phi = fminf(fmaxf(wi_p.y * inv, -1.f), 1.f), cos_phi = fminf(fmaxf(wi_p.x * inv, -1.f), 1.f);
        if (fabsf(sin_theta_2) <= 4.f * kEpsilon) { sin_phi = 0.f; cos_phi = 1.f; }
        float slx, sly; sample_visible_11(wi_p.z, sx, sy, slx, sly);
        float rx = fmaf(cos_phi, slx, -(sin_phi * sly)) * au, ry = fmaf(sin_phi, slx, cos_phi * sly) * av;
        float3 m = normalize(mk3(-rx, -ry, 1.f));
        pdf_out = eval(m) * smith_g1(wi, m) * fabsf(dot(wi, m)) / wi.z;
        return m;
    }
};

// ---- rough plastic tables (roughplastic.cpp: parameters_changed; microfacet.h:454-552 eval_reflectance / eval_transmittance) ----
// One entry: the hemispherical transmittance (or reflectance) of a rough dielectric boundary for incidence cosine mu, by a
// res x res Gauss-Legendre rule over the visible-normal sample square.  nodes / weights: the rule on [-1, 1].
constexpr int kRoughTransRes = 64;       // MTS_ROUGH_TRANSMITTANCE_RES
HD float rough_boundary_integral(int distr_type, float alpha, float eta, float mu, const float *nodes, const float *weights, int res, bool transmit) {
    Microfacet distr(distr_type, alpha, alpha, true);
    float3 wi = mk3(sqrtf(1.f - mu * mu), 0.f, mu);
    float accum = 0.f;
    for (int a = 0; a < res; ++a)
        for (int b = 0; b < res; ++b) {
            float pdf; float3 m = distr.sample(wi, fmaf(nodes[b], 0.5f, 0.5f), fmaf(nodes[a], 0.5f, 0.5f), pdf);
            float f, cos_t, eta_it, eta_ti, wim = dot(wi, m);
            fresnel_dielectric(wim, eta, f, cos_t, eta_it, eta_ti);
            float smith;
            if (transmit) {
                float3 wo = m * fmaf(wim, eta_ti, cos_t) - wi * eta_ti;
                smith = wo.z * wi.z >= 0.f ? 0.f : distr.smith_g1(wo, m) * (1.f - f);
            } else {
                float3 wo = m * (2.f * wim) - wi;
                smith = (wo.z <= 0.f || wi.z <= 0.f) ? 0.f : distr.smith_g1(wo, m) * f;
            }
            accum += smith * (weights[a] * weights[b]);
        }
    return accum * 0.25f;
}
HD int rough_table_res(float eta) { return eta > 1.f ? 32 : 128; }          // microfacet.h:468-472 (already multiples of the packet size)
HD float rough_table_mu(int i) { return fmaxf(1e-6f, (float) i / (float) (kRoughTransRes - 1)); }
// roughplastic.cpp: lerp_gather over the 64-entry table
HD float rough_lerp(const float *data, float x) {
    x *= (float) (kRoughTransRes - 1);
    uint32_t index = min((uint32_t) x, (uint32_t) kRoughTransRes - 2u);
    return lerpf(data[index], data[index + 1], x - (float) index);
}

// ---- BSDF::sample / eval / pdf --------------------------------------------------------------------
// The textures of a BSDF at one surface point, looked up once per path vertex and shared by eval (emitter sample) and
// sample: diffuse a = reflectance; roughconductor a = eta, b = k, c = specular_reflectance (1 if absent);
// dielectric a = specular_reflectance, b = specular_transmittance (1 if absent); conductor like roughconductor;
// plastic a = diffuse_reflectance, b = specular_reflectance (1 if absent).
template <typename M> struct BsdfTexT { typename M::Spec a, b, c; float opacity; /* `mask` adapter: clamp(opacity texture), else 1 */ };
// fresnel_diffuse_reflectance (include/mitsuba/render/fresnel.h:324-358)
HD float fresnel_diffuse_reflectance(float eta) {
    if (eta < 1.f) return -1.4399f * (eta * eta) + 0.7099f * eta + 0.6681f + 0.0636f / eta;
    float i1 = 1.f / eta, i2 = i1 * i1, i3 = i2 * i1, i4 = i3 * i1, i5 = i4 * i1;
    return 0.919317f - 3.4793f * i1 + 6.75335f * i2 - 7.80989f * i3 + 4.98554f * i4 - 1.36881f * i5;
}
// plastic.cpp:236,262: diffuse reflectance divided by 1 - fdr_int (times the reflectance itself when `nonlinear`)
HD float3 plastic_div(float3 d, float f, bool nl) { return mk3(d.x / (1.f - (nl ? d.x * f : f)), d.y / (1.f - (nl ? d.y * f : f)), d.z / (1.f - (nl ? d.z * f : f))); }
HD float4 plastic_div(float4 d, float f, bool nl) { return make_float4(d.x / (1.f - (nl ? d.x * f : f)), d.y / (1.f - (nl ? d.y * f : f)), d.z / (1.f - (nl ? d.z * f : f)), d.w / (1.f - (nl ? d.w * f : f))); }
template <typename M> HD typename M::Spec plastic_diffuse(const BsdfDev &b, typename M::Spec d) { return plastic_div(d, fresnel_diffuse_reflectance(1.f / b.eta_f), b.nonlinear != 0); }
template <typename M>
HD BsdfTexT<M> bsdf_textures(const SceneDev &sc, const BsdfDev &b, const SI &si, float4 wav) {
    BsdfTexT<M> t; t.a = t.b = t.c = M::splat(1.f); t.opacity = 1.f;
    if (M::ext && b.opacity >= 0) t.opacity = fminf(fmaxf(tex_fetch<M>(sc, b.opacity, si, wav).x, 0.f), 1.f);      // mask.cpp: eval_opacity
    if (b.type == 0) t.a = tex_fetch<M>(sc, b.refl, si, wav);
    else if (M::ext && (b.type == 4 || b.type == 6)) { t.a = tex_fetch<M>(sc, b.refl, si, wav); if (b.spec_r >= 0) t.b = tex_fetch<M>(sc, b.spec_r, si, wav); }
    else if (b.type == 1 || (M::ext && b.type == 3)) {
        t.a = tex_fetch<M>(sc, b.eta, si, wav); t.b = tex_fetch<M>(sc, b.k, si, wav);
        if (b.spec_r >= 0) t.c = tex_fetch<M>(sc, b.spec_r, si, wav);
    } else {                      // dielectric, rough dielectric
        if (b.spec_r >= 0) t.a = tex_fetch<M>(sc, b.spec_r, si, wav);
        if (b.spec_t >= 0) t.b = tex_fetch<M>(sc, b.spec_t, si, wav);
    }
    return t;
}
template <typename M>
HD void bsdf_eval_pdf_one_sided(const SceneDev &sc, const BsdfDev &b, const SI &si, float3 wo, typename M::Spec &val, float &pdf, const BsdfTexT<M> &tx);
template <typename M>
HD BsdfSampleT<M> bsdf_sample_one_sided(const SceneDev &sc, const BsdfDev &b, const SI &si, float s1, float sx, float sy, const BsdfTexT<M> &tx) {
    typedef typename M::Spec Spec;
    BsdfSampleT<M> bs; bs.wo = mk3(0.f); bs.pdf = 0.f; bs.eta = 1.f; bs.delta = false; bs.weight = M::splat(0.f);
    float cos_i = si.wi.z;
    if (b.type == 0) {            // src/bsdfs/diffuse.cpp:78-103
        if (!(cos_i > 0.f)) return bs;
        bs.wo = square_to_cosine_hemisphere(sx, sy);
        bs.pdf = kInvPi * bs.wo.z;
        if (bs.pdf > 0.f) bs.weight = tx.a;
        return bs;
    }
    if (b.type == 1) {            // src/bsdfs/roughconductor.cpp:196-276
        if (!(cos_i > 0.f)) return bs;
        Microfacet distr(b.distr, b.alpha_u, b.alpha_v, b.sample_visible != 0);
        float3 m = distr.sample(si.wi, sx, sy, bs.pdf);
        bs.wo = m * (2.f * dot(si.wi, m)) - si.wi;
        bool active = bs.pdf != 0.f && bs.wo.z > 0.f;
        float weight = b.sample_visible ? distr.smith_g1(bs.wo, m)
                                        : distr.G(si.wi, bs.wo, m) * dot(si.wi, m) / (cos_i * m.z);
        bs.pdf /= 4.f * dot(bs.wo, m);
        Spec F = fresnel_conductor(dot(si.wi, m), tx.a, tx.b);
        if (b.spec_r >= 0) F *= tx.c;
        bs.weight = active ? F * weight : M::splat(0.f);
        return bs;
    }
    if (M::ext && b.type == 3) {  // src/bsdfs/conductor.cpp (sample): delta reflection, weight = reflectance * F(cos_i)
        if (!(cos_i > 0.f)) return bs;
        bs.wo = mk3(-si.wi.x, -si.wi.y, si.wi.z); bs.pdf = 1.f; bs.delta = true;
        Spec F = fresnel_conductor(cos_i, tx.a, tx.b);
        if (b.spec_r >= 0) F = tx.c * F;
        bs.weight = F;
        return bs;
    }
    if (M::ext && b.type == 4) {  // src/bsdfs/plastic.cpp:178-245
        if (!(cos_i > 0.f)) return bs;
        float f_i, ct, e1, e2; fresnel_dielectric(cos_i, b.eta_f, f_i, ct, e1, e2);
        float prob_specular = f_i * b.spec_weight, prob_diffuse = (1.f - f_i) * (1.f - b.spec_weight);
        prob_specular = prob_specular / (prob_specular + prob_diffuse); prob_diffuse = 1.f - prob_specular;
        if (s1 < prob_specular) {
            bs.wo = mk3(-si.wi.x, -si.wi.y, si.wi.z); bs.pdf = prob_specular; bs.delta = true;
            Spec value = M::splat(f_i / bs.pdf);
            if (b.spec_r >= 0) value *= tx.b;
            bs.weight = value;
        } else {
            bs.wo = square_to_cosine_hemisphere(sx, sy); bs.pdf = prob_diffuse * (kInvPi * bs.wo.z);
            float f_o; fresnel_dielectric(bs.wo.z, b.eta_f, f_o, ct, e1, e2);
            bs.weight = plastic_diffuse<M>(b, tx.a) * ((1.f / (b.eta_f * b.eta_f)) * (1.f - f_i) * (1.f - f_o) / prob_diffuse);
        }
        return bs;
    }
    if (M::ext && b.type == 6) {  // src/bsdfs/roughplastic.cpp (sample): pick a lobe, then weight = eval / pdf of BOTH lobes
        if (!(cos_i > 0.f)) return bs;
        float t_i = rough_lerp(b.rough_t, cos_i);
        float prob_specular = (1.f - t_i) * b.spec_weight, prob_diffuse = t_i * (1.f - b.spec_weight);
        prob_specular = prob_specular / (prob_specular + prob_diffuse);
        if (s1 < prob_specular) {
            Microfacet distr(b.distr, b.alpha_u, b.alpha_u, b.sample_visible != 0);
            float unused; float3 m = distr.sample(si.wi, sx, sy, unused);
            bs.wo = m * (2.f * dot(si.wi, m)) - si.wi;
        } else
            bs.wo = square_to_cosine_hemisphere(sx, sy);
        Spec val; bsdf_eval_pdf_one_sided<M>(sc, b, si, bs.wo, val, bs.pdf, tx);
        bs.weight = bs.pdf > 0.f ? val * (1.f / bs.pdf) : M::splat(0.f);
        if (!(bs.pdf > 0.f)) bs.pdf = 0.f;
        return bs;
    }
    if (M::ext && b.type == 5) {  // src/bsdfs/roughdielectric.cpp (sample): microfacet normal, then Fresnel lobe choice
        if (cos_i == 0.f) return bs;
        Microfacet distr(b.distr, b.alpha_u, b.alpha_v, b.sample_visible != 0), sample_distr = distr;
        if (!b.sample_visible) { float sc_ = 1.2f - 0.2f * sqrtf(fabsf(cos_i)); sample_distr.au *= sc_; sample_distr.av *= sc_; }
        float3 wi_up = mk3(mulsign(si.wi.x, cos_i), mulsign(si.wi.y, cos_i), mulsign(si.wi.z, cos_i));
        float3 m = sample_distr.sample(wi_up, sx, sy, bs.pdf);
        if (bs.pdf == 0.f) return bs;
        float F, cos_t, eta_it, eta_ti, wim = dot(si.wi, m);
        fresnel_dielectric(wim, b.eta_f, F, cos_t, eta_it, eta_ti);
        bool sel_r = s1 <= F;
        Spec weight = M::splat(1.f);
        bs.pdf *= sel_r ? F : 1.f - F;
        bs.eta = sel_r ? 1.f : eta_it;
        float dwh_dwo;
        if (sel_r) {
            bs.wo = m * (2.f * wim) - si.wi;
            if (b.spec_r >= 0) weight *= tx.a;
            dwh_dwo = 1.f / (4.f * dot(bs.wo, m));
        } else {
            bs.wo = m * fmaf(wim, eta_ti, cos_t) - si.wi * eta_ti;
            Spec factor = M::splat(sqr(eta_ti));
            if (b.spec_t >= 0) factor *= tx.b;
            weight *= factor;
            dwh_dwo = (sqr(bs.eta) * dot(bs.wo, m)) / sqr(wim + bs.eta * dot(bs.wo, m));
        }
        if (b.sample_visible) weight *= distr.smith_g1(bs.wo, m);
        else weight *= distr.G(si.wi, bs.wo, m) * wim / (cos_i * m.z);
        bs.pdf *= fabsf(dwh_dwo);
        bs.weight = weight;
        return bs;
    }
    // src/bsdfs/dielectric.cpp:201-316
    float r_i, cos_t, eta_it, eta_ti;
    fresnel_dielectric(cos_i, b.eta_f, r_i, cos_t, eta_it, eta_ti);
    float t_i = 1.f - r_i;
    bool sel_r = s1 <= r_i;
    bs.pdf = sel_r ? r_i : t_i;
    bs.delta = true;
    bs.wo = sel_r ? mk3(-si.wi.x, -si.wi.y, si.wi.z) : mk3(-eta_ti * si.wi.x, -eta_ti * si.wi.y, cos_t);
    bs.eta = sel_r ? 1.f : eta_it;
    Spec w = M::splat(1.f);
    if (sel_r) { if (b.spec_r >= 0) w *= tx.a; }
    else { if (b.spec_t >= 0) w *= tx.b; w *= sqr(eta_ti); }
    bs.weight = w;
    return bs;
}

template <typename M>
HD void bsdf_eval_pdf_one_sided(const SceneDev &sc, const BsdfDev &b, const SI &si, float3 wo, typename M::Spec &val, float &pdf, const BsdfTexT<M> &tx) {
    typedef typename M::Spec Spec;
    val = M::splat(0.f); pdf = 0.f;
    float ci = si.wi.z, co = wo.z;
    if (b.type == 0) {            // src/bsdfs/diffuse.cpp:105-135
        if (ci > 0.f && co > 0.f) { val = tx.a * kInvPi * co; pdf = kInvPi * co; }
        return;
    }
    if (b.type == 1) {            // src/bsdfs/roughconductor.cpp:278-394
        if (!(ci > 0.f && co > 0.f)) return;
        float3 H = normalize(wo + si.wi);
        Microfacet distr(b.distr, b.alpha_u, b.alpha_v, b.sample_visible != 0);
        float D = distr.eval(H);
        if (D != 0.f) {
            float G = distr.G(si.wi, wo, H);
            float res = D * G / (4.f * ci);
            Spec F = fresnel_conductor(dot(si.wi, H), tx.a, tx.b);
            if (b.spec_r >= 0) F *= tx.c;
            val = F * res;
        }
        if (dot(si.wi, H) > 0.f && dot(wo, H) > 0.f)
            pdf = b.sample_visible ? D * distr.smith_g1(si.wi, H) / (4.f * ci) : distr.pdf(si.wi, H) / (4.f * dot(wo, H));
        return;
    }
    if (M::ext && b.type == 5) {  // src/bsdfs/roughdielectric.cpp (eval, pdf)
        if (ci == 0.f) return;
        bool reflect = ci * co > 0.f;
        float eta = ci > 0.f ? b.eta_f : 1.f / b.eta_f, inv_eta = ci > 0.f ? 1.f / b.eta_f : b.eta_f;
        float3 m = normalize(si.wi + wo * (reflect ? 1.f : eta));
        m = mk3(mulsign(m.x, m.z), mulsign(m.y, m.z), mulsign(m.z, m.z));
        Microfacet distr(b.distr, b.alpha_u, b.alpha_v, b.sample_visible != 0), sample_distr = distr;
        float wim = dot(si.wi, m), wom = dot(wo, m);
        float D = distr.eval(m), F, ct, e1, e2;
        fresnel_dielectric(wim, b.eta_f, F, ct, e1, e2);
        float G = distr.G(si.wi, wo, m);
        if (reflect) {
            val = M::splat(F * D * G / (4.f * fabsf(ci)));
            if (b.spec_r >= 0) val *= tx.a;
        } else {
            val = M::splat(fabsf((sqr(inv_eta) * (1.f - F) * D * G * eta * eta * wim * wom) / (ci * sqr(wim + eta * wom))));
            if (b.spec_t >= 0) val *= tx.b;
        }
        if (!(wim * ci > 0.f && wom * co > 0.f)) return;          // pdf only: micro / macro surface disagree on the side
        float dwh_dwo = reflect ? 1.f / (4.f * wom) : (eta * eta * wom) / sqr(wim + eta * wom);
        if (!b.sample_visible) { float sc_ = 1.2f - 0.2f * sqrtf(fabsf(ci)); sample_distr.au *= sc_; sample_distr.av *= sc_; }
        float prob = sample_distr.pdf(mk3(mulsign(si.wi.x, ci), mulsign(si.wi.y, ci), mulsign(si.wi.z, ci)), m);
        prob *= reflect ? F : 1.f - F;
        pdf = prob * fabsf(dwh_dwo);
        return;
    }
    if (M::ext && b.type == 6) {  // src/bsdfs/roughplastic.cpp (eval, pdf)
        if (!(ci > 0.f && co > 0.f)) return;
        Microfacet distr(b.distr, b.alpha_u, b.alpha_u, b.sample_visible != 0);
        float3 H = normalize(wo + si.wi);
        float D = distr.eval(H), F, ct, e1, e2;
        fresnel_dielectric(dot(si.wi, H), b.eta_f, F, ct, e1, e2);
        Spec value = M::splat(F * D * distr.G(si.wi, wo, H) / (4.f * ci));
        if (b.spec_r >= 0) value *= tx.b;
        float t_i = rough_lerp(b.rough_t, ci), t_o = rough_lerp(b.rough_t, co);
        val = value + plastic_div(tx.a, b.internal_refl, b.nonlinear != 0) * (kInvPi * (1.f / (b.eta_f * b.eta_f)) * co * t_i * t_o);
        float prob_specular = (1.f - t_i) * b.spec_weight, prob_diffuse = t_i * (1.f - b.spec_weight);
        prob_specular = prob_specular / (prob_specular + prob_diffuse); prob_diffuse = 1.f - prob_specular;
        float r = b.sample_visible ? D * distr.smith_g1(si.wi, H) / (4.f * ci) : distr.pdf(si.wi, H) / (4.f * dot(wo, H));
        pdf = r * prob_specular + prob_diffuse * (kInvPi * co);
        return;
    }
    if (M::ext && b.type == 4) {  // src/bsdfs/plastic.cpp:247-296: the diffuse lobe (the specular one is a delta)
        if (!(ci > 0.f && co > 0.f)) return;
        float f_i, f_o, ct, e1, e2; fresnel_dielectric(ci, b.eta_f, f_i, ct, e1, e2); fresnel_dielectric(co, b.eta_f, f_o, ct, e1, e2);
        val = plastic_diffuse<M>(b, tx.a) * ((kInvPi * co) * (1.f / (b.eta_f * b.eta_f)) * (1.f - f_i) * (1.f - f_o));
        float prob_specular = f_i * b.spec_weight, prob_diffuse = (1.f - f_i) * (1.f - b.spec_weight);
        pdf = (kInvPi * co) * (prob_diffuse / (prob_specular + prob_diffuse));
    }
}

// `twosided` (src/bsdfs/twosided.cpp:98-200): on the back side mirror wi (and wo) about the surface and use the
// nested BRDF; cos(wi) == 0 belongs to neither side.  One call site each, so the nested code is not duplicated.
template <typename M = Rgb>
HD BsdfSampleT<M> bsdf_sample(const SceneDev &sc, const BsdfDev &b, const SI &si, float s1, float sx, float sy, const BsdfTexT<M> &tx) {
    if (M::ext && b.opacity >= 0) { // src/bsdfs/mask.cpp:95-128: null lobe with probability 1 - opacity, else the nested BSDF with
        if (!(s1 < tx.opacity)) { // sample1 rescaled; the nested sample's pdf is NOT multiplied by the opacity (as in the reference)
            BsdfSampleT<M> bs; bs.wo = -si.wi; bs.eta = 1.f; bs.pdf = 1.f - tx.opacity; bs.delta = true; bs.weight = M::splat(1.f);
            return bs;
        }
        s1 /= tx.opacity;
    }
    bool back = b.two_sided && si.wi.z < 0.f;
    SI s2 = si;
    if (back) s2.wi.z = -s2.wi.z;
    BsdfSampleT<M> bs = bsdf_sample_one_sided<M>(sc, b, s2, s1, sx, sy, tx);
    if (back) bs.wo.z = -bs.wo.z;
    return bs;
}
template <typename M = Rgb>
HD void bsdf_eval_pdf(const SceneDev &sc, const BsdfDev &b, const SI &si, float3 wo, typename M::Spec &val, float &pdf, const BsdfTexT<M> &tx) {
    bool back = b.two_sided && si.wi.z < 0.f;
    SI s2 = si;
    if (back) { s2.wi.z = -s2.wi.z; wo.z = -wo.z; }
    bsdf_eval_pdf_one_sided<M>(sc, b, s2, wo, val, pdf, tx);
    if (M::ext && b.opacity >= 0) { val *= tx.opacity; pdf *= tx.opacity; }      // mask.cpp:130-155
}
// stand-alone forms (parity hooks, adjoint): look the textures up themselves
template <typename M = RgbX>
HD BsdfSampleT<M> bsdf_sample(const SceneDev &sc, const BsdfDev &b, const SI &si, float s1, float sx, float sy, float4 wav = float4()) {
    return bsdf_sample<M>(sc, b, si, s1, sx, sy, bsdf_textures<M>(sc, b, si, wav));
}
template <typename M = RgbX>
HD void bsdf_eval_pdf(const SceneDev &sc, const BsdfDev &b, const SI &si, float3 wo, typename M::Spec &val, float &pdf, float4 wav = float4()) {
    bsdf_eval_pdf<M>(sc, b, si, wo, val, pdf, bsdf_textures<M>(sc, b, si, wav));
}
HD bool bsdf_is_smooth(const BsdfDev &b) { return b.type != 2 && b.type != 3; }      // dielectric and smooth conductor are all-delta

} // namespace mtsb
