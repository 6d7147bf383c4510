// adjoint.h -- vertex-position adjoint of the path tracer (HD bodies).
//
// What the reference does: src/python/python/autodiff.py:6-194 records the whole render on the Enoki tape
// (ext/enoki/src/autodiff/autodiff.cpp:838-988) and back-propagates through it; vertex positions are exposed by
// Mesh::traverse / parameters_changed (src/librender/mesh.cpp:781-804).  `pathreparam` is not in the tree.
// What this file does instead (no tape, O(pool) memory), in the wavefront replay of grad.cuh:
//
//   dI/dv = interior term + boundary term on the film (primary visibility) + boundary term seen from x1 (shadows).
//
// * Interior term.  The pixel sample keeps its film position (x1 = camera ray ∩ moving triangle slides along the ray);
//   every later vertex keeps its barycentric coordinates on the triangle it hit ("material" parameterisation), so the
//   path integral is taken over triangle areas.  With sampling densities, MIS weights and Russian roulette detached
//   (they are functions of the material coordinates only), a path's radiance is a sum of products of LOCAL factors
//       S_k = f_s(x_{k-1}, x_k, x_{k+1}) * |N_{k+1} . D| / (2 |D|^3),     D = x_{k+1} - x_k,  N = e1 x e2 (unnormalised),
//   (BSDF value incl. the shading cosine, times the geometry term incl. the triangle-area Jacobian), and
//       dL/dv = sum_k (L - P_k) * dlog S_k/dv,
//   where P_k is the radiance gathered before S_k is used -- the same bookkeeping as texture_adjoint() in path.h, so
//   the replay only needs the two previous vertices of a path (PathQueue::prev1/prev2).  dlog f_s is evaluated with
//   forward-mode dual numbers over (wi, wo); everything else is closed form.
// * Boundary term (edge_* bodies): Reynolds transport over the film plane.  Mesh edges are sampled proportionally to
//   their projected length, the radiance on both sides of the edge is path-traced with common random numbers, and
//   (L+ - L-) * dL/dI(pixel) * d(alpha)/dv is scattered to the two edge vertices (Li et al. 2018, primary visibility).
//   Every edge is a candidate, so shading discontinuities of flat-shaded meshes are covered as well as silhouettes.
// * Boundary term seen from the first path vertex (edge2_* bodies, end of this file): silhouette edges as seen from x1.
// Not covered (documented in DESIGN.md): visibility changes seen from the second and later vertices, paths beyond a delta
// (dielectric) vertex, mesh-mesh intersections, emitter meshes.
#pragma once
#include "path.h"

namespace mtsb {

// ---- forward-mode dual numbers over N inputs ------------------------------------------------------------------------
template <int N> struct Dual { float v; float d[N]; };
template <int N> HD Dual<N> dconst(float v) { Dual<N> r; r.v = v; for (int i = 0; i < N; ++i) r.d[i] = 0.f; return r; }
template <int N> HD Dual<N> dvar(float v, int k) { Dual<N> r = dconst<N>(v); r.d[k] = 1.f; return r; }
template <int N> HD Dual<N> operator+(Dual<N> a, Dual<N> b) { a.v += b.v; for (int i = 0; i < N; ++i) a.d[i] += b.d[i]; return a; }
template <int N> HD Dual<N> operator-(Dual<N> a, Dual<N> b) { a.v -= b.v; for (int i = 0; i < N; ++i) a.d[i] -= b.d[i]; return a; }
template <int N> HD Dual<N> operator-(Dual<N> a) { a.v = -a.v; for (int i = 0; i < N; ++i) a.d[i] = -a.d[i]; return a; }
template <int N> HD Dual<N> operator*(Dual<N> a, Dual<N> b) { Dual<N> r; r.v = a.v * b.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i]; return r; }
template <int N> HD Dual<N> operator/(Dual<N> a, Dual<N> b) {
    Dual<N> r; float inv = 1.f / b.v; r.v = a.v * inv;
    for (int i = 0; i < N; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * inv;
    return r;
}
template <int N> HD Dual<N> operator+(Dual<N> a, float b) { a.v += b; return a; }
template <int N> HD Dual<N> operator+(float b, Dual<N> a) { a.v += b; return a; }
template <int N> HD Dual<N> operator-(Dual<N> a, float b) { a.v -= b; return a; }
template <int N> HD Dual<N> operator-(float b, Dual<N> a) { return -a + b; }
template <int N> HD Dual<N> operator*(Dual<N> a, float b) { a.v *= b; for (int i = 0; i < N; ++i) a.d[i] *= b; return a; }
template <int N> HD Dual<N> operator*(float b, Dual<N> a) { return a * b; }
template <int N> HD Dual<N> operator/(Dual<N> a, float b) { return a * (1.f / b); }
template <int N> HD Dual<N> operator/(float b, Dual<N> a) { return dconst<N>(b) / a; }
template <int N> HD Dual<N> dsqr(Dual<N> a) { return a * a; }
template <int N> HD Dual<N> dsqrt(Dual<N> a) {          // safe_sqrt: clamps at 0 with a zero derivative
    if (!(a.v > 0.f)) return dconst<N>(0.f);
    Dual<N> r; r.v = sqrtf(a.v); float h = 0.5f / r.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * h; return r;
}
template <int N> HD Dual<N> dexp(Dual<N> a) { Dual<N> r; r.v = expf(a.v); for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * r.v; return r; }

typedef Dual<6> D6;      // inputs: wi.xyz (0..2), wo.xyz (3..5), both in the local shading frame

// MicrofacetDistribution::smith_g1 (include/mitsuba/render/microfacet.h:289-321), same branches as Microfacet::smith_g1
HD D6 rc_smith_g1(const D6 v[3], D6 dot_vm, float au, float av, int type) {
    D6 xy2 = dsqr(v[0] * au) + dsqr(v[1] * av);
    if (xy2.v == 0.f) return dconst<6>(1.f);
    if (dot_vm.v * v[2].v <= 0.f) return dconst<6>(0.f);
    D6 tan2 = xy2 / dsqr(v[2]);
    if (type == 0) {
        D6 a = 1.f / dsqrt(tan2);
        if (a.v >= 1.6f) return dconst<6>(1.f);
        D6 a2 = a * a;
        return (3.535f * a + 2.181f * a2) / (1.f + 2.276f * a + 2.577f * a2);
    }
    return 2.f / (1.f + dsqrt(1.f + tan2));
}
// fresnel_conductor (include/mitsuba/render/fresnel.h:92-116) on dual numbers
HD D6 rc_fresnel_conductor(D6 cos_i, float eta_r, float eta_i) {
    D6 c2 = cos_i * cos_i, s2 = 1.f - c2, s4 = s2 * s2;
    D6 temp_1 = (eta_r * eta_r - eta_i * eta_i) - s2;
    D6 a2pb2 = dsqrt(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r);
    D6 a = dsqrt(0.5f * (a2pb2 + temp_1));
    D6 term_1 = a2pb2 + c2, term_2 = 2.f * cos_i * a;
    D6 r_s = (term_1 - term_2) / (term_1 + term_2);
    D6 term_3 = a2pb2 * c2 + s4, term_4 = term_2 * s2;
    D6 r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return 0.5f * (r_s + r_p);
}

// sum_c w_c * dlog f_c / d(wi, wo) for the smooth BSDFs, f = BSDF value * cos(wo) as BSDF::eval returns it
// (src/bsdfs/diffuse.cpp:105-121, src/bsdfs/roughconductor.cpp:278-340, twosided.cpp:144-170).  false: f = 0.
HD bool bsdf_log_grad(const SceneDev &sc, const BsdfDev &b, float u, float v, float3 wi, float3 wo, float3 w,
                      float3 &g_wi, float3 &g_wo) {
    g_wi = mk3(0.f); g_wo = mk3(0.f);
    bool back = b.two_sided && wi.z < 0.f;
    if (back) { wi.z = -wi.z; wo.z = -wo.z; }
    if (!(wi.z > 0.f && wo.z > 0.f)) return false;
    float q = w.x + w.y + w.z;
    if (b.type == 0) {
        g_wo.z = q / wo.z;
    } else if (b.type == 1) {
        D6 i[3] = { dvar<6>(wi.x, 0), dvar<6>(wi.y, 1), dvar<6>(wi.z, 2) }, o[3] = { dvar<6>(wo.x, 3), dvar<6>(wo.y, 4), dvar<6>(wo.z, 5) };
        D6 h[3] = { i[0] + o[0], i[1] + o[1], i[2] + o[2] };
        D6 inv = 1.f / dsqrt(dsqr(h[0]) + dsqr(h[1]) + dsqr(h[2]));
        for (int k = 0; k < 3; ++k) h[k] = h[k] * inv;
        float au = fmaxf(b.alpha_u, 1e-4f), av = fmaxf(b.alpha_v, 1e-4f);
        D6 ct2 = dsqr(h[2]), D;
        if (b.distr == 0) D = dexp(-(dsqr(h[0] / au) + dsqr(h[1] / av)) / ct2) / (kPi * au * av * dsqr(ct2));
        else D = 1.f / (kPi * au * av * dsqr(dsqr(h[0] / au) + dsqr(h[1] / av) + ct2));
        if (!(D.v * h[2].v > 1e-20f)) return false;
        D6 cih = i[0] * h[0] + i[1] * h[1] + i[2] * h[2], coh = o[0] * h[0] + o[1] * h[1] + o[2] * h[2];
        D6 res = D * rc_smith_g1(i, cih, au, av, b.distr) * rc_smith_g1(o, coh, au, av, b.distr) / (4.f * i[2]);
        if (!(res.v > 0.f)) return false;
        float3 eta = tex_eval(sc, b.eta, u, v), k = tex_eval(sc, b.k, u, v);
        float etas[3] = { eta.x, eta.y, eta.z }, ks[3] = { k.x, k.y, k.z }, ws[3] = { w.x, w.y, w.z };
        float acc[6]; float rinv = 1.f / res.v;
        for (int j = 0; j < 6; ++j) acc[j] = q * res.d[j] * rinv;
        for (int c = 0; c < 3; ++c) {
            if (ws[c] == 0.f) continue;
            D6 F = rc_fresnel_conductor(cih, etas[c], ks[c]);
            if (!(F.v > 0.f)) continue;
            float s = ws[c] / F.v;
            for (int j = 0; j < 6; ++j) acc[j] += s * F.d[j];
        }
        g_wi = mk3(acc[0], acc[1], acc[2]); g_wo = mk3(acc[3], acc[4], acc[5]);
    } else
        return false;
    if (back) { g_wi.z = -g_wi.z; g_wo.z = -g_wo.z; }
    return true;
}

// ---- geometry helpers -------------------------------------------------------------------------------------------------
struct VtxRec { uint32_t prim; float b1, b2; };           // a path vertex: triangle + barycentrics; prim ~0 = the camera
HD VtxRec vtx_rec(float4 r) { VtxRec v; v.prim = f2u(r.x) & kPrimMask; if (f2u(r.x) == 0xffffffffu) v.prim = 0xffffffffu; v.b1 = r.y; v.b2 = r.z; return v; }
struct TriGeo { uint32_t i0, i1, i2; int mesh; float3 p0, e1, e2, N; };
HD TriGeo tri_geo(const SceneDev &sc, uint32_t prim) {
    TriGeo t; uint4 f = sc.faces[prim];
    t.i0 = f.x; t.i1 = f.y; t.i2 = f.z; t.mesh = (int) f.w;
    t.p0 = ld3(sc.pos, f.x); t.e1 = ld3(sc.pos, f.y) - t.p0; t.e2 = ld3(sc.pos, f.z) - t.p0;
    t.N = cross(t.e1, t.e2);
    return t;
}
HD float3 tri_point(const TriGeo &t, const VtxRec &r) { return t.p0 + t.e1 * r.b1 + t.e2 * r.b2; }

template <typename Add> HD void add_vertex(const GradCtx &g, uint32_t vi, float3 v, Add add) {
    uint32_t k = vi - g.vert_begin;
    if (k < g.vert_count) { add(g.grad_pos + 3 * (size_t) k, v.x); add(g.grad_pos + 3 * (size_t) k + 1, v.y); add(g.grad_pos + 3 * (size_t) k + 2, v.z); }
}
// adjoint of a path-vertex position.  Attached vertex: x = sum b_i v_i.  Primary vertex: x = o + t d slides along the camera
// ray, dx = d (N . sum b_i dv_i) / (N . d).
template <typename Add>
HD void scatter_point(const GradCtx &g, const TriGeo &t, const VtxRec &r, bool primary, float3 x, float3 gx, Add add) {
    if (t.mesh != g.mesh_id) return;
    if (primary) {
        float3 d = x - g.cam_o; float dn = dot(d, t.N);
        if (!(fabsf(dn) > 1e-20f)) return;
        gx = t.N * (dot(d, gx) / dn);
    }
    add_vertex(g, t.i0, gx * (1.f - r.b1 - r.b2), add); add_vertex(g, t.i1, gx * r.b1, add); add_vertex(g, t.i2, gx * r.b2, add);
}
// adjoint of the unnormalised normal N = e1 x e2
template <typename Add> HD void scatter_normal(const GradCtx &g, const TriGeo &t, float3 gN, Add add) {
    if (t.mesh != g.mesh_id) return;
    float3 ge1 = cross(t.e2, gN), ge2 = cross(gN, t.e1);
    add_vertex(g, t.i1, ge1, add); add_vertex(g, t.i2, ge2, add); add_vertex(g, t.i0, -(ge1 + ge2), add);
}

// value and uv-gradient of a bitmap texture lookup (tex_eval_rgb in shade.h: repeat, bilinear); false for other textures
HD bool tex_grad_rgb(const SceneDev &sc, int id, float u, float v, float3 &a, float3 &da_du, float3 &da_dv) {
    const TexDev &t = sc.tex[id];
    if (t.type != 1) return false;
    float tu = t.to_uv[0] * u + t.to_uv[1] * v + t.to_uv[2], tv = t.to_uv[3] * u + t.to_uv[4] * v + t.to_uv[5];
    tu -= floorf(tu); tv -= floorf(tv);
    tu *= (float) (t.w - 1); tv *= (float) (t.h - 1);
    uint32_t px = min((uint32_t) tu, t.w - 2), py = min((uint32_t) tv, t.h - 2);
    float w1x = tu - (float) px, w1y = tv - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    uint32_t i = px + py * t.w;
    float3 v00, v10, v01, v11;
    if (t.ch == 1) { v00 = mk3(t.data[i]); v10 = mk3(t.data[i + 1]); v01 = mk3(t.data[i + t.w]); v11 = mk3(t.data[i + t.w + 1]); }
    else { v00 = ld3(t.data, i); v10 = ld3(t.data, i + 1); v01 = ld3(t.data, i + t.w); v11 = ld3(t.data, i + t.w + 1); }
    a = (v00 * w0x + v10 * w1x) * w0y + (v01 * w0x + v11 * w1x) * w1y;
    float3 dx = ((v10 - v00) * w0y + (v11 - v01) * w1y) * (float) (t.w - 1), dy = ((v01 - v00) * w0x + (v11 - v10) * w1x) * (float) (t.h - 1);
    da_du = dx * t.to_uv[0] + dy * t.to_uv[3]; da_dv = dx * t.to_uv[1] + dy * t.to_uv[4];
    return true;
}

// ---- one local factor S = f_s(A, B, C) * G~(B, C) --------------------------------------------------------------------
enum { kEndSurface = 0, kEndDirection = 1, kEndFixedArea = 2, kEndFixedPoint = 3 };
// A = previous vertex (position xA; recA.prim ~0: camera), B = the vertex whose BSDF is used (siB reconstructed by
// make_si), C = what the outgoing direction ends on: an attached surface point (xC, recC), a fixed direction (xC holds
// it; environment), a fixed point with normal nC on an emitter that is not differentiated, or a fixed point light.
// w = per-channel weight dL/dI * (radiance carried through this factor).
template <typename Add>
HD void vertex_triple(const SceneDev &sc, const GradCtx &g, float3 xA, VtxRec recA, bool A_primary, const SI &siB, VtxRec recB,
                      bool B_primary, int kindC, float3 xC, VtxRec recC, float3 nC, float3 w, Add add) {
    if (w.x == 0.f && w.y == 0.f && w.z == 0.f) return;
    const BsdfDev bsdf = sc.bsdfs[sc.meshes[siB.mesh].bsdf];
    float3 din = xA - siB.p; float rin = norm(din);
    if (!(rin > 0.f)) return;
    float3 win = din / rin, dout = mk3(0.f), wout = xC; float rout = 1.f;
    if (kindC != kEndDirection) { dout = xC - siB.p; rout = norm(dout); if (!(rout > 0.f)) return; wout = dout / rout; }
    float3 wi = siB.sh.to_local(win), wo = siB.sh.to_local(wout), g_wi, g_wo;
    if (!bsdf_log_grad(sc, bsdf, siB.u, siB.v, wi, wo, w, g_wi, g_wo)) return;
    float q = w.x + w.y + w.z;
    float3 n = siB.sh.n;
    float3 g_win = siB.sh.to_world(g_wi), g_wout = siB.sh.to_world(g_wo);
    // shading normal; the tangents follow it rigidly (exact for isotropic BSDFs): ds = -n (s . dn), dt = -n (t . dn)
    float3 g_n = win * g_wi.z + wout * g_wo.z - (siB.sh.s * g_wi.x + siB.sh.t * g_wi.y) * dot(n, win)
                                              - (siB.sh.s * g_wo.x + siB.sh.t * g_wo.y) * dot(n, wout);
    float3 g_din = (g_win - win * dot(win, g_win)) / rin, g_dout = mk3(0.f);
    if (kindC != kEndDirection) g_dout = (g_wout - wout * dot(wout, g_wout)) / rout;
    TriGeo tC; float3 g_NC = mk3(0.f);
    if (kindC == kEndSurface) {
        tC = tri_geo(sc, recC.prim);
        float nd = dot(tC.N, dout);
        if (fabsf(nd) > 1e-20f) { g_dout += (tC.N * (1.f / nd) - dout * (3.f / (rout * rout))) * q; g_NC = dout * (q / nd); }
    } else if (kindC == kEndFixedArea) {
        float nd = dot(nC, dout);
        if (fabsf(nd) > 1e-20f) g_dout += (nC * (1.f / nd) - dout * (3.f / (rout * rout))) * q;
    } else if (kindC == kEndFixedPoint)
        g_dout += dout * (-2.f * q / (rout * rout));
    // scatter
    if (recA.prim != 0xffffffffu) { TriGeo tA = tri_geo(sc, recA.prim); scatter_point(g, tA, recA, A_primary, xA, g_din, add); }
    TriGeo tB = tri_geo(sc, recB.prim);
    scatter_point(g, tB, recB, B_primary, siB.p, -(g_din + g_dout), add);
    float lN = norm(tB.N);
    if (g.grad_vnrm != nullptr && tB.mesh == g.mesh_id && sc.meshes[tB.mesh].has_normals) {
        // smooth shading: n_s = normalize(sum b_i n_i); hand the adjoint to the three vertex normals (their dependence on
        // the positions -- Mesh::recompute_vertex_normals, mesh.cpp:199-276 -- is back-propagated once, after the replay)
        float3 n0 = ld3(sc.nrm, tB.i0), n1 = ld3(sc.nrm, tB.i1), n2 = ld3(sc.nrm, tB.i2);
        float b0 = 1.f - recB.b1 - recB.b2;
        float L = norm(n0 * b0 + n1 * recB.b1 + n2 * recB.b2);
        if (L > 0.f) {
            float3 gm = (g_n - n * dot(n, g_n)) / L;
            GradCtx gv = g; gv.grad_pos = g.grad_vnrm;
            add_vertex(gv, tB.i0, gm * b0, add); add_vertex(gv, tB.i1, gm * recB.b1, add); add_vertex(gv, tB.i2, gm * recB.b2, add);
        }
    } else if (lN > 0.f) {
        // dn_g = P dN / |N|; the shading normal turns with the face: dn_s = (n_g x dn_g) x n_s (= dn_g for flat shading)
        float3 ng = tB.N / lN;
        float3 g_ng = g_n * dot(ng, n) - n * dot(g_n, ng);
        scatter_normal(g, tB, (g_ng - ng * dot(ng, g_ng)) / lN, add);
    }
    if (B_primary && tB.mesh == g.mesh_id && lN > 0.f && (sc.meshes[tB.mesh].has_normals || sc.meshes[tB.mesh].has_uv)) {
        // x1 slides over the triangle when the vertices move under a fixed camera ray: r = dt d - u (in-plane),
        // db1 = r . (e2 x N)/|N|^2, db2 = r . (N x e1)/|N|^2.  Through the barycentrics this changes the INTERPOLATED normal
        // and the texture coordinates (a bitmap reflectance slides under the pixel).
        float b0 = 1.f - recB.b1 - recB.b2, gb1 = 0.f, gb2 = 0.f;
        if (sc.meshes[tB.mesh].has_normals) {
            float3 n0 = ld3(sc.nrm, tB.i0), n1 = ld3(sc.nrm, tB.i1), n2 = ld3(sc.nrm, tB.i2);
            float L = norm(n0 * b0 + n1 * recB.b1 + n2 * recB.b2);
            if (L > 0.f) { float3 gm = (g_n - n * dot(n, g_n)) / L; gb1 += dot(gm, n1 - n0); gb2 += dot(gm, n2 - n0); }
        }
        int tid = bsdf.type == 0 ? bsdf.refl : (bsdf.type == 1 ? bsdf.spec_r : -1);
        if (sc.meshes[tB.mesh].has_uv && tid >= 0) {
            float3 a, da_du, da_dv;
            if (tex_grad_rgb(sc, tid, siB.u, siB.v, a, da_du, da_dv)) {      // f_c is proportional to the texel mix a_c: dlog f_c = da_c / a_c
                float3 ga = mk3(a.x > 1e-8f ? w.x / a.x : 0.f, a.y > 1e-8f ? w.y / a.y : 0.f, a.z > 1e-8f ? w.z / a.z : 0.f);
                float gu_ = dot(ga, da_du), gv_ = dot(ga, da_dv);
                float u0 = sc.uv[2 * tB.i0], v0 = sc.uv[2 * tB.i0 + 1];
                gb1 += gu_ * (sc.uv[2 * tB.i1] - u0) + gv_ * (sc.uv[2 * tB.i1 + 1] - v0);
                gb2 += gu_ * (sc.uv[2 * tB.i2] - u0) + gv_ * (sc.uv[2 * tB.i2 + 1] - v0);
            }
        }
        float3 d = siB.p - g.cam_o; float nd = dot(tB.N, d);
        if ((gb1 != 0.f || gb2 != 0.f) && fabsf(nd) > 1e-20f) {
            float3 gr = (cross(tB.e2, tB.N) * gb1 + cross(tB.N, tB.e1) * gb2) / (lN * lN);
            float3 gu = tB.N * (dot(d, gr) / nd) - gr;
            add_vertex(g, tB.i0, gu * b0, add); add_vertex(g, tB.i1, gu * recB.b1, add); add_vertex(g, tB.i2, gu * recB.b2, add);
        }
    }
    if (kindC == kEndSurface) { scatter_point(g, tC, recC, false, xC, g_dout, add); scatter_normal(g, tC, g_NC, add); }
}

// ---- vertex normals: n_v = normalize(a_v), a_v = sum over the faces around v of (corner angle) x (unit face normal) ----------
// (Mesh::recompute_vertex_normals, src/librender/mesh.cpp:199-276, run by parameters_changed after a position update).
// vnormal_accum_body rebuilds a_v; vnormal_adjoint_body takes the adjoint of the vertex normals gathered during the replay
// back to the positions: through the normalisation, the unit face normal and the corner angles.
HD float corner_angle(float3 d0, float3 d1) {          // enoki unit_angle
    float dp = dot(d0, d1);
    float temp = 2.f * asinf(fminf(1.f, 0.5f * norm(d1 - (dp < 0.f ? -d0 : d0))));
    return dp >= 0.f ? temp : kPi - temp;
}
template <typename Add> HD void vnormal_accum_body(const SceneDev &sc, const GradCtx &g, uint32_t prim, float *acc, Add add) {
    TriGeo t = tri_geo(sc, prim);
    float l = norm(t.N);
    if (!(l > 0.f)) return;
    float3 nf = t.N / l, v[3] = { t.p0, t.p0 + t.e1, t.p0 + t.e2 };
    uint32_t idx[3] = { t.i0, t.i1, t.i2 };
    GradCtx ga = g; ga.grad_pos = acc;
    for (int k = 0; k < 3; ++k)
        add_vertex(ga, idx[k], nf * corner_angle(normalize(v[(k + 1) % 3] - v[k]), normalize(v[(k + 2) % 3] - v[k])), add);
}
template <typename Add> HD void vnormal_adjoint_body(const SceneDev &sc, const GradCtx &g, const float *acc, uint32_t prim, Add add) {
    TriGeo t = tri_geo(sc, prim);
    float l = norm(t.N);
    if (!(l > 0.f)) return;
    float3 nf = t.N / l, v[3] = { t.p0, t.p0 + t.e1, t.p0 + t.e2 }, g_nf = mk3(0.f);
    uint32_t idx[3] = { t.i0, t.i1, t.i2 };
    for (int k = 0; k < 3; ++k) {
        uint32_t vi = idx[k] - g.vert_begin;
        float3 a = ld3(acc, vi), gn = ld3(g.grad_vnrm, vi);
        float la = norm(a);
        if (!(la > 0.f) || (gn.x == 0.f && gn.y == 0.f && gn.z == 0.f)) continue;
        float3 nv = a / la, ga = (gn - nv * dot(nv, gn)) / la;          // through n_v = a / |a|
        uint32_t k1 = (k + 1) % 3, k2 = (k + 2) % 3;
        float3 e0 = v[k1] - v[k], e1 = v[k2] - v[k];
        float l0 = norm(e0), l1 = norm(e1);
        float3 d0 = e0 / l0, d1 = e1 / l1;
        float theta = corner_angle(d0, d1), c = dot(d0, d1);
        g_nf += ga * theta;                                               // a += theta * n_f
        float s2 = 1.f - c * c;
        if (s2 > 1e-12f) {
            float gc = -dot(ga, nf) / sqrtf(s2);                          // theta = acos(c)
            float3 ge0 = (d1 - d0 * c) * (gc / l0), ge1 = (d0 - d1 * c) * (gc / l1);
            add_vertex(g, idx[k1], ge0, add); add_vertex(g, idx[k2], ge1, add); add_vertex(g, idx[k], -(ge0 + ge1), add);
        }
    }
    scatter_normal(g, t, (g_nf - nf * dot(nf, g_nf)) / l, add);           // n_f = N / |N|
}

// ---- replay hooks (plugged into shade_body, path.h) ----------------------------------------------------------------
// Add = functor (float *dst, float v) accumulating into a gradient buffer (atomicAdd on the device).
template <typename Add, bool Vtx = true> struct ReplayGrad {
    static constexpr bool kVertex = Vtx;
    GradCtx g; Add add;
    HD float3 weight(const RenderBatch &b, uint32_t path_id, float3 radiance) const {
        return image_adjoint(g, b, path_id) * radiance;
    }
    // factor S_{k-1} = f_s(x_{k-2}, x_{k-1}, x_k) G~(x_{k-1}, x_k), complete now that x_k (or a miss) is known
    HD void on_vertex(const SceneDev &sc, const PathQueue &q, const RenderBatch &b, uint32_t i, uint32_t path_id, uint32_t depth,
                      float3 o, float3 d, const SI &si) const {
        if (g.grad_pos == nullptr || depth < 2) return;
        float4 fin = g.final_radiance[path_id], run = b.result[path_id];
        float3 w = weight(b, path_id, mk3(fin.x - run.x, fin.y - run.y, fin.z - run.z));
        VtxRec recB = vtx_rec(q.prev1[i]), recA = vtx_rec(q.prev2[i]), recC; recC.prim = si.prim; recC.b1 = si.u; recC.b2 = si.v;
        float3 xA = g.cam_o;
        if (recA.prim != 0xffffffffu) xA = tri_point(tri_geo(sc, recA.prim), recA);
        SI siB = make_si(sc, xA, mk3(0.f, 0.f, 1.f), make_float4(1.f, recB.b1, recB.b2, u2f(recB.prim)));
        if (si.valid) {
            float4 h = q.hit[i]; recC.b1 = h.y; recC.b2 = h.z;
            vertex_triple(sc, g, xA, recA, depth == 3, siB, recB, depth == 2, kEndSurface, si.p, recC, mk3(0.f), w, add);
        } else
            vertex_triple(sc, g, xA, recA, depth == 3, siB, recB, depth == 2, kEndDirection, d, recC, mk3(0.f), w, add);
        (void) o;
    }
    HD void on_bsdf(const SceneDev &sc, const RenderBatch &b, uint32_t path_id, const SI &si, const BsdfDev &bsdf) const {
        if (g.grad_tex != nullptr) texture_adjoint(sc, b, g, path_id, si, bsdf, add);
    }
};

// emitter sample at x_k, differentiated once its shadow ray is known to be free (k_nee_grad after k_shadow): factor
// The emitter sample of (sx, sy) as a material point: which area light, and when that light's mesh is `mesh_id` (a triangle mesh)
// the triangle and the barycentrics of Mesh::sample_position (emit.h area_sample_direction, same arithmetic).  False otherwise.
HD bool nee_sample_on_mesh(const SceneDev &sc, int mesh_id, float sx, float sy, VtxRec &rec) {
    if (sc.n_emitters == 0) return false;
    uint32_t index = 0;
    if (sc.n_emitters > 1) {
        float ne = (float) sc.n_emitters;
        index = min((uint32_t) (sx * ne), (uint32_t) sc.n_emitters - 1);
        sx = (sx - (float) index * (1.f / ne)) * ne;
    }
    const EmitterDev &e = sc.emitters[index];
    if (e.type != 0 || e.mesh != mesh_id || e.cdf == nullptr) return false;
    const MeshDev m = sc.meshes[e.mesh];
    float value = sy * e.cdf_sum;
    const float *cdf = e.cdf;
    uint32_t idx = binary_search(e.valid_lo, e.valid_hi, [&](uint32_t i) { return cdf[i] < value; });
    uint4 f = sc.faces[m.face_begin + idx];
    float3 p0 = ld3(sc.pos, f.x), e0 = ld3(sc.pos, f.y) - p0, e1 = ld3(sc.pos, f.z) - p0;
    float pmf = 0.5f * norm(cross(e0, e1)) * e.normalization;
    float cdf_prev = idx > 0 ? cdf[idx - 1] * e.normalization : 0.f;
    sy = (sy - cdf_prev) / pmf;
    float t = safe_sqrt(1.f - sx);
    rec.prim = m.face_begin + idx; rec.b1 = 1.f - t; rec.b2 = t * sy;
    return true;
}

// f_s(x_{k-1}, x_k, y) G~(x_k, y), weighted by the sample's own contribution.  `rec` = ShadowQueue::nee_rec (queue index of
// the vertex, the two sampling numbers, depth; depth 0 = path already passed a delta vertex), `contrib` = the shadow
// record's contribution and path id.  The queue entry of the vertex is still intact at that point of the batch.
template <typename Add>
HD void nee_adjoint(const SceneDev &sc, const GradCtx &g, const PathQueue &q, const RenderBatch &b, float4 rec, float4 contrib, Add add) {
    uint32_t i = f2u(rec.x), depth = f2u(rec.w), path_id = f2u(contrib.w);
    if (depth == 0u || g.grad_pos == nullptr) return;
    float4 o4 = q.o_mint[i], d4 = q.d_maxt[i], hc = q.hit[i];
    SI si = make_si(sc, xyz(o4), xyz(d4), hc);
    DirSample ds = sample_emitter_direction(sc, si.p, rec.y, rec.z);
    VtxRec recA; recA.prim = 0xffffffffu; recA.b1 = recA.b2 = 0.f;
    float3 xA = g.cam_o;
    if (depth > 1) { recA = vtx_rec(q.prev1[i]); xA = tri_point(tri_geo(sc, recA.prim), recA); }
    VtxRec recB; recB.prim = si.prim; recB.b1 = hc.y; recB.b2 = hc.z;
    // area light: fixed point with a normal; envmap / constant background: fixed direction; point light: fixed point.
    // An area light on the differentiated mesh: the sample is a material point of that mesh (triangle and barycentrics are what
    // the sampler fixed; the triangle's probability and the uniform barycentric density are detached like every sampling
    // decision), so the factor is the attached-surface one: |N . d| / r^3 with the UNNORMALISED normal carries the area Jacobian.
    int kind = ds.etype == 0 ? kEndFixedArea : (ds.etype == 3 ? kEndFixedPoint : kEndDirection);
    VtxRec recC = recB;
    if (ds.etype == 0 && nee_sample_on_mesh(sc, g.mesh_id, rec.y, rec.z, recC)) kind = kEndSurface;
    float3 xC = kind == kEndDirection ? ds.d : ds.p;
    float3 w = image_adjoint(g, b, path_id) * xyz(contrib);
    vertex_triple(sc, g, xA, recA, depth == 2, si, recB, depth == 1, kind, xC, recC, ds.n, w, add);
}

// ---- boundary term: edge sampling on the film plane ---------------------------------------------------------------
struct EdgeCtx {
    const uint4 *edges; uint32_t n_edges;   // unique edges of the differentiated mesh: scene-global vertex indices a, b and
                                            // the primitive ids of the two adjacent faces (second = ~0 on a border)
    const float *cdf;                       // inclusive prefix sum of the projected edge lengths (pixels)
    float total_len;
    float w2s[16];                          // world -> sample space (x, y in [0,1]^2 after the divide by w), row-major
    uint32_t width, height;
    uint64_t seed; uint64_t first;          // global index of this call's first boundary sample
    float weight;                           // 1 / (boundary samples of the whole job)
    float eps;                              // half distance (pixels) between the two radiance samples
    int rim_override;                       // 1: silhouette pairs take the mesh side as black where the dark rim reaches the edge (default)
};
HD bool edge_project(const EdgeCtx &e, float3 v, float &x, float &y, float &hw) {
    const float *m = e.w2s;
    float hx = m[0] * v.x + m[1] * v.y + m[2] * v.z + m[3], hy = m[4] * v.x + m[5] * v.y + m[6] * v.z + m[7];
    hw = m[12] * v.x + m[13] * v.y + m[14] * v.z + m[15];
    if (!(hw > 1e-6f)) return false;
    x = hx / hw * (float) e.width; y = hy / hw * (float) e.height;
    return true;
}
// ---- shading-normal terminator of a smooth-shaded face --------------------------------------------------------------------
// A one-sided BSDF is black where the INTERPOLATED normal looks away from the camera (Frame::cos_theta(si.wi) <= 0 in the
// shading frame, e.g. diffuse.cpp:95-103), which happens on front-facing triangles next to every silhouette and on the flanks
// of bumps.  The curve  g(b) = n_s(b) . (p(b) - cam) = 0  inside such a triangle is a jump of the image (lit | exactly black)
// that moves with the mesh: a film-space boundary like a silhouette edge, but not a mesh edge.  Entry k of the edge list with
// edges[k].y == kTermTag stands for the terminator of face edges[k].x.  Its end points are the roots of g (a quadratic) on
// two of the triangle's edges -- the same function the silhouette pairs use to decide whether the rim reaches a mesh edge, so
// the two kinds of samples partition the jumps exactly -- the chord between them is sampled like an edge, the jump is the
// radiance of the lit side evaluated AT the point (path.h: the face is shaded as if lit; no pair of offset rays, so the
// chord's distance from the true curve does not matter), and the curve moves with the material (barycentric mix of the
// three vertex displacements; its slip over the material is along the line of sight and does not show on the film).
constexpr uint32_t kTermTag = 0xfffffffeu;
HD bool term_bsdf(const SceneDev &sc, const MeshDev &m) {      // does this mesh have a terminator at all?
    if (!m.has_normals || m.emitter >= 0) return false;
    const BsdfDev b = sc.bsdfs[m.bsdf];
    return !b.two_sided && b.type != 2 && b.type != 5 && b.opacity < 0;
}
struct TermSeg { float3 qa, qb, lit; float wa[3], wb[3]; uint4 f; bool valid; };      // lit = a vertex on the lit side
HD TermSeg term_segment(const SceneDev &sc, float3 cam_o, uint32_t prim) {
    TermSeg t; t.valid = false; t.qa = t.qb = t.lit = mk3(0.f);
    for (int i = 0; i < 3; ++i) t.wa[i] = t.wb[i] = 0.f;
    uint4 f = sc.faces[prim]; t.f = f;
    const MeshDev m = sc.meshes[f.w];
    if (!term_bsdf(sc, m)) return t;
    float3 v[3] = { ld3(sc.pos, f.x), ld3(sc.pos, f.y), ld3(sc.pos, f.z) };
    float3 n[3] = { ld3(sc.nrm, f.x), ld3(sc.nrm, f.y), ld3(sc.nrm, f.z) };
    if (!(dot(cross(v[1] - v[0], v[2] - v[0]), v[0] - cam_o) < 0.f)) return t;      // back-facing triangles are not seen
    float c[3] = { dot(n[0], v[0] - cam_o), dot(n[1], v[1] - cam_o), dot(n[2], v[2] - cam_o) };
    int found = 0;
    for (int i = 0; i < 3 && found < 2; ++i) {
        int j = (i + 1) % 3;
        if ((c[i] < 0.f) == (c[j] < 0.f)) continue;
        // g(s) = (n_i + s dn) . (P_i + s dv) on the edge i -> j: root near the linear estimate, two Newton steps
        float3 dn = n[j] - n[i], dv = v[j] - v[i], Pi = v[i] - cam_o;
        float A = dot(dn, dv), B = dot(n[i], dv) + dot(dn, Pi), C = c[i];
        float sroot = C / (C - c[j]);
        for (int it = 0; it < 2; ++it) { float gs = (A * sroot + B) * sroot + C, dg = 2.f * A * sroot + B; if (dg != 0.f) sroot -= gs / dg; }
        sroot = fminf(fmaxf(sroot, 0.f), 1.f);
        float *w = found == 0 ? t.wa : t.wb;
        w[i] = 1.f - sroot; w[j] = sroot;
        if (found == 0) t.qa = v[i] + dv * sroot; else t.qb = v[i] + dv * sroot;
        found++;
    }
    t.lit = c[0] < 0.f ? v[0] : (c[1] < 0.f ? v[1] : v[2]);
    t.valid = found == 2;
    return t;
}
// does the dark rim reach the point a + t (b - a) of a mesh edge?  (same g as above, on that edge)
HD bool edge_point_dark(const SceneDev &sc, float3 cam_o, uint4 ed, float t) {
    float3 a = ld3(sc.pos, ed.x), b = ld3(sc.pos, ed.y), na = ld3(sc.nrm, ed.x), nb = ld3(sc.nrm, ed.y);
    return dot(na + (nb - na) * t, a + (b - a) * t - cam_o) >= 0.f;
}

// sampling weight of an edge = projected length.  A smooth-shaded mesh (vertex normals) is discontinuous on the film only
// along its silhouette -- edges whose two faces look opposite ways from the camera, and border edges -- so the others get
// weight 0; a flat-shaded mesh also jumps across every edge that has a front-facing side.
HD float edge_len_body(const SceneDev &sc, const EdgeCtx &e, float3 cam_o, uint32_t k) {
    uint4 ed = e.edges[k]; float ax, ay, aw, bx, by, bw;
    if (ed.y == kTermTag) {
        TermSeg ts = term_segment(sc, cam_o, ed.x);
        if (!ts.valid || !edge_project(e, ts.qa, ax, ay, aw) || !edge_project(e, ts.qb, bx, by, bw)) return 0.f;
        float lt = sqrtf(sqr(ax - bx) + sqr(ay - by));
        return lt < 1e30f ? lt : 0.f;
    }
    float3 va = ld3(sc.pos, ed.x);
    if (!edge_project(e, va, ax, ay, aw) || !edge_project(e, ld3(sc.pos, ed.y), bx, by, bw)) return 0.f;
    if (ed.w != 0xffffffffu) {
        float s0 = dot(tri_geo(sc, ed.z).N, va - cam_o), s1 = dot(tri_geo(sc, ed.w).N, va - cam_o);
        bool smooth = sc.meshes[sc.faces[ed.z].w].has_normals != 0;
        if (smooth ? (s0 < 0.f) == (s1 < 0.f) : (s0 >= 0.f && s1 >= 0.f)) return 0.f;
    }
    float l = sqrtf(sqr(ax - bx) + sqr(ay - by));
    return l < 1e30f ? l : 0.f;
}
struct EdgeSample { uint32_t edge; float t, px, py, nx, ny, len; float3 va, vb; bool valid; Pcg32 rng;
                    bool term, lit_plus; uint32_t vid[3]; float vw[3];         // terminator sample: lit side at +n?, the face's vertices, weights
                    int rim_side; uint32_t front; };      // silhouette pair whose edge point lies in the dark rim: +1 / -1 = mesh side at +n / -n, front face
HD EdgeSample edge_sample(const SceneDev &sc, const EdgeCtx &e, uint32_t j, float3 cam_o) {
    EdgeSample s; s.valid = false; s.edge = 0; s.t = 0.f; s.px = s.py = s.nx = s.ny = s.len = 0.f; s.va = s.vb = mk3(0.f);
    s.term = s.lit_plus = false; for (int i = 0; i < 3; ++i) { s.vid[i] = 0; s.vw[i] = 0.f; }
    s.rim_side = 0; s.front = 0;
    float3 lit = mk3(0.f);
    s.rng = sampler_seed(e.seed ^ 0x9e3779b97f4a7c15ull, (uint32_t) (e.first + j));
    float u1 = s.rng.next_f32(), u2 = s.rng.next_f32(); (void) s.rng.next_f32();
    if (!(e.total_len > 0.f) || e.n_edges == 0) return s;
    float value = u1 * e.total_len; const float *cdf = e.cdf;
    s.edge = min(binary_search(0u, e.n_edges, [&](uint32_t i) { return cdf[i] <= value; }), e.n_edges - 1);
    uint4 ed = e.edges[s.edge];
    if (ed.y == kTermTag) {
        TermSeg ts = term_segment(sc, cam_o, ed.x);
        if (!ts.valid) return s;
        s.term = true; s.va = ts.qa; s.vb = ts.qb; lit = ts.lit;
        s.vid[0] = ts.f.x; s.vid[1] = ts.f.y; s.vid[2] = ts.f.z;
        for (int i = 0; i < 3; ++i) s.vw[i] = (1.f - u2) * ts.wa[i] + u2 * ts.wb[i];
    } else { s.va = ld3(sc.pos, ed.x); s.vb = ld3(sc.pos, ed.y); }
    float ax, ay, aw, bx, by, bw;
    if (!edge_project(e, s.va, ax, ay, aw) || !edge_project(e, s.vb, bx, by, bw)) return s;
    s.len = sqrtf(sqr(ax - bx) + sqr(ay - by));
    if (!(s.len > 0.f)) return s;
    s.t = u2; s.px = ax + u2 * (bx - ax); s.py = ay + u2 * (by - ay);
    s.nx = (ay - by) / s.len; s.ny = (bx - ax) / s.len;            // grad_xy(alpha) / |grad_xy(alpha)|
    if (s.term) { float lx, ly, lw; if (!edge_project(e, lit, lx, ly, lw)) return s; s.lit_plus = (lx - ax) * s.nx + (ly - ay) * s.ny > 0.f; }
    else if (e.rim_override && term_bsdf(sc, sc.meshes[sc.faces[ed.z].w]) && edge_point_dark(sc, cam_o, ed, u2)) {
        // Smooth-shaded, one-sided: the dark rim reaches this point of the edge, so the mesh side of the jump AT the edge is
        // exactly 0 -- whatever the offset ray 0.004 px inside sees of the front face (it may already be beyond a rim thinner
        // than that, and that jump belongs to the terminator samples).  edge_raygen_body tags the ray on the mesh side as
        // black (path.h) -- and both rays when the edge point is hidden from the camera, so that hidden edges still cancel.
        // Mesh side = the side of the front face's third vertex.
        uint32_t front = ed.z;
        if (ed.w != 0xffffffffu && !(dot(tri_geo(sc, ed.z).N, s.va - cam_o) < 0.f)) front = ed.w;
        uint4 ff = sc.faces[front];
        uint32_t third = (ff.x != ed.x && ff.x != ed.y) ? ff.x : ((ff.y != ed.x && ff.y != ed.y) ? ff.y : ff.z);
        float tx, ty, tw;
        if (edge_project(e, ld3(sc.pos, third), tx, ty, tw)) { s.rim_side = (tx - ax) * s.nx + (ty - ay) * s.ny > 0.f ? 1 : -1; s.front = front; }
    }
    s.valid = s.px >= 0.f && s.py >= 0.f && s.px < (float) e.width && s.py < (float) e.height;
    return s;
}
// two primary rays per boundary sample: 2j on the alpha > 0 side, 2j+1 on the other, same random-number stream
HD void edge_raygen_body(const SceneDev &sc, const SensorDev &sn, const EdgeCtx &e, const RenderBatch &b, const PathQueue &q, uint32_t j) {
    EdgeSample s = edge_sample(sc, e, j, mk3(sn.to_world[3], sn.to_world[7], sn.to_world[11]));
    if (!s.valid) { s.px = 0.5f * (float) e.width; s.py = 0.5f * (float) e.height; s.nx = 1.f; s.ny = 0.f; s.term = false; }
    bool black[2] = { false, false };
    if (!s.term && s.valid && s.rim_side != 0) {
        float3 cam = mk3(sn.to_world[3], sn.to_world[7], sn.to_world[11]);
        float3 dir = s.va + (s.vb - s.va) * s.t - cam; float dist = norm(dir); dir = dir / dist;
        uint2 fast[8]; TravStack<8, 56> st; st.fast = fast; st.stride = 1; Hit h;
        bool hidden = bvh_trace<true>(sc.bvh, cam, dir, 1e-4f * dist, dist * (1.f - 1e-4f), st, h);
        black[0] = hidden || s.rim_side > 0; black[1] = hidden || s.rim_side < 0;
    }
    for (uint32_t side = 0; side < 2; ++side) {
        uint32_t i = 2 * j + side; float sg = s.term ? 0.f : (side == 0 ? e.eps : -e.eps);
        float4 o, d;
        sensor_sample_ray(sn, (s.px + sg * s.nx) / (float) sn.width, (s.py + sg * s.ny) / (float) sn.height, o, d);
        // terminator sample: ONE path through the point itself, tagged with the face it must see (path.h shades it as lit);
        // the second slot of the pair is not used (a ray of zero length)
        uint32_t tag = s.term && side == 0 ? e.edges[s.edge].x + 1u : 0u;
        if (black[side]) tag = 0xffffffffu;
        if (s.term && side == 1) d.w = 0.f;
        q.o_mint[i] = o; q.d_maxt[i] = d; q.thr_eta[i] = make_float4(1.f, 1.f, 1.f, 1.f);
        q.misc[i] = make_uint4(i, 1u, f2u(0.f), tag);
        rng_store(&q.rng[i], s.rng);
        b.result[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
// dPhi = (1/N) sum  dL/dI(pixel) . (L+ - L-) * grad_v(alpha) * total_len / len,  alpha = edge function through a, b
template <typename Add>
HD void edge_scatter_body(const SceneDev &sc, const EdgeCtx &e, const GradCtx &g, const float4 *result, uint32_t j, Add add) {
    EdgeSample s = edge_sample(sc, e, j, g.cam_o);
    if (!s.valid) return;
    float4 lp = result[2 * j], lm = result[2 * j + 1];
    uint4 ed = e.edges[s.edge];
    const float *m = e.w2s;
    if (s.term) {       // jump = radiance of the lit side (slot 0) against exactly 0
        if (s.lit_plus) lm = make_float4(0.f, 0.f, 0.f, 0.f); else { lm = lp; lp = make_float4(0.f, 0.f, 0.f, 0.f); }
    }
    float3 dI;
    if (g.inv_w != nullptr) dI = film_gather(g.filt, g.dL_dimage, g.inv_w, s.px, s.py);       // wide reconstruction filter
    else dI = ld3(g.dL_dimage, (uint32_t) s.py * e.width + (uint32_t) s.px);
    float sc_ = (dI.x * (lp.x - lm.x) + dI.y * (lp.y - lm.y) + dI.z * (lp.z - lm.z)) * e.weight * e.total_len;
    if (sc_ == 0.f || !(fabsf(sc_) < 1e30f)) return;
    if (s.term) {
        // the sample point p = sum_i w_i v_i moves with the material: displacement along the chord's film normal =
        // n . J(p) . sum_i w_i dv_i; sc_ carries total_len = (chord length) / (pick probability per unit t)
        float3 p = s.va + (s.vb - s.va) * s.t; float x, y, hw;
        if (!edge_project(e, p, x, y, hw)) return;
        float sx = x / (float) e.width, sy = y / (float) e.height, W = (float) e.width / hw, H = (float) e.height / hw;
        float gx = -sc_ * s.nx, gy = -sc_ * s.ny;
        float3 gp = mk3(gx * W * (m[0] - sx * m[12]) + gy * H * (m[4] - sy * m[12]),
                        gx * W * (m[1] - sx * m[13]) + gy * H * (m[5] - sy * m[13]),
                        gx * W * (m[2] - sx * m[14]) + gy * H * (m[6] - sy * m[14]));
        for (int i = 0; i < 3; ++i) add_vertex(g, s.vid[i], gp * s.vw[i], add);
        return;
    }
    // grad_a(alpha) = -(1 - t) len n, grad_b(alpha) = -t len n; the 1/len of the edge pick cancels len
    for (int k = 0; k < 2; ++k) {
        float3 v = k == 0 ? s.va : s.vb; float wgt = -(k == 0 ? 1.f - s.t : s.t) * sc_;
        float gx = wgt * s.nx, gy = wgt * s.ny, x, y, hw;
        edge_project(e, v, x, y, hw);
        float sx = x / (float) e.width, sy = y / (float) e.height, W = (float) e.width / hw, H = (float) e.height / hw;
        float3 gv = mk3(gx * W * (m[0] - sx * m[12]) + gy * H * (m[4] - sy * m[12]),
                        gx * W * (m[1] - sx * m[13]) + gy * H * (m[5] - sy * m[13]),
                        gx * W * (m[2] - sx * m[14]) + gy * H * (m[6] - sy * m[14]));
        add_vertex(g, k == 0 ? ed.x : ed.y, gv, add);
    }
}

// ---- boundary term seen from the first path vertex: edge sampling in the direction domain of x1 ---------------------------
// The interior term integrates x2.. over triangle areas with fixed barycentrics, so what is left at a visibility boundary is
// the far side only:  at a silhouette edge (as seen from x1) of the differentiated mesh the region of directions that see
// the surface BEHIND the edge shrinks or grows, with the edge's angular velocity measured relative to the material point
// x2 seen behind it (the occluder's own side moves with its material -- no term):
//     dPhi += - dL/dI . f_s(x1, wi, w) cos . L_far(w) * n_out . [ (dp - dx1)/r - (dx2 - dx1)/r2 ] * |dw/dt| / P(edge, t)
// w = direction to the edge point p = a + t (b - a), r = |p - x1|, n_out = unit normal of the plane (x1, a, b) pointing away
// from the occluding face, x2 = first hit of the ray just outside the edge (r2 = its distance; infinite for a miss),
// L_far = full incident radiance along that ray (path-traced by the normal wavefront from depth 2, emitter hits unweighted).
// Edges are drawn by resampling: `candidates` edges proportional to their 3D length, one of those that are silhouettes from
// x1 kept with probability proportional to its angular length / r, P = (len / total) * candidates * w_chosen / sum(w).  Covers moving shadow boundaries and occlusion
// of indirect light as seen directly by the camera.
// Deeper vertices (Edge2Ctx::depth = k > 1): the same term with x_k in the place of x1.  The prefix x1 .. x_k is an ordinary
// sampled path (the primal shade kernel runs for k - 1 bounces), its throughput rides in the queue and multiplies L_far; x_k
// is a material point (fixed barycentrics, as the interior term parametrises it) instead of a point pinned to a pixel.
struct Edge2Ctx {
    const uint4 *edges; uint32_t n_edges; const float *cdf; float total_len;   // cdf over 3D edge lengths
    int candidates; float eps;                                                 // eps: angular offset of the far ray
    int moving_mesh;                                                           // the differentiated mesh (see edge2_raygen_body)
    float self_rmin;                                                           // x1 on that mesh: its own edges nearer than this many edge lengths are skipped
    float4 *aux;                     // 5 x float4 per PATH: (edge, t, |dw/dt| / P, valid) | far hit (prim, b1, b2, r2) | d | hit of x_k | o
    uint32_t mov_begin, mov_end;     // the differentiated mesh's edges in `edges` (candidates: half by length over all edges, half over these)
    int depth;                       // k: the path vertex the edges are seen from (1 = the camera's hit; deeper: mtsb_grad_config boundary 3)
};
HD float edge_len3_body(const SceneDev &sc, const Edge2Ctx &e, uint32_t k) {
    uint4 ed = e.edges[k];
    float l = norm(ld3(sc.pos, ed.y) - ld3(sc.pos, ed.x));
    return l < 1e30f ? l : 0.f;
}
struct Edge2Geo { float3 a, b, p, w, n_out; float r, dwdt; bool silhouette; };
// geometry of edge `ed` at parameter t seen from x; silhouette = its faces look opposite ways from x (or it is a border)
HD Edge2Geo edge2_geo(const SceneDev &sc, uint4 ed, float t, float3 x) {
    Edge2Geo g; g.a = ld3(sc.pos, ed.x); g.b = ld3(sc.pos, ed.y);
    TriGeo f0 = tri_geo(sc, ed.z);
    float s0 = dot(f0.N, g.a - x);
    g.silhouette = true;
    uint32_t front = ed.z;                  // the face that looks at x occludes; the far region is on the other side of the edge
    if (ed.w != 0xffffffffu) {
        float s1 = dot(tri_geo(sc, ed.w).N, g.a - x);
        g.silhouette = (s0 < 0.f) != (s1 < 0.f);
        if (!(s0 < 0.f)) front = ed.w;
    }
    g.p = g.a + (g.b - g.a) * t;
    float3 dp = g.p - x; g.r = norm(dp); g.w = dp / g.r;
    float3 ne = cross(g.a - x, g.b - x); float ln = norm(ne);
    if (!(ln > 0.f) || !(g.r > 0.f)) { g.silhouette = false; g.n_out = mk3(0.f); g.dwdt = 0.f; return g; }
    ne = ne / ln;
    uint4 ff = sc.faces[front];             // third vertex of the occluding face tells its side
    uint32_t third = (ff.x != ed.x && ff.x != ed.y) ? ff.x : ((ff.y != ed.x && ff.y != ed.y) ? ff.y : ff.z);
    g.n_out = dot(ne, ld3(sc.pos, third) - g.a) > 0.f ? -ne : ne;
    float3 e = g.b - g.a; float3 pe = e - g.w * dot(g.w, e);
    g.dwdt = norm(pe) / g.r;
    return g;
}
// stage B: queue entry j holds the ray into x_k, its hit and the sampler state; writes the far ray as a depth-(k+1) path
HD void edge2_raygen_body(const SceneDev &sc, const Edge2Ctx &e, const RenderBatch &b, const PathQueue &qin, const PathQueue &qout, uint32_t j) {
    float4 o4 = qin.o_mint[j], d4 = qin.d_maxt[j], h4 = qin.hit[j];
    const uint32_t pid = qin.misc[j].x;
    Pcg32 rng = rng_load(qin.rng[j]);
    float4 rec0 = make_float4(0.f, 0.f, 0.f, 0.f), rec1 = make_float4(u2f(0xffffffffu), 0.f, 0.f, INFINITY);
    float4 ro = make_float4(0.f, 0.f, 0.f, INFINITY), rd = make_float4(0.f, 0.f, 1.f, 0.f);
    SI si = make_si(sc, xyz(o4), xyz(d4), h4);
    bool ok = si.valid && e.n_edges > 0 && e.total_len > 0.f;
    BsdfDev bsdf;
    if (ok) { bsdf = sc.bsdfs[sc.meshes[si.mesh].bsdf]; ok = bsdf.type != 2 && (bsdf.two_sided ? si.wi.z != 0.f : si.wi.z > 0.f); }
    // Resampled importance sampling over the candidates: a silhouette candidate is kept with probability proportional to
    // w = (angular length of the edge seen from x1) / r -- the factor the integrand grows with for near edges -- and the
    // estimate is weighted by sum(w) / (candidates * w_chosen).
    // Edges of the mesh x1 itself lies on are skipped when that mesh is the differentiated one: the self-occlusion of a
    // bumpy surface seen from a point ON it has edges at every distance down to 0, the integrand grows like len / r^2 and a
    // length-proportional edge pick has infinite variance there (measured at C4: the standard error of the sphere's own
    // pixels went from 0.03 to 4 with this term).  Documented bias: the change of a mesh's shadow on ITSELF is not sampled.
    // Candidates come from a mixture: half of them proportional to length over ALL edges, half over the edges of the differentiated
    // mesh (static edges only matter where the moving mesh is x_k or lies behind them; in a scene of big static polygons they would
    // otherwise take nearly every candidate).  q = pdf per unit length relative to the plain one.
    uint32_t chosen = 0, n_sil = 0; float w_sum = 0.f, w_chosen = 0.f;
    const bool x1_on_moving = ok && si.mesh == e.moving_mesh;
    const float *cdf = e.cdf;
    float mov_lo = 0.f, mov_total = 0.f;
    if (ok && e.mov_end > e.mov_begin) { mov_lo = e.mov_begin ? cdf[e.mov_begin - 1] : 0.f; mov_total = cdf[e.mov_end - 1] - mov_lo; }
    const bool mix = mov_total > 0.f && mov_total < e.total_len;
    for (int c = 0; c < e.candidates; ++c) {          // same number of draws whatever happens: the stream stays aligned
        float u = rng.next_f32(), pick = rng.next_f32();
        if (!ok) continue;
        uint32_t k;
        if (mix && u >= 0.5f) {
            float value = fmaf(2.f * u - 1.f, mov_total, mov_lo);
            k = min(max(binary_search(e.mov_begin, e.mov_end, [&](uint32_t i) { return cdf[i] <= value; }), e.mov_begin), e.mov_end - 1);
        } else {
            float value = (mix ? 2.f * u : u) * e.total_len;
            k = min(binary_search(0u, e.n_edges, [&](uint32_t i) { return cdf[i] <= value; }), e.n_edges - 1);
        }
        const float q = mix ? 0.5f + (k >= e.mov_begin && k < e.mov_end ? 0.5f * e.total_len / mov_total : 0.f) : 1.f;
        uint4 ek = e.edges[k];
        if (ek.z == si.prim || ek.w == si.prim) continue;                 // edges of x1's own face lie in its tangent plane
        Edge2Geo gk = edge2_geo(sc, ek, 0.5f, si.p);
        if (!gk.silhouette) continue;
        if (x1_on_moving && (int) sc.faces[ek.z].w == si.mesh && gk.r < e.self_rmin * norm(gk.b - gk.a)) continue;
        // resampling target: angular length / r, times a prior of 0.1 for a static edge seen from a static x_k (it only
        // contributes when the moving mesh happens to lie behind it, which is not known before the far ray is traced)
        float target = gk.dwdt / gk.r * ((k >= e.mov_begin && k < e.mov_end) || x1_on_moving ? 1.f : 0.1f), w = target / q;
        if (!(w > 0.f) || !(w < 1e30f)) continue;
        n_sil++; w_sum += w;
        if (pick * w_sum < w) { chosen = k; w_chosen = target; }      // weighted reservoir
    }
    float t = rng.next_f32();
    if (ok && n_sil > 0 && w_chosen > 0.f) {
        uint4 ed = e.edges[chosen];
        Edge2Geo g = edge2_geo(sc, ed, t, si.p);
        float len = norm(g.b - g.a);
        float mint = kRayEpsilon * (1.f + hmax(abs3(si.p)));
        uint2 fast[8]; TravStack<8, 56> st; st.fast = fast; st.stride = 1; Hit h;
        bool vis = g.silhouette && len > 0.f && dot(g.w, si.sh.n) * si.wi.z > 0.f &&
                   !bvh_trace<true>(sc.bvh, si.p, g.w, mint, g.r * (1.f - 1e-3f), st, h);
        if (vis) {
            float3 wf = normalize(g.w + g.n_out * e.eps);
            bool hit2 = bvh_trace<false>(sc.bvh, si.p, wf, mint, INFINITY, st, h);
            if (!hit2 || h.t > g.r * (1.f - 1e-3f)) {     // the ray really passes the edge
                float inv_p = (e.total_len / len) * (w_sum / ((float) e.candidates * w_chosen));
                rec0 = make_float4(u2f(chosen), t, g.dwdt * inv_p, 1.f);
                if (hit2) rec1 = make_float4(u2f(h.prim), h.u, h.v, h.t);
                ro = mk4(si.p, mint); rd = mk4(wf, INFINITY);
            }
        }
    }
    float4 *aux = e.aux + 5 * (size_t) pid;
    aux[0] = rec0; aux[1] = rec1; aux[2] = d4; aux[3] = h4; aux[4] = o4;
    qout.o_mint[j] = ro; qout.d_maxt[j] = rd; qout.thr_eta[j] = qin.thr_eta[j];       // prefix throughput (1 at the camera's hit)
    qout.misc[j] = make_uint4(pid, (uint32_t) (e.depth + 1) | 0x80000000u, f2u(1.f), 0u);      // next depth after a "delta" bounce: emitter hits count fully
    rng_store(&qout.rng[j], rng);
    b.result[pid] = make_float4(0.f, 0.f, 0.f, 0.f);      // whatever the prefix gathered is not part of this term
}
template <typename Add>
HD void edge2_scatter_body(const SceneDev &sc, const Edge2Ctx &e, const GradCtx &g, const RenderBatch &b, uint32_t j, Add add) {
    const float4 *aux = e.aux + 5 * (size_t) j;        // j = path id
    float4 rec0 = aux[0];
    if (rec0.w == 0.f) return;
    float4 rec1 = aux[1], d4 = aux[2], h4 = aux[3];
    SI si = make_si(sc, xyz(aux[4]), xyz(d4), h4);
    uint4 ed = e.edges[f2u(rec0.x)];
    Edge2Geo eg = edge2_geo(sc, ed, rec0.y, si.p);
    const BsdfDev bsdf = sc.bsdfs[sc.meshes[si.mesh].bsdf];
    float3 val; float pdf;
    bsdf_eval_pdf(sc, bsdf, si, si.sh.to_local(eg.w), val, pdf);
    float4 L = b.result[j];
    float3 dI = image_adjoint(g, b, j);
    float c = -(dI.x * val.x * L.x + dI.y * val.y * L.y + dI.z * val.z * L.z) * rec0.z;
    if (c == 0.f || !(fabsf(c) < 1e30f)) return;
    float inv_r = 1.f / eg.r, inv_r2 = rec1.w < INFINITY ? 1.f / rec1.w : 0.f;
    float3 gp = eg.n_out * (c * inv_r);
    add_vertex(g, ed.x, gp * (1.f - rec0.y), add); add_vertex(g, ed.y, gp * rec0.y, add);
    VtxRec r1; r1.prim = si.prim; r1.b1 = h4.y; r1.b2 = h4.z;
    scatter_point(g, tri_geo(sc, r1.prim), r1, e.depth == 1, si.p, eg.n_out * (c * (inv_r2 - inv_r)), add);
    if (inv_r2 > 0.f) {
        VtxRec r2; r2.prim = f2u(rec1.x); r2.b1 = rec1.y; r2.b2 = rec1.z;
        TriGeo t2 = tri_geo(sc, r2.prim);
        scatter_point(g, t2, r2, false, tri_point(t2, r2), eg.n_out * (-c * inv_r2), add);
    }
}

} // namespace mtsb
