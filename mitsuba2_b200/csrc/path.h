// path.h -- wavefront path-tracer stage bodies (HD): raygen, shade (emission + RR + emitter sampling with
// MIS + BSDF sampling), shadow accumulate, film splat.  State machine = PathIntegrator::sample
// (src/integrators/path.cpp:100-211) unrolled across kernel launches; per-sample prefix =
// SamplingIntegrator::render_sample (src/librender/integrator.cpp:224-271).
#pragma once
#include "emit.h"

namespace mtsb {

// Compacted per-segment queue (ping-pong). 96 B per live path.
struct PathQueue {
    float4 *o_mint, *d_maxt;   // extension ray
    float4 *hit;               // closest-hit result for that ray (t,u,v,prim)
    float4 *thr_eta;           // throughput rgb, eta
    uint4  *misc;              // path_id, depth | (prev_delta << 31), prev bsdf pdf (bits), depth 1 only: terminator face + 1 (else 0)
    uint4  *rng;               // PCG32 state lo/hi, inc lo/hi
    float2 *wav_eta;           // spectral mode only (+8 B): wavelength sample of the path, eta; thr_eta then holds 4 channels
    float4 *prev1, *prev2;     // vertex-gradient replay only (else null): the two previous path vertices
                               // (prim bits, b1, b2, chain flag); prim 0xffffffff = the camera
};
struct ShadowQueue {
    float4 *o_mint, *d_maxt;   // shadow ray
    float4 *contrib_id;        // mis * throughput * bsdf * Le/pdf (rgb; spectral mode: already converted to XYZ), path_id (bits)
    // vertex-gradient replay only (else null): what k_nee_grad needs to differentiate the emitter sample once the shadow ray
    // is known to be free -- (queue index of the vertex, the two emitter-sampling random numbers, depth) and the any-hit result
    float4 *nee_rec; uint32_t *occluded;
};
struct RenderBatch {
    uint32_t n_paths;          // paths in this batch
    uint64_t offset;           // global path index of the batch's first path (index = pixel*spp_batch + s)
    uint32_t spp_begin, spp_batch, spp_total;
    uint64_t seed;
    float4 *result;            // per path: radiance rgb (spectral mode: XYZ), alpha
};

HD void rng_store(uint4 *dst, const Pcg32 &r) {
    *dst = make_uint4((uint32_t) r.state, (uint32_t) (r.state >> 32), (uint32_t) r.inc, (uint32_t) (r.inc >> 32));
}
HD Pcg32 rng_load(uint4 v) {
    Pcg32 r; r.state = (uint64_t) v.x | ((uint64_t) v.y << 32); r.inc = (uint64_t) v.z | ((uint64_t) v.w << 32);
    return r;
}

// Sample stream of path i = pixel * spp_total + sample index, independent of how the samples are split into batches and
// ranks.  The sampler takes a 32-bit index (TEA pair, integrator.cpp:144-169); a job with more than 2^32 samples folds the
// high half of the 64-bit index into the seed, so indices below 2^32 seed exactly as before.
HD void path_ids(const RenderBatch &b, uint32_t i, uint32_t &pixel, uint32_t &stream, uint64_t &seed) {
    uint64_t g = b.offset + i;
    pixel = (uint32_t) (g / b.spp_batch);
    uint32_t s = b.spp_begin + (uint32_t) (g % b.spp_batch);
    uint64_t idx = (uint64_t) pixel * b.spp_total + s;
    stream = (uint32_t) idx; seed = b.seed + (idx >> 32);
}

// PerspectiveCamera::sample_ray (src/sensors/perspective.cpp:151-188); pos in [0,1]^2
HD void sensor_sample_ray(const SensorDev &sn, float px, float py, float4 &o_mint, float4 &d_maxt) {
    float3 near_p = xform_point_persp(sn.sample_to_camera, mk3(px, py, 0.f));
    float3 d = normalize(near_p);
    float inv_z = 1.f / d.z;
    float3 o = mk3(sn.to_world[3], sn.to_world[7], sn.to_world[11]);
    float3 dw = xform_vec(sn.to_world, d);
    o_mint = mk4(o, sn.near_clip * inv_z); d_maxt = mk4(dw, sn.far_clip * inv_z);
}

template <typename M = Rgb>
HD void raygen_body(const SensorDev &sn, const RenderBatch &b, const PathQueue &q, uint32_t i) {
    uint32_t pixel, stream; uint64_t seed; path_ids(b, i, pixel, stream, seed);
    Pcg32 rng = sampler_seed(seed, stream);
    float jx = rng.next_f32(), jy = rng.next_f32();
    float wavelength_sample = rng.next_f32();    // integrator.cpp:237; drawn in every mode, used by the spectral one
    if (M::spectral) q.wav_eta[i] = make_float2(wavelength_sample, 1.f);
    float fx = (float) (pixel % sn.width) + jx, fy = (float) (pixel / sn.width) + jy;
    float4 o, d;
    sensor_sample_ray(sn, fx / (float) sn.width, fy / (float) sn.height, o, d);
    q.o_mint[i] = o; q.d_maxt[i] = d;
    q.thr_eta[i] = make_float4(1.f, 1.f, 1.f, 1.f);
    q.misc[i] = make_uint4(i, 1u, f2u(0.f), 0u);
    rng_store(&q.rng[i], rng);
    b.result[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

HD float mis_weight(float a, float b) { a *= a; b *= b; return a > 0.f ? a / (a + b) : 0.f; }

struct ShadeOut {
    bool extend, shadow;
    float4 o_mint, d_maxt, thr_eta; uint4 misc, rng; float2 wav_eta;   // extension record
    float4 s_o_mint, s_d_maxt, s_contrib_id;                  // shadow record
    float4 prev1, prev2;                                       // vertex-gradient replay: records carried to the next vertex
    float4 s_nee_rec;                                          // vertex-gradient replay: ShadowQueue::nee_rec of the shadow record
    // adjoint bookkeeping (filled always; cheap): first-hit info of this vertex
    float3 emitted;                                            // radiance added at this vertex (already weighted)
};

// per-mode packing of (throughput, eta): rgb -> one float4; spectral -> 4 channels + (wavelength sample, eta) in wav_eta
HD void load_throughput(const PathQueue &, uint32_t, float4 te, float3 &thr, float &eta, float2 &) { thr = xyz(te); eta = te.w; }
HD void load_throughput(const PathQueue &q, uint32_t i, float4 te, float4 &thr, float &eta, float2 &we) { thr = te; we = q.wav_eta[i]; eta = we.y; }
HD void store_throughput(ShadeOut &out, float3 thr, float eta, float2) { out.thr_eta = mk4(thr, eta); }
HD void store_throughput(ShadeOut &out, float4 thr, float eta, float2 we) { out.thr_eta = thr; out.wav_eta = make_float2(we.x, eta); }
// what a contribution adds to the per-path result: rgb radiance as is; spectral mode applies the sensor's wavelength
// weight and converts to XYZ right away (integrator.cpp:249-262 does both once per sample; the conversion is linear)
HD float3 to_result(const SceneDev &, float3 c, float4, float4) { return c; }
HD float3 to_result(const SceneDev &sc, float4 c, float4 wav, float4 ray_weight) { return spectrum_to_xyz(sc, ray_weight * c, wav); }

// ---- adjoint of the path w.r.t. a bitmap texture that enters a BSDF multiplicatively ------------------------------
// (diffuse `reflectance`, src/bsdfs/diffuse.cpp:78-135; roughconductor `specular_reflectance`, roughconductor.cpp:270)
// L = P_k + a_k * R_k at every vertex k that looks the texture up (a_k = bilinear texel mix, P_k = radiance gathered
// before the vertex's own BSDF is used, a_k R_k = everything after).  The replay pass knows the final radiance of the
// path (from the primal pass) and the running prefix, so  dL/da_k = (L - P_k) / a_k  and the texel gradient is that
// times the bilinear weights.  Sampling decisions and Russian roulette are treated as constants (detached), which
// keeps the estimator unbiased.  Replaces the Enoki tape of src/python/python/autodiff.py:6-194 for this parameter.
struct GradCtx {
    const float4 *final_radiance;   // per path, from the primal pass
    const float *dL_dimage;         // width*height*3, dL/d(developed RGB)
    float *grad_tex;                // texture-sized accumulator
    int tex_id;
    float scale;                    // d(developed pixel)/d(sample radiance) = 1/spp_total (box filter)
    // reconstruction filters wider than a pixel: the developed image is sum_s w_s(p) L_s / W[p], so a sample's adjoint weight
    // is a gather over its footprint (film_gather below), done once per path and batch by k_path_adjoint; null = box filter
    const float4 *path_dl = nullptr;
    const float *inv_w = nullptr;   // per pixel: spp_total / W[p], W = the film's weight channel over ALL spp_total samples (0 where W = 0)
    FilmFilter filt = { 0, 0.f, 0.f, 0.f, 0u, 0u };
    // vertex-position adjoint (adjoint.h); grad_pos == nullptr: texture gradient only
    float *grad_pos;                // 3 floats per vertex of mesh `mesh_id`
    float *grad_vnrm;               // same indexing: adjoint of the mesh's VERTEX NORMALS (smooth-shaded mesh), turned into
                                    // position gradients after the replay by vnormal_adjoint_body (adjoint.h); may be null
    int mesh_id; uint32_t vert_begin, vert_count;
    float3 cam_o;                   // sensor origin (x0 of every path)
};
// dL/d(radiance of path `path_id`), per channel: box filter = the path's pixel / spp_total; wider filters = the gathered value
HD float3 image_adjoint(const GradCtx &g, const RenderBatch &b, uint32_t path_id) {
    if (g.path_dl != nullptr) return xyz(g.path_dl[path_id]);
    uint32_t pixel = (uint32_t) ((b.offset + path_id) / b.spp_batch);
    return ld3(g.dL_dimage, pixel) * g.scale;
}
// Hooks the adjoint replays plug into shade_body (all empty for the primal render):
//   on_vertex: a path vertex was reconstructed (before its emission is added to the result);
//   on_bsdf:   the vertex's BSDF is going to be used (after emission, Russian roulette and the depth cut).
// (Emitter samples are differentiated after the shadow pass, from ShadowQueue::nee_rec -- nee_adjoint() in adjoint.h.)
struct NoGrad {
    static constexpr bool kVertex = false;
    HD void on_vertex(const SceneDev &, const PathQueue &, const RenderBatch &, uint32_t, uint32_t, uint32_t, float3, float3, const SI &) const { }
    HD void on_bsdf(const SceneDev &, const RenderBatch &, uint32_t, const SI &, const BsdfDev &) const { }
};

template <typename AddFn>
HD void texture_adjoint(const SceneDev &sc, const RenderBatch &b, const GradCtx &g, uint32_t path_id, const SI &si,
                        const BsdfDev &bsdf, AddFn add) {
    int tid = bsdf.type == 0 ? bsdf.refl : (bsdf.type == 1 ? bsdf.spec_r : -1);
    if (tid != g.tex_id || tid < 0) return;
    if (bsdf.two_sided ? si.wi.z == 0.f : !(si.wi.z > 0.f)) return;   // no light is carried past this vertex
    const TexDev &t = sc.tex[tid];
    uint32_t idx[4]; float w[4];
    if (!tex_footprint(t, si.u, si.v, idx, w)) return;
    float3 a = tex_eval(sc, tid, si.u, si.v);
    float4 fin = g.final_radiance[path_id], run = b.result[path_id];
    float3 dI = image_adjoint(g, b, path_id);
    const float tiny = 1e-8f;
    float3 ga = mk3(a.x > tiny ? dI.x * (fin.x - run.x) / a.x : 0.f, a.y > tiny ? dI.y * (fin.y - run.y) / a.y : 0.f,
                    a.z > tiny ? dI.z * (fin.z - run.z) / a.z : 0.f);
    for (int k = 0; k < 4; ++k) {
        if (t.ch == 1) add(g.grad_tex + idx[k], (ga.x + ga.y + ga.z) * w[k]);
        else { add(g.grad_tex + 3 * (size_t) idx[k], ga.x * w[k]); add(g.grad_tex + 3 * (size_t) idx[k] + 1, ga.y * w[k]); add(g.grad_tex + 3 * (size_t) idx[k] + 2, ga.z * w[k]); }
    }
}

// One iteration of the loop in path.cpp:124-208 for queue entry i.  `result` accumulates emission.  `grad` is called
// once the vertex's BSDF is known to be used (after emission, Russian roulette and the depth cut).
template <typename M = Rgb, typename GradFn = NoGrad>
HD ShadeOut shade_body(const SceneDev &sc, const PathQueue &q, const RenderBatch &b, uint32_t i, const GradFn &grad = GradFn()) {
    typedef typename M::Spec Spec;
    ShadeOut out; out.extend = false; out.shadow = false; out.emitted = mk3(0.f);
    float4 o4 = q.o_mint[i], d4 = q.d_maxt[i], h4 = q.hit[i], te = q.thr_eta[i];
    uint4 misc = q.misc[i];
    uint32_t path_id = misc.x, depth = misc.y & 0x7fffffffu; bool prev_delta = (misc.y >> 31) != 0;
    float prev_pdf = u2f(misc.z);
    Spec thr; float eta; float4 wav = float4(), ray_weight = float4(); float2 we = float2();
    load_throughput(q, i, te, thr, eta, we);
    if (M::spectral) sample_wavelengths(we.x, wav, ray_weight);
    float3 o = xyz(o4), d = xyz(d4);
    SI si = make_si<M::ext>(sc, o, d, h4);
    if constexpr (M::ext) {
        // boundary sample on a shading-normal terminator (adjoint.h, kTermTag): misc.w = face + 1.  The path measures the
        // radiance of the LIT side of that curve: it counts only if the camera ray really sees the face, and the face is
        // shaded as if its interpolated normal just looked at the camera (the dark side is exactly 0 for a one-sided BSDF)
        // misc.w = ~0: a silhouette pair's ray that counts as black (the mesh side where the dark rim reaches the edge).
        if (depth == 1 && misc.w != 0u) {
            if (misc.w == 0xffffffffu || !si.valid || si.prim + 1u != misc.w) return out;
            si.wi.z = fmaxf(fabsf(si.wi.z), 1e-4f);
        }
    }
    bool chain_ok = true;                     // vertex-gradient replay: no delta vertex on the path so far
    if (GradFn::kVertex) {
        float4 p1 = depth > 1 ? q.prev1[i] : make_float4(u2f(0xffffffffu), 0.f, 0.f, u2f(1u));
        chain_ok = f2u(p1.w) != 0u;
        out.prev2 = p1;
        if (chain_ok) grad.on_vertex(sc, q, b, i, path_id, depth, o, d, si);
    }

    // ---- emission (path.cpp:128-129) with the MIS weight of the previous BSDF sample (:194-205)
    int em = si.valid ? sc.meshes[si.mesh].emitter : sc.env_emitter;
    if (em >= 0) {
        const EmitterDev &e = sc.emitters[em];
        Spec Le; float emitter_pdf = 0.f;
        if (e.type == 0) {
            Le = si.wi.z > 0.f ? emitter_radiance<M>(sc, e, wav) : M::splat(0.f);
            if (depth > 1 && !prev_delta) {
                float3 dd = si.p - o; float dist = norm(dd); dd = dd / dist;
                emitter_pdf = area_pdf_direction<M::ext>(sc, e, o, dd, si.sh.n, dist);
            }
        } else if (e.type == 1) {
            Le = env_eval<M>(sc, e, d, wav);
            if (depth > 1 && !prev_delta) emitter_pdf = env_pdf_direction(e, d);
        } else {                                                  // constant background (constant.cpp:105-108)
            Le = emitter_radiance<M>(sc, e, wav);
            if (depth > 1 && !prev_delta) emitter_pdf = 0.25f * kInvPi;
        }
        if (sc.n_emitters > 1) emitter_pdf *= 1.f / (float) sc.n_emitters;
        float w = depth > 1 ? mis_weight(prev_pdf, emitter_pdf) : 1.f;
        float3 add = to_result(sc, thr * Le * w, wav, ray_weight);
        out.emitted = add;
        float4 r = b.result[path_id];
        r.x += add.x; r.y += add.y; r.z += add.z;
        if (depth == 1) r.w = si.valid ? 1.f : 0.f;
        b.result[path_id] = r;
    } else if (depth == 1) {
        float4 r = b.result[path_id]; r.w = si.valid ? 1.f : 0.f; b.result[path_id] = r;
    }
    if (!si.valid) return out;

    Pcg32 rng = rng_load(q.rng[i]);
    // ---- Russian roulette (path.cpp:137-141)
    if ((int) depth > sc.rr_depth) {
        float qq = fminf(hmax(thr) * sqr(eta), 0.95f);
        bool survive = rng.next_f32() < qq;
        thr *= 1.f / qq;
        if (!survive) return out;
    }
    if ((uint32_t) depth >= (uint32_t) sc.max_depth) return out;

    const BsdfDev bsdf = sc.bsdfs[sc.meshes[si.mesh].bsdf];
    grad.on_bsdf(sc, b, path_id, si, bsdf);
    // spectral mode: one texture lookup shared by eval and sample (spectrum evaluations are expensive); the rgb kernel
    // re-reads the few floats at each use instead of keeping them live (it is compiled for 48 registers)
    BsdfTexT<M> btex;
    if (M::spectral) btex = bsdf_textures<M>(sc, bsdf, si, wav);
    // ---- emitter sampling (path.cpp:155-172, scene.cpp:141-189)
    if (bsdf_is_smooth(bsdf)) {
        float sx = rng.next_f32(), sy = rng.next_f32();
        if (sc.n_emitters > 0) {
            DirSampleT<M> ds = sample_emitter_direction<M>(sc, si.p, sx, sy, wav);
            if (ds.pdf != 0.f) {
                float3 wo = si.sh.to_local(ds.d);
                Spec bval; float bpdf;
                if (!M::spectral) btex = bsdf_textures<M>(sc, bsdf, si, wav);
                bsdf_eval_pdf<M>(sc, bsdf, si, wo, bval, bpdf, btex);
                float mis = ds.delta ? 1.f : mis_weight(ds.pdf, bpdf);          // path.cpp:169
                Spec c = thr * bval * ds.value * mis;
                if (!is_black(c)) {
                    if (GradFn::kVertex) out.s_nee_rec = make_float4(u2f(i), sx, sy, u2f(chain_ok ? depth : 0u));
                    out.shadow = true;
                    out.s_o_mint = mk4(si.p, kRayEpsilon * (1.f + hmax(abs3(si.p))));
                    out.s_d_maxt = mk4(ds.d, ds.dist * (1.f - kShadowEpsilon));
                    out.s_contrib_id = mk4(to_result(sc, c, wav, ray_weight), u2f(path_id));
                }
            }
        }
    }
    // ---- BSDF sampling (path.cpp:177-190)
    float s1 = rng.next_f32(), s2x = rng.next_f32(), s2y = rng.next_f32();
    if (!M::spectral) btex = bsdf_textures<M>(sc, bsdf, si, wav);
    BsdfSampleT<M> bs = bsdf_sample<M>(sc, bsdf, si, s1, s2x, s2y, btex);
    thr *= bs.weight;
    if (is_black(thr)) return out;
    eta *= bs.eta;
    float3 wo_w = si.sh.to_world(bs.wo);
    out.extend = true;
    out.o_mint = mk4(si.p, (1.f + hmax(abs3(si.p))) * kRayEpsilon);
    out.d_maxt = mk4(wo_w, INFINITY);
    store_throughput(out, thr, eta, we);
    out.misc = make_uint4(path_id, (depth + 1u) | (bs.delta ? 0x80000000u : 0u), f2u(bs.pdf), 0u);
    rng_store(&out.rng, rng);
    if (GradFn::kVertex) out.prev1 = make_float4(u2f(si.prim), h4.y, h4.z, u2f(chain_ok && !bs.delta ? 1u : 0u));
    return out;
}

// ---- film: ImageBlock::put (src/librender/imageblock.cpp:79-172) on a border-less full-film block ----------
template <typename S> HD float rfilter_eval(const S &sn, float x) {
    if (sn.rfilter == 0) return fabsf(x) <= sn.radius ? 1.f : 0.f;                 // src/rfilters/box.cpp:34-36
    if (sn.rfilter == 1) return fmaxf(0.f, 1.f - fabsf(x / sn.radius));            // src/rfilters/tent.cpp:33-35
    if (sn.rfilter == 3 || sn.rfilter == 4) {                                      // mitchell.cpp:40-53, catmullrom.cpp:29-43
        float B = sn.rfilter == 3 ? sn.stddev : 0.f, C = sn.rfilter == 3 ? sn.filter_c : 0.5f;
        x = fabsf(x); float x2 = x * x, x3 = x2 * x;
        float r = (1.f / 6.f) * (x < 1.f ? (12.f - 9.f * B - 6.f * C) * x3 + (-18.f + 12.f * B + 6.f * C) * x2 + (6.f - 2.f * B)
                                         : (-B - 6.f * C) * x3 + (6.f * B + 30.f * C) * x2 + (-12.f * B - 48.f * C) * x + (8.f * B + 24.f * C));
        return x < 2.f ? r : 0.f;
    }
    if (sn.rfilter == 5) {                                                         // lanczos.cpp:44-54
        x = fabsf(x); float x1 = kPi * x, x2 = x1 / sn.radius;
        return x < kEpsilon ? 1.f : (x > sn.radius ? 0.f : (sinf(x1) * sinf(x2)) / (x1 * x2));
    }
    float alpha = -1.f / (2.f * sn.stddev * sn.stddev);                            // src/rfilters/gaussian.cpp:34-47
    return fmaxf(0.f, expf(alpha * x * x) - expf(alpha * sn.radius * sn.radius));
}

// ImageBlock::put drops a sample with a non-finite or negative (< -1e-5) channel, weight included, instead of
// accumulating it (src/librender/imageblock.cpp:84-108; warn_negative / warn_invalid are on for the path integrator's blocks)
HD bool film_sample_valid(float4 xyza, float wgt) {
    float lo = fminf(fminf(fminf(xyza.x, xyza.y), fminf(xyza.z, xyza.w)), wgt), sum = xyza.x + xyza.y + xyza.z + xyza.w + wgt;
    return lo >= -1e-5f && fabsf(sum) <= FLT_MAX && xyza.x == xyza.x && xyza.y == xyza.y && xyza.z == xyza.z && xyza.w == xyza.w;
}
template <typename AddFn>
HD void film_put(const SensorDev &sn, float px, float py, float4 xyza, float wgt, AddFn add) {
    if (!film_sample_valid(xyza, wgt)) return;
    // full-film block: offset 0, border 0 => pos = pos_ - 0.5
    float x = px - 0.5f, y = py - 0.5f;
    if (sn.radius > 1.f) {
        int lox = max((int) ceilf(x - sn.radius), 0), loy = max((int) ceilf(y - sn.radius), 0);
        int hix = min((int) floorf(x + sn.radius), (int) sn.width - 1), hiy = min((int) floorf(y + sn.radius), (int) sn.height - 1);
        int n = (int) ceilf((sn.radius - 2.f * kRayEpsilon) * 2.f);
        for (int yr = 0; yr < n; ++yr) {
            int yy = loy + yr; if (yy > hiy) break;
            float wy = rfilter_eval(sn, (float) loy - y + (float) yr);
            for (int xr = 0; xr < n; ++xr) {
                int xx = lox + xr; if (xx > hix) break;
                float w = wy * rfilter_eval(sn, (float) lox - x + (float) xr);
                add((uint32_t) (yy * (int) sn.width + xx), make_float4(xyza.x * w, xyza.y * w, xyza.z * w, xyza.w * w), wgt * w);
            }
        }
    } else {
        int lx = (int) ceilf(x - 0.5f), ly = (int) ceilf(y - 0.5f);
        if (lx >= 0 && ly >= 0 && lx < (int) sn.width && ly < (int) sn.height)
            add((uint32_t) (ly * (int) sn.width + lx), xyza, wgt);
    }
}

// Adjoint of film_put followed by the film's division by its weight channel (hdrfilm.cpp develop): a sample at (px, py)
// contributes w(p - x) L / W[p] to every pixel p of its footprint, so dL/dL_sample = sum_p dL/dI[p] w(p - x) inv_w[p] / spp_total
// (inv_w = spp_total / W; the caller applies 1 / spp_total).  Same footprint arithmetic as film_put.
HD float3 film_gather(const FilmFilter &sn, const float *dl, const float *inv_w, float px, float py) {
    float x = px - 0.5f, y = py - 0.5f;
    float3 acc = mk3(0.f);
    int lox = max((int) ceilf(x - sn.radius), 0), loy = max((int) ceilf(y - sn.radius), 0);
    int hix = min((int) floorf(x + sn.radius), (int) sn.width - 1), hiy = min((int) floorf(y + sn.radius), (int) sn.height - 1);
    int n = (int) ceilf((sn.radius - 2.f * kRayEpsilon) * 2.f);
    for (int yr = 0; yr < n; ++yr) {
        int yy = loy + yr; if (yy > hiy) break;
        float wy = rfilter_eval(sn, (float) loy - y + (float) yr);
        for (int xr = 0; xr < n; ++xr) {
            int xx = lox + xr; if (xx > hix) break;
            uint32_t pix = (uint32_t) (yy * (int) sn.width + xx);
            acc = acc + ld3(dl, pix) * (wy * rfilter_eval(sn, (float) lox - x + (float) xr) * inv_w[pix]);
        }
    }
    return acc;
}

// src/librender/integrator.cpp:254-270: rgb -> XYZ (core/spectrum.h:220-227), aovs = {X,Y,Z,alpha,1}
HD float4 radiance_to_xyza(float4 r);
template <typename M> HD float4 result_to_xyza(float4 r) { if constexpr (M::spectral) return r; else return radiance_to_xyza(r); }
HD float4 radiance_to_xyza(float4 r) {
    return make_float4(0.412453f * r.x + 0.357580f * r.y + 0.180423f * r.z,
                       0.212671f * r.x + 0.715160f * r.y + 0.072169f * r.z,
                       0.019334f * r.x + 0.119193f * r.y + 0.950227f * r.z, r.w);
}
// position sample of path i (re-derived from the RNG stream instead of stored)
HD void path_film_pos(const SensorDev &sn, const RenderBatch &b, uint32_t i, float &fx, float &fy) {
    uint32_t pixel, stream; uint64_t seed; path_ids(b, i, pixel, stream, seed);
    Pcg32 rng = sampler_seed(seed, stream);
    float jx = rng.next_f32(), jy = rng.next_f32();
    fx = (float) (pixel % sn.width) + jx; fy = (float) (pixel / sn.width) + jy;
}

} // namespace mtsb
