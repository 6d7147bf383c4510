// hostsim.cu -- TEST-ONLY host simulator: runs the product's per-thread stage bodies (mitsuba2_b200/csrc/*.h,
// the HD functions the CUDA kernels wrap) sequentially on the CPU, so the BVH builder, the compressed-node
// traversal and the wavefront state machine can be checked against the oracle in the `-m "not gpu"` suite.
// It is NOT part of the product library and makes no CUDA runtime calls.
#define MTSB_TRAV_STATS 1
#include "../../include/mtsb200.h"
#include "../../mitsuba2_b200/csrc/adjoint.h"
#include "../../mitsuba2_b200/csrc/host_tables.h"
#include "../../mitsuba2_b200/csrc/cie_tables.h"
#include <vector>
#include <algorithm>
#include <numeric>
#include <cstring>
#include <memory>

using namespace mtsb;

struct hs_scene {
    std::vector<MeshDev> meshes; std::vector<TexDev> tex; std::vector<BsdfDev> bsdfs; std::vector<EmitterDev> emitters;
    std::vector<SensorDev> sensors;
    std::vector<float> pos, nrm, uv; std::vector<uint4> faces; std::vector<ShapeDev> shapes;
    std::vector<std::vector<float>> tables; std::vector<float4> env_rgba; std::vector<float> cie; bool spectral = false, ext = false;
    std::vector<uint4> nodes; std::vector<float4> tris;
    std::vector<std::vector<float4>> film_xyza; std::vector<std::vector<float>> film_w;
    SceneDev dev; uint32_t bvh_nodes = 0, bvh_depth = 0;
    std::vector<uint32_t> bvh_levels;       // as mtsb_scene::bvh_levels (api.cu)
    float3 lo, hi;
};

// EXPERIMENT (HS_SAH=1): top-down binned-SAH binary tree in the node numbering the collapse expects, to measure how much
// a higher-quality tree would save over the Morton-order LBVH.
struct SahCtx { BvhBuild *b; std::vector<float3> clo, chi, cen; int next_internal; };
static int sah_build(SahCtx &c, uint32_t a, uint32_t e) {      // range [a, e) of b.order
    BvhBuild &b = *c.b; int n = (int) b.n;
    if (e - a == 1) { int id = n - 1 + (int) a; uint32_t t = b.order[a]; b.nlo[id] = mk4(c.clo[t], 0.f); b.nhi[id] = mk4(c.chi[t], 0.f); return id; }
    int id = c.next_internal++;
    float3 lo = mk3(INFINITY), hi = mk3(-INFINITY), klo = mk3(INFINITY), khi = mk3(-INFINITY);
    for (uint32_t i = a; i < e; ++i) { uint32_t t = b.order[i]; lo = fmin3(lo, c.clo[t]); hi = fmax3(hi, c.chi[t]); klo = fmin3(klo, c.cen[t]); khi = fmax3(khi, c.cen[t]); }
    const int NB = 16; float best = INFINITY; int best_axis = -1, best_bin = -1;
    for (int ax = 0; ax < 3; ++ax) {
        float k0 = ax == 0 ? klo.x : ax == 1 ? klo.y : klo.z, k1 = ax == 0 ? khi.x : ax == 1 ? khi.y : khi.z;
        if (!(k1 > k0)) continue;
        float3 blo[NB], bhi[NB]; uint32_t cnt[NB];
        for (int k = 0; k < NB; ++k) { blo[k] = mk3(INFINITY); bhi[k] = mk3(-INFINITY); cnt[k] = 0; }
        float scale = (float) NB / (k1 - k0);
        for (uint32_t i = a; i < e; ++i) { uint32_t t = b.order[i]; float cv = ax == 0 ? c.cen[t].x : ax == 1 ? c.cen[t].y : c.cen[t].z;
            int k = std::min(NB - 1, std::max(0, (int) ((cv - k0) * scale))); cnt[k]++; blo[k] = fmin3(blo[k], c.clo[t]); bhi[k] = fmax3(bhi[k], c.chi[t]); }
        float ra[NB]; uint32_t rc[NB]; float3 l2 = mk3(INFINITY), h2 = mk3(-INFINITY); uint32_t cc = 0;
        for (int k = NB - 1; k > 0; --k) { l2 = fmin3(l2, blo[k]); h2 = fmax3(h2, bhi[k]); cc += cnt[k]; ra[k] = cc ? half_area(l2, h2) : 0.f; rc[k] = cc; }
        l2 = mk3(INFINITY); h2 = mk3(-INFINITY); cc = 0;
        for (int k = 0; k + 1 < NB; ++k) { l2 = fmin3(l2, blo[k]); h2 = fmax3(h2, bhi[k]); cc += cnt[k];
            if (cc == 0 || rc[k + 1] == 0) continue;
            float cost = half_area(l2, h2) * (float) cc + ra[k + 1] * (float) rc[k + 1];
            if (cost < best) { best = cost; best_axis = ax; best_bin = k; } }
    }
    uint32_t mid;
    if (best_axis < 0) mid = (a + e) / 2;
    else {
        float k0 = best_axis == 0 ? klo.x : best_axis == 1 ? klo.y : klo.z, k1 = best_axis == 0 ? khi.x : best_axis == 1 ? khi.y : khi.z;
        float scale = (float) NB / (k1 - k0);
        auto it = std::partition(b.order + a, b.order + e, [&](uint32_t t) { float cv = best_axis == 0 ? c.cen[t].x : best_axis == 1 ? c.cen[t].y : c.cen[t].z;
            return std::min(NB - 1, std::max(0, (int) ((cv - k0) * scale))) <= best_bin; });
        mid = (uint32_t) (it - b.order);
        if (mid == a || mid == e) mid = (a + e) / 2;
    }
    b.rfirst[id] = a; b.rlast[id] = e - 1; b.nlo[id] = mk4(lo, 0.f); b.nhi[id] = mk4(hi, 0.f);
    int l = sah_build(c, a, mid), r = sah_build(c, mid, e);
    b.left[id] = l; b.right[id] = r; b.parent[l] = id; b.parent[r] = id;
    return id;
}

static void hs_build_bvh(hs_scene *s) {
    uint32_t n = (uint32_t) s->faces.size();
    s->dev.bvh.n_tris = n; s->dev.bvh.nodes = nullptr; s->dev.bvh.tris = nullptr;
    s->lo = mk3(INFINITY); s->hi = mk3(-INFINITY);
    if (n == 0) return;
    std::vector<float4> tlo(n), thi(n), nlo(2 * (size_t) n), nhi(2 * (size_t) n);
    std::vector<uint64_t> keys(n), keys2(n); std::vector<uint32_t> order(n), order2(n), rfirst(n), rlast(n), flags(n, 0), counters(2);
    std::vector<int> left(n), right(n), parent(2 * (size_t) n), wide_bin(n + 1);
    s->nodes.assign(5 * ((size_t) n + 1), make_uint4(0, 0, 0, 0)); s->tris.assign(3 * (size_t) n, make_float4(0, 0, 0, 0));
    BvhBuild b;
    b.pos = s->pos.data(); b.faces = s->faces.data(); b.shapes = s->shapes.data(); b.n = n;
    b.tlo = tlo.data(); b.thi = thi.data(); b.keys = keys.data(); b.order = order.data(); b.left = left.data(); b.right = right.data();
    b.parent = parent.data(); b.rfirst = rfirst.data(); b.rlast = rlast.data(); b.nlo = nlo.data(); b.nhi = nhi.data(); b.flags = flags.data();
    b.nodes = s->nodes.data(); b.tris = s->tris.data(); b.wide_bin = wide_bin.data(); b.counters = counters.data();
    for (uint32_t i = 0; i < n; ++i) { float3 lo, hi; tri_bounds_body(b, i, lo, hi); s->lo = fmin3(s->lo, lo); s->hi = fmax3(s->hi, hi); }
    b.scene_lo = s->lo; b.scene_hi = s->hi;
    for (uint32_t i = 0; i < n; ++i) morton_body(b, i);
    std::vector<uint32_t> perm(n); std::iota(perm.begin(), perm.end(), 0u);
    std::stable_sort(perm.begin(), perm.end(), [&](uint32_t a, uint32_t c) { return keys[a] < keys[c]; });
    for (uint32_t i = 0; i < n; ++i) { keys2[i] = keys[perm[i]]; order2[i] = order[perm[i]]; }
    b.keys = keys2.data(); b.order = order2.data();
    const char *mode = getenv("HS_BVH");           // default: PLOC (the product default); "lbvh" = Morton-split tree
    if (!getenv("HS_SAH") && n > 1 && !(mode && !strcmp(mode, "lbvh"))) {
        // same sequence as build_bvh() in api.cu
        std::vector<int> cidA(n), cidB(n), nn(n), leaf_new(n); std::vector<uint32_t> order3(n);
        counters[0] = 0;
        for (uint32_t i = 0; i < n; ++i) ploc_init_body(b, cidA.data(), i);
        uint32_t m = n; int *cur = cidA.data(), *nxt = cidB.data();
        while (m > 1) {
            for (uint32_t i = 0; i < m; ++i) ploc_nearest_body(b, cur, m, nn.data(), i);
            for (uint32_t i = 0; i < m; ++i) ploc_merge_body(b, cur, nn.data(), nxt, i);
            uint32_t k = 0;
            for (uint32_t i = 0; i < m; ++i) if (nxt[i] >= 0) cur[k++] = nxt[i];
            m = k;
        }
        for (uint32_t i = 0; i + 1 < n; ++i) ploc_finalize_internal_body(b, (int) i);
        for (uint32_t i = 0; i < n; ++i) ploc_finalize_leaf_body(b, order3.data(), leaf_new.data(), i);
        for (uint32_t i = 0; i < n; ++i) ploc_patch_leaf_body(b, leaf_new.data(), i);
        for (uint32_t i = 0; i < n; ++i) ploc_leaf_bounds_body(b, order3.data(), i);
        order2 = order3; b.order = order2.data();
    } else if (getenv("HS_SAH") && n > 1) {
        SahCtx c; c.b = &b; c.next_internal = 0; c.clo.resize(n); c.chi.resize(n); c.cen.resize(n);
        for (uint32_t i = 0; i < n; ++i) { float3 lo, hi; tri_bounds_body(b, i, lo, hi); c.clo[i] = lo; c.chi[i] = hi; c.cen[i] = (lo + hi) * 0.5f; }
        sah_build(c, 0, n);
    } else {
    for (uint32_t i = 0; i + 1 < n; ++i) lbvh_node_body(b, (int) i);
    for (uint32_t i = 0; i < n; ++i) lbvh_fit_body(b, (int) i);
    }
    counters[0] = 1; counters[1] = 0; wide_bin[0] = 0;
    uint32_t begin = 0, end = 1, depth = 0;
    s->bvh_levels.clear();
    while (begin < end) { s->bvh_levels.push_back(begin); for (uint32_t w = begin; w < end; ++w) collapse_body(b, w); begin = end; end = counters[0]; depth++; }
    s->bvh_levels.push_back(end);
    s->bvh_nodes = end; s->bvh_depth = depth;
    s->dev.bvh.nodes = s->nodes.data(); s->dev.bvh.tris = s->tris.data(); s->dev.bvh.shapes = s->shapes.empty() ? nullptr : s->shapes.data();
}

extern "C" void hs_scene_destroy(hs_scene *s) { delete s; }

// refit_bvh() of api.cu: new positions of one mesh under the topology of the last build
extern "C" void hs_refit_vertices(hs_scene *s, uint32_t mesh_id, const float *positions) {
    const MeshDev &m = s->meshes[mesh_id];
    std::copy(positions, positions + 3 * (size_t) m.vert_count, s->pos.begin() + 3 * (size_t) m.vert_begin);
    uint32_t n = (uint32_t) s->faces.size();
    std::vector<float4> box_lo(n + 1), box_hi(n + 1);
    BvhRefit r; r.nodes = s->nodes.data(); r.tris = s->tris.data(); r.pos = s->pos.data(); r.faces = s->faces.data(); r.shapes = s->shapes.data();
    r.box_lo = box_lo.data(); r.box_hi = box_hi.data();
    for (uint32_t t = 0; t < n; ++t) refit_tri_body(r, t);
    for (size_t l = s->bvh_levels.size() - 1; l-- > 0;)
        for (uint32_t w = s->bvh_levels[l]; w < s->bvh_levels[l + 1]; ++w) refit_node_body(r, w);
    s->lo = xyz(box_lo[0]); s->hi = xyz(box_hi[0]);
}

extern "C" hs_scene *hs_scene_create(const mtsb_scene_desc *d) {
    hs_scene *s = new hs_scene();
    memset(&s->dev, 0, sizeof(SceneDev));
    uint32_t vbase = 0, fbase = 0;
    std::vector<std::vector<uint32_t>> mesh_faces;
    for (uint32_t i = 0; i < d->n_meshes; ++i) {
        const mtsb_mesh_desc &m = d->meshes[i];
        MeshDev md; md.face_begin = fbase; md.face_count = m.face_count; md.vert_begin = vbase; md.vert_count = m.vertex_count;
        md.bsdf = m.bsdf_id; md.emitter = m.emitter_id; md.has_normals = m.normals != nullptr; md.has_uv = m.texcoords != nullptr;
        md.shape = -1;
        s->meshes.push_back(md);
        s->pos.insert(s->pos.end(), m.positions, m.positions + 3 * (size_t) m.vertex_count);
        if (m.normals) s->nrm.insert(s->nrm.end(), m.normals, m.normals + 3 * (size_t) m.vertex_count); else s->nrm.insert(s->nrm.end(), 3 * (size_t) m.vertex_count, 0.f);
        if (m.texcoords) s->uv.insert(s->uv.end(), m.texcoords, m.texcoords + 2 * (size_t) m.vertex_count); else s->uv.insert(s->uv.end(), 2 * (size_t) m.vertex_count, 0.f);
        mesh_faces.emplace_back(m.faces, m.faces + 3 * (size_t) m.face_count);
        for (uint32_t f = 0; f < m.face_count; ++f) s->faces.push_back(make_uint4(m.faces[3 * f] + vbase, m.faces[3 * f + 1] + vbase, m.faces[3 * f + 2] + vbase, i));
        vbase += m.vertex_count; fbase += m.face_count;
    }
    for (uint32_t k = 0; k < d->n_shapes; ++k) {       // analytic shapes: same layout as mtsb_scene_create
        const mtsb_shape_desc &sh = d->shapes[k];
        ShapeDev sd; const char *err = nullptr;
        if (!make_shape_dev(sh.type, sh.flip_normals, sh.to_world, sd, &err)) { delete s; return nullptr; }
        s->shapes.push_back(sd); s->ext = true;
        MeshDev md; md.face_begin = fbase + k; md.face_count = 1; md.vert_begin = 0; md.vert_count = 0;
        md.bsdf = sh.bsdf_id; md.emitter = sh.emitter_id; md.has_normals = 0; md.has_uv = 1; md.shape = (int) k;
        s->meshes.push_back(md);
        s->faces.push_back(make_uint4(k, kAnalyticTag, (uint32_t) sh.type, d->n_meshes + k));
    }
    s->tables.reserve(64 + d->n_textures + d->n_emitters);
    s->spectral = d->mode == MTSB_MODE_SPECTRAL;
    if (s->spectral) {
        s->cie.resize(4 * 95);
        for (int k = 0; k < 3 * 95; ++k) s->cie[k] = kCie1931[k];
        for (int k = 0; k < 95; ++k) s->cie[3 * 95 + k] = kCieD65[k] * (1.f / 10568.f);
        s->dev.cie = s->cie.data();
    }
    for (uint32_t i = 0; i < d->n_textures; ++i) {
        const mtsb_texture_desc &t = d->textures[i]; TexDev td; memset(&td, 0, sizeof(td));
        td.type = t.type; td.value = mk3(t.value[0], t.value[1], t.value[2]);
        if (t.type == MTSB_TEX_BITMAP) { td.w = t.width; td.h = t.height; td.ch = t.channels; memcpy(td.to_uv, t.to_uv, sizeof(td.to_uv));
            s->tables.emplace_back(t.data, t.data + (size_t) t.width * t.height * t.channels); td.data = s->tables.back().data(); }
        if (t.type == MTSB_TEX_REGULAR || t.type == MTSB_TEX_IRREGULAR) { td.w = t.width; td.lambda_min = t.lambda_min; td.lambda_max = t.lambda_max;
            if (t.type == MTSB_TEX_IRREGULAR) { s->tables.emplace_back(irregular_with_buckets(t.data, t.width)); td.h = kIrregularBuckets; }
            else s->tables.emplace_back(t.data, t.data + (size_t) t.width);
            td.data = s->tables.back().data(); }
        td.d65_scale = t.d65_scale;
        s->tex.push_back(td);
    }
    for (uint32_t i = 0; i < d->n_bsdfs; ++i) {
        const mtsb_bsdf_desc &b = d->bsdfs[i]; BsdfDev bd; memset(&bd, 0, sizeof(bd));
        bd.type = b.type; bd.refl = b.reflectance_tex; bd.eta = b.eta_tex; bd.k = b.k_tex; bd.spec_r = b.specular_reflectance_tex; bd.spec_t = b.specular_transmittance_tex;
        bd.alpha_u = b.alpha_u; bd.alpha_v = b.alpha_v; bd.distr = b.distribution; bd.sample_visible = b.sample_visible; bd.eta_f = b.eta; bd.two_sided = b.two_sided ? 1 : 0;
        bd.nonlinear = b.nonlinear; bd.spec_weight = b.specular_sampling_weight; bd.opacity = b.opacity_tex_plus1 - 1;
        if (b.type >= MTSB_BSDF_CONDUCTOR || bd.opacity >= 0) s->ext = true;
        if (b.type == MTSB_BSDF_ROUGHPLASTIC) {      // k_rough_tables + the host reduction of mtsb_scene_create
            std::vector<float> n32, w32, n128, w128; gauss_legendre_rule(32, n32, w32); gauss_legendre_rule(128, n128, w128);
            s->tables.emplace_back(2 * kRoughTransRes);
            std::vector<float> &tab = s->tables.back();
            for (int k = 0; k < 2 * kRoughTransRes; ++k) { bool tr = k < kRoughTransRes; float e = tr ? b.eta : 1.f / b.eta; int res = rough_table_res(e);
                tab[k] = rough_boundary_integral(b.distribution, b.alpha_u, e, rough_table_mu(k % kRoughTransRes), res == 32 ? n32.data() : n128.data(), res == 32 ? w32.data() : w128.data(), res, tr); }
            double acc = 0; for (int k = 0; k < kRoughTransRes; ++k) acc += (double) tab[kRoughTransRes + k] * rough_table_mu(k);
            bd.rough_t = tab.data(); bd.internal_refl = (float) (acc / kRoughTransRes) * 2.f;
        }
        s->bsdfs.push_back(bd);
    }
    hs_build_bvh(s);
    s->dev.env_emitter = -1;
    for (uint32_t i = 0; i < d->n_emitters; ++i) {
        const mtsb_emitter_desc &e = d->emitters[i]; EmitterDev ed; memset(&ed, 0, sizeof(ed)); ed.type = e.type; ed.radiance_tex = e.radiance_tex;
        if (e.type == MTSB_EMITTER_AREA) {
            ed.radiance = mk3(e.radiance[0], e.radiance[1], e.radiance[2]); ed.mesh = e.mesh_id;
            const MeshDev &m = s->meshes[e.mesh_id];
            if (m.shape >= 0) { ed.normalization = s->shapes[m.shape].inv_area; s->emitters.push_back(ed); continue; }
            AreaTable t = build_area_table(s->pos.data() + 3 * (size_t) m.vert_begin, mesh_faces[e.mesh_id].data(), m.face_count);
            s->tables.push_back(t.cdf); ed.cdf = s->tables.back().data(); ed.cdf_sum = t.sum; ed.normalization = t.normalization; ed.valid_lo = t.valid_lo; ed.valid_hi = t.valid_hi;
        } else if (e.type == MTSB_EMITTER_CONSTANT || e.type == MTSB_EMITTER_POINT) {
            ed.radiance = mk3(e.radiance[0], e.radiance[1], e.radiance[2]); ed.mesh = -1;
            ed.position = mk3(e.position[0], e.position[1], e.position[2]);
            if (e.type == MTSB_EMITTER_CONSTANT) {
                s->dev.env_emitter = (int) i;
                float3 c = (s->lo + s->hi) * 0.5f; float r = s->faces.size() ? norm(s->hi - c) : 0.f;
                ed.bsphere_radius = fmaxf(kRayEpsilon, r * (1.f + kRayEpsilon));
            }
        } else {
            s->dev.env_emitter = (int) i; ed.w = e.width; ed.h = e.height; ed.scale = e.scale;
            memcpy(ed.to_world, e.to_world, sizeof(ed.to_world)); invert4x4(e.to_world, ed.to_local);
            size_t npx = (size_t) e.width * e.height; s->env_rgba.resize(npx);
            for (size_t k = 0; k < npx; ++k) s->env_rgba[k] = s->spectral ? make_float4(e.coeff_scale[4 * k], e.coeff_scale[4 * k + 1], e.coeff_scale[4 * k + 2], e.coeff_scale[4 * k + 3])
                                                             : make_float4(e.rgba[4 * k], e.rgba[4 * k + 1], e.rgba[4 * k + 2], 1.f);
            ed.rgba = s->env_rgba.data();
            std::vector<float> lum = env_importance(e.rgba, e.width, e.height);
            H2DTable h = build_h2d(lum.data(), e.width, e.height);
            ed.n_levels = (int) h.levels.size();
            for (size_t l = 0; l < h.levels.size(); ++l) { s->tables.push_back(h.levels[l]); ed.level_data[l] = s->tables.back().data(); ed.level_width[l] = h.widths[l]; }
            float3 c = (s->lo + s->hi) * 0.5f; float r = s->faces.size() ? norm(s->hi - c) : 0.f;
            ed.bsphere_radius = fmaxf(kRayEpsilon, r * (1.f + kRayEpsilon));
        }
        s->emitters.push_back(ed);
    }
    s->dev.n_emitters = (int) s->emitters.size();
    for (uint32_t i = 0; i < d->n_sensors; ++i) {
        const mtsb_sensor_desc &sd = d->sensors[i]; SensorDev sn; memset(&sn, 0, sizeof(sn));
        memcpy(sn.to_world, sd.to_world, sizeof(sn.to_world)); memcpy(sn.sample_to_camera, sd.sample_to_camera, sizeof(sn.sample_to_camera));
        sn.near_clip = sd.near_clip; sn.far_clip = sd.far_clip; sn.width = sd.film_width; sn.height = sd.film_height; sn.rfilter = sd.rfilter;
        sn.stddev = sd.rfilter_stddev; sn.filter_c = sd.rfilter_c; sn.radius = sd.rfilter == MTSB_RFILTER_BOX ? sd.rfilter_radius + kRayEpsilon : sd.rfilter_radius;
        s->film_xyza.emplace_back((size_t) sn.width * sn.height, make_float4(0, 0, 0, 0)); s->film_w.emplace_back((size_t) sn.width * sn.height, 0.f);
        s->sensors.push_back(sn);
    }
    for (size_t i = 0; i < s->sensors.size(); ++i) { s->sensors[i].film_xyza = s->film_xyza[i].data(); s->sensors[i].film_w = s->film_w[i].data(); }
    s->dev.max_depth = d->integrator.max_depth < 0 ? (1 << 30) : d->integrator.max_depth; s->dev.rr_depth = d->integrator.rr_depth;
    s->dev.pos = s->pos.data(); s->dev.nrm = s->nrm.data(); s->dev.uv = s->uv.data(); s->dev.faces = s->faces.data(); s->dev.shapes = s->shapes.data();
    s->dev.meshes = s->meshes.data(); s->dev.tex = s->tex.data(); s->dev.bsdfs = s->bsdfs.data(); s->dev.emitters = s->emitters.data();
    return s;
}

extern "C" void hs_trav_stats(unsigned long long *out, int reset) { out[0] = g_trav_nodes; out[1] = g_trav_tris; if (reset) g_trav_nodes = g_trav_tris = 0; }
// histogram of the wide nodes: out[0..8] = nodes with k used slots, out[9..33] = nodes with k leaf triangles, out[34] = internal children
extern "C" void hs_node_stats(hs_scene *s, unsigned long long *out) {
    for (int i = 0; i < 35; ++i) out[i] = 0;
    for (uint32_t w = 0; w < s->bvh_nodes; ++w) {
        uint4 n0 = s->nodes[5 * (size_t) w], n1 = s->nodes[5 * (size_t) w + 1];
        uint32_t imask = n0.w >> 24, valid = n1.z;
        int slots = popc32(imask), tris = popc32(valid);
        for (int k = 0; k < 8; ++k) if ((valid >> (3 * k)) & 7u) slots++;
        out[slots]++; out[9 + tris]++; out[34] += (unsigned long long) popc32(imask);
    }
}
extern "C" void hs_bvh_info(hs_scene *s, uint32_t *nodes, uint32_t *depth) { *nodes = s->bvh_nodes; *depth = s->bvh_depth; }

extern "C" void hs_intersect(hs_scene *s, uint64_t n, const float *o, const float *d, float *hits, uint32_t *occ, int any) {
    std::vector<uint2> fast(8);
    for (uint64_t i = 0; i < n; ++i) {
        TravStack<8, 56> st; st.fast = fast.data(); st.stride = 1;
        Hit h;
        if (any) occ[i] = bvh_trace<true>(s->dev.bvh, mk3(o[4 * i], o[4 * i + 1], o[4 * i + 2]), mk3(d[4 * i], d[4 * i + 1], d[4 * i + 2]), o[4 * i + 3], d[4 * i + 3], st, h) ? 1u : 0u;
        else { bvh_trace<false>(s->dev.bvh, mk3(o[4 * i], o[4 * i + 1], o[4 * i + 2]), mk3(d[4 * i], d[4 * i + 1], d[4 * i + 2]), o[4 * i + 3], d[4 * i + 3], st, h);
            hits[4 * i] = h.t; hits[4 * i + 1] = h.u; hits[4 * i + 2] = h.v; hits[4 * i + 3] = u2f(h.prim); }
    }
}

// The wavefront loop of mtsb_render (run_batch in api.cu), stage by stage, with sequential "threads".
struct HsPool {
    std::vector<float4> qo[2], qd[2], qh[2], qt[2], so, sdv, scv, snee, result;
    std::vector<uint4> qm[2], qr[2]; std::vector<float2> qw[2]; std::vector<float4> p1[2], p2[2];
    explicit HsPool(uint32_t pool) : so(pool), sdv(pool), scv(pool), snee(pool), result(pool) {
        for (int k = 0; k < 2; ++k) { qo[k].resize(pool); qd[k].resize(pool); qh[k].resize(pool); qt[k].resize(pool); qm[k].resize(pool); qr[k].resize(pool); qw[k].resize(pool);
                                      p1[k].resize(pool); p2[k].resize(pool); }
    }
    PathQueue queue(int k) { PathQueue q; q.o_mint = qo[k].data(); q.d_maxt = qd[k].data(); q.hit = qh[k].data(); q.thr_eta = qt[k].data(); q.misc = qm[k].data(); q.rng = qr[k].data(); q.wav_eta = qw[k].data();
                             q.prev1 = p1[k].data(); q.prev2 = p2[k].data(); return q; }
};
// k_nee_grad: vertex adjoint of an unoccluded emitter sample (no-op for the primal render)
static void hs_nee(const NoGrad &, const SceneDev &, const PathQueue &, const RenderBatch &, float4, float4) { }
template <typename G> static void hs_nee(const G &grad, const SceneDev &sc, const PathQueue &q, const RenderBatch &b, float4 rec, float4 c) {
    nee_adjoint(sc, grad.g, q, b, rec, c, grad.add);
}
template <typename M, typename GradFn>
static void hs_run_batch(hs_scene *s, const SensorDev &sn, const RenderBatch &b, HsPool &P, uint64_t *cnt, GradFn grad, bool raygen = true,
                         const Edge2Ctx *edge2 = nullptr) {
    std::vector<uint2> fast(8);
    uint32_t n = b.n_paths; int cur = 0, depth0 = 1; uint32_t count = n;
    if (raygen) for (uint32_t i = 0; i < n; ++i) raygen_body<M>(sn, b, P.queue(cur), i);
    for (int depth = depth0; count > 0 && depth <= s->dev.max_depth; ++depth) {
        PathQueue qin = P.queue(cur), qout = P.queue(cur ^ 1);
        for (uint32_t i = 0; i < count; ++i) {
            TravStack<8, 56> st; st.fast = fast.data(); st.stride = 1; Hit h;
            bvh_trace<false>(s->dev.bvh, xyz(qin.o_mint[i]), xyz(qin.d_maxt[i]), qin.o_mint[i].w, qin.d_maxt[i].w, st, h);
            qin.hit[i] = make_float4(h.t, h.u, h.v, u2f(h.prim));
        }
        cnt[0] += count; cnt[2] += count;
        if (edge2 && depth == edge2->depth) {      // run_batch(edge2) in api.cu: the far rays are the next depth's entries, 1:1
            for (uint32_t i = 0; i < count; ++i) edge2_raygen_body(s->dev, *edge2, b, qin, qout, i);
            cur ^= 1;
            continue;
        }
        uint32_t next = 0, nshadow = 0;
        for (uint32_t i = 0; i < count; ++i) {
            ShadeOut out = shade_body<M>(s->dev, qin, b, i, grad);
            if (out.extend) { qout.o_mint[next] = out.o_mint; qout.d_maxt[next] = out.d_maxt; qout.thr_eta[next] = out.thr_eta; qout.misc[next] = out.misc; qout.rng[next] = out.rng; if (M::spectral) qout.wav_eta[next] = out.wav_eta;
                              if (GradFn::kVertex) { qout.prev1[next] = out.prev1; qout.prev2[next] = out.prev2; } next++; }
            if (out.shadow) { P.so[nshadow] = out.s_o_mint; P.sdv[nshadow] = out.s_d_maxt; P.scv[nshadow] = out.s_contrib_id;
                              if (GradFn::kVertex) P.snee[nshadow] = out.s_nee_rec; nshadow++; }
        }
        for (uint32_t i = 0; i < nshadow; ++i) {
            TravStack<8, 56> st; st.fast = fast.data(); st.stride = 1; Hit h;
            if (!bvh_trace<true>(s->dev.bvh, xyz(P.so[i]), xyz(P.sdv[i]), P.so[i].w, P.sdv[i].w, st, h)) {
                uint32_t pid = f2u(P.scv[i].w); P.result[pid].x += P.scv[i].x; P.result[pid].y += P.scv[i].y; P.result[pid].z += P.scv[i].z;
                hs_nee(grad, s->dev, qin, b, P.snee[i], P.scv[i]); }      // k_nee_grad
        }
        cnt[1] += nshadow;
        count = next; cur ^= 1;
    }
}

extern "C" void hs_render(hs_scene *s, uint32_t sensor_id, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total, uint64_t seed,
                          float *xyzaw, uint64_t *counts) {
    SensorDev sn = s->sensors[sensor_id];
    uint64_t total = (uint64_t) sn.width * sn.height * spp_count;
    uint32_t pool = (uint32_t) std::min<uint64_t>(total, 1u << 20);
    HsPool P(pool);
    uint64_t cnt[3] = { 0, 0, 0 };
    for (uint64_t off = 0; off < total; off += pool) {
        uint32_t n = (uint32_t) std::min<uint64_t>(pool, total - off);
        RenderBatch b; b.n_paths = n; b.offset = off; b.spp_begin = spp_begin; b.spp_batch = spp_count; b.spp_total = spp_total; b.seed = seed; b.result = P.result.data();
        if (s->spectral) { if (s->ext) hs_run_batch<SpectralX>(s, sn, b, P, cnt, NoGrad()); else hs_run_batch<Spectral>(s, sn, b, P, cnt, NoGrad()); }
        else { if (s->ext) hs_run_batch<RgbX>(s, sn, b, P, cnt, NoGrad()); else hs_run_batch<Rgb>(s, sn, b, P, cnt, NoGrad()); }
        for (uint32_t i = 0; i < n; ++i) {
            float fx, fy; path_film_pos(sn, b, i, fx, fy);
            film_put(sn, fx, fy, s->spectral ? result_to_xyza<Spectral>(P.result[i]) : result_to_xyza<Rgb>(P.result[i]), 1.f, [&](uint32_t pix, float4 v, float w) {
                xyzaw[5 * (size_t) pix] += v.x; xyzaw[5 * (size_t) pix + 1] += v.y; xyzaw[5 * (size_t) pix + 2] += v.z; xyzaw[5 * (size_t) pix + 3] += v.w; xyzaw[5 * (size_t) pix + 4] += w; });
        }
    }
    if (counts) { counts[0] = total; counts[1] = cnt[0]; counts[2] = cnt[1]; counts[3] = cnt[2]; }
}

// mtsb_render_grad (grad.cuh) on the host: primal pass -> final radiance, replay with the texel / vertex scatter, then
// the boundary (edge-sampling) pass for vertex gradients
struct HostAdd { HD void operator()(float *dst, float v) const { *dst += v; } };
static void invert4(const float *m, double *out) {
    double a[4][8];
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { a[r][c] = m[4 * r + c]; a[r][4 + c] = r == c ? 1.0 : 0.0; }
    for (int c = 0; c < 4; ++c) {
        int piv = c; for (int r = c + 1; r < 4; ++r) if (fabs(a[r][c]) > fabs(a[piv][c])) piv = r;
        for (int k = 0; k < 8; ++k) std::swap(a[c][k], a[piv][k]);
        double inv = 1.0 / a[c][c]; for (int k = 0; k < 8; ++k) a[c][k] *= inv;
        for (int r = 0; r < 4; ++r) if (r != c) { double f = a[r][c]; for (int k = 0; k < 8; ++k) a[r][k] -= f * a[c][k]; }
    }
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out[4 * r + c] = a[r][4 + c];
}
extern "C" void hs_render_grad(hs_scene *s, uint32_t sensor_id, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total, uint64_t seed,
                               const float *dL_dimage, int32_t tex_id, float *grad_texture, int32_t mesh_id, float *grad_vertices,
                               int boundary, float edge_fraction) {
    SensorDev sn = s->sensors[sensor_id];
    uint64_t npix = (uint64_t) sn.width * sn.height, total = npix * spp_count;
    uint32_t pool = (uint32_t) std::min<uint64_t>(total, 1u << 20);
    HsPool P(pool);
    std::vector<float4> fin(pool);
    uint64_t cnt[3] = { 0, 0, 0 };
    GradCtx g; g.final_radiance = fin.data(); g.dL_dimage = dL_dimage; g.grad_tex = grad_texture; g.tex_id = grad_texture ? tex_id : -1; g.scale = 1.f / (float) spp_total;
    g.grad_pos = grad_vertices; g.grad_vnrm = nullptr; g.mesh_id = mesh_id; g.vert_begin = g.vert_count = 0;
    g.cam_o = mk3(sn.to_world[3], sn.to_world[7], sn.to_world[11]);
    if (grad_vertices) { g.vert_begin = s->meshes[mesh_id].vert_begin; g.vert_count = s->meshes[mesh_id].vert_count; }
    std::vector<float> vnrm_adj, vnrm_acc;
    if (grad_vertices && s->meshes[mesh_id].has_normals) { vnrm_adj.assign(3 * (size_t) g.vert_count, 0.f); vnrm_acc.assign(3 * (size_t) g.vert_count, 0.f); g.grad_vnrm = vnrm_adj.data(); }
    // reconstruction filter wider than a pixel: k_film_weights, k_film_weights_invert, k_path_adjoint of kernels.cuh
    std::vector<float> inv_w; std::vector<float4> path_dl;
    const bool wide = sn.radius > 1.f;
    auto fill_path_dl = [&](const RenderBatch &b, const GradCtx &gc) {
        for (uint32_t i = 0; i < b.n_paths; ++i) { float fx, fy; path_film_pos(sn, b, i, fx, fy); path_dl[i] = mk4(film_gather(gc.filt, gc.dL_dimage, gc.inv_w, fx, fy) * gc.scale, 0.f); }
    };
    if (wide) {
        inv_w.assign(npix, 0.f); path_dl.resize(pool);
        RenderBatch wb; wb.n_paths = (uint32_t) (npix * spp_total); wb.offset = 0; wb.spp_begin = 0; wb.spp_batch = spp_total; wb.spp_total = spp_total; wb.seed = seed; wb.result = nullptr;
        for (uint32_t i = 0; i < wb.n_paths; ++i) { float fx, fy; path_film_pos(sn, wb, i, fx, fy);
            film_put(sn, fx, fy, make_float4(0.f, 0.f, 0.f, 0.f), 1.f, [&](uint32_t pix, float4, float w) { inv_w[pix] += w; }); }
        for (float &w : inv_w) w = w != 0.f ? (float) spp_total / w : 0.f;
        g.inv_w = inv_w.data(); g.path_dl = path_dl.data(); g.filt = FilmFilter{ sn.rfilter, sn.radius, sn.stddev, sn.filter_c, sn.width, sn.height };
    }
    for (uint64_t off = 0; off < total; off += pool) {
        uint32_t n = (uint32_t) std::min<uint64_t>(pool, total - off);
        RenderBatch b; b.n_paths = n; b.offset = off; b.spp_begin = spp_begin; b.spp_batch = spp_count; b.spp_total = spp_total; b.seed = seed; b.result = P.result.data();
        hs_run_batch<RgbX>(s, sn, b, P, cnt, NoGrad());
        std::copy(P.result.begin(), P.result.begin() + n, fin.begin());
        if (wide) fill_path_dl(b, g);
        ReplayGrad<HostAdd> rg; rg.g = g;
        hs_run_batch<RgbX>(s, sn, b, P, cnt, rg);
    }
    if (g.grad_vnrm != nullptr) {        // k_vnormal_accum, k_vnormal_adjoint
        const MeshDev &mm = s->meshes[mesh_id];
        for (uint32_t f = 0; f < mm.face_count; ++f) vnormal_accum_body(s->dev, g, mm.face_begin + f, vnrm_acc.data(), HostAdd());
        for (uint32_t f = 0; f < mm.face_count; ++f) vnormal_adjoint_body(s->dev, g, vnrm_acc.data(), mm.face_begin + f, HostAdd());
    }
    if (!grad_vertices || !boundary) return;
    // ---- boundary term: unique edges of the mesh, projected-length cdf, two paths per edge sample
    const MeshDev &m = s->meshes[mesh_id];
    // same construction as build_edge_list() in grad.cuh
    std::vector<std::pair<uint64_t, uint32_t>> keys;
    for (uint32_t f = 0; f < m.face_count; ++f) { uint4 fc = s->faces[m.face_begin + f]; uint32_t v[3] = { fc.x, fc.y, fc.z };
        for (int k = 0; k < 3; ++k) { uint32_t a = v[k], c = v[(k + 1) % 3]; if (a > c) std::swap(a, c); keys.emplace_back(((uint64_t) a << 32) | c, m.face_begin + f); } }
    std::sort(keys.begin(), keys.end());
    std::vector<uint4> edges;
    for (size_t i = 0; i < keys.size();) {
        size_t j = i + 1; while (j < keys.size() && keys[j].first == keys[i].first) ++j;
        edges.push_back(make_uint4((uint32_t) (keys[i].first >> 32), (uint32_t) keys[i].first, keys[i].second, j - i > 1 ? keys[i + 1].second : 0xffffffffu));
        i = j;
    }
    if (m.has_normals && !getenv("HS_NO_TERM")) for (uint32_t f = 0; f < m.face_count; ++f) edges.push_back(make_uint4(m.face_begin + f, kTermTag, m.face_begin + f, m.face_begin + f));   // mesh_edge_list()
    EdgeCtx e; e.edges = edges.data(); e.n_edges = (uint32_t) edges.size(); e.width = sn.width; e.height = sn.height; e.seed = seed; e.eps = 0.004f; e.rim_override = getenv("HS_NO_TERM") ? 0 : 1;
    double c2w_inv[16], s2c_inv[16]; invert4(sn.to_world, c2w_inv); invert4(sn.sample_to_camera, s2c_inv);
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { double a = 0; for (int k = 0; k < 4; ++k) a += s2c_inv[4 * r + k] * c2w_inv[4 * k + c]; e.w2s[4 * r + c] = (float) a; }
    std::vector<float> cdf(edges.size()); double acc = 0;
    for (uint32_t k = 0; k < e.n_edges; ++k) { acc += edge_len_body(s->dev, e, g.cam_o, k); cdf[k] = (float) acc; }
    e.cdf = cdf.data(); e.total_len = (float) acc;
    uint64_t nb_total = std::max<uint64_t>(1, (uint64_t) ((double) npix * spp_total * edge_fraction));
    uint64_t nb_first = (uint64_t) ((double) npix * spp_begin * edge_fraction), nb_end = (uint64_t) ((double) npix * (spp_begin + spp_count) * edge_fraction);
    if (spp_begin + spp_count == spp_total) nb_end = nb_total;
    e.weight = 1.f / (float) nb_total;
    uint32_t half = pool / 2;
    for (uint64_t off = nb_first; off < nb_end; off += half) {
        uint32_t nb = (uint32_t) std::min<uint64_t>(half, nb_end - off);
        e.first = off;
        RenderBatch b; b.n_paths = 2 * nb; b.offset = 0; b.spp_begin = 0; b.spp_batch = 1; b.spp_total = 1; b.seed = seed; b.result = P.result.data();
        for (uint32_t j = 0; j < nb; ++j) edge_raygen_body(s->dev, sn, e, b, P.queue(0), j);
        hs_run_batch<RgbX>(s, sn, b, P, cnt, NoGrad(), false);
        for (uint32_t j = 0; j < nb; ++j) edge_scatter_body(s->dev, e, g, P.result.data(), j, HostAdd());
    }
    if (boundary < 2) return;
    // ---- boundary term seen from the first path vertex (same sequence as mtsb_render_grad)
    // edges of ALL meshes, mesh by mesh (all_edge_list() in grad.cuh): static occluders sweep over moving material too
    std::vector<uint4> all_edges; uint32_t mov_begin = 0, mov_end = 0;
    for (size_t mi = 0; mi < s->meshes.size(); ++mi) {
        const MeshDev &mm = s->meshes[mi];
        if ((int) mi == mesh_id) mov_begin = (uint32_t) all_edges.size();
        std::vector<std::pair<uint64_t, uint32_t>> k2;
        for (uint32_t f = 0; f < mm.face_count; ++f) { uint4 fc = s->faces[mm.face_begin + f]; uint32_t v[3] = { fc.x, fc.y, fc.z };
            for (int k = 0; k < 3; ++k) { uint32_t a = v[k], c = v[(k + 1) % 3]; if (a > c) std::swap(a, c); k2.emplace_back(((uint64_t) a << 32) | c, mm.face_begin + f); } }
        std::sort(k2.begin(), k2.end());
        for (size_t i = 0; i < k2.size();) {
            size_t j = i + 1; while (j < k2.size() && k2[j].first == k2[i].first) ++j;
            all_edges.push_back(make_uint4((uint32_t) (k2[i].first >> 32), (uint32_t) k2[i].first, k2[i].second, j - i > 1 ? k2[i + 1].second : 0xffffffffu));
            i = j;
        }
        if ((int) mi == mesh_id) mov_end = (uint32_t) all_edges.size();
    }
    Edge2Ctx e2; e2.edges = all_edges.data(); e2.n_edges = (uint32_t) all_edges.size(); e2.candidates = 32; e2.eps = 1e-3f; e2.moving_mesh = mesh_id; e2.self_rmin = 1e30f;
    e2.mov_begin = mov_begin; e2.mov_end = mov_end; e2.depth = 1;
    std::vector<float> cdf3(all_edges.size()); acc = 0;
    for (uint32_t k = 0; k < e2.n_edges; ++k) { acc += edge_len3_body(s->dev, e2, k); cdf3[k] = (float) acc; }
    e2.cdf = cdf3.data(); e2.total_len = (float) acc;
    std::vector<float4> aux(5 * (size_t) pool); e2.aux = aux.data();
    const int k_last = boundary >= 3 ? std::min(s->dev.max_depth - 1, 8) : 1;
    for (int k = 1; k <= k_last; ++k) {
        e2.depth = k;
        uint32_t s2_total = (uint32_t) ceil((double) spp_total * edge_fraction / (double) (1u << (k - 1))); if (s2_total < 1) s2_total = 1;
        uint32_t s2_begin = (uint32_t) ((uint64_t) spp_begin * s2_total / spp_total), s2_end = (uint32_t) ((uint64_t) (spp_begin + spp_count) * s2_total / spp_total);
        GradCtx g2 = g; g2.scale = 1.f / (float) s2_total;
        uint64_t total2 = npix * (uint64_t) (s2_end - s2_begin);
        for (uint64_t off = 0; off < total2; off += pool) {
            uint32_t n2 = (uint32_t) std::min<uint64_t>(pool, total2 - off);
            RenderBatch b; b.n_paths = n2; b.offset = off; b.spp_begin = s2_begin; b.spp_batch = s2_end - s2_begin; b.spp_total = s2_total;
            b.seed = (seed ^ 0x51ed270b7f4a7c15ull) + 0x9e3779b97f4a7c15ull * (uint64_t) (k - 1); b.result = P.result.data();
            if (k > 1) std::fill(aux.begin(), aux.begin() + 5 * (size_t) n2, make_float4(0.f, 0.f, 0.f, 0.f));
            hs_run_batch<RgbX>(s, sn, b, P, cnt, NoGrad(), true, &e2);
            if (wide) fill_path_dl(b, g2);
            for (uint32_t j = 0; j < n2; ++j) edge2_scatter_body(s->dev, e2, g2, b, j, HostAdd());
        }
    }
}

extern "C" void hs_bsdf_sample(hs_scene *s, uint32_t id, uint64_t n, const float *wi, const float *uv, const float *s1, const float *s2, float *out9) {
    for (uint64_t i = 0; i < n; ++i) {
        SI si; si.p = mk3(0.f); si.n = mk3(0, 0, 1); si.sh.s = mk3(1, 0, 0); si.sh.t = mk3(0, 1, 0); si.sh.n = mk3(0, 0, 1);
        si.u = uv[2 * i]; si.v = uv[2 * i + 1]; si.wi = ld3(wi, (uint32_t) i); si.t = 1.f; si.prim = 0; si.mesh = 0; si.valid = true;
        BsdfSample bs = bsdf_sample(s->dev, s->bsdfs[id], si, s1[i], s2[2 * i], s2[2 * i + 1]); float *o = out9 + 9 * i;
        o[0] = bs.wo.x; o[1] = bs.wo.y; o[2] = bs.wo.z; o[3] = bs.pdf; o[4] = bs.eta; o[5] = bs.weight.x; o[6] = bs.weight.y; o[7] = bs.weight.z; o[8] = bs.delta ? 1.f : 0.f;
    }
}
extern "C" void hs_sample_emitter_direction(hs_scene *s, uint64_t n, const float *ref_p, const float *s2, float *out11) {
    for (uint64_t i = 0; i < n; ++i) { DirSampleT<RgbX> ds = sample_emitter_direction<RgbX>(s->dev, ld3(ref_p, (uint32_t) i), s2[2 * i], s2[2 * i + 1]); float *o = out11 + 11 * i;
        o[0] = ds.d.x; o[1] = ds.d.y; o[2] = ds.d.z; o[3] = ds.dist; o[4] = ds.pdf; o[5] = ds.value.x; o[6] = ds.value.y; o[7] = ds.value.z; o[8] = ds.n.x; o[9] = ds.n.y; o[10] = ds.n.z; }
}
