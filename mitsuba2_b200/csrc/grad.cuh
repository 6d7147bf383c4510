// grad.cuh -- adjoint pass entry point (included at the end of api.cu).
//
// mtsb_render_grad = the backward pass of mtsb_render for one bitmap texture (texture_adjoint() in path.h) and / or the
// vertex positions of one mesh (adjoint.h).
// Per batch: (1) primal wavefront without the splat -> per-path final radiance; (2) replay with identical RNG streams,
// k_shade_grad scatters dL/dtexel and the interior term of dL/dvertex with float atomics.  Vertex gradients then add the
// primary-visibility boundary term: edges sampled by projected length, two paths per edge sample (k_edge_*).  Memory is O(pool), independent of spp and path length, unlike the
// reference's tape (src/python/python/autodiff.py:6-194, ext/enoki/src/autodiff/autodiff.cpp:838-988).
#pragma once



// row-major 4x4 inverse (double), for world -> sample space of a sensor
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
// unique edges of a mesh (scene-global vertex indices + the primitive ids of the two adjacent faces); host sort once per
// mesh: the list is topology, not positions
static void build_edge_list(const uint32_t *faces, size_t n_faces, uint32_t vert_begin, uint32_t face_begin, std::vector<uint4> &out) {
    std::vector<std::pair<uint64_t, uint32_t>> keys; keys.reserve(3 * n_faces);
    for (size_t t = 0; t < n_faces; ++t)
        for (int k = 0; k < 3; ++k) {
            uint32_t a = faces[3 * t + k] + vert_begin, c = faces[3 * t + (k + 1) % 3] + vert_begin;
            if (a > c) std::swap(a, c);
            keys.emplace_back(((uint64_t) a << 32) | c, face_begin + (uint32_t) t);
        }
    std::sort(keys.begin(), keys.end());
    out.clear();
    for (size_t i = 0; i < keys.size();) {
        size_t j = i + 1; while (j < keys.size() && keys[j].first == keys[i].first) ++j;
        out.push_back(make_uint4((uint32_t) (keys[i].first >> 32), (uint32_t) keys[i].first, keys[i].second, j - i > 1 ? keys[i + 1].second : 0xffffffffu));
        i = j;
    }
}
static int mesh_edge_list(mtsb_scene *s, int mesh_id, const uint4 **edges, uint32_t *n_edges) {
    auto it = s->mesh_edges.find(mesh_id);
    if (it == s->mesh_edges.end()) {
        const MeshDev &m = s->meshes[mesh_id];
        std::vector<uint4> e;
        build_edge_list(s->mesh_faces[mesh_id].data(), m.face_count, m.vert_begin, m.face_begin, e);
        // a smooth-shaded mesh also jumps along the shading-normal terminator inside its faces (adjoint.h, kTermTag): one
        // candidate entry per face, given a length only where the curve exists from the current viewpoint
        if (m.has_normals) for (uint32_t f = 0; f < m.face_count; ++f) e.push_back(make_uint4(m.face_begin + f, kTermTag, m.face_begin + f, m.face_begin + f));
        std::unique_ptr<DevBuf<uint4>> buf(new DevBuf<uint4>());
        CU(buf->upload(e.data(), e.size(), s->ctx->stream));
        CU(cudaStreamSynchronize(s->ctx->stream));
        it = s->mesh_edges.emplace(mesh_id, std::move(buf)).first;
    }
    *edges = it->second->p; *n_edges = (uint32_t) it->second->n;
    return MTSB_OK;
}

// unique edges of ALL meshes (key -1 of the cache).  The boundary term seen from x1 needs them: with x2 parameterised on its
// own triangle, a STATIC occluder's silhouette also sweeps over the material behind it when x1 or x2 sit on the moving mesh
// (the shadow a fixed box casts onto the moving object; the part of the moving object seen past a fixed edge).
static int all_edge_list(mtsb_scene *s, const uint4 **edges, uint32_t *n_edges) {
    auto it = s->mesh_edges.find(-1);
    if (it == s->mesh_edges.end()) {
        std::vector<uint4> all, e;
        s->edge_offsets.assign(1, 0u);
        for (size_t k = 0; k < s->meshes.size(); ++k) {
            const MeshDev &m = s->meshes[k];
            build_edge_list(s->mesh_faces[k].data(), m.face_count, m.vert_begin, m.face_begin, e);
            all.insert(all.end(), e.begin(), e.end());
            s->edge_offsets.push_back((uint32_t) all.size());      // mesh k's edges = [edge_offsets[k], edge_offsets[k + 1])
        }
        std::unique_ptr<DevBuf<uint4>> buf(new DevBuf<uint4>());
        CU(buf->upload(all.data(), all.size(), s->ctx->stream));
        CU(cudaStreamSynchronize(s->ctx->stream));
        it = s->mesh_edges.emplace(-1, std::move(buf)).first;
    }
    *edges = it->second->p; *n_edges = (uint32_t) it->second->n;
    return MTSB_OK;
}

/* boundary: 0 = interior term only, 1 = + primary-visibility edge sampling, 2 = + edge sampling seen from the first path
 * vertex (shadow / occlusion boundaries); edge_fraction = edge samples per path sample */
extern "C" int mtsb_grad_config(mtsb_scene *s, int boundary, float edge_fraction) {
    CHECK_ARG(s && edge_fraction >= 0.f && edge_fraction <= 64.f);
    s->grad_boundary = boundary < 0 ? 0 : boundary; s->grad_edge_fraction = edge_fraction;
    return MTSB_OK;
}

extern "C" int mtsb_render_grad(mtsb_scene *s, uint32_t sensor_id, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total,
                                uint64_t seed, const float *dL_dimage, int32_t tex_id, float *grad_texture, int32_t mesh_id,
                                float *grad_vertices, int mem, mtsb_stats *stats) {
    CHECK_ARG(s && sensor_id < s->sensors.size() && spp_count > 0 && spp_begin + spp_count <= spp_total && dL_dimage);
    CHECK_ARG(grad_texture || grad_vertices);
    if (s->spectral) return fail(MTSB_E_INVALID, "mtsb_render_grad: the adjoint pass is implemented for MTSB_MODE_RGB only");
    if (!s->shapes.empty()) return fail(MTSB_E_INVALID, "mtsb_render_grad: scenes with analytic shapes are not supported (triangle meshes only)");
    size_t ntex = 0, nvert = 0;
    if (grad_texture) {
        CHECK_ARG(tex_id >= 0 && (size_t) tex_id < s->tex.size());
        const TexDev td = s->tex[tex_id];
        if (td.type != 1) return fail(MTSB_E_INVALID, "mtsb_render_grad: texture is not a bitmap");
        ntex = (size_t) td.w * td.h * td.ch;
    }
    if (grad_vertices) {
        CHECK_ARG(mesh_id >= 0 && (size_t) mesh_id < s->n_tri_meshes);
        if (s->meshes[mesh_id].emitter >= 0 && s->meshes[mesh_id].has_normals)      // the sampler's pdf uses the interpolated normal there
            return fail(MTSB_E_INVALID, "mtsb_render_grad: vertex gradients of an emitter mesh with vertex normals are not supported");
        nvert = 3 * (size_t) s->meshes[mesh_id].vert_count;
    }
    CU(cudaSetDevice(s->ctx->device));
    cudaStream_t st = s->ctx->stream;
    SensorDev sn = s->sensors[sensor_id]->dev;
    const bool wide = sn.radius > 1.f;           // else: single-pixel splat of weight 1 (imageblock.cpp:162-169)
    uint64_t npix = (uint64_t) sn.width * sn.height, total = npix * spp_count;
    uint32_t pool = (uint32_t) (total < pool_limit() ? total : pool_limit());
    if (grad_vertices && pool < 2) pool = 2;
    RC(ensure_pool(s, pool));
    CU(s->final_radiance.alloc(s->pool));
    if (grad_vertices) { for (int k = 0; k < 2; ++k) { CU(s->q_p1[k].alloc(s->pool)); CU(s->q_p2[k].alloc(s->pool)); }
                         CU(s->s_nee.alloc(s->pool)); CU(s->s_occ.alloc(s->pool)); }
    Staging sg; const float *d_dl;
    RC(sg.in(dL_dimage, 3 * npix, mem, st, &d_dl));
    DevBuf<float> gbuf, vbuf;
    float *d_g = nullptr, *d_v = nullptr;
    if (grad_texture) {
        if (mem == MTSB_MEM_DEVICE) d_g = grad_texture;
        else { CU(gbuf.alloc(ntex)); d_g = gbuf.p; CU(cudaMemsetAsync(d_g, 0, ntex * sizeof(float), st)); }
    }
    if (grad_vertices) {
        if (mem == MTSB_MEM_DEVICE) d_v = grad_vertices;
        else { CU(vbuf.alloc(nvert)); d_v = vbuf.p; CU(cudaMemsetAsync(d_v, 0, nvert * sizeof(float), st)); }
    }
    GradCtx g; g.final_radiance = s->final_radiance.p; g.dL_dimage = d_dl; g.grad_tex = d_g; g.tex_id = grad_texture ? tex_id : -1;
    g.scale = 1.f / (float) spp_total;
    g.grad_pos = d_v; g.grad_vnrm = nullptr; g.mesh_id = grad_vertices ? mesh_id : -1; g.vert_begin = g.vert_count = 0;
    DevBuf<float> vnrm_adj, vnrm_acc;        // smooth-shaded mesh: adjoint of its vertex normals, and the unnormalised normals
    if (grad_vertices && s->meshes[mesh_id].has_normals) {
        CU(vnrm_adj.alloc(nvert)); CU(vnrm_acc.alloc(nvert));
        CU(cudaMemsetAsync(vnrm_adj.p, 0, nvert * sizeof(float), st)); CU(cudaMemsetAsync(vnrm_acc.p, 0, nvert * sizeof(float), st));
        g.grad_vnrm = vnrm_adj.p;
    }
    if (grad_vertices) { g.vert_begin = s->meshes[mesh_id].vert_begin; g.vert_count = s->meshes[mesh_id].vert_count; }
    g.cam_o = mk3(sn.to_world[3], sn.to_world[7], sn.to_world[11]);
    DevBuf<float> inv_w;
    if (wide) {
        // the developed image divides by the film's weight channel: accumulate it over ALL spp_total samples of the job
        // (every rank of a sharded job derives the same positions from the sampler streams), invert, gather per path
        CU(inv_w.alloc(npix)); CU(cudaMemsetAsync(inv_w.p, 0, npix * sizeof(float), st));
        uint64_t all = npix * spp_total;
        for (uint64_t off = 0; off < all; off += 1u << 30) {
            RenderBatch b; b.n_paths = (uint32_t) ((all - off) < (1u << 30) ? (all - off) : (1u << 30)); b.offset = off;
            b.spp_begin = 0; b.spp_batch = spp_total; b.spp_total = spp_total; b.seed = seed; b.result = nullptr;
            k_film_weights<<<grid_for(b.n_paths), kBlock, 0, st>>>(sn, b, inv_w.p);
        }
        k_film_weights_invert<<<grid_for(npix), kBlock, 0, st>>>(inv_w.p, (uint32_t) npix, (float) spp_total);
        CU(s->path_dl.alloc(s->pool));
        g.inv_w = inv_w.p; g.path_dl = s->path_dl.p;
        g.filt = FilmFilter{ sn.rfilter, sn.radius, sn.stddev, sn.filter_c, sn.width, sn.height };
    }
    CU(cudaMemsetAsync(s->dstats.p, 0, 4 * sizeof(unsigned long long), st));
    CU(cudaEventRecord(s->ctx->ev0, st));
    uint64_t launches = 0;
    s->ctx->cancel = 0;      // a cancel that arrived before this call does not abort it (integrator.cpp:54)
    for (uint64_t off = 0; off < total; off += pool) {
        if (s->ctx->cancel) { s->ctx->cancel = 0; cudaStreamSynchronize(st); return fail(MTSB_E_CANCELLED, "render cancelled"); }
        uint32_t n = (uint32_t) ((total - off) < pool ? (total - off) : pool);
        RenderBatch b; b.n_paths = n; b.offset = off; b.spp_begin = spp_begin; b.spp_batch = spp_count; b.spp_total = spp_total;
        b.seed = seed; b.result = s->result.p;
        RC(run_batch(s, sn, b, nullptr, false, launches));
        CU(cudaMemcpyAsync(s->final_radiance.p, s->result.p, (size_t) n * sizeof(float4), cudaMemcpyDeviceToDevice, st));
        if (wide) { k_path_adjoint<<<grid_for(n), kBlock, 0, st>>>(sn, b, g, s->path_dl.p); launches++; }
        RC(run_batch(s, sn, b, &g, false, launches));
    }
    if (g.grad_vnrm != nullptr) {        // vertex normals -> positions (Mesh::recompute_vertex_normals), once per call
        const MeshDev &m = s->meshes[mesh_id];
        k_vnormal_accum<<<grid_for(m.face_count), kBlock, 0, st>>>(s->dev, g, m.face_begin, m.face_count, vnrm_acc.p);
        k_vnormal_adjoint<<<grid_for(m.face_count), kBlock, 0, st>>>(s->dev, g, m.face_begin, m.face_count, vnrm_acc.p);
        launches += 2;
    }
    uint64_t edge_samples = 0;
    if (grad_vertices && s->grad_boundary && s->grad_edge_fraction > 0.f) {
        // ---- boundary term: sum over edge samples of dL/dI (L+ - L-) d(alpha)/dv (adjoint.h)
        EdgeLaunch el; EdgeCtx &e = el.e;
        RC(mesh_edge_list(s, mesh_id, &e.edges, &e.n_edges));
        if (e.n_edges > 0) {
            e.width = sn.width; e.height = sn.height; e.seed = seed; e.eps = 0.004f; e.total_len = 0.f; e.rim_override = 1;
            double c2w_inv[16], s2c_inv[16]; invert4(sn.to_world, c2w_inv); invert4(sn.sample_to_camera, s2c_inv);
            for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { double a = 0; for (int k = 0; k < 4; ++k) a += s2c_inv[4 * r + k] * c2w_inv[4 * k + c]; e.w2s[4 * r + c] = (float) a; }
            DevBuf<float> len, cdf; DevBuf<uint8_t> tmp; size_t tmp_bytes = 0;
            CU(len.alloc(e.n_edges)); CU(cdf.alloc(e.n_edges));
            k_edge_len<<<grid_for(e.n_edges), kBlock, 0, st>>>(s->dev, e, g.cam_o, len.p); launches++;
            CU(cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, len.p, cdf.p, (int) e.n_edges, st));
            CU(tmp.alloc(tmp_bytes));
            CU(cub::DeviceScan::InclusiveSum(tmp.p, tmp_bytes, len.p, cdf.p, (int) e.n_edges, st)); launches++;
            e.cdf = cdf.p; el.total_len = cdf.p + (e.n_edges - 1);
            double frac = s->grad_edge_fraction;
            uint64_t nb_total = (uint64_t) ((double) npix * spp_total * frac); if (nb_total < 1) nb_total = 1;
            uint64_t nb_first = (uint64_t) ((double) npix * spp_begin * frac), nb_end = (uint64_t) ((double) npix * (spp_begin + spp_count) * frac);
            if (spp_begin + spp_count == spp_total) nb_end = nb_total;
            e.weight = 1.f / (float) nb_total;
            uint32_t half = s->pool / 2;
            for (uint64_t off = nb_first; off < nb_end; off += half) {
                if (s->ctx->cancel) { s->ctx->cancel = 0; cudaStreamSynchronize(st); return fail(MTSB_E_CANCELLED, "render cancelled"); }
                uint32_t nb = (uint32_t) ((nb_end - off) < half ? (nb_end - off) : half);
                e.first = off;
                RenderBatch b; b.n_paths = 2 * nb; b.offset = 0; b.spp_begin = 0; b.spp_batch = 1; b.spp_total = 1; b.seed = seed; b.result = s->result.p;
                RC(run_batch(s, sn, b, nullptr, false, launches, &el));
                k_edge_scatter<<<grid_for(nb), kBlock, 0, st>>>(s->dev, e, el.total_len, g, s->result.p, nb); launches++;
                edge_samples += nb;
            }
            // ---- boundary term seen from the first path vertex (moving shadow / occlusion boundaries)
            if (s->grad_boundary >= 2) {
                Edge2Launch l2; Edge2Ctx &e2 = l2.e;
                RC(all_edge_list(s, &e2.edges, &e2.n_edges));
                e2.candidates = 32; e2.eps = 1e-3f; e2.total_len = 0.f; e2.moving_mesh = mesh_id; e2.self_rmin = 1e30f;      // all of them (3 and 8 edge lengths were measured: more noise, same bias); candidates 8 / 16 / 32 at C4: 69.0 / 70.6 / 76.0 ms per pass, variance x time of the three fields lowest at 32
                e2.mov_begin = s->edge_offsets[mesh_id]; e2.mov_end = s->edge_offsets[mesh_id + 1]; e2.depth = 1;
                DevBuf<float> len3, cdf3; DevBuf<unsigned char> tmp2;
                CU(len3.alloc(e2.n_edges)); CU(cdf3.alloc(e2.n_edges));
                k_edge_len3<<<grid_for(e2.n_edges), kBlock, 0, st>>>(s->dev, e2, len3.p); launches++;
                size_t tmp2_bytes = 0;
                CU(cub::DeviceScan::InclusiveSum(nullptr, tmp2_bytes, len3.p, cdf3.p, (int) e2.n_edges, st)); CU(tmp2.alloc(tmp2_bytes));
                CU(cub::DeviceScan::InclusiveSum(tmp2.p, tmp2_bytes, len3.p, cdf3.p, (int) e2.n_edges, st)); launches++;
                e2.cdf = cdf3.p; l2.total_len = cdf3.p + (e2.n_edges - 1);
                CU(s->edge_aux.alloc(5 * (size_t) s->pool)); e2.aux = s->edge_aux.p;
                // the vertex the edges are seen from: the camera's hit (boundary 2), or every vertex that can still be followed by
                // another one (boundary 3), each depth with its own samples -- half as many as the one before
                const int k_last = s->grad_boundary >= 3 ? (s->dev.max_depth - 1 < 8 ? s->dev.max_depth - 1 : 8) : 1;
                for (int k = 1; k <= k_last; ++k) {
                    e2.depth = k;
                    // sample counts per pixel follow the path samples: the same (begin, count, total) split, scaled
                    uint32_t s2_total = (uint32_t) ceil((double) spp_total * frac / (double) (1u << (k - 1))); if (s2_total < 1) s2_total = 1;
                    uint32_t s2_begin = (uint32_t) ((uint64_t) spp_begin * s2_total / spp_total), s2_end = (uint32_t) ((uint64_t) (spp_begin + spp_count) * s2_total / spp_total);
                    GradCtx g2 = g; g2.scale = 1.f / (float) s2_total;
                    uint64_t total2 = npix * (uint64_t) (s2_end - s2_begin);
                    for (uint64_t off = 0; off < total2; off += s->pool) {
                        uint32_t n2 = (uint32_t) ((total2 - off) < s->pool ? (total2 - off) : s->pool);
                        RenderBatch b; b.n_paths = n2; b.offset = off; b.spp_begin = s2_begin; b.spp_batch = s2_end - s2_begin; b.spp_total = s2_total;
                        b.seed = (seed ^ 0x51ed270b7f4a7c15ull) + 0x9e3779b97f4a7c15ull * (uint64_t) (k - 1); b.result = s->result.p;
                        if (k > 1) CU(cudaMemsetAsync(e2.aux, 0, 5 * (size_t) n2 * sizeof(float4), st));      // paths that end before vertex k carry no sample
                        RC(run_batch(s, sn, b, nullptr, false, launches, nullptr, &l2));
                        if (wide) { k_path_adjoint<<<grid_for(n2), kBlock, 0, st>>>(sn, b, g2, s->path_dl.p); launches++; }
                        k_edge2_scatter<<<grid_for(n2), kBlock, 0, st>>>(s->dev, e2, g2, b); launches++;
                        edge_samples += n2 / 2;
                    }
                }
                CU(cudaStreamSynchronize(st));
            }
            CU(cudaStreamSynchronize(st));       // len / cdf / tmp are released at scope exit
        }
    }
    CU(cudaEventRecord(s->ctx->ev1, st));
    unsigned long long hs[4];
    CU(cudaMemcpyAsync(hs, s->dstats.p, sizeof(hs), cudaMemcpyDeviceToHost, st));
    if (mem != MTSB_MEM_DEVICE) {
        std::vector<float> h(ntex > nvert ? ntex : nvert);
        if (grad_texture) {
            CU(cudaMemcpyAsync(h.data(), d_g, ntex * sizeof(float), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            for (size_t i = 0; i < ntex; ++i) grad_texture[i] += h[i];
        }
        if (grad_vertices) {
            CU(cudaMemcpyAsync(h.data(), d_v, nvert * sizeof(float), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            for (size_t i = 0; i < nvert; ++i) grad_vertices[i] += h[i];
        }
    }
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->samples = total + 2 * edge_samples; stats->closest_rays = hs[0]; stats->shadow_rays = hs[1]; stats->segments = hs[2];
        stats->kernel_launches = launches;
        CU(cudaEventElapsedTime(&stats->ms_total, s->ctx->ev0, s->ctx->ev1));
    }
    collect_stage_times(s->ctx, stats);
    return MTSB_OK;
}
