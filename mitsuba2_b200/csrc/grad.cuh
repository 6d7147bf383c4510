// grad.cuh -- adjoint pass entry point (included at the end of api.cu).
//
// mtsb_render_grad = the backward pass of mtsb_render for one bitmap texture (see texture_adjoint() in path.h).
// Per batch: (1) primal wavefront without the splat -> per-path final radiance; (2) replay with identical RNG streams,
// k_shade_grad scatters dL/dtexel with float atomics.  Memory is O(pool), independent of spp and path length, unlike the
// reference's tape (src/python/python/autodiff.py:6-194, ext/enoki/src/autodiff/autodiff.cpp:838-988).
#pragma once

__global__ void k_axpy(size_t n, const float *__restrict__ x, float *__restrict__ y) {
    size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x;
    if (i < n) y[i] += x[i];
}

extern "C" int mtsb_render_grad(mtsb_scene *s, uint32_t sensor_id, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total,
                                uint64_t seed, const float *dL_dimage, int32_t tex_id, float *grad_texture, int32_t mesh_id,
                                float *grad_vertices, int mem, mtsb_stats *stats) {
    CHECK_ARG(s && sensor_id < s->sensors.size() && spp_count > 0 && spp_begin + spp_count <= spp_total && dL_dimage);
    if (grad_vertices || mesh_id >= 0)
        return fail(MTSB_E_INVALID, "mtsb_render_grad: vertex-position gradients are not implemented (texture gradients only)");
    CHECK_ARG(grad_texture && tex_id >= 0 && (size_t) tex_id < s->tex.size());
    const TexDev td = s->tex[tex_id];
    if (td.type != 1) return fail(MTSB_E_INVALID, "mtsb_render_grad: texture is not a bitmap");
    if (s->spectral) return fail(MTSB_E_INVALID, "mtsb_render_grad: the adjoint pass is implemented for MTSB_MODE_RGB only");
    CU(cudaSetDevice(s->ctx->device));
    cudaStream_t st = s->ctx->stream;
    SensorDev sn = s->sensors[sensor_id]->dev;
    if (sn.rfilter != 0 || sn.radius > 1.f)      // single-pixel splat of weight 1 (imageblock.cpp:162-169)
        return fail(MTSB_E_INVALID, "mtsb_render_grad: only the box reconstruction filter is supported");
    uint64_t npix = (uint64_t) sn.width * sn.height, total = npix * spp_count;
    if (npix * spp_total > 0xffffffffull) return fail(MTSB_E_INVALID, "pixels * spp_total exceeds the 32-bit sample-stream index");
    uint32_t pool = (uint32_t) (total < pool_limit() ? total : pool_limit());
    RC(ensure_pool(s, pool));
    CU(s->final_radiance.alloc(s->pool));
    size_t ntex = (size_t) td.w * td.h * td.ch;
    Staging sg; const float *d_dl; float *d_g = nullptr;
    RC(sg.in(dL_dimage, 3 * npix, mem, st, &d_dl));
    DevBuf<float> gbuf;
    if (mem == MTSB_MEM_DEVICE) d_g = grad_texture;
    else { CU(gbuf.alloc(ntex)); d_g = gbuf.p; CU(cudaMemsetAsync(d_g, 0, ntex * sizeof(float), st)); }
    GradCtx g; g.final_radiance = s->final_radiance.p; g.dL_dimage = d_dl; g.grad_tex = d_g; g.tex_id = tex_id;
    g.scale = 1.f / (float) spp_total;
    CU(cudaMemsetAsync(s->dstats.p, 0, 4 * sizeof(unsigned long long), st));
    CU(cudaEventRecord(s->ctx->ev0, st));
    uint64_t launches = 0;
    for (uint64_t off = 0; off < total; off += pool) {
        if (s->ctx->cancel) { s->ctx->cancel = 0; cudaStreamSynchronize(st); return fail(MTSB_E_CANCELLED, "render cancelled"); }
        uint32_t n = (uint32_t) ((total - off) < pool ? (total - off) : pool);
        RenderBatch b; b.n_paths = n; b.offset = off; b.spp_begin = spp_begin; b.spp_batch = spp_count; b.spp_total = spp_total;
        b.seed = seed; b.result = s->result.p;
        RC(run_batch(s, sn, b, nullptr, false, launches));
        CU(cudaMemcpyAsync(s->final_radiance.p, s->result.p, (size_t) n * sizeof(float4), cudaMemcpyDeviceToDevice, st));
        RC(run_batch(s, sn, b, &g, false, launches));
    }
    CU(cudaEventRecord(s->ctx->ev1, st));
    unsigned long long hs[4];
    CU(cudaMemcpyAsync(hs, s->dstats.p, sizeof(hs), cudaMemcpyDeviceToHost, st));
    if (mem != MTSB_MEM_DEVICE) {
        std::vector<float> h(ntex);
        CU(cudaMemcpyAsync(h.data(), d_g, ntex * sizeof(float), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        for (size_t i = 0; i < ntex; ++i) grad_texture[i] += h[i];
    }
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->samples = total; stats->closest_rays = hs[0]; stats->shadow_rays = hs[1]; stats->segments = hs[2];
        stats->kernel_launches = launches;
        CU(cudaEventElapsedTime(&stats->ms_total, s->ctx->ev0, s->ctx->ev1));
    }
    collect_stage_times(s->ctx, stats);
    return MTSB_OK;
}
