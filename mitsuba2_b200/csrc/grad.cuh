// grad.cuh -- adjoint pass entry point (included at the end of api.cu).
#pragma once
extern "C" int mtsb_render_grad(mtsb_scene *s, uint32_t sensor_id, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total,
                                uint64_t seed, const float *dL_dimage, int32_t tex_id, float *grad_texture, int32_t mesh_id,
                                float *grad_vertices, int mem, mtsb_stats *stats) {
    (void) s; (void) sensor_id; (void) spp_begin; (void) spp_count; (void) spp_total; (void) seed; (void) dL_dimage;
    (void) tex_id; (void) grad_texture; (void) mesh_id; (void) grad_vertices; (void) mem; (void) stats;
    return fail(MTSB_E_INVALID, "mtsb_render_grad: not implemented yet");
}
