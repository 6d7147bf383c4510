// host_render.cpp -- a C++ host driving the C ABI directly (no Python): builds a small Cornell-box-like scene from
// quads, renders it with mtsb_render and writes the developed film as a PFM.  This is the call sequence the adapter
// plugin in INTEGRATION.md makes from Mitsuba's `Integrator::render`.
//
//   g++ -std=c++17 -O2 -Iinclude examples/host_render.cpp -Lmitsuba2_b200 -lmtsb200 -Wl,-rpath,$PWD/mitsuba2_b200 -o host_render
//   ./host_render out.pfm [spp]
#include <mtsb200.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CHECK(x) do { if ((x) != 0) { fprintf(stderr, "%s failed: %s\n", #x, mtsb_last_error()); return 1; } } while (0)

struct Quad { float p[12]; int bsdf; int emitter; };

int main(int argc, char **argv) {
    const char *out = argc > 1 ? argv[1] : "host_render.pfm";
    uint32_t spp = argc > 2 ? (uint32_t) atoi(argv[2]) : 32;
    const uint32_t W = 128, H = 128;
    // textures: 0 white, 1 red, 2 green
    mtsb_texture_desc tex[3]; memset(tex, 0, sizeof(tex));
    const float cols[3][3] = { { .73f, .73f, .73f }, { .65f, .05f, .05f }, { .12f, .45f, .15f } };
    for (int i = 0; i < 3; ++i) { tex[i].type = MTSB_TEX_CONSTANT; memcpy(tex[i].value, cols[i], sizeof(cols[i])); }
    mtsb_bsdf_desc bsdf[3]; memset(bsdf, 0, sizeof(bsdf));
    for (int i = 0; i < 3; ++i) { bsdf[i].type = MTSB_BSDF_DIFFUSE; bsdf[i].reflectance_tex = i; bsdf[i].eta_tex = bsdf[i].k_tex = -1;
        bsdf[i].specular_reflectance_tex = bsdf[i].specular_transmittance_tex = -1; }
    // unit box [-1,1]^3 open towards -z, light quad under the ceiling
    std::vector<Quad> quads = {
        { { -1, -1, -1,  1, -1, -1,  1, -1, 1,  -1, -1, 1 }, 0, -1 },      // floor  (normal +y)
        { { -1, 1, -1,  -1, 1, 1,  1, 1, 1,  1, 1, -1 }, 0, -1 },          // ceiling (normal -y)
        { { -1, -1, 1,  1, -1, 1,  1, 1, 1,  -1, 1, 1 }, 0, -1 },          // back   (normal -z)
        { { -1, -1, -1,  -1, -1, 1,  -1, 1, 1,  -1, 1, -1 }, 1, -1 },      // left   (normal +x)
        { { 1, -1, -1,  1, 1, -1,  1, 1, 1,  1, -1, 1 }, 2, -1 },          // right  (normal -x)
        { { -.3f, .98f, -.3f,  -.3f, .98f, .3f,  .3f, .98f, .3f,  .3f, .98f, -.3f }, 0, 0 },   // light (normal -y)
    };
    static const uint32_t faces[6] = { 0, 2, 1, 0, 3, 2 };      // winding chosen so the normals listed above point into the box
    std::vector<mtsb_mesh_desc> meshes(quads.size());
    for (size_t i = 0; i < quads.size(); ++i) {
        memset(&meshes[i], 0, sizeof(mtsb_mesh_desc));
        meshes[i].faces = faces; meshes[i].positions = quads[i].p; meshes[i].face_count = 2; meshes[i].vertex_count = 4;
        meshes[i].bsdf_id = quads[i].bsdf; meshes[i].emitter_id = quads[i].emitter;
    }
    mtsb_emitter_desc em; memset(&em, 0, sizeof(em));
    em.type = MTSB_EMITTER_AREA; em.radiance[0] = 18.f; em.radiance[1] = 14.f; em.radiance[2] = 8.f; em.mesh_id = 5;
    // perspective camera at (0,0,-3.4) looking down +z, 40 degrees: sample_to_camera as in perspective.cpp:125-133
    mtsb_sensor_desc sn; memset(&sn, 0, sizeof(sn));
    const float to_world[16] = { 1, 0, 0, 0,  0, 1, 0, 0,  0, 0, 1, -3.4f,  0, 0, 0, 1 };       // look_at((0,0,-3.4) -> origin, up +y)
    memcpy(sn.to_world, to_world, sizeof(to_world));
    float tanv = tanf(40.f * 0.5f * 3.14159265f / 180.f), nearc = 0.01f, farc = 100.f, aspect = (float) W / H;
    // inverse of scale(-.5,-.5*aspect,1) * translate(-1,-1/aspect,0) * perspective(fov,near,far)
    const float s2c[16] = { -2 * tanv, 0, 0, tanv,   0, -2 * tanv / aspect, 0, tanv / aspect,   0, 0, 0, 1,
                            0, 0, (nearc - farc) / (farc * nearc), 1 / nearc };
    memcpy(sn.sample_to_camera, s2c, sizeof(s2c));
    sn.near_clip = nearc; sn.far_clip = farc; sn.film_width = W; sn.film_height = H;
    sn.rfilter = MTSB_RFILTER_BOX; sn.rfilter_radius = 0.5f;

    mtsb_scene_desc sd; memset(&sd, 0, sizeof(sd));
    sd.n_meshes = (uint32_t) meshes.size(); sd.meshes = meshes.data();
    sd.n_bsdfs = 3; sd.bsdfs = bsdf; sd.n_textures = 3; sd.textures = tex; sd.n_emitters = 1; sd.emitters = &em;
    sd.n_sensors = 1; sd.sensors = &sn; sd.integrator.max_depth = 6; sd.integrator.rr_depth = 5;

    mtsb_ctx *ctx = nullptr; mtsb_scene *scene = nullptr; mtsb_stats st;
    CHECK(mtsb_ctx_create(0, &ctx));
    CHECK(mtsb_scene_create(ctx, &sd, &scene));
    CHECK(mtsb_film_clear(scene, 0));
    CHECK(mtsb_render(scene, 0, 0, spp, spp, /*seed*/ 1, &st));
    std::vector<float> rgb((size_t) W * H * 3);
    CHECK(mtsb_film_develop(scene, 0, rgb.data(), MTSB_MEM_HOST));
    double mean = 0; for (float v : rgb) mean += v; mean /= (double) rgb.size();
    printf("rendered %ux%u x %u spp in %.3f ms (%.1f Msamples/s), %llu launches, mean %.4f\n", W, H, spp, st.ms_total,
           (double) st.samples / (st.ms_total * 1e3), (unsigned long long) st.kernel_launches, mean);
    FILE *f = fopen(out, "wb");
    if (!f) { perror(out); return 1; }
    fprintf(f, "PF\n%u %u\n-1.0\n", W, H);
    for (int y = (int) H - 1; y >= 0; --y) fwrite(rgb.data() + (size_t) y * W * 3, sizeof(float), (size_t) W * 3, f);
    fclose(f);
    mtsb_scene_destroy(scene); mtsb_ctx_destroy(ctx);
    return mean > 0.05 ? 0 : 2;
}
