// ref_query.cpp -- TEST INFRASTRUCTURE.  A small driver linked against the UNMODIFIED reference libraries built by
// oracle/build_reference.sh (libmitsuba-core / libmitsuba-render, scalar_rgb).  It loads a Mitsuba XML scene with the
// reference's own loader and answers batched queries through the reference's own virtual interfaces, so that the C
// oracle (oracle/mts_oracle.c) and the CUDA path can be compared with the real thing on identical inputs:
//   intersect : Scene::ray_intersect  (src/librender/scene.cpp:108-116 -> kd-tree, kdtree.h:2080-2175)
//   naive     : Scene::ray_intersect_naive (scene.cpp:118-129)
//   test      : Scene::ray_test       (scene.cpp:131-139)
//   bsdf      : BSDF::sample / eval / pdf of the BSDF of shape k (include/mitsuba/render/bsdf.h)
//   emitter   : Scene::sample_emitter_direction without the visibility test (scene.cpp:141-176)
//   sensor    : Sensor::sample_ray (perspective.cpp:151-188)
//   render    : Integrator::render repeated on the loaded scene (the timed CPU baseline of bench.py)
// Usage: ref_query <scene.xml> <cmd> <in.bin> <out.bin> [threads] [arg]
// Binary files are raw little-endian float32 / uint32 arrays; layouts are documented per command below and mirrored in
// tests/helpers.py (class RefBuild).  Timing of the intersect commands (TBB parallel_for over all rays, grain 4096) is
// printed as "query_seconds <s>" so it can serve as the CPU intersect baseline.
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/logger.h>
#include <mitsuba/core/thread.h>
#include <mitsuba/core/xml.h>
#include <mitsuba/core/bitmap.h>
#include <mitsuba/core/jit.h>
#include <mitsuba/core/profiler.h>
#include <mitsuba/core/util.h>
#include <mitsuba/core/filesystem.h>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/integrator.h>
#include <mitsuba/render/bsdf.h>
#include <mitsuba/render/emitter.h>
#include <mitsuba/render/sensor.h>
#include <mitsuba/render/shape.h>
#include <mitsuba/render/mesh.h>
#include <mitsuba/render/records.h>
#include <mitsuba/render/interaction.h>
#include <tbb/tbb.h>
#include <unistd.h>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <vector>

using namespace mitsuba;

static std::vector<float> read_floats(const char *path) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
    size_t bytes = (size_t) f.tellg(); f.seekg(0);
    std::vector<float> v(bytes / 4);
    f.read((char *) v.data(), (std::streamsize) (v.size() * 4));
    return v;
}
static void write_floats(const char *path, const std::vector<float> &v) {
    std::ofstream f(path, std::ios::binary);
    f.write((const char *) v.data(), (std::streamsize) (v.size() * 4));
}
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

template <typename Float, typename Spectrum>
static int run(int argc, char **argv, int threads, int arg) {
    MTS_IMPORT_TYPES(Scene, Shape, Mesh, BSDF, Emitter, Sensor)
    (void) argc; (void) threads;

    int rc = 0;
    try {
        ref<Thread> thread = Thread::thread();
        ref<FileResolver> fr = thread->file_resolver();
        fs::path base = util::library_path().parent_path();
        if (!fr->contains(base)) fr->append(base);
        fs::path filename(argv[1]);
        if (!fr->contains(filename.parent_path())) fr->append(filename.parent_path());
        ref<Object> parsed = xml::load_file(filename, "scalar_rgb");
        Scene *scene = dynamic_cast<Scene *>(parsed.get());
        if (!scene) Throw("root element is not a scene");
        std::string cmd = argv[2];
        std::vector<float> in = read_floats(argv[3]), out;
        // global primitive index = prefix sum of primitive counts over Scene::shapes() (kdtree.h:2382-2400)
        std::vector<uint32_t> prefix; uint32_t acc = 0;
        for (auto &s : scene->shapes()) { prefix.push_back(acc); acc += (uint32_t) s->primitive_count(); }
        auto shape_index = [&](const Shape *sh) -> int {
            for (size_t i = 0; i < scene->shapes().size(); ++i) if (scene->shapes()[i].get() == sh) return (int) i;
            return -1;
        };
        auto t0 = std::chrono::steady_clock::now();
        if (cmd == "intersect" || cmd == "naive") {
            // in: n x 8 = (o.xyz, mint, d.xyz, maxt); out: n x 16 = (t, uv.xy, bitcast global prim, p.xyz, n.xyz, sh_n.xyz, wi.xyz)
            size_t n = in.size() / 8; out.assign(n * 16, 0.f);
            bool naive = cmd == "naive";
            tbb::parallel_for(tbb::blocked_range<size_t>(0, n, 4096), [&](const tbb::blocked_range<size_t> &r) {
                for (size_t i = r.begin(); i < r.end(); ++i) {
                    const float *q = &in[8 * i];
                    Ray3f ray(Point3f(q[0], q[1], q[2]), Vector3f(q[4], q[5], q[6]), q[3], q[7], 0.f, Wavelength());
                    SurfaceInteraction3f si = naive ? scene->ray_intersect_naive(ray) : scene->ray_intersect(ray);
                    float *o = &out[16 * i];
                    if (si.is_valid()) {
                        int k = shape_index(si.shape);
                        o[0] = si.t; o[1] = si.uv.x(); o[2] = si.uv.y(); o[3] = u2f(prefix[k] + (uint32_t) si.prim_index);
                        for (int c = 0; c < 3; ++c) { o[4 + c] = si.p[c]; o[7 + c] = si.n[c]; o[10 + c] = si.sh_frame.n[c]; o[13 + c] = si.wi[c]; }
                    } else { o[0] = std::numeric_limits<float>::infinity(); o[3] = u2f(0xffffffffu); }
                }
            });
        } else if (cmd == "test") {
            size_t n = in.size() / 8; out.assign(n, 0.f);
            tbb::parallel_for(tbb::blocked_range<size_t>(0, n, 4096), [&](const tbb::blocked_range<size_t> &r) {
                for (size_t i = r.begin(); i < r.end(); ++i) {
                    const float *q = &in[8 * i];
                    Ray3f ray(Point3f(q[0], q[1], q[2]), Vector3f(q[4], q[5], q[6]), q[3], q[7], 0.f, Wavelength());
                    out[i] = scene->ray_test(ray) ? 1.f : 0.f;
                }
            });
        } else if (cmd == "bsdf") {
            // arg = shape index whose BSDF is queried.  in: n x 11 = (wi.xyz, uv.xy, sample1, sample2.xy, wo_eval.xyz)
            // out: n x 13 = sample: (wo.xyz, pdf, eta, weight.rgb, delta flag) ; eval(wo_eval).rgb ; pdf(wo_eval)
            const BSDF *bsdf = scene->shapes()[(size_t) arg]->bsdf();
            size_t n = in.size() / 11; out.assign(n * 13, 0.f);
            BSDFContext ctx;
            for (size_t i = 0; i < n; ++i) {
                const float *q = &in[11 * i];
                SurfaceInteraction3f si;
                si.t = 1.f; si.p = Point3f(0.f); si.n = Normal3f(0.f, 0.f, 1.f); si.sh_frame = Frame3f(Normal3f(0.f, 0.f, 1.f));
                si.sh_frame.s = Vector3f(1.f, 0.f, 0.f); si.sh_frame.t = Vector3f(0.f, 1.f, 0.f);
                si.uv = Point2f(q[3], q[4]); si.wi = Vector3f(q[0], q[1], q[2]); si.wavelengths = Wavelength();
                si.dp_du = Vector3f(1.f, 0.f, 0.f); si.dp_dv = Vector3f(0.f, 1.f, 0.f); si.duv_dx = si.duv_dy = Vector2f(0.f);
                si.time = 0.f; si.prim_index = 0;
                auto [bs, w] = bsdf->sample(ctx, si, q[5], Point2f(q[6], q[7]));
                float *o = &out[13 * i];
                o[0] = bs.wo.x(); o[1] = bs.wo.y(); o[2] = bs.wo.z(); o[3] = bs.pdf; o[4] = bs.eta; o[5] = w[0]; o[6] = w[1]; o[7] = w[2];
                o[8] = has_flag(bs.sampled_type, BSDFFlags::Delta) ? 1.f : 0.f;
                Vector3f wo(q[8], q[9], q[10]);
                Spectrum ev = bsdf->eval(ctx, si, wo);
                o[9] = ev[0]; o[10] = ev[1]; o[11] = ev[2]; o[12] = bsdf->pdf(ctx, si, wo);
            }
        } else if (cmd == "emitter") {
            // in: n x 5 = (ref.xyz, sample.xy); out: n x 11 = (d.xyz, dist, pdf, value.rgb, n.xyz)
            size_t n = in.size() / 5; out.assign(n * 11, 0.f);
            for (size_t i = 0; i < n; ++i) {
                const float *q = &in[5 * i];
                Interaction3f it; it.t = 0.f; it.time = 0.f; it.p = Point3f(q[0], q[1], q[2]); it.wavelengths = Wavelength();
                auto [ds, val] = scene->sample_emitter_direction(it, Point2f(q[3], q[4]), false);
                float *o = &out[11 * i];
                o[0] = ds.d.x(); o[1] = ds.d.y(); o[2] = ds.d.z(); o[3] = ds.dist; o[4] = ds.pdf; o[5] = val[0]; o[6] = val[1]; o[7] = val[2];
                o[8] = ds.n.x(); o[9] = ds.n.y(); o[10] = ds.n.z();
            }
        } else if (cmd == "sensor") {
            // in: n x 2 = position sample in [0,1]^2; out: n x 8 = (o.xyz, mint, d.xyz, maxt)
            size_t n = in.size() / 2; out.assign(n * 8, 0.f);
            const Sensor *sensor = scene->sensors()[(size_t) arg].get();
            for (size_t i = 0; i < n; ++i) {
                auto [ray, w] = sensor->sample_ray(0.f, 0.5f, Point2f(in[2 * i], in[2 * i + 1]), Point2f(0.5f));
                float *o = &out[8 * i];
                o[0] = ray.o.x(); o[1] = ray.o.y(); o[2] = ray.o.z(); o[3] = ray.mint; o[4] = ray.d.x(); o[5] = ray.d.y(); o[6] = ray.d.z(); o[7] = ray.maxt;
            }
        } else if (cmd == "render") {
            // arg = number of repetitions of Integrator::render(scene, sensor 0) (src/librender/integrator.cpp:52-176) on the
            // loaded scene; prints one "render_seconds" line per repetition (scene loading and kd-tree build excluded, like
            // the CLI's "Rendering finished. (took ...)" line)
            Sensor *sensor = scene->sensors()[0].get();
            for (int r = 0; r < (arg < 1 ? 1 : arg); ++r) {
                auto r0 = std::chrono::steady_clock::now();
                if (!scene->integrator()->render(scene, sensor)) Throw("render failed");
                printf("render_seconds %.6f\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - r0).count());
                fflush(stdout);
            }
        } else if (cmd == "info") {
            printf("shapes %zu emitters %zu primitives %u\n", scene->shapes().size(), scene->emitters().size(), acc);
            for (size_t i = 0; i < scene->shapes().size(); ++i) printf("shape %zu prims %zu first %u\n", i, (size_t) scene->shapes()[i]->primitive_count(), prefix[i]);
        } else Throw("unknown command %s", cmd);
        double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        printf("query_seconds %.6f\n", dt);
        write_floats(argv[4], out);
    } catch (const std::exception &e) {
        fprintf(stderr, "ref_query: %s\n", e.what());
        rc = 1;
    }
    return rc;
}

int main(int argc, char **argv) {
    if (argc < 5) { fprintf(stderr, "usage: ref_query scene.xml cmd in.bin out.bin [threads] [arg]\n"); return 2; }
    Jit::static_initialization();
    Class::static_initialization();
    Thread::static_initialization();
    Logger::static_initialization();
    Bitmap::static_initialization();
    Profiler::static_initialization();
    librender_nop();
    Thread::thread()->logger()->set_log_level(Warn);
    int threads = argc > 5 ? atoi(argv[5]) : 1;
    int arg = argc > 6 ? atoi(argv[6]) : 0;
    __global_thread_count = (size_t) (threads < 1 ? 1 : threads);
    tbb::task_scheduler_init init((int) __global_thread_count);
    int rc = run<float, Color<float, 3>>(argc, argv, threads, arg);
    // results are on disk; skip the static shutdown sequence (it can dead-lock against idle TBB workers) and leave
    fflush(stdout); fflush(stderr);
    _exit(rc);
}
