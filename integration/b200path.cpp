// b200path.cpp -- Mitsuba 2 integrator plugin `b200path`: drop-in for `path` (src/integrators/path.cpp) that hands the
// already-loaded scene to libmtsb200.so through its C ABI (include/mtsb200.h) and delivers the result through
// Film::put(const ImageBlock*) (include/mitsuba/render/film.h:29), so `hdrfilm` develops / writes the image unchanged.
//
//     <integrator type="b200path"> <integer name="max_depth" value="6"/> </integrator>
//     mitsuba -m scalar_rgb scene.xml
//
// Compiled against the reference headers, outside the product library (which never includes a Mitsuba header):
// integration/Makefile builds it into <reference build>/dist/plugins/b200path.so.  It only uses public reference
// interfaces: Integrator::render / cancel (include/mitsuba/render/integrator.h:42,51), Scene::shapes / emitters / bbox
// (render/scene.h), Mesh buffers (render/mesh.h:43-65,119-122), Object::traverse + TraversalCallback (core/object.h:75,263-279)
// for plugin parameters, Class::name (core/class.h) for plugin kinds, Sensor / Film accessors, ImageBlock::data.
// Not exposed by traverse() in this snapshot and therefore read from the plugins' to_string(): the micro-facet
// distribution type and `sample_visible` of roughconductor (src/bsdfs/roughconductor.cpp:409-427).
#include <mitsuba/core/properties.h>
#include <mitsuba/core/rfilter.h>
#include <mitsuba/core/transform.h>
#include <mitsuba/render/bsdf.h>
#include <mitsuba/render/emitter.h>
#include <mitsuba/render/film.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/render/integrator.h>
#include <mitsuba/render/mesh.h>
#include <mitsuba/render/sampler.h>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/sensor.h>
#include <mitsuba/render/texture.h>
#include <map>
#include <typeinfo>
#include <vector>
#include "../include/mtsb200.h"

NAMESPACE_BEGIN(mitsuba)

namespace b200 {
/// What Object::traverse() reports about one plugin instance: typed parameter pointers and child objects, by name.
struct Params : TraversalCallback {
    struct P { const std::type_info *type; void *ptr; };
    std::map<std::string, P> params;
    std::map<std::string, Object *> objects;
    explicit Params(Object *o) { o->traverse(this); }
    void put_object(const std::string &name, Object *obj) override { objects[name] = obj; }
    void put_parameter_impl(const std::string &name, const std::type_info &type, void *ptr) override { params[name] = P{ &type, ptr }; }
    template <typename T> T *get(const std::string &name) const {
        auto it = params.find(name);
        return (it != params.end() && *it->second.type == typeid(T)) ? (T *) it->second.ptr : nullptr;
    }
    Object *object(const std::string &name) const { auto it = objects.find(name); return it == objects.end() ? nullptr : it->second; }
};
/// value of "key = <token>" in a plugin's to_string()
inline std::string printed(const Object *o, const std::string &key) {
    std::string s = o->to_string(), k = key + " = ";
    size_t i = s.find(k);
    if (i == std::string::npos) return "";
    size_t b = i + k.size(), e = s.find_first_of(",\n", b);
    return s.substr(b, e == std::string::npos ? std::string::npos : e - b);
}
}  // namespace b200

template <typename Float, typename Spectrum>
class B200PathIntegrator final : public Integrator<Float, Spectrum> {
public:
    MTS_IMPORT_BASE(Integrator)
    MTS_IMPORT_TYPES(Scene, Sensor, Film, ImageBlock, Mesh, Shape, BSDF, Emitter, Texture, Sampler, ReconstructionFilter, ProjectiveCamera)

    B200PathIntegrator(const Properties &props) : Base(props) {
        // MonteCarloIntegrator (src/librender/integrator.cpp:285-298)
        m_max_depth = props.int_("max_depth", -1);
        m_rr_depth = props.int_("rr_depth", 5);
        m_device = props.int_("device", 0);
        m_seed = (uint64_t) props.size_("seed", 0);
        if (m_max_depth < 0 && m_max_depth != -1) Throw("\"max_depth\" must be set to -1 (infinite) or a value >= 0");
        if (m_rr_depth <= 0) Throw("\"rr_depth\" must be set to a value greater than zero!");
    }
    ~B200PathIntegrator() {
        if (m_scene) mtsb_scene_destroy(m_scene);
        if (m_ctx) mtsb_ctx_destroy(m_ctx);
    }

    bool render(Scene *scene, Sensor *sensor) override {
        if constexpr (is_spectral_v<Spectrum> || is_array_v<Float> || !std::is_same_v<scalar_t<Float>, float>)
            Throw("b200path: this adapter flattens scalar_rgb scenes (float, rgb); load the scene with -m scalar_rgb");
        else {
            if (!m_ctx && mtsb_ctx_create(m_device, &m_ctx)) Throw("b200path: %s", mtsb_last_error());
            upload(scene, sensor);
            ref<Film> film = sensor->film();
            ScalarVector2i size = film->crop_size();
            uint32_t spp = (uint32_t) sensor->sampler()->sample_count();
            film->prepare({ "X", "Y", "Z", "A", "W" });                       // integrator.cpp:72-74
            if (mtsb_film_clear(m_scene, 0)) Throw("b200path: %s", mtsb_last_error());
            mtsb_stats st;
            int rc = mtsb_render(m_scene, 0, 0, spp, spp, m_seed, &st);
            if (rc == MTSB_E_CANCELLED) return false;                          // Integrator::cancel() arrived (integrator.h:51)
            if (rc) Throw("b200path: %s", mtsb_last_error());
            ref<ImageBlock> block = new ImageBlock(size, 5);                   // no filter: border 0, the device film is already filtered
            block->set_offset(film->crop_offset());
            if (mtsb_film_read(m_scene, 0, block->data().data(), MTSB_MEM_HOST)) Throw("b200path: %s", mtsb_last_error());
            film->put(block);                                                   // Film::put(const ImageBlock *), film.h:29
            Log(Info, "b200path: %u samples/pixel, %.1f ms on the device (%.1f Msamples/s), %u BVH8 nodes",
                spp, st.ms_total, st.samples / (st.ms_total * 1e-3) / 1e6, m_bvh_nodes);
            return true;
        }
        return false;
    }

    void cancel() override { if (m_ctx) mtsb_cancel(m_ctx); }

    std::string to_string() const override {
        return tfm::format("B200PathIntegrator[\n  max_depth = %i,\n  rr_depth = %i,\n  device = %i\n]", m_max_depth, m_rr_depth, m_device);
    }

    MTS_DECLARE_CLASS()
private:
    // ---- scene flattening (what Scene::Scene collected, src/librender/scene.cpp:22-99 -> mtsb_scene_desc) --------------
    struct Flat {
        std::vector<mtsb_mesh_desc> meshes; std::vector<mtsb_bsdf_desc> bsdfs; std::vector<mtsb_texture_desc> textures;
        std::vector<mtsb_emitter_desc> emitters; std::map<const Object *, int> tex_id, bsdf_id;
        std::vector<mtsb_shape_desc> shapes;
        std::vector<std::vector<float>> keep;
    };
    // Shape::m_to_world is protected (render/shape.h:419); a never-instantiated subclass reads it
    struct ShapeAccess : Shape {
        static const Transform<Point<float, 4>> &to_world(const Shape *s) { return static_cast<const ShapeAccess *>(s)->m_to_world; }
    };
    static void identity9(float *m) { for (int i = 0; i < 9; ++i) m[i] = (i % 4 == 0) ? 1.f : 0.f; }

    int texture(Flat &f, Object *o) {
        if (!o) return -1;
        auto it = f.tex_id.find(o);
        if (it != f.tex_id.end()) return it->second;
        mtsb_texture_desc t; memset(&t, 0, sizeof(t)); identity9(t.to_uv);
        b200::Params p(o);
        std::string cls = o->class_()->name();
        if (cls == "SRGBReflectanceSpectrum" || cls == "SRGBEmitterSpectrum") {           // src/spectra/srgb.cpp, srgb_d65.cpp
            auto *v = p.get<Color<float, 3>>("value");
            if (!v) Throw("b200path: %s without an rgb value", cls);
            t.type = MTSB_TEX_CONSTANT; for (int c = 0; c < 3; ++c) t.value[c] = (*v)[c];
        } else if (cls == "UniformSpectrum") {                                              // src/spectra/uniform.cpp
            auto *v = p.get<float>("value");
            if (!v) Throw("b200path: uniform spectrum without a value");
            t.type = MTSB_TEX_CONSTANT; t.value[0] = t.value[1] = t.value[2] = *v;
        } else if (cls.rfind("BitmapTextureImpl", 0) == 0) {                                // src/textures/bitmap.cpp:295-299
            auto *data = p.get<DynamicBuffer<float>>("data");
            auto *res = p.get<Vector<uint32_t, 2>>("resolution");
            auto *tr = p.get<Transform<Point<float, 3>>>("transform");
            if (!data || !res) Throw("b200path: bitmap texture does not expose data / resolution");
            t.type = MTSB_TEX_BITMAP; t.width = res->x(); t.height = res->y();
            t.channels = (uint32_t) (data->size() / ((size_t) res->x() * res->y()));
            t.data = data->data();
            if (tr) for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) t.to_uv[3 * r + c] = tr->matrix(r, c);
        } else
            Throw("b200path: texture plugin %s is not supported by the backend", cls);
        f.textures.push_back(t);
        return f.tex_id[o] = (int) f.textures.size() - 1;
    }
    float constant_of(Flat &f, Object *o, float fallback) {
        if (!o) return fallback;
        int id = texture(f, o);
        if (f.textures[id].type != MTSB_TEX_CONSTANT) Throw("b200path: textured roughness is not supported");
        return f.textures[id].value[0];
    }

    int bsdf(Flat &f, Object *o) {
        auto it = f.bsdf_id.find(o);
        if (it != f.bsdf_id.end()) return it->second;
        mtsb_bsdf_desc b; memset(&b, 0, sizeof(b));
        b.reflectance_tex = b.eta_tex = b.k_tex = b.specular_reflectance_tex = b.specular_transmittance_tex = -1;
        b.alpha_u = b.alpha_v = 0.1f; b.sample_visible = 1; b.eta = 1.f; b.specular_sampling_weight = 0.5f;
        std::string cls = o->class_()->name();
        b200::Params p(o);
        if (cls == "TwoSidedBRDF") {                                        // src/bsdfs/twosided.cpp:183-186
            Object *b0 = p.object("brdf_0"), *b1 = p.object("brdf_1");
            if (!b0 || (b1 && b1 != b0)) Throw("b200path: twosided with two different BRDFs is not supported");
            int inner = bsdf(f, b0);
            b = f.bsdfs[inner]; b.two_sided = 1;
        } else if (cls == "SmoothDiffuse") {                                // src/bsdfs/diffuse.cpp:137-139
            b.type = MTSB_BSDF_DIFFUSE; b.reflectance_tex = texture(f, p.object("reflectance"));
        } else if (cls == "RoughConductor" || cls == "SmoothConductor") {   // roughconductor.cpp:396-407, conductor.cpp
            b.type = cls == "RoughConductor" ? MTSB_BSDF_ROUGHCONDUCTOR : MTSB_BSDF_CONDUCTOR;
            b.eta_tex = texture(f, p.object("eta")); b.k_tex = texture(f, p.object("k"));
            b.specular_reflectance_tex = texture(f, p.object("specular_reflectance"));
            if (cls == "RoughConductor") {
                if (p.object("alpha")) b.alpha_u = b.alpha_v = constant_of(f, p.object("alpha"), 0.1f);
                else { b.alpha_u = constant_of(f, p.object("alpha_u"), 0.1f); b.alpha_v = constant_of(f, p.object("alpha_v"), 0.1f); }
                b.distribution = b200::printed(o, "distribution") == "ggx" ? MTSB_MF_GGX : MTSB_MF_BECKMANN;
                b.sample_visible = b200::printed(o, "sample_visible") != "0";
            }
        } else if (cls == "SmoothDielectric") {                             // src/bsdfs/dielectric.cpp:328-334
            auto *eta = p.get<float>("eta");
            if (!eta) Throw("b200path: dielectric does not expose eta");
            b.type = MTSB_BSDF_DIELECTRIC; b.eta = *eta;
            b.specular_reflectance_tex = texture(f, p.object("specular_reflectance"));
            b.specular_transmittance_tex = texture(f, p.object("specular_transmittance"));
        } else
            Throw("b200path: BSDF plugin %s is not supported by the backend", cls);
        f.bsdfs.push_back(b);
        return f.bsdf_id[o] = (int) f.bsdfs.size() - 1;
    }

    static void matrix16(const Matrix<float, 4> &m, float *out) { for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out[4 * r + c] = m(r, c); }

    void upload(Scene *scene, Sensor *sensor) {
        if (m_scene) { mtsb_scene_destroy(m_scene); m_scene = nullptr; }
        Flat f;
        std::map<const Object *, int> emitter_index;
        for (size_t i = 0; i < scene->emitters().size(); ++i) emitter_index[scene->emitters()[i].get()] = (int) i;
        f.emitters.resize(scene->emitters().size());
        // shapes in Scene::shapes() order: the global primitive index is the prefix sum over them (kdtree.h:2382-2400)
        std::vector<std::pair<int, int>> shape_emitters;      // (emitter index, analytic shape index): mesh_id is fixed up below
        for (auto &sh : scene->shapes()) {
            Mesh *mesh = dynamic_cast<Mesh *>(sh.get());
            if (!mesh) {
                // analytic shapes (src/shapes/{sphere,disk,cylinder,rectangle}.cpp) become BVH leaf primitives of the backend.
                // The backend lists them after the meshes; only the order of primitive ids differs, not the image.
                std::string cls = sh->class_()->name();
                mtsb_shape_desc d; memset(&d, 0, sizeof(d));
                d.type = cls == "Sphere" ? MTSB_SHAPE_SPHERE : cls == "Disk" ? MTSB_SHAPE_DISK : cls == "Cylinder" ? MTSB_SHAPE_CYLINDER
                       : cls == "Rectangle" ? MTSB_SHAPE_RECTANGLE : 0;
                if (!d.type) Throw("b200path: shape plugin %s is not supported by the backend", cls);
                matrix16(ShapeAccess::to_world(sh.get()).matrix, d.to_world);       // the plugin's state after update()
                // flip_normals is private to Sphere / Cylinder: read it off a sampled position (disk / rectangle fold it into to_world)
                if (d.type == MTSB_SHAPE_SPHERE) {
                    auto ps = sh->sample_position(0.f, Point<float, 2>(0.3f, 0.3f));
                    d.flip_normals = dot(ps.n, ps.p - sh->bbox().center()) < 0.f;
                } else if (d.type == MTSB_SHAPE_CYLINDER) {
                    d.flip_normals = sh->sample_position(0.f, Point<float, 2>(0.5f, 0.f)).n.x() < 0.f;     // cylinder.cpp:222-227: object-space normal
                    // after update() the cylinder's to_world is rigid and radius / length are private (cylinder.cpp:116-123): measure
                    // them on sampled positions (sample_position: (r cos 2 pi v, r sin 2 pi v, u length)) and fold them back in
                    auto p00 = sh->sample_position(0.f, Point<float, 2>(0.f, 0.f)).p, p05 = sh->sample_position(0.f, Point<float, 2>(0.f, 0.5f)).p,
                         p10 = sh->sample_position(0.f, Point<float, 2>(1.f, 0.f)).p;
                    float radius = 0.5f * norm(p00 - p05), length = norm(p10 - p00);
                    for (int r = 0; r < 3; ++r) { d.to_world[4 * r] *= radius; d.to_world[4 * r + 1] *= radius; d.to_world[4 * r + 2] *= length; }
                }
                d.bsdf_id = bsdf(f, const_cast<BSDF *>(sh->bsdf()));
                d.emitter_id = -1;
                if (sh->is_emitter()) {
                    Emitter *em = const_cast<Emitter *>(sh->emitter());
                    int ei = emitter_index.at(em);
                    mtsb_emitter_desc e; memset(&e, 0, sizeof(e));
                    e.type = MTSB_EMITTER_AREA; e.radiance_tex = -1;
                    int t = texture(f, b200::Params(em).object("radiance"));
                    if (t < 0 || f.textures[t].type != MTSB_TEX_CONSTANT) Throw("b200path: textured area lights are not supported");
                    memcpy(e.radiance, f.textures[t].value, sizeof(e.radiance));
                    f.emitters[ei] = e; d.emitter_id = ei;
                    shape_emitters.emplace_back(ei, (int) f.shapes.size());
                }
                f.shapes.push_back(d);
                continue;
            }
            mtsb_mesh_desc m; memset(&m, 0, sizeof(m));
            m.faces = (const uint32_t *) mesh->faces_buffer().data();        // mesh.h:63-65
            m.positions = mesh->vertex_positions_buffer().data();            // mesh.h:48-50 (world space: loaders bake to_world)
            m.normals = mesh->has_vertex_normals() ? mesh->vertex_normals_buffer().data() : nullptr;
            m.texcoords = mesh->has_vertex_texcoords() ? mesh->vertex_texcoords_buffer().data() : nullptr;
            m.face_count = (uint32_t) mesh->face_count(); m.vertex_count = (uint32_t) mesh->vertex_count();
            m.bsdf_id = bsdf(f, const_cast<BSDF *>(sh->bsdf()));
            m.emitter_id = -1;
            if (sh->is_emitter()) {
                Emitter *em = const_cast<Emitter *>(sh->emitter());
                int ei = emitter_index.at(em);
                mtsb_emitter_desc e; memset(&e, 0, sizeof(e));
                e.type = MTSB_EMITTER_AREA; e.mesh_id = (int) f.meshes.size(); e.radiance_tex = -1;
                int t = texture(f, b200::Params(em).object("radiance"));     // src/emitters/area.cpp:129-131
                if (t < 0 || f.textures[t].type != MTSB_TEX_CONSTANT) Throw("b200path: textured area lights are not supported");
                memcpy(e.radiance, f.textures[t].value, sizeof(e.radiance));
                f.emitters[ei] = e; m.emitter_id = ei;
            }
            f.meshes.push_back(m);
        }
        for (auto &se : shape_emitters) f.emitters[se.first].mesh_id = (int) f.meshes.size() + se.second;     // mtsb200.h: n_meshes + shape index
        // the remaining emitters, in Scene::emitters() order (scene.cpp:23-48; decides `index` in sample_emitter_direction)
        for (size_t i = 0; i < scene->emitters().size(); ++i) {
            Emitter *em = scene->emitters()[i].get();
            std::string cls = em->class_()->name();
            if (cls == "AreaLight") continue;
            mtsb_emitter_desc e; memset(&e, 0, sizeof(e)); e.mesh_id = -1; e.radiance_tex = -1;
            b200::Params p(em);
            if (cls == "EnvironmentMapEmitter") {                            // src/emitters/envmap.cpp:214-218
                auto *data = p.get<DynamicBuffer<float>>("data"); auto *res = p.get<Vector<uint32_t, 2>>("resolution"); auto *scale = p.get<float>("scale");
                if (!data || !res || !scale) Throw("b200path: envmap does not expose data / resolution / scale");
                e.type = MTSB_EMITTER_ENVMAP; e.width = res->x(); e.height = res->y(); e.rgba = data->data(); e.scale = *scale;
                matrix16(em->world_transform()->eval(0.f).matrix, e.to_world);
            } else if (cls == "ConstantBackgroundEmitter" || cls == "PointLight") {
                int t = texture(f, p.object(cls == "PointLight" ? "intensity" : "radiance"));
                if (t < 0 || f.textures[t].type != MTSB_TEX_CONSTANT) Throw("b200path: %s needs a constant colour", cls);
                e.type = cls == "PointLight" ? MTSB_EMITTER_POINT : MTSB_EMITTER_CONSTANT;
                memcpy(e.radiance, f.textures[t].value, sizeof(e.radiance));
                if (cls == "PointLight") { auto pos = em->world_transform()->eval(0.f) * Point<float, 3>(0.f); for (int c = 0; c < 3; ++c) e.position[c] = pos[c]; }
            } else
                Throw("b200path: emitter plugin %s is not supported by the backend", cls);
            f.emitters[i] = e;
        }
        mtsb_sensor_desc sd = sensor_desc(sensor);
        mtsb_scene_desc d; memset(&d, 0, sizeof(d));
        d.n_meshes = (uint32_t) f.meshes.size(); d.n_bsdfs = (uint32_t) f.bsdfs.size(); d.n_textures = (uint32_t) f.textures.size();
        d.n_emitters = (uint32_t) f.emitters.size(); d.n_sensors = 1;
        d.meshes = f.meshes.data(); d.bsdfs = f.bsdfs.data(); d.textures = f.textures.data(); d.emitters = f.emitters.data(); d.sensors = &sd;
        d.n_shapes = (uint32_t) f.shapes.size(); d.shapes = f.shapes.data();
        d.integrator.max_depth = m_max_depth; d.integrator.rr_depth = m_rr_depth; d.mode = MTSB_MODE_RGB;
        if (mtsb_scene_create(m_ctx, &d, &m_scene)) Throw("b200path: %s", mtsb_last_error());
        mtsb_stats st; mtsb_scene_stats(m_scene, &st); m_bvh_nodes = st.bvh_nodes;
    }

    // ---- sensor + film (src/sensors/perspective.cpp:106-222, src/librender/film.cpp:16-64, rfilters) ---------------------
    static void mul4(const double *a, const double *b, double *out) {
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { double s = 0; for (int k = 0; k < 4; ++k) s += a[4 * r + k] * b[4 * k + c]; out[4 * r + c] = s; }
    }
    static float number_after(const std::string &s, const std::string &key, float fallback) {
        size_t i = s.find(key);
        return i == std::string::npos ? fallback : (float) atof(s.c_str() + i + key.size());
    }
    mtsb_sensor_desc sensor_desc(Sensor *sensor) {
        if (std::string(sensor->class_()->name()) != "PerspectiveCamera") Throw("b200path: sensor plugin %s is not supported", sensor->class_()->name());
        auto *cam = dynamic_cast<ProjectiveCamera *>(sensor);
        Film *film = sensor->film();
        mtsb_sensor_desc sd; memset(&sd, 0, sizeof(sd));
        auto to_world = sensor->world_transform()->eval(0.f);
        matrix16(to_world.matrix, sd.to_world);
        sd.near_clip = cam->near_clip(); sd.far_clip = cam->far_clip();
        ScalarVector2i size = film->size(), crop = film->crop_size(); ScalarPoint2i off = film->crop_offset();
        sd.film_width = (uint32_t) crop.x(); sd.film_height = (uint32_t) crop.y();
        double rsx = (double) crop.x() / size.x(), rsy = (double) crop.y() / size.y(), rox = (double) off.x() / size.x(), roy = (double) off.y() / size.y();
        // x_fov is not exposed (perspective.cpp:237-240 "TODO x_fov"): recover tan(fov_x / 2) from two rays of the sensor itself.
        // camera space: d.x / d.z = tan * (1 - 2 x_rel), x_rel = crop offset + sample * crop size relative to the full film
        auto inv = to_world.inverse();
        double q[2];
        for (int k = 0; k < 2; ++k) {
            auto [ray, w] = sensor->sample_ray(0.f, 0.5f, Point<float, 2>((float) k, 0.5f), Point<float, 2>(0.5f));
            auto d = inv * ray.d;
            q[k] = (double) d.x() / (double) d.z();
        }
        double tan_half = (q[0] - q[1]) / (2.0 * rsx), aspect = (double) size.x() / size.y(), n = sd.near_clip, fr = sd.far_clip;
        // inverse of scale(1/rel_size) translate(-rel_offset) scale(-.5,-.5 aspect,1) translate(-1,-1/aspect,0) perspective(fov,near,far)
        double ip[16] = { tan_half, 0, 0, 0, 0, tan_half, 0, 0, 0, 0, 0, 1, 0, 0, (n - fr) / (fr * n), 1.0 / n };
        double it[16] = { 1, 0, 0, 1, 0, 1, 0, 1.0 / aspect, 0, 0, 1, 0, 0, 0, 0, 1 };
        double is[16] = { -2, 0, 0, 0, 0, -2.0 / aspect, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
        double ic[16] = { rsx, 0, 0, rox, 0, rsy, 0, roy, 0, 0, 1, 0, 0, 0, 0, 1 };
        double a[16], b[16], c[16];
        mul4(ip, it, a); mul4(a, is, b); mul4(b, ic, c);
        for (int k = 0; k < 16; ++k) sd.sample_to_camera[k] = (float) c[k];
        const ReconstructionFilter *rf = film->reconstruction_filter();
        std::string fc = rf->class_()->name(), fs = rf->to_string();
        sd.rfilter_radius = rf->radius(); sd.rfilter_c = 1.f / 3.f;
        if (fc == "BoxFilter") { sd.rfilter = MTSB_RFILTER_BOX; sd.rfilter_radius = 0.5f; }      // the library adds RayEpsilon itself (box.cpp:30)
        else if (fc == "TentFilter") sd.rfilter = MTSB_RFILTER_TENT;
        else if (fc == "GaussianFilter") { sd.rfilter = MTSB_RFILTER_GAUSSIAN; sd.rfilter_stddev = rf->radius() * 0.25f; }      // gaussian.cpp:31-32
        else if (fc == "MitchellNetravaliFilter") { sd.rfilter = MTSB_RFILTER_MITCHELL; sd.rfilter_stddev = number_after(fs, "B=", 1.f / 3.f); sd.rfilter_c = number_after(fs, "C=", 1.f / 3.f); }
        else if (fc == "CatmullRomFilter") sd.rfilter = MTSB_RFILTER_CATMULLROM;
        else if (fc == "LanczosSincFilter") sd.rfilter = MTSB_RFILTER_LANCZOS;
        else Throw("b200path: reconstruction filter %s is not supported", fc);
        return sd;
    }

    int m_max_depth, m_rr_depth, m_device; uint64_t m_seed;
    mtsb_ctx *m_ctx = nullptr; mtsb_scene *m_scene = nullptr; uint32_t m_bvh_nodes = 0;
};

MTS_IMPLEMENT_CLASS_VARIANT(B200PathIntegrator, Integrator)
MTS_EXPORT_PLUGIN(B200PathIntegrator, "B200 path integrator (libmtsb200 adapter)");
NAMESPACE_END(mitsuba)
