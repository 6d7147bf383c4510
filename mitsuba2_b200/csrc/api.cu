// api.cu -- the C ABI (include/mtsb200.h): scene snapshot upload, GPU BVH build, batched intersection,
// wavefront render loop, film export.  No torch types, no Mitsuba headers.
#include "../../include/mtsb200.h"
#include "kernels.cuh"
#include "host_tables.h"
#include "cie_tables.h"
#include <cub/cub.cuh>
#include <string>
#include <vector>
#include <map>
#include <algorithm>
#include <cmath>
#include <memory>
#include <cstring>
#include <cstdio>
#include <chrono>

using namespace mtsb;

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(MTSB_E_CUDA, std::string(#call ": ") + cudaGetErrorString(e_)); } while (0)
#define CHECK_ARG(c) do { if (!(c)) return fail(MTSB_E_INVALID, "invalid argument: " #c); } while (0)

template <typename T> struct DevBuf {
    T *p = nullptr; size_t n = 0;
    DevBuf() {}
    DevBuf(const DevBuf &) = delete; DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    cudaError_t alloc(size_t count) {
        if (count <= n && p) return cudaSuccess;
        release();
        cudaError_t e = cudaMalloc((void **) &p, (count ? count : 1) * sizeof(T));
        if (e == cudaSuccess) n = count;
        return e;
    }
    cudaError_t upload(const T *src, size_t count, cudaStream_t s) {
        cudaError_t e = alloc(count); if (e != cudaSuccess) return e;
        if (count == 0) return cudaSuccess;
        return cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, s);
    }
};

struct mtsb_ctx {
    int device; cudaStream_t stream; cudaEvent_t ev0, ev1; volatile int cancel;
    cudaEvent_t ring[2];                     // one event per in-flight render batch (mtsb_render keeps at most two queued)
    float timeout = -1.f;                    // seconds; <= 0: none (mtsb_set_timeout)
    int profiling = 0;                       // per-stage CUDA-event timing (mtsb_set_profiling)
    int bin_materials = 1;                   // sort queue entries by material class before shading (MTSB_BIN=0 disables)
    int bin_emitters = 1;                    // ... and by the emitter the path is going to sample (MTSB_BIN_EMITTER=0 disables)
    int bin_depth1 = 1;                      // primary hits too: coherent in material, but not in the emitter pick (MTSB_BIN_DEPTH1=0: identity lists there)
    unsigned trace_grid = 148 * 8, refill = 22; // persistent traversal launch shape (set in mtsb_ctx_create)
    std::vector<cudaEvent_t> pool;           // event pairs, grown on demand
    std::vector<int> pool_stage; size_t pool_used = 0;
};
// stage ids for the profiling pool
enum { ST_RAYGEN = 0, ST_CLOSEST = 1, ST_SHADE = 2, ST_ANY = 3, ST_FILM = 4, ST_COUNT = 5 };
struct StageTimer {      // records an event pair around one launch when profiling is on
    mtsb_ctx *c; bool on;
    StageTimer(mtsb_ctx *ctx, int stage) : c(ctx), on(ctx->profiling != 0) {
        if (!on) return;
        if (c->pool_used + 2 > c->pool.size()) { for (int k = 0; k < 2; ++k) { cudaEvent_t e; cudaEventCreate(&e); c->pool.push_back(e); } c->pool_stage.resize(c->pool.size() / 2); }
        c->pool_stage[c->pool_used / 2] = stage;
        cudaEventRecord(c->pool[c->pool_used], c->stream);
    }
    ~StageTimer() { if (on) { cudaEventRecord(c->pool[c->pool_used + 1], c->stream); c->pool_used += 2; } }
};
static void collect_stage_times(mtsb_ctx *c, mtsb_stats *st) {   // call after the stream is synchronised
    float acc[ST_COUNT] = { 0, 0, 0, 0, 0 };
    for (size_t i = 0; i + 1 < c->pool_used; i += 2) { float ms = 0.f; cudaEventElapsedTime(&ms, c->pool[i], c->pool[i + 1]); acc[c->pool_stage[i / 2]] += ms; }
    c->pool_used = 0;
    if (st) { st->ms_raygen = acc[ST_RAYGEN]; st->ms_closest = acc[ST_CLOSEST]; st->ms_shade = acc[ST_SHADE]; st->ms_any = acc[ST_ANY]; st->ms_film = acc[ST_FILM]; }
}

struct SensorHost { SensorDev dev; DevBuf<float4> xyza, m2; DevBuf<float> w; DevBuf<float> packed; };

struct BuildScratch {
    DevBuf<float4> tlo, thi, nlo, nhi; DevBuf<uint64_t> keys, keys2; DevBuf<uint32_t> order, order2, rfirst, rlast, flags, counters;
    DevBuf<int> left, right, parent, wide_bin, d_count; DevBuf<float> bounds; DevBuf<unsigned char> sort_tmp, select_tmp;
};
struct mtsb_scene {
    mtsb_ctx *ctx;
    BuildScratch build_scratch;
    std::vector<uint32_t> bvh_levels;     // wide-node index where every level of the last build starts (+ the total): the refit walks them backwards
    int update_refit = 0; uint32_t rebuild_every = 0, refits_since_build = 0;      // mtsb_scene_update_config
    // host copies needed for updates
    std::vector<MeshDev> meshes; std::vector<std::vector<uint32_t>> mesh_faces; std::vector<bool> computed_normals;
    std::vector<TexDev> tex; std::vector<BsdfDev> bsdfs; std::vector<EmitterDev> emitters;
    std::map<int, std::vector<float>> emitter_pos;   // host positions of EMITTER meshes only (area CDF is built on the host in double)
    // device
    DevBuf<float> pos, nrm, uv; DevBuf<uint4> faces;
    std::vector<ShapeDev> shapes; DevBuf<ShapeDev> d_shapes;   // analytic shapes; meshes[n_tri_meshes + k] is shape k
    uint32_t n_tri_meshes = 0;
    DevBuf<MeshDev> d_meshes; DevBuf<TexDev> d_tex; DevBuf<BsdfDev> d_bsdfs; DevBuf<EmitterDev> d_emitters;
    std::vector<std::unique_ptr<DevBuf<float>>> tables;       // texture data, cdfs, envmap levels
    std::vector<int> tex_table;                               // texture id -> tables index (-1 constant)
    std::vector<int> emitter_cdf_table;
    DevBuf<float4> env_rgba;
    bool spectral = false;                   // MTSB_MODE_SPECTRAL: 4 wavelengths per path
    bool ext = false;                        // some BSDF is a conductor / plastic / rough dielectric or masked: RgbX / SpectralX kernels
    DevBuf<float> cie;                       // spectral mode: CIE 1931 x/y/z + D65/10568 (cie_tables.h)
    DevBuf<float2> q_we[2];                  // spectral mode: wavelength sample + eta per queue entry
    DevBuf<uint4> nodes; DevBuf<float4> tris;
    SceneDev dev;
    std::vector<std::unique_ptr<SensorHost>> sensors;
    // wavefront pool
    uint32_t pool = 0;
    DevBuf<float4> q_o[2], q_d[2], q_hit[2], q_te[2], s_o, s_d, s_c, result, final_radiance, path_dl;
    DevBuf<uint8_t> prim_class;              // per primitive: 1 + BSDF type (shade class), see k_classify
    DevBuf<uint32_t> class_lists;            // material bins of the shade stage: kShadeClasses x pool queue indices
    DevBuf<uint4> q_misc[2], q_rng[2];
    DevBuf<float4> q_p1[2], q_p2[2];         // vertex-gradient replay: previous two path vertices per queue entry
    std::map<int, std::unique_ptr<DevBuf<uint4>>> mesh_edges;   // unique edges per mesh (boundary term), built on first use
    int grad_boundary = 2; float grad_edge_fraction = 0.25f;    // mtsb_grad_config
    std::vector<uint32_t> edge_offsets;                         // all_edge_list(): where every mesh's edges start
    DevBuf<float4> s_nee; DevBuf<uint32_t> s_occ;     // vertex-gradient replay: ShadowQueue::nee_rec / occluded
    DevBuf<float4> edge_aux;                 // boundary term seen from x1: 4 float4 per sample
    DevBuf<uint32_t> counters; DevBuf<unsigned long long> dstats;
    mtsb_stats build_stats;
    float3 bbox_lo, bbox_hi;
};

extern "C" const char *mtsb_last_error(void) { return g_err.c_str(); }
extern "C" int mtsb_abi_version(void) { return MTSB_ABI_VERSION; }

extern "C" int mtsb_ctx_create(int device, mtsb_ctx **out) {
    CHECK_ARG(out);
    int n = 0;
    CU(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail(MTSB_E_INVALID, "no such CUDA device");
    CU(cudaSetDevice(device));
    std::unique_ptr<mtsb_ctx> holder(new mtsb_ctx());       // released on any early return below
    mtsb_ctx *c = holder.get();
    c->device = device; c->cancel = 0;
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&c->ev0)); CU(cudaEventCreate(&c->ev1));
    CU(cudaEventCreateWithFlags(&c->ring[0], cudaEventDisableTiming)); CU(cudaEventCreateWithFlags(&c->ring[1], cudaEventDisableTiming));
    {   // persistent traversal grid: resident blocks per SM x SM count (tunable for experiments)
        int sms = 148, per_sm = 8, occ = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_trace<false>, kBlock, 0) == cudaSuccess && occ > 0) per_sm = occ;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shadow<false>, kBlock, 0) == cudaSuccess && occ > 0 && occ < per_sm) per_sm = occ;
        if (const char *e = getenv("MTSB_TRACE_BLOCKS_PER_SM")) { int v = atoi(e); if (v > 0 && v <= 32) per_sm = v; }
        if (const char *e = getenv("MTSB_REFILL")) { int v = atoi(e); if (v >= 0 && v <= 32) c->refill = (unsigned) v; }
        c->trace_grid = (unsigned) (sms * per_sm);
        if (const char *e = getenv("MTSB_BIN")) c->bin_materials = atoi(e) != 0;
        if (const char *e = getenv("MTSB_BIN_EMITTER")) c->bin_emitters = atoi(e) != 0;
        if (const char *e = getenv("MTSB_BIN_DEPTH1")) c->bin_depth1 = atoi(e) != 0;
    }
    *out = holder.release();
    return MTSB_OK;
}
extern "C" int mtsb_ctx_destroy(mtsb_ctx *c) {
    if (!c) return MTSB_OK;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaEventDestroy(c->ev0); cudaEventDestroy(c->ev1); cudaEventDestroy(c->ring[0]); cudaEventDestroy(c->ring[1]); cudaStreamDestroy(c->stream);
    delete c;
    return MTSB_OK;
}
extern "C" int mtsb_cancel(mtsb_ctx *c) { if (c) c->cancel = 1; return MTSB_OK; }
extern "C" int mtsb_set_timeout(mtsb_ctx *c, float seconds) { if (!c) return fail(MTSB_E_INVALID, "null context"); c->timeout = seconds; return MTSB_OK; }
extern "C" int mtsb_set_profiling(mtsb_ctx *c, int on) { if (!c) return fail(MTSB_E_INVALID, "null context"); c->profiling = on; return MTSB_OK; }

static inline unsigned grid_for(size_t n) { return (unsigned) ((n + kBlock - 1) / kBlock); }
// counters buffer (uint32 words): [0,1024) live paths per depth, [1024,2048) shadow rays per depth,
// [2048,3072) work-fetch cursor of the closest-hit launch of that depth, [3072,4096) same for the shadow launch
// [4096, 8192) shade-class counters, 4 per depth
constexpr uint32_t kCounterWords = 12288, kWorkClosest = 2048, kWorkShadow = 3072, kClassCounts = 4096;      // class counters: 8 per depth
static inline unsigned trace_grid_for(const mtsb_ctx *c, size_t n) {
    unsigned g = grid_for(n);
    return g < c->trace_grid ? (g ? g : 1u) : c->trace_grid;
}

// ---------------------------------------------------------------------------------------------------
// GPU BVH build
// ---------------------------------------------------------------------------------------------------
static int build_bvh(mtsb_scene *s) {
    cudaStream_t st = s->ctx->stream;
    uint32_t n = (uint32_t) s->faces.n;
    s->dev.bvh.n_tris = n;
    s->build_stats.bvh_nodes = 0; s->build_stats.bvh_depth = 0;
    if (n == 0) { s->dev.bvh.nodes = nullptr; s->dev.bvh.tris = nullptr; return MTSB_OK; }
    CU(cudaEventRecord(s->ctx->ev0, st));
    // build scratch lives in a per-scene arena: a rebuild after `parameters_changed` allocates nothing (DevBuf::alloc keeps a large-enough buffer)
    BuildScratch &bs = s->build_scratch;
    DevBuf<float4> &tlo = bs.tlo, &thi = bs.thi, &nlo = bs.nlo, &nhi = bs.nhi; DevBuf<uint64_t> &keys = bs.keys, &keys2 = bs.keys2;
    DevBuf<uint32_t> &order = bs.order, &order2 = bs.order2, &rfirst = bs.rfirst, &rlast = bs.rlast, &flags = bs.flags, &counters = bs.counters;
    DevBuf<int> &left = bs.left, &right = bs.right, &parent = bs.parent, &wide_bin = bs.wide_bin; DevBuf<float> &bounds = bs.bounds;
    CU(tlo.alloc(n)); CU(thi.alloc(n)); CU(nlo.alloc(2 * (size_t) n)); CU(nhi.alloc(2 * (size_t) n));
    CU(keys.alloc(n)); CU(keys2.alloc(n)); CU(order.alloc(n)); CU(order2.alloc(n));
    CU(rfirst.alloc(n)); CU(rlast.alloc(n)); CU(flags.alloc(n)); CU(counters.alloc(2));
    CU(left.alloc(n)); CU(right.alloc(n)); CU(parent.alloc(2 * (size_t) n)); CU(wide_bin.alloc(n + 1));
    CU(bounds.alloc(6));
    CU(s->nodes.alloc(5 * ((size_t) n + 1))); CU(s->tris.alloc(3 * (size_t) n));
    BvhBuild b;
    b.pos = s->pos.p; b.faces = s->faces.p; b.shapes = s->d_shapes.p; b.n = n;
    b.tlo = tlo.p; b.thi = thi.p; b.keys = keys.p; b.order = order.p; b.left = left.p; b.right = right.p; b.parent = parent.p;
    b.rfirst = rfirst.p; b.rlast = rlast.p; b.nlo = nlo.p; b.nhi = nhi.p; b.flags = flags.p;
    b.nodes = s->nodes.p; b.tris = s->tris.p; b.wide_bin = wide_bin.p; b.counters = counters.p;
    float init[6] = { INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY };
    CU(cudaMemcpyAsync(bounds.p, init, sizeof(init), cudaMemcpyHostToDevice, st));
    k_tri_bounds<<<grid_for(n), kBlock, 0, st>>>(b, bounds.p);
    float hb[6];
    CU(cudaMemcpyAsync(hb, bounds.p, sizeof(hb), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    b.scene_lo = mk3(hb[0], hb[1], hb[2]); b.scene_hi = mk3(hb[3], hb[4], hb[5]);
    s->bbox_lo = b.scene_lo; s->bbox_hi = b.scene_hi;
    k_morton<<<grid_for(n), kBlock, 0, st>>>(b);
    {   // radix sort (library primitive; the build is not on the per-ray hot path)
        size_t tmp_bytes = 0;
        CU(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys.p, keys2.p, order.p, order2.p, (int) n, 0, 63, st));
        DevBuf<unsigned char> &tmp = bs.sort_tmp; CU(tmp.alloc(tmp_bytes));
        CU(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, keys.p, keys2.p, order.p, order2.p, (int) n, 0, 63, st));
        CU(cudaStreamSynchronize(st));
        b.keys = keys2.p; b.order = order2.p;
    }
    CU(cudaMemsetAsync(flags.p, 0, n * sizeof(uint32_t), st));
    const char *bvh_mode = getenv("MTSB_BVH");
    if (n > 1 && !(bvh_mode && !strcmp(bvh_mode, "lbvh"))) {
        // PLOC agglomeration over the Morton order (default).  The Morton keys are dead after the sort, so their two
        // buffers serve as scratch: cluster arrays (ping/pong), nearest-neighbour indices, new leaf ids.
        int *cur = (int *) keys.p, *nxt = cur + n, *nn = (int *) keys2.p, *leaf_new = nn + n;
        uint32_t *order_out = order.p;            // the buffer the sort did not leave the result in
        uint32_t zero = 0;
        CU(cudaMemcpyAsync(counters.p, &zero, sizeof(zero), cudaMemcpyHostToDevice, st));
        k_ploc_init<<<grid_for(n), kBlock, 0, st>>>(b, cur);
        DevBuf<unsigned char> &tmp = bs.select_tmp; DevBuf<int> &d_count = bs.d_count; CU(d_count.alloc(1));
        size_t tmp_bytes = 0;
        CU(cub::DeviceSelect::If(nullptr, tmp_bytes, nxt, cur, d_count.p, (int) n, PlocAlive(), st));
        CU(tmp.alloc(tmp_bytes));
        uint32_t m = n; int rounds = 0;
        while (m > 1) {
            k_ploc_nearest<<<grid_for(m), kBlock, 0, st>>>(b, cur, m, nn);
            k_ploc_merge<<<grid_for(m), kBlock, 0, st>>>(b, cur, nn, m, nxt);
            CU(cub::DeviceSelect::If(tmp.p, tmp_bytes, nxt, cur, d_count.p, (int) m, PlocAlive(), st));
            int hm = 0;
            CU(cudaMemcpyAsync(&hm, d_count.p, sizeof(int), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            if (hm <= 0 || (uint32_t) hm >= m) return fail(MTSB_E_INVALID, "PLOC made no progress");
            m = (uint32_t) hm;
            if (++rounds > 4096) return fail(MTSB_E_INVALID, "PLOC did not terminate");
        }
        k_ploc_finalize_internal<<<grid_for(n), kBlock, 0, st>>>(b);
        k_ploc_finalize_leaf<<<grid_for(n), kBlock, 0, st>>>(b, order_out, leaf_new);
        k_ploc_patch_leaf<<<grid_for(n), kBlock, 0, st>>>(b, leaf_new);
        k_ploc_leaf_bounds<<<grid_for(n), kBlock, 0, st>>>(b, order_out);
        b.order = order_out;
    } else {
        if (n > 1) k_lbvh_nodes<<<grid_for(n - 1), kBlock, 0, st>>>(b);
        k_lbvh_fit<<<grid_for(n), kBlock, 0, st>>>(b);
    }
    uint32_t c0[2] = { 1u, 0u };
    int root = n > 1 ? 0 : 0;     // internal node 0, or leaf id (n-1)+0 = 0 when n == 1
    CU(cudaMemcpyAsync(counters.p, c0, sizeof(c0), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(wide_bin.p, &root, sizeof(int), cudaMemcpyHostToDevice, st));
    uint32_t begin = 0, end = 1, depth = 0;
    s->bvh_levels.clear(); s->refits_since_build = 0;
    while (begin < end) {
        s->bvh_levels.push_back(begin);
        k_collapse<<<grid_for(end - begin), kBlock, 0, st>>>(b, begin, end);
        uint32_t hc[2];
        CU(cudaMemcpyAsync(hc, counters.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        begin = end; end = hc[0]; depth++;
        if (depth > 4096) return fail(MTSB_E_INVALID, "BVH collapse did not terminate");
    }
    s->bvh_levels.push_back(end);
    CU(cudaGetLastError());
    if (depth > (uint32_t) (kStackFast + kStackSlow)) return fail(MTSB_E_INVALID, "BVH too deep for the traversal stack");
    CU(cudaEventRecord(s->ctx->ev1, st));
    CU(cudaEventSynchronize(s->ctx->ev1));
    CU(cudaEventElapsedTime(&s->build_stats.ms_build, s->ctx->ev0, s->ctx->ev1));
    s->build_stats.bvh_nodes = end; s->build_stats.bvh_depth = depth;
    s->dev.bvh.nodes = s->nodes.p; s->dev.bvh.tris = s->tris.p; s->dev.bvh.shapes = s->shapes.empty() ? nullptr : s->d_shapes.p;
    return MTSB_OK;
}

// Refit of the BVH of the last build to the current position array (bvh.h refit_*_body): one pass over the leaf slots, one
// launch per level from the deepest up, one read-back (the scene box, for the environment emitters' bounding sphere).
static int refit_bvh(mtsb_scene *s) {
    cudaStream_t st = s->ctx->stream;
    uint32_t n = (uint32_t) s->faces.n;
    if (n == 0 || s->bvh_levels.size() < 2) return build_bvh(s);
    CU(cudaEventRecord(s->ctx->ev0, st));
    BuildScratch &bs = s->build_scratch;
    BvhRefit r; r.nodes = s->nodes.p; r.tris = s->tris.p; r.pos = s->pos.p; r.faces = s->faces.p; r.shapes = s->d_shapes.p;
    r.box_lo = bs.nlo.p; r.box_hi = bs.nhi.p;      // 2n entries each from the build; there are at most n wide nodes
    k_refit_tris<<<grid_for(n), kBlock, 0, st>>>(r, n);
    for (size_t l = s->bvh_levels.size() - 1; l-- > 0;) {
        uint32_t begin = s->bvh_levels[l], end = s->bvh_levels[l + 1];
        if (end > begin) k_refit_level<<<grid_for(end - begin), kBlock, 0, st>>>(r, begin, end);
    }
    float4 hb[2];
    CU(cudaMemcpyAsync(&hb[0], r.box_lo, sizeof(float4), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&hb[1], r.box_hi, sizeof(float4), cudaMemcpyDeviceToHost, st));
    CU(cudaEventRecord(s->ctx->ev1, st));
    CU(cudaEventSynchronize(s->ctx->ev1));
    CU(cudaGetLastError());
    CU(cudaEventElapsedTime(&s->build_stats.ms_build, s->ctx->ev0, s->ctx->ev1));
    s->bbox_lo = xyz(hb[0]); s->bbox_hi = xyz(hb[1]);
    s->refits_since_build++;
    return MTSB_OK;
}

// ---------------------------------------------------------------------------------------------------
// Scene snapshot
// ---------------------------------------------------------------------------------------------------
static int upload_struct_arrays(mtsb_scene *s) {
    cudaStream_t st = s->ctx->stream;
    CU(s->d_meshes.upload(s->meshes.data(), s->meshes.size(), st));
    CU(s->d_tex.upload(s->tex.data(), s->tex.size(), st));
    CU(s->d_bsdfs.upload(s->bsdfs.data(), s->bsdfs.size(), st));
    CU(s->d_emitters.upload(s->emitters.data(), s->emitters.size(), st));
    {   // shade class of every primitive (1 + BSDF type), looked up by the traversal kernel when a ray finishes
        size_t np = 0; for (const MeshDev &m : s->meshes) np += m.face_count;
        std::vector<uint8_t> cls(np ? np : 1, 0);
        for (const MeshDev &m : s->meshes)
            for (uint32_t f = 0; f < m.face_count; ++f) cls[m.face_begin + f] = (uint8_t) (1 + (s->bsdfs[m.bsdf].type < 2 ? s->bsdfs[m.bsdf].type : 2));   // 2 bits: the rest share a class
        CU(s->prim_class.upload(cls.data(), cls.size(), st));
    }
    CU(cudaStreamSynchronize(st));
    s->dev.pos = s->pos.p; s->dev.nrm = s->nrm.p; s->dev.uv = s->uv.p; s->dev.faces = s->faces.p; s->dev.shapes = s->d_shapes.p;
    s->dev.meshes = s->d_meshes.p; s->dev.tex = s->d_tex.p; s->dev.bsdfs = s->d_bsdfs.p; s->dev.emitters = s->d_emitters.p;
    return MTSB_OK;
}

static int build_area_tables(mtsb_scene *s) {
    cudaStream_t st = s->ctx->stream;
    for (size_t e = 0; e < s->emitters.size(); ++e) {
        EmitterDev &em = s->emitters[e];
        if (em.type != MTSB_EMITTER_AREA) continue;
        const MeshDev &m = s->meshes[em.mesh];
        if (m.shape >= 0) { em.cdf = nullptr; em.normalization = s->shapes[m.shape].inv_area; continue; }      // analytic: sampled in closed form
        AreaTable t;
        try { t = build_area_table(s->emitter_pos.at(em.mesh).data(), s->mesh_faces[em.mesh].data(), m.face_count); }
        catch (const std::exception &ex) { return fail(MTSB_E_INVALID, ex.what()); }
        int ti = s->emitter_cdf_table[e];
        if (ti < 0) { s->tables.emplace_back(new DevBuf<float>()); ti = (int) s->tables.size() - 1; s->emitter_cdf_table[e] = ti; }
        CU(s->tables[ti]->upload(t.cdf.data(), t.cdf.size(), st));
        CU(cudaStreamSynchronize(st));
        em.cdf = s->tables[ti]->p; em.cdf_sum = t.sum; em.normalization = t.normalization; em.valid_lo = t.valid_lo; em.valid_hi = t.valid_hi;
    }
    return MTSB_OK;
}

static void update_env_bsphere(mtsb_scene *s) {
    // Scene bbox -> bounding sphere (envmap.cpp:126-130): radius = |max - center| * (1 + RayEpsilon)
    float3 c = (s->bbox_lo + s->bbox_hi) * 0.5f;
    float r = s->faces.n ? norm(s->bbox_hi - c) : 0.f;
    r = fmaxf(kRayEpsilon, r * (1.f + kRayEpsilon));
    for (auto &em : s->emitters) if (em.type == MTSB_EMITTER_ENVMAP || em.type == MTSB_EMITTER_CONSTANT) em.bsphere_radius = r;   // constant.cpp:45-49
}

extern "C" int mtsb_scene_destroy(mtsb_scene *s) {
    if (!s) return MTSB_OK;
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
    delete s;
    return MTSB_OK;
}

extern "C" int mtsb_scene_create(mtsb_ctx *ctx, const mtsb_scene_desc *d, mtsb_scene **out) {
    CHECK_ARG(ctx); CHECK_ARG(d); CHECK_ARG(out);
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    std::unique_ptr<mtsb_scene> s(new mtsb_scene());
    s->ctx = ctx;
    memset(&s->dev, 0, sizeof(SceneDev)); memset(&s->build_stats, 0, sizeof(mtsb_stats));
    // ---- meshes
    // Direct-to-device upload: vertex data goes from the caller's buffers (host or device, mtsb_mesh_desc.mem) straight into
    // the concatenated device arrays, faces are expanded to (i0, i1, i2, mesh) with global vertex indices by a kernel.
    // Host copies are kept only where the host needs them: face indices (edge lists, area tables) and emitter positions.
    size_t nv_total = 0, nf_total = 0;
    for (uint32_t i = 0; i < d->n_meshes; ++i) { nv_total += d->meshes[i].vertex_count; nf_total += d->meshes[i].face_count; }
    CHECK_ARG(d->n_shapes == 0 || d->shapes != nullptr);
    if (nf_total + d->n_shapes > (size_t) kPrimMask) return fail(MTSB_E_INVALID, "more than 2^30 - 1 primitives");
    CU(s->pos.alloc(3 * nv_total)); CU(s->nrm.alloc(3 * nv_total)); CU(s->uv.alloc(2 * nv_total)); CU(s->faces.alloc(nf_total + d->n_shapes));
    DevBuf<uint32_t> raw_faces;
    uint32_t vbase = 0, fbase = 0;
    for (uint32_t i = 0; i < d->n_meshes; ++i) {
        const mtsb_mesh_desc &m = d->meshes[i];
        CHECK_ARG(m.positions && (m.faces || m.face_count == 0));
        CHECK_ARG(m.bsdf_id >= 0 && (uint32_t) m.bsdf_id < d->n_bsdfs);
        CHECK_ARG(m.emitter_id < (int32_t) d->n_emitters);
        CHECK_ARG(m.mem == MTSB_MEM_HOST || m.mem == MTSB_MEM_DEVICE);
        const cudaMemcpyKind kind = m.mem == MTSB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
        MeshDev md; md.face_begin = fbase; md.face_count = m.face_count; md.vert_begin = vbase; md.vert_count = m.vertex_count;
        md.bsdf = m.bsdf_id; md.emitter = m.emitter_id; md.has_normals = m.normals != nullptr; md.has_uv = m.texcoords != nullptr;
        md.shape = -1;
        s->meshes.push_back(md);
        const size_t nv = m.vertex_count, nf = m.face_count;
        CU(cudaMemcpyAsync(s->pos.p + 3 * (size_t) vbase, m.positions, 3 * nv * sizeof(float), kind, st));
        if (m.normals) CU(cudaMemcpyAsync(s->nrm.p + 3 * (size_t) vbase, m.normals, 3 * nv * sizeof(float), kind, st));
        else CU(cudaMemsetAsync(s->nrm.p + 3 * (size_t) vbase, 0, 3 * nv * sizeof(float), st));
        if (m.texcoords) CU(cudaMemcpyAsync(s->uv.p + 2 * (size_t) vbase, m.texcoords, 2 * nv * sizeof(float), kind, st));
        else CU(cudaMemsetAsync(s->uv.p + 2 * (size_t) vbase, 0, 2 * nv * sizeof(float), st));
        // host copy of the indices (validated here; the device copy is expanded from the same values)
        s->mesh_faces.emplace_back(3 * nf);
        std::vector<uint32_t> &hf = s->mesh_faces.back();
        if (m.mem == MTSB_MEM_DEVICE) { CU(cudaMemcpyAsync(hf.data(), m.faces, 3 * nf * sizeof(uint32_t), cudaMemcpyDeviceToHost, st)); CU(cudaStreamSynchronize(st)); }
        else if (nf) memcpy(hf.data(), m.faces, 3 * nf * sizeof(uint32_t));
        for (size_t k = 0; k < 3 * nf; ++k) if (hf[k] >= m.vertex_count) return fail(MTSB_E_INVALID, "face index out of range");
        if (nf) {
            const uint32_t *src = m.faces;
            if (m.mem == MTSB_MEM_HOST) { CU(raw_faces.alloc(3 * nf)); CU(cudaMemcpyAsync(raw_faces.p, hf.data(), 3 * nf * sizeof(uint32_t), cudaMemcpyHostToDevice, st)); src = raw_faces.p; }
            k_faces_expand<<<grid_for(nf), kBlock, 0, st>>>(src, (uint32_t) nf, vbase, i, s->faces.p + fbase);
            if (m.mem == MTSB_MEM_HOST) CU(cudaStreamSynchronize(st));       // raw_faces is reused by the next mesh
        }
        if (m.emitter_id >= 0) {
            std::vector<float> &hp = s->emitter_pos[(int) i]; hp.resize(3 * nv);
            if (m.mem == MTSB_MEM_DEVICE) { CU(cudaMemcpyAsync(hp.data(), m.positions, 3 * nv * sizeof(float), cudaMemcpyDeviceToHost, st)); CU(cudaStreamSynchronize(st)); }
            else memcpy(hp.data(), m.positions, 3 * nv * sizeof(float));
        }
        s->computed_normals.push_back(false);
        vbase += m.vertex_count; fbase += m.face_count;
    }
    s->n_tri_meshes = d->n_meshes;
    // ---- analytic shapes: one primitive and one MeshDev entry each, after the triangle meshes
    if (d->n_shapes) {
        std::vector<uint4> sf(d->n_shapes);
        for (uint32_t k = 0; k < d->n_shapes; ++k) {
            const mtsb_shape_desc &sh = d->shapes[k];
            CHECK_ARG(sh.bsdf_id >= 0 && (uint32_t) sh.bsdf_id < d->n_bsdfs);
            CHECK_ARG(sh.emitter_id < (int32_t) d->n_emitters);
            ShapeDev sd; const char *err = nullptr;
            if (!make_shape_dev(sh.type, sh.flip_normals, sh.to_world, sd, &err)) return fail(MTSB_E_INVALID, err);
            s->shapes.push_back(sd);
            MeshDev md; md.face_begin = fbase + k; md.face_count = 1; md.vert_begin = 0; md.vert_count = 0;
            md.bsdf = sh.bsdf_id; md.emitter = sh.emitter_id; md.has_normals = 0; md.has_uv = 1; md.shape = (int) k;
            s->meshes.push_back(md);
            sf[k] = make_uint4(k, kAnalyticTag, (uint32_t) sh.type, d->n_meshes + k);
        }
        s->ext = true;                   // analytic surface interactions / light sampling live in the RgbX / SpectralX shade instances
        CU(s->d_shapes.upload(s->shapes.data(), s->shapes.size(), st));
        CU(cudaMemcpyAsync(s->faces.p + fbase, sf.data(), sf.size() * sizeof(uint4), cudaMemcpyHostToDevice, st));
        CU(cudaStreamSynchronize(st));
    }
    CU(cudaStreamSynchronize(st));
    // ---- colour mode
    CHECK_ARG(d->mode == MTSB_MODE_RGB || d->mode == MTSB_MODE_SPECTRAL);
    s->spectral = d->mode == MTSB_MODE_SPECTRAL;
    if (s->spectral) {
        std::vector<float> tbl(4 * 95);
        for (int k = 0; k < 3 * 95; ++k) tbl[k] = kCie1931[k];
        for (int k = 0; k < 95; ++k) tbl[3 * 95 + k] = kCieD65[k] * (1.f / 10568.f);      // D65Spectrum: scale / 10568 (d65.cpp:63)
        CU(s->cie.upload(tbl.data(), tbl.size(), st));
        CU(cudaStreamSynchronize(st));
        s->dev.cie = s->cie.p;
    }
    // ---- textures
    for (uint32_t i = 0; i < d->n_textures; ++i) {
        const mtsb_texture_desc &t = d->textures[i];
        TexDev td; memset(&td, 0, sizeof(td));
        td.type = t.type; td.value = mk3(t.value[0], t.value[1], t.value[2]);
        int ti = -1;
        if (t.type == MTSB_TEX_BITMAP) {
            CHECK_ARG(t.data && t.width >= 2 && t.height >= 2 && (t.channels == 1 || t.channels == 3));
            td.w = t.width; td.h = t.height; td.ch = t.channels; memcpy(td.to_uv, t.to_uv, sizeof(td.to_uv));
            s->tables.emplace_back(new DevBuf<float>()); ti = (int) s->tables.size() - 1;
            CU(s->tables[ti]->upload(t.data, (size_t) t.width * t.height * t.channels, st));
            CU(cudaStreamSynchronize(st));
            td.data = s->tables[ti]->p;
        } else if (t.type == MTSB_TEX_REGULAR || t.type == MTSB_TEX_IRREGULAR) {
            if (!s->spectral) return fail(MTSB_E_INVALID, "tabulated spectra need MTSB_MODE_SPECTRAL (rgb variants pre-integrate them to a colour on the host)");
            CHECK_ARG(t.data && t.width >= 2);
            if (t.type == MTSB_TEX_REGULAR) CHECK_ARG(t.lambda_max > t.lambda_min);
            else for (uint32_t k = 1; k < t.width; ++k) CHECK_ARG(t.data[k] > t.data[k - 1]);     // IrregularContinuousDistribution: nodes strictly increasing
            td.w = t.width; td.lambda_min = t.lambda_min; td.lambda_max = t.lambda_max;
            s->tables.emplace_back(new DevBuf<float>()); ti = (int) s->tables.size() - 1;
            if (t.type == MTSB_TEX_IRREGULAR) {       // nodes, values + the bucket index of host_tables.h
                std::vector<float> packed = irregular_with_buckets(t.data, t.width);
                td.h = kIrregularBuckets;
                CU(s->tables[ti]->upload(packed.data(), packed.size(), st));
                CU(cudaStreamSynchronize(st));
            } else CU(s->tables[ti]->upload(t.data, (size_t) t.width, st));
            CU(cudaStreamSynchronize(st));
            td.data = s->tables[ti]->p;
        } else if (t.type == MTSB_TEX_MESH_ATTRIBUTE) {
            CHECK_ARG(t.data && t.width >= 1 && (t.channels == 1 || t.channels == 3) && t.height <= 1);
            td.w = t.width; td.h = t.height; td.ch = t.channels;
            s->tables.emplace_back(new DevBuf<float>()); ti = (int) s->tables.size() - 1;
            CU(s->tables[ti]->upload(t.data, (size_t) t.width * t.channels, st));
            CU(cudaStreamSynchronize(st));
            td.data = s->tables[ti]->p;
            s->ext = true;                   // evaluated by the RgbX / SpectralX shade instances only
        } else if (t.type == MTSB_TEX_UNIFORM) {
            if (!s->spectral) return fail(MTSB_E_INVALID, "uniform spectra need MTSB_MODE_SPECTRAL");
        } else CHECK_ARG(t.type == MTSB_TEX_CONSTANT);
        if (s->spectral) td.d65_scale = t.d65_scale; else CHECK_ARG(t.d65_scale == 0.f);
        s->tex.push_back(td); s->tex_table.push_back(ti);
    }
    // ---- bsdfs
    auto tex_ok = [&](int32_t id, bool optional) { return (optional && id < 0) || (id >= 0 && (uint32_t) id < d->n_textures); };
    for (uint32_t i = 0; i < d->n_bsdfs; ++i) {
        const mtsb_bsdf_desc &b = d->bsdfs[i];
        BsdfDev bd; memset(&bd, 0, sizeof(bd));
        bd.type = b.type; bd.refl = b.reflectance_tex; bd.eta = b.eta_tex; bd.k = b.k_tex; bd.spec_r = b.specular_reflectance_tex;
        bd.spec_t = b.specular_transmittance_tex; bd.alpha_u = b.alpha_u; bd.alpha_v = b.alpha_v; bd.distr = b.distribution;
        bd.sample_visible = b.sample_visible; bd.eta_f = b.eta; bd.two_sided = b.two_sided ? 1 : 0;
        bd.nonlinear = b.nonlinear; bd.spec_weight = b.specular_sampling_weight; bd.opacity = b.opacity_tex_plus1 - 1;
        if (b.type >= MTSB_BSDF_CONDUCTOR || bd.opacity >= 0) s->ext = true;
        CHECK_ARG(tex_ok(b.opacity_tex_plus1 - 1, true));
        if (b.two_sided && b.type == MTSB_BSDF_DIELECTRIC) return fail(MTSB_E_INVALID, "twosided: only materials without a transmission component can be nested");
        if (b.type == MTSB_BSDF_DIFFUSE) CHECK_ARG(tex_ok(b.reflectance_tex, false));
        else if (b.type == MTSB_BSDF_ROUGHCONDUCTOR || b.type == MTSB_BSDF_CONDUCTOR) CHECK_ARG(tex_ok(b.eta_tex, false) && tex_ok(b.k_tex, false) && tex_ok(b.specular_reflectance_tex, true));
        else if (b.type == MTSB_BSDF_PLASTIC) CHECK_ARG(b.eta > 0.f && tex_ok(b.reflectance_tex, false) && tex_ok(b.specular_reflectance_tex, true) &&
                                                         b.specular_sampling_weight >= 0.f && b.specular_sampling_weight <= 1.f);
        else if (b.type == MTSB_BSDF_DIELECTRIC || b.type == MTSB_BSDF_ROUGHDIELECTRIC) CHECK_ARG(b.eta > 0.f && tex_ok(b.specular_reflectance_tex, true) && tex_ok(b.specular_transmittance_tex, true));
        else if (b.type == MTSB_BSDF_ROUGHPLASTIC) {
            CHECK_ARG(b.eta > 0.f && b.alpha_u > 0.f && tex_ok(b.reflectance_tex, false) && tex_ok(b.specular_reflectance_tex, true) &&
                      b.specular_sampling_weight >= 0.f && b.specular_sampling_weight <= 1.f);
            // roughplastic.cpp parameters_changed: external transmittance at 64 cosines + mean internal reflectance, integrated on the GPU
            std::vector<float> n32, w32, n128, w128;
            gauss_legendre_rule(32, n32, w32); gauss_legendre_rule(128, n128, w128);
            DevBuf<float> d_n32, d_w32, d_n128, d_w128;
            CU(d_n32.upload(n32.data(), 32, st)); CU(d_w32.upload(w32.data(), 32, st)); CU(d_n128.upload(n128.data(), 128, st)); CU(d_w128.upload(w128.data(), 128, st));
            s->tables.emplace_back(new DevBuf<float>());
            DevBuf<float> &tab = *s->tables.back();
            CU(tab.alloc(2 * kRoughTransRes));
            k_rough_tables<<<1, 2 * kRoughTransRes, 0, st>>>(b.distribution, b.alpha_u, b.eta, d_n32.p, d_w32.p, d_n128.p, d_w128.p, tab.p);
            float h[2 * kRoughTransRes];
            CU(cudaMemcpyAsync(h, tab.p, sizeof(h), cudaMemcpyDeviceToHost, st)); CU(cudaStreamSynchronize(st));
            double acc = 0; for (int k = 0; k < kRoughTransRes; ++k) acc += (double) h[kRoughTransRes + k] * rough_table_mu(k);
            bd.rough_t = tab.p; bd.internal_refl = (float) (acc / kRoughTransRes) * 2.f;
        }
        else return fail(MTSB_E_INVALID, "unknown BSDF type");
        s->bsdfs.push_back(bd);
    }
    // a mesh_attribute buffer must cover every mesh whose BSDF looks it up (Mesh::eval_attribute indexes by face / vertex)
    for (const MeshDev &m : s->meshes) {
        const BsdfDev &b = s->bsdfs[m.bsdf];
        const int ids[6] = { b.refl, b.eta, b.k, b.spec_r, b.spec_t, b.opacity };
        for (int id : ids)
            if (id >= 0 && s->tex[id].type == MTSB_TEX_MESH_ATTRIBUTE && m.shape >= 0)
                return fail(MTSB_E_INVALID, "mesh_attribute texture on an analytic shape");
        for (int id : ids)
            if (id >= 0 && s->tex[id].type == MTSB_TEX_MESH_ATTRIBUTE && s->tex[id].w != (s->tex[id].h == 1 ? m.face_count : m.vert_count))
                return fail(MTSB_E_INVALID, "mesh_attribute texture: element count does not match the mesh that uses it");
    }
    // ---- BVH (also yields the scene bbox)
    int rc = build_bvh(s.get());
    if (rc) return rc;
    // ---- emitters
    s->dev.env_emitter = -1;
    for (uint32_t i = 0; i < d->n_emitters; ++i) {
        const mtsb_emitter_desc &e = d->emitters[i];
        EmitterDev ed; memset(&ed, 0, sizeof(ed));
        ed.type = e.type; ed.radiance_tex = -1;
        s->emitter_cdf_table.push_back(-1);
        if (s->spectral && e.type != MTSB_EMITTER_ENVMAP) {
            CHECK_ARG(e.radiance_tex >= 0 && (uint32_t) e.radiance_tex < d->n_textures && d->textures[e.radiance_tex].type != MTSB_TEX_BITMAP);
            ed.radiance_tex = e.radiance_tex;
        }
        if (e.type == MTSB_EMITTER_AREA) {
            CHECK_ARG(e.mesh_id >= 0 && (uint32_t) e.mesh_id < d->n_meshes + d->n_shapes);
            CHECK_ARG(((uint32_t) e.mesh_id < d->n_meshes ? d->meshes[e.mesh_id].emitter_id : d->shapes[e.mesh_id - d->n_meshes].emitter_id) == (int32_t) i);
            ed.radiance = mk3(e.radiance[0], e.radiance[1], e.radiance[2]); ed.mesh = e.mesh_id;
        } else if (e.type == MTSB_EMITTER_CONSTANT) {
            if (s->dev.env_emitter >= 0) return fail(MTSB_E_INVALID, "only one environment emitter is allowed");
            s->dev.env_emitter = (int) i;
            ed.radiance = mk3(e.radiance[0], e.radiance[1], e.radiance[2]); ed.mesh = -1;
        } else if (e.type == MTSB_EMITTER_POINT) {
            ed.radiance = mk3(e.radiance[0], e.radiance[1], e.radiance[2]); ed.mesh = -1;
            ed.position = mk3(e.position[0], e.position[1], e.position[2]);
        } else if (e.type == MTSB_EMITTER_ENVMAP) {
            CHECK_ARG(e.rgba && e.width >= 2 && e.height >= 2);
            if (s->dev.env_emitter >= 0) return fail(MTSB_E_INVALID, "only one environment emitter is allowed");
            s->dev.env_emitter = (int) i;
            ed.w = e.width; ed.h = e.height; ed.scale = e.scale;
            memcpy(ed.to_world, e.to_world, sizeof(ed.to_world));
            if (!invert4x4(e.to_world, ed.to_local)) return fail(MTSB_E_INVALID, "envmap to_world is singular");
            size_t npx = (size_t) e.width * e.height;
            std::vector<float4> rgba(npx);
            if (s->spectral) {       // texels = sRGB-model coefficients + scale (envmap.cpp:98-110); importance still from the rgb image
                CHECK_ARG(e.coeff_scale != nullptr);
                for (size_t k = 0; k < npx; ++k) rgba[k] = make_float4(e.coeff_scale[4 * k], e.coeff_scale[4 * k + 1], e.coeff_scale[4 * k + 2], e.coeff_scale[4 * k + 3]);
            } else
            for (size_t k = 0; k < npx; ++k) rgba[k] = make_float4(e.rgba[4 * k], e.rgba[4 * k + 1], e.rgba[4 * k + 2], 1.f);
            CU(s->env_rgba.upload(rgba.data(), npx, st));
            ed.rgba = s->env_rgba.p;
            std::vector<float> lum = env_importance(e.rgba, e.width, e.height);
            H2DTable h = build_h2d(lum.data(), e.width, e.height);
            if ((int) h.levels.size() > kMaxEnvLevels) return fail(MTSB_E_INVALID, "envmap too large");
            ed.n_levels = (int) h.levels.size();
            for (size_t l = 0; l < h.levels.size(); ++l) {
                s->tables.emplace_back(new DevBuf<float>());
                CU(s->tables.back()->upload(h.levels[l].data(), h.levels[l].size(), st));
                ed.level_data[l] = s->tables.back()->p; ed.level_width[l] = h.widths[l];
            }
            CU(cudaStreamSynchronize(st));
        } else return fail(MTSB_E_INVALID, "unknown emitter type");
        s->emitters.push_back(ed);
    }
    s->dev.n_emitters = (int) s->emitters.size();
    rc = build_area_tables(s.get()); if (rc) return rc;
    update_env_bsphere(s.get());
    // ---- sensors + films
    for (uint32_t i = 0; i < d->n_sensors; ++i) {
        const mtsb_sensor_desc &sd = d->sensors[i];
        CHECK_ARG(sd.film_width > 0 && sd.film_height > 0);
        std::unique_ptr<SensorHost> sh(new SensorHost());
        SensorDev &sn = sh->dev; memset(&sn, 0, sizeof(sn));
        memcpy(sn.to_world, sd.to_world, sizeof(sn.to_world)); memcpy(sn.sample_to_camera, sd.sample_to_camera, sizeof(sn.sample_to_camera));
        sn.near_clip = sd.near_clip; sn.far_clip = sd.far_clip; sn.width = sd.film_width; sn.height = sd.film_height;
        sn.rfilter = sd.rfilter; sn.stddev = sd.rfilter_stddev; sn.filter_c = sd.rfilter_c;
        sn.radius = sd.rfilter == MTSB_RFILTER_BOX ? sd.rfilter_radius + kRayEpsilon : sd.rfilter_radius;
        size_t npx = (size_t) sn.width * sn.height;
        CU(sh->xyza.alloc(npx)); CU(sh->w.alloc(npx)); CU(sh->packed.alloc(5 * npx));
        CU(cudaMemsetAsync(sh->xyza.p, 0, npx * sizeof(float4), st)); CU(cudaMemsetAsync(sh->w.p, 0, npx * sizeof(float), st));
        sn.film_xyza = sh->xyza.p; sn.film_w = sh->w.p;
        s->sensors.push_back(std::move(sh));
    }
    s->dev.max_depth = d->integrator.max_depth < 0 ? (1 << 30) : d->integrator.max_depth;
    s->dev.rr_depth = d->integrator.rr_depth;
    rc = upload_struct_arrays(s.get()); if (rc) return rc;
    CU(s->counters.alloc(kCounterWords)); CU(s->dstats.alloc(4));
    *out = s.release();
    return MTSB_OK;
}

extern "C" int mtsb_scene_stats(mtsb_scene *s, mtsb_stats *out) {
    CHECK_ARG(s && out);
    *out = s->build_stats;
    return MTSB_OK;
}

// ---------------------------------------------------------------------------------------------------
// Batched intersection: Scene::ray_intersect / ray_test / ray_intersect_naive
// ---------------------------------------------------------------------------------------------------
struct Staging {   // host<->device staging for MTSB_MEM_HOST callers
    std::vector<std::unique_ptr<DevBuf<unsigned char>>> bufs;
    template <typename T> int in(const T *src, size_t count, int mem, cudaStream_t st, const T **out) {
        if (mem == MTSB_MEM_DEVICE) { *out = src; return MTSB_OK; }
        bufs.emplace_back(new DevBuf<unsigned char>());
        CU(bufs.back()->alloc(count * sizeof(T)));
        CU(cudaMemcpyAsync(bufs.back()->p, src, count * sizeof(T), cudaMemcpyHostToDevice, st));
        *out = (const T *) bufs.back()->p; return MTSB_OK;
    }
    template <typename T> int out(T *dst, size_t count, int mem, T **dev) {
        if (mem == MTSB_MEM_DEVICE) { *dev = dst; return MTSB_OK; }
        bufs.emplace_back(new DevBuf<unsigned char>());
        CU(bufs.back()->alloc(count * sizeof(T)));
        *dev = (T *) bufs.back()->p; return MTSB_OK;
    }
    template <typename T> int back(T *dst, const T *dev, size_t count, int mem, cudaStream_t st) {
        if (mem == MTSB_MEM_DEVICE) return MTSB_OK;
        CU(cudaMemcpyAsync(dst, dev, count * sizeof(T), cudaMemcpyDeviceToHost, st));
        return MTSB_OK;
    }
};
#define RC(x) do { int rc_ = (x); if (rc_) return rc_; } while (0)

static int intersect_common(mtsb_scene *s, uint64_t n, const float *o_mint, const float *d_maxt, float *hits, uint32_t *occ,
                            int mem, int mode, mtsb_stats *stats) {
    CHECK_ARG(s && (n == 0 || (o_mint && d_maxt)) && n < (1ull << 32));
    CU(cudaSetDevice(s->ctx->device));
    cudaStream_t st = s->ctx->stream;
    if (stats) memset(stats, 0, sizeof(*stats));
    if (n == 0) return MTSB_OK;
    Staging sg; const float *d_o, *d_d; float *d_h = nullptr; uint32_t *d_occ = nullptr;
    RC(sg.in(o_mint, 4 * n, mem, st, &d_o)); RC(sg.in(d_maxt, 4 * n, mem, st, &d_d));
    if (mode == 1) RC(sg.out(occ, n, mem, &d_occ)); else RC(sg.out(hits, 4 * n, mem, &d_h));
    CU(cudaEventRecord(s->ctx->ev0, st));
    uint32_t *work = s->counters.p + kWorkClosest;
    if (mode != 2) CU(cudaMemsetAsync(work, 0, sizeof(uint32_t), st));
    unsigned tg = trace_grid_for(s->ctx, n);
    const bool shp = !s->shapes.empty();      // scenes with analytic leaf primitives run the <.., true> instances
    if (mode == 0 && !shp) k_trace<false, false><<<tg, kBlock, 0, st>>>(s->dev.bvh, nullptr, (uint32_t) n, (const float4 *) d_o, (const float4 *) d_d, (float4 *) d_h, nullptr, work, s->ctx->refill, nullptr);
    else if (mode == 0) k_trace<false, true><<<tg, kBlock, 0, st>>>(s->dev.bvh, nullptr, (uint32_t) n, (const float4 *) d_o, (const float4 *) d_d, (float4 *) d_h, nullptr, work, s->ctx->refill, nullptr);
    else if (mode == 1 && !shp) k_trace<true, false><<<tg, kBlock, 0, st>>>(s->dev.bvh, nullptr, (uint32_t) n, (const float4 *) d_o, (const float4 *) d_d, nullptr, d_occ, work, s->ctx->refill, nullptr);
    else if (mode == 1) k_trace<true, true><<<tg, kBlock, 0, st>>>(s->dev.bvh, nullptr, (uint32_t) n, (const float4 *) d_o, (const float4 *) d_d, nullptr, d_occ, work, s->ctx->refill, nullptr);
    else k_trace_naive<<<grid_for(n), kBlock, 0, st>>>(s->dev.bvh, (uint32_t) n, (const float4 *) d_o, (const float4 *) d_d, (float4 *) d_h);
    CU(cudaEventRecord(s->ctx->ev1, st));
    CU(cudaGetLastError());
    if (mode == 1) RC(sg.back(occ, d_occ, n, mem, st)); else RC(sg.back(hits, d_h, 4 * n, mem, st));
    CU(cudaStreamSynchronize(st));
    if (stats) {
        CU(cudaEventElapsedTime(&stats->ms_total, s->ctx->ev0, s->ctx->ev1));
        stats->kernel_launches = 1;
        if (mode == 1) { stats->shadow_rays = n; stats->ms_any = stats->ms_total; } else { stats->closest_rays = n; stats->ms_closest = stats->ms_total; }
    }
    return MTSB_OK;
}
extern "C" int mtsb_intersect_closest(mtsb_scene *s, uint64_t n, const float *o, const float *d, float *hits, int mem, mtsb_stats *stats) {
    CHECK_ARG(n == 0 || hits); return intersect_common(s, n, o, d, hits, nullptr, mem, 0, stats);
}
extern "C" int mtsb_intersect_any(mtsb_scene *s, uint64_t n, const float *o, const float *d, uint32_t *occ, int mem, mtsb_stats *stats) {
    CHECK_ARG(n == 0 || occ); return intersect_common(s, n, o, d, nullptr, occ, mem, 1, stats);
}
extern "C" int mtsb_intersect_naive(mtsb_scene *s, uint64_t n, const float *o, const float *d, float *hits, int mem) {
    CHECK_ARG(n == 0 || hits); return intersect_common(s, n, o, d, hits, nullptr, mem, 2, nullptr);
}

// ---------------------------------------------------------------------------------------------------
// Wavefront render
// ---------------------------------------------------------------------------------------------------
// Paths per wavefront batch.  Big batches amortise the tail of every persistent traversal launch (measured on C3:
// 2^23 -> 673, 2^24 -> 740, 2^25 -> 777 Msamples/s); 2^25 paths cost ~10 GB of queues out of 180 GB.
static uint32_t pool_limit() {
    const char *e = getenv("MTSB_POOL");
    if (e) { long v = atol(e); if (v >= 1024 && v <= (1l << 28)) return (uint32_t) v; }
    return 1u << 25;
}
static int ensure_pool(mtsb_scene *s, uint32_t n) {
    if (n <= s->pool) return MTSB_OK;
    s->pool = 0;                 // a failed allocation below must not leave a stale size next to released buffers
    for (int k = 0; k < 2; ++k) {
        CU(s->q_o[k].alloc(n)); CU(s->q_d[k].alloc(n)); CU(s->q_hit[k].alloc(n)); CU(s->q_te[k].alloc(n));
        CU(s->q_misc[k].alloc(n)); CU(s->q_rng[k].alloc(n));
        if (s->spectral) CU(s->q_we[k].alloc(n));
    }
    CU(s->s_o.alloc(n)); CU(s->s_d.alloc(n)); CU(s->s_c.alloc(n)); CU(s->result.alloc(n)); CU(s->class_lists.alloc((size_t) kShadeClasses * n)); 
    s->pool = n;
    return MTSB_OK;
}
static PathQueue make_queue(mtsb_scene *s, int k) {
    PathQueue q; q.o_mint = s->q_o[k].p; q.d_maxt = s->q_d[k].p; q.hit = s->q_hit[k].p; q.thr_eta = s->q_te[k].p;
    q.misc = s->q_misc[k].p; q.rng = s->q_rng[k].p; q.wav_eta = s->q_we[k].p; q.prev1 = s->q_p1[k].p; q.prev2 = s->q_p2[k].p; return q;
}

// One batch of the wavefront: raygen -> [closest -> shade -> shadow]* .  `grad` != null selects the adjoint replay
// (k_shade_grad scatters texel gradients); `splat` adds the finished paths to the film.
// the shade kernel has four instances: colour mode x (lean | with the N4 plugins: conductor, plastic, rough dielectric, mask)
// (boundary passes of the gradient, `edge` != null, always take the X instance: it honours the terminator tag of path.h)
#define LAUNCH_SHADE(grid, ...) do { \
    if (s->spectral) { if (s->ext) k_shade<SpectralX><<<(grid), kBlock, 0, st>>>(__VA_ARGS__); else k_shade<Spectral><<<(grid), kBlock, 0, st>>>(__VA_ARGS__); } \
    else { if (s->ext || edge != nullptr) k_shade<RgbX><<<(grid), kBlock, 0, st>>>(__VA_ARGS__); else k_shade<Rgb><<<(grid), kBlock, 0, st>>>(__VA_ARGS__); } } while (0)
struct EdgeLaunch { EdgeCtx e; const float *total_len; };     // primary rays come from edge samples instead of the film
struct Edge2Launch { Edge2Ctx e; const float *total_len; };   // depth-2 rays come from edge samples seen from the primary hits
static int run_batch(mtsb_scene *s, const SensorDev &sn, const RenderBatch &b, const GradCtx *grad, bool splat, uint64_t &launches,
                     const EdgeLaunch *edge = nullptr, const Edge2Launch *edge2 = nullptr) {
    cudaStream_t st = s->ctx->stream;
    const uint32_t n = b.n_paths;
    ShadowQueue sq; sq.o_mint = s->s_o.p; sq.d_maxt = s->s_d.p; sq.contrib_id = s->s_c.p;
    const bool vgrad = grad && grad->grad_pos;
    sq.nee_rec = vgrad ? s->s_nee.p : nullptr; sq.occluded = vgrad ? s->s_occ.p : nullptr;
    int max_iter = s->dev.max_depth < 1000 ? s->dev.max_depth : 1000;
    bool poll = s->dev.max_depth > 16;
    CU(cudaMemsetAsync(s->counters.p, 0, kCounterWords * sizeof(uint32_t), st));
    int cur = 0;
    const unsigned tg = trace_grid_for(s->ctx, n);
    if (edge) { StageTimer t(s->ctx, ST_RAYGEN); k_edge_raygen<<<grid_for(n / 2), kBlock, 0, st>>>(s->dev, sn, edge->e, edge->total_len, b, make_queue(s, cur), s->counters.p); launches++; }
    else
    { StageTimer t(s->ctx, ST_RAYGEN); if (s->spectral) k_raygen<Spectral><<<grid_for(n), kBlock, 0, st>>>(sn, b, make_queue(s, cur), s->counters.p);
      else k_raygen<Rgb><<<grid_for(n), kBlock, 0, st>>>(sn, b, make_queue(s, cur), s->counters.p); } launches++;
    const int depth0 = 1;
    for (int depth = depth0; depth <= max_iter; ++depth) {
        PathQueue qin = make_queue(s, cur), qout = make_queue(s, cur ^ 1);
        { StageTimer t(s->ctx, ST_CLOSEST); if (s->shapes.empty()) k_trace<false, false><<<tg, kBlock, 0, st>>>(s->dev.bvh, s->counters.p + depth, n, qin.o_mint, qin.d_maxt, qin.hit, nullptr, s->counters.p + kWorkClosest + depth, s->ctx->refill, s->prim_class.p);
          else k_trace<false, true><<<tg, kBlock, 0, st>>>(s->dev.bvh, s->counters.p + depth, n, qin.o_mint, qin.d_maxt, qin.hit, nullptr, s->counters.p + kWorkClosest + depth, s->ctx->refill, s->prim_class.p); }
        if (edge2 && depth == edge2->e.depth) {      // the edges are seen from this vertex: one far ray per live path as the next depth's entry
            { StageTimer t(s->ctx, ST_SHADE); k_edge2_raygen<<<grid_for(n), kBlock, 0, st>>>(s->dev, edge2->e, edge2->total_len, b, qin, qout, s->counters.p); }
            launches += 2; cur ^= 1;
            continue;
        }
        { StageTimer t(s->ctx, ST_SHADE);
          if (grad && grad->grad_pos) { if (s->ext) k_shade_grad<true, RgbX><<<grid_for(n), kBlock, 0, st>>>(s->dev, b, qin, qout, sq, s->counters.p, s->dstats.p, (uint32_t) depth, *grad);
                                        else k_shade_grad<true, Rgb><<<grid_for(n), kBlock, 0, st>>>(s->dev, b, qin, qout, sq, s->counters.p, s->dstats.p, (uint32_t) depth, *grad); }
          else if (grad) { if (s->ext) k_shade_grad<false, RgbX><<<grid_for(n), kBlock, 0, st>>>(s->dev, b, qin, qout, sq, s->counters.p, s->dstats.p, (uint32_t) depth, *grad);
                           else k_shade_grad<false, Rgb><<<grid_for(n), kBlock, 0, st>>>(s->dev, b, qin, qout, sq, s->counters.p, s->dstats.p, (uint32_t) depth, *grad); }
          else if ((depth == 1 && !s->spectral && !s->ctx->bin_depth1) || !s->ctx->bin_materials) {
              // identity lists.  (Round 1: binning the coherent primary hits of the rgb kernel by material alone cost 10.85 -> 11.0 ms of
              // shade per C3 step; with the emitter pick in the class it pays: 10.45 -> 10.11 ms, so depth 1 is binned by default now.)
              ShadeBins bins; bins.lists = nullptr; bins.counts = s->counters.p + depth; bins.pool = 0;
              LAUNCH_SHADE(grid_for(n), s->dev, b, qin, qout, sq, s->counters.p, s->dstats.p, (uint32_t) depth, bins);
          } else {
              ShadeBins bins; bins.lists = s->class_lists.p; bins.counts = s->counters.p + kClassCounts + kShadeClasses * depth; bins.pool = s->pool;
              k_classify<<<grid_for(n), kBlock, 0, st>>>(s->counters.p + depth, qin.hit, bins, s->ctx->bin_emitters ? qin.rng : nullptr, depth > s->dev.rr_depth ? 1 : 0, s->dev.n_emitters);
              LAUNCH_SHADE(grid_for((size_t) n + 32 * kShadeClasses), s->dev, b, qin, qout, sq, s->counters.p, s->dstats.p, (uint32_t) depth, bins);
              launches++;
          } }
        launches += 2;
        if (depth < s->dev.max_depth) {
            { StageTimer t(s->ctx, ST_ANY); if (s->shapes.empty()) k_shadow<false><<<tg, kBlock, 0, st>>>(s->dev.bvh, sq, s->counters.p, s->dstats.p, (uint32_t) depth, n, s->result.p, s->counters.p + kWorkShadow + depth, s->ctx->refill);
              else k_shadow<true><<<tg, kBlock, 0, st>>>(s->dev.bvh, sq, s->counters.p, s->dstats.p, (uint32_t) depth, n, s->result.p, s->counters.p + kWorkShadow + depth, s->ctx->refill); }
            launches++;
            if (vgrad) { StageTimer t(s->ctx, ST_SHADE); k_nee_grad<<<grid_for(n), kBlock, 0, st>>>(s->dev, b, qin, sq, s->counters.p, (uint32_t) depth, n, *grad); launches++; }
        }
        cur ^= 1;
        if (poll) {
            uint32_t next = 0;
            CU(cudaMemcpyAsync(&next, s->counters.p + depth + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            if (next == 0) break;
        }
    }
    if (splat) { StageTimer t(s->ctx, ST_FILM); if (s->spectral) k_splat<Spectral><<<grid_for(n), kBlock, 0, st>>>(sn, b); else k_splat<Rgb><<<grid_for(n), kBlock, 0, st>>>(sn, b); launches++; }
    return MTSB_OK;
}

extern "C" int mtsb_render(mtsb_scene *s, uint32_t sensor_id, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total,
                           uint64_t seed, mtsb_stats *stats) {
    CHECK_ARG(s && sensor_id < s->sensors.size() && spp_count > 0 && (uint64_t) spp_begin + spp_count <= (uint64_t) spp_total);
    CU(cudaSetDevice(s->ctx->device));
    s->ctx->cancel = 0;      // like `m_stop = false` at the top of SamplingIntegrator::render (integrator.cpp:54): a stale cancel does not abort this render
    cudaStream_t st = s->ctx->stream;
    SensorDev sn = s->sensors[sensor_id]->dev;
    uint64_t npix = (uint64_t) sn.width * sn.height, total = npix * spp_count;
    uint32_t pool = (uint32_t) (total < pool_limit() ? total : pool_limit());
    RC(ensure_pool(s, pool));
    CU(cudaMemsetAsync(s->dstats.p, 0, 4 * sizeof(unsigned long long), st));
    CU(cudaEventRecord(s->ctx->ev0, st));
    uint64_t launches = 0, started = 0; int nb = 0;
    const auto t_start = std::chrono::steady_clock::now();
    for (uint64_t off = 0; off < total; off += pool) {
        // at most two batches are queued ahead of the GPU, so `cancel` and `timeout` (Integrator::should_stop,
        // include/mitsuba/render/integrator.h:143-146) act within two batch times instead of after the whole job was enqueued
        if (nb >= 2) CU(cudaEventSynchronize(s->ctx->ring[nb & 1]));
        if (s->ctx->cancel) { s->ctx->cancel = 0; cudaStreamSynchronize(st); return fail(MTSB_E_CANCELLED, "render cancelled"); }
        if (s->ctx->timeout > 0.f && std::chrono::duration<float>(std::chrono::steady_clock::now() - t_start).count() > s->ctx->timeout) break;
        uint32_t n = (uint32_t) ((total - off) < pool ? (total - off) : pool);
        RenderBatch b; b.n_paths = n; b.offset = off; b.spp_begin = spp_begin; b.spp_batch = spp_count; b.spp_total = spp_total;
        b.seed = seed; b.result = s->result.p;
        RC(run_batch(s, sn, b, nullptr, true, launches));
        CU(cudaEventRecord(s->ctx->ring[nb & 1], st)); nb++; started += n;
    }
    CU(cudaEventRecord(s->ctx->ev1, st));
    unsigned long long hs[4];
    CU(cudaMemcpyAsync(hs, s->dstats.p, sizeof(hs), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->samples = started; stats->closest_rays = hs[0];      // < total when the timeout stopped the job (the film holds what was rendered) stats->shadow_rays = hs[1]; stats->segments = hs[2];
        stats->kernel_launches = launches;
        CU(cudaEventElapsedTime(&stats->ms_total, s->ctx->ev0, s->ctx->ev1));
    }
    collect_stage_times(s->ctx, stats);
    return MTSB_OK;
}

// ---------------------------------------------------------------------------------------------------
// Film
// ---------------------------------------------------------------------------------------------------
extern "C" int mtsb_film_clear(mtsb_scene *s, uint32_t sensor_id) {
    CHECK_ARG(s && sensor_id < s->sensors.size());
    CU(cudaSetDevice(s->ctx->device));
    SensorHost &sh = *s->sensors[sensor_id]; size_t npx = (size_t) sh.dev.width * sh.dev.height;
    CU(cudaMemsetAsync(sh.xyza.p, 0, npx * sizeof(float4), s->ctx->stream)); CU(cudaMemsetAsync(sh.w.p, 0, npx * sizeof(float), s->ctx->stream));
    if (sh.dev.film_m2) CU(cudaMemsetAsync(sh.m2.p, 0, npx * sizeof(float4), s->ctx->stream));
    CU(cudaStreamSynchronize(s->ctx->stream));
    return MTSB_OK;
}
// `moment` integrator (src/integrators/moment.cpp): second-moment AOVs of the nested integrator's XYZ value
extern "C" int mtsb_film_enable_moments(mtsb_scene *s, uint32_t sensor_id, int on) {
    CHECK_ARG(s && sensor_id < s->sensors.size());
    CU(cudaSetDevice(s->ctx->device));
    SensorHost &sh = *s->sensors[sensor_id]; size_t npx = (size_t) sh.dev.width * sh.dev.height;
    if (on) {
        CU(sh.m2.alloc(npx));
        CU(cudaMemsetAsync(sh.m2.p, 0, npx * sizeof(float4), s->ctx->stream)); CU(cudaStreamSynchronize(s->ctx->stream));
        sh.dev.film_m2 = sh.m2.p;
    } else sh.dev.film_m2 = nullptr;
    return MTSB_OK;
}
extern "C" int mtsb_film_read_moments(mtsb_scene *s, uint32_t sensor_id, float *m2_xyz, int mem) {
    CHECK_ARG(s && sensor_id < s->sensors.size() && m2_xyz);
    CU(cudaSetDevice(s->ctx->device));
    SensorHost &sh = *s->sensors[sensor_id]; size_t npx = (size_t) sh.dev.width * sh.dev.height;
    if (!sh.dev.film_m2) return fail(MTSB_E_INVALID, "second moments are not enabled for this sensor (mtsb_film_enable_moments)");
    Staging sg; float *d_out; RC(sg.out(m2_xyz, 3 * npx, mem, &d_out));
    k_film_pack_m2<<<grid_for(npx), kBlock, 0, s->ctx->stream>>>(sh.dev, d_out);
    RC(sg.back(m2_xyz, d_out, 3 * npx, mem, s->ctx->stream));
    CU(cudaStreamSynchronize(s->ctx->stream));
    return MTSB_OK;
}
extern "C" int mtsb_film_device_ptr(mtsb_scene *s, uint32_t sensor_id, void **ptr, uint64_t *n_floats) {
    CHECK_ARG(s && sensor_id < s->sensors.size() && ptr && n_floats);
    CU(cudaSetDevice(s->ctx->device));
    SensorHost &sh = *s->sensors[sensor_id]; size_t npx = (size_t) sh.dev.width * sh.dev.height;
    k_film_pack<<<grid_for(npx), kBlock, 0, s->ctx->stream>>>(sh.dev, sh.packed.p);
    CU(cudaStreamSynchronize(s->ctx->stream));
    *ptr = sh.packed.p; *n_floats = 5 * npx;
    return MTSB_OK;
}
extern "C" int mtsb_film_read(mtsb_scene *s, uint32_t sensor_id, float *xyzaw, int mem) {
    CHECK_ARG(s && sensor_id < s->sensors.size() && xyzaw);
    CU(cudaSetDevice(s->ctx->device));
    SensorHost &sh = *s->sensors[sensor_id]; size_t npx = (size_t) sh.dev.width * sh.dev.height;
    if (mem == MTSB_MEM_DEVICE) k_film_pack<<<grid_for(npx), kBlock, 0, s->ctx->stream>>>(sh.dev, xyzaw);      // straight into the caller's buffer
    else {
        k_film_pack<<<grid_for(npx), kBlock, 0, s->ctx->stream>>>(sh.dev, sh.packed.p);
        CU(cudaMemcpyAsync(xyzaw, sh.packed.p, 5 * npx * sizeof(float), cudaMemcpyDeviceToHost, s->ctx->stream));
    }
    CU(cudaStreamSynchronize(s->ctx->stream));
    return MTSB_OK;
}
static int film_write(mtsb_scene *s, uint32_t sensor_id, const float *xyzaw, int mem, int replace) {
    CHECK_ARG(s && sensor_id < s->sensors.size() && xyzaw);
    CU(cudaSetDevice(s->ctx->device));
    SensorHost &sh = *s->sensors[sensor_id]; size_t npx = (size_t) sh.dev.width * sh.dev.height;
    if ((const void *) xyzaw != (const void *) sh.packed.p)
        CU(cudaMemcpyAsync(sh.packed.p, xyzaw, 5 * npx * sizeof(float), mem == MTSB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, s->ctx->stream));
    k_film_unpack_add<<<grid_for(npx), kBlock, 0, s->ctx->stream>>>(sh.dev, sh.packed.p, replace);
    CU(cudaStreamSynchronize(s->ctx->stream));
    return MTSB_OK;
}
extern "C" int mtsb_film_accumulate(mtsb_scene *s, uint32_t sensor_id, const float *xyzaw, int mem) { return film_write(s, sensor_id, xyzaw, mem, 0); }
extern "C" int mtsb_film_develop(mtsb_scene *s, uint32_t sensor_id, float *rgb, int mem) {
    CHECK_ARG(s && sensor_id < s->sensors.size() && rgb);
    CU(cudaSetDevice(s->ctx->device));
    SensorHost &sh = *s->sensors[sensor_id]; size_t npx = (size_t) sh.dev.width * sh.dev.height;
    Staging sg; float *d_rgb; RC(sg.out(rgb, 3 * npx, mem, &d_rgb));
    k_film_develop<<<grid_for(npx), kBlock, 0, s->ctx->stream>>>(sh.dev, d_rgb);
    RC(sg.back(rgb, d_rgb, 3 * npx, mem, s->ctx->stream));
    CU(cudaStreamSynchronize(s->ctx->stream));
    return MTSB_OK;
}
extern "C" int mtsb_film_put(mtsb_scene *s, uint32_t sensor_id, uint64_t n, const float *pos2, const float *value5) {
    CHECK_ARG(s && sensor_id < s->sensors.size() && (n == 0 || (pos2 && value5)));
    CU(cudaSetDevice(s->ctx->device));
    if (n == 0) return MTSB_OK;
    cudaStream_t st = s->ctx->stream; Staging sg; const float *dp, *dv;
    RC(sg.in(pos2, 2 * n, MTSB_MEM_HOST, st, &dp)); RC(sg.in(value5, 5 * n, MTSB_MEM_HOST, st, &dv));
    k_film_put<<<grid_for(n), kBlock, 0, st>>>(s->sensors[sensor_id]->dev, (uint32_t) n, dp, dv);
    CU(cudaStreamSynchronize(st)); CU(cudaGetLastError());
    return MTSB_OK;
}

// ---------------------------------------------------------------------------------------------------
// Parameter updates (Mesh::parameters_changed, BitmapTexture::parameters_changed)
// ---------------------------------------------------------------------------------------------------
extern "C" int mtsb_scene_update_vertices(mtsb_scene *s, uint32_t mesh_id, const float *positions, int mem) {
    CHECK_ARG(s && mesh_id < s->n_tri_meshes && positions);
    CU(cudaSetDevice(s->ctx->device));
    cudaStream_t st = s->ctx->stream;
    const MeshDev &m = s->meshes[mesh_id];
    const size_t nf = 3 * (size_t) m.vert_count;
    float *dp = s->pos.p + 3 * (size_t) m.vert_begin, *hp = m.emitter >= 0 ? s->emitter_pos[(int) mesh_id].data() : nullptr;
    // positions go straight to the device array (device callers: one D2D copy, no host round trip); the host copy is only
    // refreshed for emitter meshes, whose area CDF is rebuilt on the host in double like the reference (distr_1d.h:49-88)
    CU(cudaMemcpyAsync(dp, positions, nf * sizeof(float), mem == MTSB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    if (hp && mem == MTSB_MEM_HOST) memcpy(hp, positions, nf * sizeof(float));
    else if (hp) { CU(cudaMemcpyAsync(hp, positions, nf * sizeof(float), cudaMemcpyDeviceToHost, st)); CU(cudaStreamSynchronize(st)); }
    if (m.has_normals) {      // Mesh::parameters_changed -> recompute_vertex_normals (mesh.cpp:794-804, 199-276), on the device
        float *dn = s->nrm.p + 3 * (size_t) m.vert_begin;
        CU(cudaMemsetAsync(dn, 0, nf * sizeof(float), st));
        k_vnormal_faces<<<grid_for(m.face_count), kBlock, 0, st>>>(s->pos.p, s->faces.p, m.face_begin, m.face_count, s->nrm.p);
        k_vnormal_normalize<<<grid_for(m.vert_count), kBlock, 0, st>>>(dn, m.vert_count);
    }
    const bool refit = s->update_refit && (s->rebuild_every == 0 || s->refits_since_build + 1 < s->rebuild_every);
    if (refit) RC(refit_bvh(s)); else RC(build_bvh(s));
    if (m.emitter >= 0) RC(build_area_tables(s));
    update_env_bsphere(s);
    if (refit && m.emitter < 0) {        // nothing but the emitters' bounding sphere changed in the tables
        CU(s->d_emitters.upload(s->emitters.data(), s->emitters.size(), st));
        CU(cudaStreamSynchronize(st));
        s->dev.emitters = s->d_emitters.p;
        return MTSB_OK;
    }
    return upload_struct_arrays(s);
}

extern "C" int mtsb_scene_update_config(mtsb_scene *s, int refit, uint32_t rebuild_every) {
    CHECK_ARG(s);
    s->update_refit = refit ? 1 : 0; s->rebuild_every = rebuild_every;
    return MTSB_OK;
}

extern "C" int mtsb_scene_update_texture(mtsb_scene *s, uint32_t tex_id, const float *data, int mem) {
    CHECK_ARG(s && tex_id < s->tex.size() && data);
    CU(cudaSetDevice(s->ctx->device));
    TexDev &t = s->tex[tex_id];
    if (t.type == MTSB_TEX_CONSTANT) {
        float v[3];
        if (mem == MTSB_MEM_DEVICE) CU(cudaMemcpy(v, data, sizeof(v), cudaMemcpyDeviceToHost)); else memcpy(v, data, sizeof(v));
        t.value = mk3(v[0], v[1], v[2]);
        return upload_struct_arrays(s);
    }
    size_t n = (size_t) t.w * t.h * t.ch;
    if (t.type == MTSB_TEX_MESH_ATTRIBUTE) n = (size_t) t.w * t.ch;
    else if (t.type != MTSB_TEX_BITMAP)      // tabulated / uniform spectra carry derived host-side tables (bucket index): recreate the scene instead
        return fail(MTSB_E_INVALID, "mtsb_scene_update_texture: only constant, bitmap and mesh_attribute textures can be updated in place");
    CU(cudaMemcpyAsync((float *) t.data, data, n * sizeof(float), mem == MTSB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, s->ctx->stream));
    CU(cudaStreamSynchronize(s->ctx->stream));
    return MTSB_OK;
}

// ---------------------------------------------------------------------------------------------------
// Parity hooks (host pointers)
// ---------------------------------------------------------------------------------------------------
extern "C" int mtsb_pcg32_floats(mtsb_ctx *c, uint64_t n, const uint64_t *initstate, const uint64_t *initseq, uint32_t n_draws, float *out) {
    CHECK_ARG(c && (n == 0 || (initstate && initseq && out)));
    CU(cudaSetDevice(c->device));
    if (n == 0 || n_draws == 0) return MTSB_OK;
    Staging sg; const uint64_t *ds, *dq; float *dout;
    RC(sg.in(initstate, n, 0, c->stream, &ds)); RC(sg.in(initseq, n, 0, c->stream, &dq)); RC(sg.out(out, n * n_draws, 0, &dout));
    k_pcg32<<<grid_for(n), kBlock, 0, c->stream>>>((uint32_t) n, ds, dq, n_draws, dout);
    RC(sg.back(out, dout, n * n_draws, 0, c->stream));
    CU(cudaStreamSynchronize(c->stream)); CU(cudaGetLastError());
    return MTSB_OK;
}
extern "C" int mtsb_sampler_seed(mtsb_ctx *c, uint64_t n, uint64_t seed, uint64_t *state, uint64_t *inc) {
    CHECK_ARG(c && (n == 0 || (state && inc)));
    CU(cudaSetDevice(c->device));
    if (n == 0) return MTSB_OK;
    Staging sg; uint64_t *ds, *di;
    RC(sg.out(state, n, 0, &ds)); RC(sg.out(inc, n, 0, &di));
    k_sampler_seed<<<grid_for(n), kBlock, 0, c->stream>>>((uint32_t) n, seed, ds, di);
    RC(sg.back(state, ds, n, 0, c->stream)); RC(sg.back(inc, di, n, 0, c->stream));
    CU(cudaStreamSynchronize(c->stream)); CU(cudaGetLastError());
    return MTSB_OK;
}
extern "C" int mtsb_sensor_sample_ray(mtsb_scene *s, uint32_t sensor_id, uint64_t n, const float *pos2, float *o_mint, float *d_maxt) {
    CHECK_ARG(s && sensor_id < s->sensors.size() && (n == 0 || (pos2 && o_mint && d_maxt)));
    CU(cudaSetDevice(s->ctx->device));
    if (n == 0) return MTSB_OK;
    cudaStream_t st = s->ctx->stream; Staging sg; const float *dp; float *d_o, *d_d;
    RC(sg.in(pos2, 2 * n, 0, st, &dp)); RC(sg.out(o_mint, 4 * n, 0, &d_o)); RC(sg.out(d_maxt, 4 * n, 0, &d_d));
    k_sensor_ray<<<grid_for(n), kBlock, 0, st>>>(s->sensors[sensor_id]->dev, (uint32_t) n, dp, (float4 *) d_o, (float4 *) d_d);
    RC(sg.back(o_mint, d_o, 4 * n, 0, st)); RC(sg.back(d_maxt, d_d, 4 * n, 0, st));
    CU(cudaStreamSynchronize(st)); CU(cudaGetLastError());
    return MTSB_OK;
}
extern "C" int mtsb_bsdf_sample(mtsb_scene *s, uint32_t bsdf_id, uint64_t n, const float *wi, const float *uv, const float *s1, const float *s2, float *out9) {
    CHECK_ARG(s && bsdf_id < s->bsdfs.size() && (n == 0 || (wi && uv && s1 && s2 && out9)));
    CU(cudaSetDevice(s->ctx->device));
    if (n == 0) return MTSB_OK;
    cudaStream_t st = s->ctx->stream; Staging sg; const float *dwi, *duv, *ds1, *ds2; float *dout;
    RC(sg.in(wi, 3 * n, 0, st, &dwi)); RC(sg.in(uv, 2 * n, 0, st, &duv)); RC(sg.in(s1, n, 0, st, &ds1)); RC(sg.in(s2, 2 * n, 0, st, &ds2));
    RC(sg.out(out9, 9 * n, 0, &dout));
    k_bsdf_sample<<<grid_for(n), kBlock, 0, st>>>(s->dev, bsdf_id, (uint32_t) n, dwi, duv, ds1, ds2, dout);
    RC(sg.back(out9, dout, 9 * n, 0, st));
    CU(cudaStreamSynchronize(st)); CU(cudaGetLastError());
    return MTSB_OK;
}
extern "C" int mtsb_bsdf_eval_pdf(mtsb_scene *s, uint32_t bsdf_id, uint64_t n, const float *wi, const float *uv, const float *wo, float *eval3, float *pdf) {
    CHECK_ARG(s && bsdf_id < s->bsdfs.size() && (n == 0 || (wi && uv && wo && eval3 && pdf)));
    CU(cudaSetDevice(s->ctx->device));
    if (n == 0) return MTSB_OK;
    cudaStream_t st = s->ctx->stream; Staging sg; const float *dwi, *duv, *dwo; float *de, *dp;
    RC(sg.in(wi, 3 * n, 0, st, &dwi)); RC(sg.in(uv, 2 * n, 0, st, &duv)); RC(sg.in(wo, 3 * n, 0, st, &dwo));
    RC(sg.out(eval3, 3 * n, 0, &de)); RC(sg.out(pdf, n, 0, &dp));
    k_bsdf_eval_pdf<<<grid_for(n), kBlock, 0, st>>>(s->dev, bsdf_id, (uint32_t) n, dwi, duv, dwo, de, dp);
    RC(sg.back(eval3, de, 3 * n, 0, st)); RC(sg.back(pdf, dp, n, 0, st));
    CU(cudaStreamSynchronize(st)); CU(cudaGetLastError());
    return MTSB_OK;
}
extern "C" int mtsb_sample_emitter_direction(mtsb_scene *s, uint64_t n, const float *ref_p, const float *s2, float *out11) {
    CHECK_ARG(s && (n == 0 || (ref_p && s2 && out11)));
    CU(cudaSetDevice(s->ctx->device));
    if (n == 0) return MTSB_OK;
    cudaStream_t st = s->ctx->stream; Staging sg; const float *dp, *ds2; float *dout;
    RC(sg.in(ref_p, 3 * n, 0, st, &dp)); RC(sg.in(s2, 2 * n, 0, st, &ds2)); RC(sg.out(out11, 11 * n, 0, &dout));
    k_sample_emitter<<<grid_for(n), kBlock, 0, st>>>(s->dev, (uint32_t) n, dp, ds2, dout);
    RC(sg.back(out11, dout, 11 * n, 0, st));
    CU(cudaStreamSynchronize(st)); CU(cudaGetLastError());
    return MTSB_OK;
}

// ---------------------------------------------------------------------------------------------------
// Allocation-free closest-hit entry point (mtsb_intersect_closest_noalloc).  No counterpart in the reference: added
// for callers that must not call cudaMalloc (CUDA graph capture, tight serving
// loops).  It runs the same k_trace<false> as mtsb_intersect_closest; everything that kernel and the host-pointer
// staging need beyond the caller's buffers lives in ONE file-scope device array, so the call makes no cudaMalloc /
// cudaFree / allocating library call.
//   layout of g_noalloc_scratch: [0, 256) the dynamic-fetch work counter of the persistent traversal (one uint32, must be
//   zero at launch: cleared by cudaMemsetAsync on the context's stream in every call); then, used for MTSB_MEM_HOST
//   callers only, o_mint | d_maxt | hits, kNoAllocMaxRays float4 each.
// Every context and scene of one device shares the array and it keeps the last call's contents: two calls that overlap
// on the GPU (different contexts / streams on one device, or different host threads) race on it.
// ---------------------------------------------------------------------------------------------------
constexpr uint64_t kNoAllocMaxRays = 1ull << 18;                      // 262 144 rays per call
constexpr size_t kNoAllocHeader = 256;
constexpr size_t kNoAllocBytes = kNoAllocHeader + 3 * kNoAllocMaxRays * sizeof(float4);      // 12 MiB + 256 B per device
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];
// In host code `g_noalloc_scratch` is only a handle of the symbol, not a device address: resolve it once per device, with
// that device current, and hand the resolved pointer to the kernel / copies.
constexpr int kNoAllocMaxDevices = 64;
static unsigned char *g_noalloc_ptr[kNoAllocMaxDevices];             // zero-initialised: not resolved yet
static int noalloc_scratch(int device, unsigned char **out) {
    if (device < 0 || device >= kNoAllocMaxDevices) return fail(MTSB_E_INVALID, "mtsb_intersect_closest_noalloc: device index beyond the scratch table");
    if (g_noalloc_ptr[device] == nullptr) {
        CU(cudaSetDevice(device));
        void *p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (p == nullptr) return fail(MTSB_E_CUDA, "cudaGetSymbolAddress returned a null scratch address");
        g_noalloc_ptr[device] = (unsigned char *) p;
    }
    *out = g_noalloc_ptr[device];
    return MTSB_OK;
}
extern "C" uint64_t mtsb_noalloc_max_rays(void) { return kNoAllocMaxRays; }
extern "C" int mtsb_intersect_closest_noalloc(mtsb_scene *s, uint64_t n, const float *o_mint, const float *d_maxt, float *hits,
                                              int mem, mtsb_stats *stats) {
    CHECK_ARG(s && (n == 0 || (o_mint && d_maxt && hits)) && (mem == MTSB_MEM_HOST || mem == MTSB_MEM_DEVICE));
    if (n > kNoAllocMaxRays)
        return fail(MTSB_E_INVALID, "mtsb_intersect_closest_noalloc: " + std::to_string(n) + " rays exceed the fixed scratch capacity of " +
                                    std::to_string(kNoAllocMaxRays) + " rays per call; split the batch or use mtsb_intersect_closest");
    CU(cudaSetDevice(s->ctx->device));
    cudaStream_t st = s->ctx->stream;
    if (stats) memset(stats, 0, sizeof(*stats));
    if (n == 0) return MTSB_OK;
    unsigned char *scratch = nullptr;
    RC(noalloc_scratch(s->ctx->device, &scratch));
    uint32_t *work = (uint32_t *) scratch;
    const float4 *d_o = (const float4 *) o_mint, *d_d = (const float4 *) d_maxt; float4 *d_h = (float4 *) hits;
    if (mem == MTSB_MEM_HOST) {        // stage through the fixed array (plain copies; nothing is allocated)
        float4 *so = (float4 *) (scratch + kNoAllocHeader), *sd = so + kNoAllocMaxRays, *sh = sd + kNoAllocMaxRays;
        CU(cudaMemcpyAsync(so, o_mint, n * sizeof(float4), cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(sd, d_maxt, n * sizeof(float4), cudaMemcpyHostToDevice, st));
        d_o = so; d_d = sd; d_h = sh;
    }
    CU(cudaMemsetAsync(work, 0, kNoAllocHeader, st));          // the fetch cursor must start at zero; the previous call left it at >= its n
    CU(cudaEventRecord(s->ctx->ev0, st));
    if (s->shapes.empty()) k_trace<false, false><<<trace_grid_for(s->ctx, n), kBlock, 0, st>>>(s->dev.bvh, nullptr, (uint32_t) n, d_o, d_d, d_h, nullptr, work, s->ctx->refill, nullptr);
    else k_trace<false, true><<<trace_grid_for(s->ctx, n), kBlock, 0, st>>>(s->dev.bvh, nullptr, (uint32_t) n, d_o, d_d, d_h, nullptr, work, s->ctx->refill, nullptr);
    CU(cudaEventRecord(s->ctx->ev1, st));
    CU(cudaGetLastError());
    if (mem == MTSB_MEM_HOST) CU(cudaMemcpyAsync(hits, d_h, n * sizeof(float4), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (stats) {
        CU(cudaEventElapsedTime(&stats->ms_total, s->ctx->ev0, s->ctx->ev1));
        stats->kernel_launches = 1; stats->closest_rays = n; stats->ms_closest = stats->ms_total;
    }
    return MTSB_OK;
}

// ---------------------------------------------------------------------------------------------------
// Gradient pass (implemented in grad.cuh)
// ---------------------------------------------------------------------------------------------------
#include "grad.cuh"
