// kernels.cuh -- __global__ wrappers around the HD stage bodies (device only).
#pragma once
#include "adjoint.h"

namespace mtsb {

constexpr int kBlock = 128;
constexpr int kStackFast = 8, kStackSlow = 56;

// ---------------- BVH build ----------------------------------------------------------------------
__device__ __forceinline__ void atomic_min_f(float *a, float v) {
    // monotone int mapping of floats
    int iv = __float_as_int(v);
    if (iv >= 0) atomicMin((int *) a, iv); else atomicMax((unsigned int *) a, (unsigned int) iv);
}
__device__ __forceinline__ void atomic_max_f(float *a, float v) {
    int iv = __float_as_int(v);
    if (iv >= 0) atomicMax((int *) a, iv); else atomicMin((unsigned int *) a, (unsigned int) iv);
}
__global__ void k_tri_bounds(BvhBuild b, float *scene_bounds /* lo xyz, hi xyz */) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float3 lo = mk3(INFINITY), hi = mk3(-INFINITY);
    if (i < b.n) tri_bounds_body(b, i, lo, hi);
    for (int o = 16; o > 0; o >>= 1) {
        lo.x = fminf(lo.x, __shfl_xor_sync(0xffffffffu, lo.x, o)); lo.y = fminf(lo.y, __shfl_xor_sync(0xffffffffu, lo.y, o));
        lo.z = fminf(lo.z, __shfl_xor_sync(0xffffffffu, lo.z, o)); hi.x = fmaxf(hi.x, __shfl_xor_sync(0xffffffffu, hi.x, o));
        hi.y = fmaxf(hi.y, __shfl_xor_sync(0xffffffffu, hi.y, o)); hi.z = fmaxf(hi.z, __shfl_xor_sync(0xffffffffu, hi.z, o));
    }
    if ((threadIdx.x & 31) == 0 && lo.x <= hi.x) {
        atomic_min_f(scene_bounds + 0, lo.x); atomic_min_f(scene_bounds + 1, lo.y); atomic_min_f(scene_bounds + 2, lo.z);
        atomic_max_f(scene_bounds + 3, hi.x); atomic_max_f(scene_bounds + 4, hi.y); atomic_max_f(scene_bounds + 5, hi.z);
    }
}
__global__ void k_morton(BvhBuild b) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < b.n) morton_body(b, i);
}
__global__ void k_lbvh_nodes(BvhBuild b) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 < b.n) lbvh_node_body(b, (int) i);
}
__global__ void k_lbvh_fit(BvhBuild b) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < b.n) lbvh_fit_body(b, (int) i);
}
// PLOC passes (bodies in bvh.h)
__global__ void k_ploc_init(BvhBuild b, int *cid) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < b.n) ploc_init_body(b, cid, i);
}
__global__ void k_ploc_nearest(BvhBuild b, const int *cid, uint32_t m, int *nn) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) ploc_nearest_body(b, cid, m, nn, i);
}
__global__ void k_ploc_merge(BvhBuild b, const int *cid, const int *nn, uint32_t m, int *cid_out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) ploc_merge_body(b, cid, nn, cid_out, i);
}
__global__ void k_ploc_finalize_internal(BvhBuild b) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 < b.n) ploc_finalize_internal_body(b, (int) i);
}
__global__ void k_ploc_finalize_leaf(BvhBuild b, uint32_t *order_out, int *leaf_new) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < b.n) ploc_finalize_leaf_body(b, order_out, leaf_new, i);
}
__global__ void k_ploc_patch_leaf(BvhBuild b, const int *leaf_new) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < b.n) ploc_patch_leaf_body(b, leaf_new, i);
}
__global__ void k_ploc_leaf_bounds(BvhBuild b, const uint32_t *order_out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < b.n) ploc_leaf_bounds_body(b, order_out, i);
}
struct PlocAlive { __host__ __device__ bool operator()(int v) const { return v >= 0; } };

__global__ void k_collapse(BvhBuild b, uint32_t begin, uint32_t end) {
    uint32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < end) collapse_body(b, i);
}

__global__ void k_refit_tris(BvhRefit r, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) refit_tri_body(r, i);
}
__global__ void k_refit_level(BvhRefit r, uint32_t begin, uint32_t end) {
    uint32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < end) refit_node_body(r, i);
}

// rough plastic tables (shade.h): thread i < 64: external transmittance at cosine mu_i; 64 <= i < 128: internal reflectance
__global__ void k_rough_tables(int distr_type, float alpha, float eta, const float *n32, const float *w32, const float *n128, const float *w128, float *out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * kRoughTransRes) return;
    bool transmit = i < (uint32_t) kRoughTransRes;
    float e = transmit ? eta : 1.f / eta;
    int res = rough_table_res(e);
    out[i] = rough_boundary_integral(distr_type, alpha, e, rough_table_mu((int) (i % kRoughTransRes)), res == 32 ? n32 : n128, res == 32 ? w32 : w128, res, transmit);
}

// faces of one mesh (3 x uint32, indices local to the mesh) -> (i0, i1, i2, mesh) with global vertex indices
__global__ void k_faces_expand(const uint32_t *__restrict__ raw, uint32_t n, uint32_t vbase, uint32_t mesh, uint4 *out) {
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < n) out[f] = make_uint4(raw[3 * f] + vbase, raw[3 * f + 1] + vbase, raw[3 * f + 2] + vbase, mesh);
}

// ---------------- Mesh::recompute_vertex_normals on the device (src/librender/mesh.cpp:199-276) ------------------------
// Angle-weighted face normals (Thuermer & Wuethrich 1998) scattered to the three corners with float atomics -- the
// reference's own vectorised branch does the same with scatter_add (:258-275); its scalar branch accumulates in float in
// face order, so the two differ by float association only (tests: <= 2e-6 per component).
__global__ void k_vnormal_faces(const float *__restrict__ pos, const uint4 *__restrict__ faces, uint32_t face_begin, uint32_t face_count, float *nrm) {
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= face_count) return;
    uint4 fi = faces[face_begin + f];
    uint32_t idx[3] = { fi.x, fi.y, fi.z };
    float3 v[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) v[k] = mk3(pos[3 * idx[k]], pos[3 * idx[k] + 1], pos[3 * idx[k] + 2]);
    float3 n = cross(v[1] - v[0], v[2] - v[0]);
    float l2 = dot(n, n);
    if (!(l2 > 0.f)) return;
    n = n * (1.f / sqrtf(l2));
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float3 d0 = normalize(v[(k + 1) % 3] - v[k]), d1 = normalize(v[(k + 2) % 3] - v[k]);
        float dp = dot(d0, d1);                                   // enoki unit_angle: robust angle between unit vectors
        float3 df = d1 - (dp < 0.f ? -d0 : d0);
        float temp = 2.f * asinf(fminf(1.f, 0.5f * norm(df)));
        float ang = dp >= 0.f ? temp : kPi - temp;
        atomicAdd(&nrm[3 * idx[k]], n.x * ang); atomicAdd(&nrm[3 * idx[k] + 1], n.y * ang); atomicAdd(&nrm[3 * idx[k] + 2], n.z * ang);
    }
}
__global__ void k_vnormal_normalize(float *nrm, uint32_t count) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float3 n = mk3(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]);
    float l = norm(n);
    n = l != 0.f ? n * (1.f / l) : mk3(1.f, 0.f, 0.f);          // "some bogus value" (mesh.cpp:247)
    nrm[3 * i] = n.x; nrm[3 * i + 1] = n.y; nrm[3 * i + 2] = n.z;
}

// ---------------- traversal ----------------------------------------------------------------------
// Persistent-warp traversal with dynamic ray fetch (after Aila & Laine 2009): the grid is sized to fill the SMs once;
// every warp claims chunks of kFetchChunk consecutive rays from a global work counter and hands them to lanes whose
// previous ray has finished as soon as fewer than `refill` lanes are still traversing, so short rays do not leave
// lanes idle until the longest ray of the warp is done.  `Sink` consumes a finished ray.
#ifndef MTSB_FETCH_CHUNK
#define MTSB_FETCH_CHUNK 64
#endif
constexpr uint32_t kFetchChunk = MTSB_FETCH_CHUNK;
#ifndef MTSB_TRACE_BLOCKS
#define MTSB_TRACE_BLOCKS 7
#endif
constexpr int kTraceBlocksPerSM = MTSB_TRACE_BLOCKS;
#ifndef MTSB_SHADE_BLOCKS
#define MTSB_SHADE_BLOCKS 8      // re-measured at the end of round 1 (ms of shade per C3 step): 6: 12.81, 7: 12.43, 8: 12.31, 9: 12.40, 10: 12.52, 12: 12.69
#endif
constexpr int kShadeBlocksPerSM = MTSB_SHADE_BLOCKS;      // resident 128-thread blocks the shade kernel is compiled for
// Inside a warp the two halves of a traversal step are scheduled differently:
//   * node test: one lane = one ray (every lane whose ray has no triangle candidates pending pops one child of its
//     current node group and intersects its 8 quantised boxes) -- ~28 of 32 lanes busy;
//   * triangle test: WARP-COOPERATIVE.  A node test leaves 0..24 candidate triangles per ray (~0.7 on average), so a
//     per-ray triangle loop runs at ~5 of 32 lanes and was half of all issued instructions (ncu, profiles/r1c_*).  Instead
//     the lanes publish their (ray lane, candidate) pairs into a per-warp list in shared memory, every lane -- idle ones
//     too -- tests ONE pair against the owner's ray (its Woop row vectors live in shared memory), the closest accepted hit
//     per ray is resolved by a 64-bit shared atomicMin on (t, pair index) and read back by the owning lane, which updates
//     tmax before its next node test (so node culling sees the same tmax as with a serial loop).
// Result of a ray = closest t among all accepted candidates, ties resolved by pair order: the same set a serial loop accepts.
constexpr int kCoopCap = 96;              // candidate pairs a warp tests per iteration (3 passes of 32); the rest stay pending
struct __align__(16) WarpCoop {
    float4 ray[4][32];                    // per lane: (o, mint) (rx, tmax) (ry, -) (rz, -)
    float4 hitrec[32];                    // (t, u, v, prim) of the hit found by each TESTING lane in the current pass
    unsigned long long best[32];          // per owning lane: min over accepted hits of (ordered t bits << 32 | testing lane)
    uint32_t items[kCoopCap];             // owner lane << 5 | candidate bit
};
__device__ __forceinline__ uint32_t float_order_bits(float t) {      // monotone float -> uint map
    uint32_t b = __float_as_uint(t);
    return b ^ ((uint32_t) ((int32_t) b >> 31) | 0x80000000u);
}
// out of line: keeps the double-precision quadratic solve out of the triangle loop's register budget
// (results by value: reference parameters of a non-inlined call would force a stack frame on the caller)
__device__ __noinline__ float4 shape_intersect_slot(const ShapeDev *sh, float3 o, float3 d, float mint, float maxt) {
    float t, u, v;
    bool ok = shape_intersect(*sh, o, d, mint, maxt, t, u, v);
    return make_float4(t, u, v, ok ? 1.f : 0.f);
}
// Shapes: the scene has analytic leaf primitives.  A template parameter rather than a run-time test: the mere presence of the
// out-of-line call in the triangle loop cost the triangle-only instance 4 % (22.3 -> 23.1 ms per C3 step, measured).
template <bool AnyHit, bool Shapes, typename Sink>
__device__ __forceinline__ void trace_persistent(const Bvh &bvh, uint32_t count, const float4 *__restrict__ o_mint,
                                                 const float4 *__restrict__ d_maxt, uint32_t *work, uint32_t refill, Sink sink) {
    __shared__ uint2 s_stack[kStackFast * kBlock];
    __shared__ WarpCoop s_coop[kBlock / 32];
    // decode tables of the node test (LSU pipe instead of ALU bit arithmetic): slot mask -> inner children in near-to-far
    // bit order for each ray octant, slot mask -> 3 candidate bits per slot
    __shared__ uint8_t s_perm[8 * 256];
    __shared__ uint32_t s_spread[256];
    for (uint32_t i = threadIdx.x; i < 8u * 256u; i += blockDim.x) s_perm[i] = (uint8_t) perm_xor8(i & 255u, i >> 8);
    for (uint32_t i = threadIdx.x; i < 256u; i += blockDim.x) s_spread[i] = spread3x8(i);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    WarpCoop &ws = s_coop[threadIdx.x >> 5];
    uint2 slow[kStackSlow];
    int sp = 0;
    float3 o = mk3(0.f), idir = mk3(0.f); float mint = 0.f, tmax = 0.f; uint32_t octinv = 0; uint2 G = make_uint2(0u, 0u);
    uint32_t tm = 0u, tb = 0u, valid = 0u;   // pending triangle candidates of the current node: bit mask, first triangle, valid24
    Hit hit; hit.t = INFINITY; hit.u = hit.v = 0.f; hit.prim = 0xffffffffu; bool found = false;
    bool active = false, exhausted = (count == 0 || bvh.n_tris == 0);
    uint32_t ray = 0, w_next = 0, w_end = 0;
    if (bvh.n_tris == 0) {      // empty scene: every ray misses
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) sink(i, false, hit);
        return;
    }
    ws.best[lane] = ~0ull;
    __syncwarp();
    while (true) {
        uint32_t idle = __ballot_sync(0xffffffffu, !active);
        if (idle && !exhausted) {
            if (w_next == w_end) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(work, kFetchChunk);
                base = __shfl_sync(0xffffffffu, base, 0);
                w_next = min(base, count); w_end = min(base + kFetchChunk, count);
                if (w_next == w_end) exhausted = true;
            }
            uint32_t take = min((uint32_t) __popc(idle), w_end - w_next);
            uint32_t r = (uint32_t) __popc(idle & lt);
            if (!active && r < take) {
                ray = w_next + r;
                float4 ro = o_mint[ray], rd = d_maxt[ray];
                const float big = 1e30f;
                float3 d = xyz(rd);
                o = xyz(ro); mint = ro.w; tmax = rd.w;
                idir = mk3(fabsf(d.x) > 1e-30f ? 1.f / d.x : copysignf(big, d.x), fabsf(d.y) > 1e-30f ? 1.f / d.y : copysignf(big, d.y),
                           fabsf(d.z) > 1e-30f ? 1.f / d.z : copysignf(big, d.z));
                octinv = 7u ^ ((d.x < 0.f ? 1u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 4u : 0u));
                RayWoop rw = woop_setup(o, d);
                ws.ray[0][lane] = mk4(o, mint); ws.ray[1][lane] = mk4(rw.rx, tmax); ws.ray[2][lane] = mk4(rw.ry, 0.f); ws.ray[3][lane] = mk4(rw.rz, 0.f);
                G = make_uint2(0u, 0x80000000u); tm = 0u; sp = 0;
                hit.t = INFINITY; hit.u = hit.v = 0.f; hit.prim = 0xffffffffu; found = false;
                active = true;
            }
            w_next += take;
            __syncwarp();
        }
        if (__ballot_sync(0xffffffffu, active) == 0u) {
            if (exhausted) break;
            continue;
        }
        while (true) {
            // ---- node phase: one child of the current group per lane (lanes with pending candidates wait)
            if (active && tm == 0u && (G.y & 0xff000000u)) {
                uint32_t bit = 31u - (uint32_t) __clz((int) G.y);
                G.y &= ~(1u << bit);
                if (G.y & 0xff000000u) { if (sp < kStackFast) s_stack[sp * kBlock + threadIdx.x] = G; else if (sp - kStackFast < kStackSlow) slow[sp - kStackFast] = G; sp++; }
                uint32_t slot = (bit - 24u) ^ octinv;
                uint32_t rel = (uint32_t) __popc(G.y & ((1u << slot) - 1u) & 0xffu);
                const uint4 *np = bvh.nodes + 5 * (size_t) (G.x + rel);
                uint4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
                uint32_t h8 = node_slots_hit(n0, n2, n3, n4, o, idir, mint, tmax), im = n0.w >> 24;
                G = make_uint2(n1.x, ((uint32_t) s_perm[octinv * 256u + (h8 & im)] << 24) | im);
                tb = n1.y; valid = n1.z; tm = s_spread[h8] & valid;
            }
            // ---- triangle phase: pool the candidates of all lanes, one (ray, triangle) pair per lane and pass
            uint32_t pend = __ballot_sync(0xffffffffu, tm != 0u);
            if (pend) {
                uint32_t n_items = 0;
                do {      // publishing rounds: every lane with candidates appends its highest one (one ballot per round)
                    if (tm) {
                        uint32_t b = 31u - (uint32_t) __clz((int) tm);
                        tm &= ~(1u << b);
                        // owner lane << 5 | rank of the candidate = offset from the node's first triangle
                        ws.items[n_items + (uint32_t) __popc(pend & lt)] = (lane << 5) | (uint32_t) __popc(valid & ((1u << b) - 1u));
                    }
                    n_items += (uint32_t) __popc(pend);
                    pend = __ballot_sync(0xffffffffu, tm != 0u);
                } while (pend && n_items <= (uint32_t) (kCoopCap - 32));      // what does not fit stays pending in tm
                __syncwarp();
                for (uint32_t base = 0; base < n_items; base += 32u) {
                    uint32_t j = base + lane;
                    bool mine = j < n_items;
                    uint32_t it = mine ? ws.items[j] : 0u, src = it >> 5;
                    uint32_t stb = __shfl_sync(0xffffffffu, tb, src);
                    // scenes with analytic shapes: the testing lane re-reads the owner's ray (the shared record holds the
                    // sheared Woop rows, not the direction)
                    uint32_t sray = 0u;
                    if constexpr (Shapes) sray = __shfl_sync(0xffffffffu, ray, src);
                    if (mine) {
                        const float4 *tp = bvh.tris + 3 * (size_t) (stb + (it & 31u));
                        float4 a0 = __ldg(tp), a1 = __ldg(tp + 1), a2 = __ldg(tp + 2);
                        float4 r0 = ws.ray[0][src], r1 = ws.ray[1][src], r2 = ws.ray[2][src], r3 = ws.ray[3][src];
                        RayWoop rw; rw.rx = xyz(r1); rw.ry = xyz(r2); rw.rz = xyz(r3);
                        float t, u, v;
                        bool ok, analytic = false;
                        if constexpr (Shapes) analytic = f2u(a1.w) != 0u;
                        if (analytic) {
                            float4 r = shape_intersect_slot(bvh.shapes + f2u(a2.w), xyz(r0), xyz(d_maxt[sray]), r0.w, r1.w);
                            t = r.x; u = r.y; v = r.z; ok = r.w != 0.f;
                        } else
                            ok = woop_intersect(rw, xyz(r0), xyz(a0), xyz(a1), xyz(a2), r0.w, r1.w, t, u, v);
                        if (ok) {
                            ws.hitrec[lane] = make_float4(t, u, v, a0.w);
                            // key = t | primitive id | testing lane: between candidates at the same t the lowest primitive id wins, whoever tested it
                            atomicMin(&ws.best[src], ((unsigned long long) float_order_bits(t) << 32) | ((unsigned long long) (f2u(a0.w) & 0x07ffffffu) << 5) | lane);
                        }
                    }
                    __syncwarp();
                    unsigned long long bk = ws.best[lane];
                    if (bk != ~0ull) {      // the closest accepted candidate of my ray was tested by lane (bk & 31)
                        float4 r = ws.hitrec[(uint32_t) bk & 31u];
                        // a later candidate at exactly the current t replaces the hit only if its primitive id is lower: the answer
                        // does not depend on which rays share the warp or in which pass a triangle was tested
                        if (AnyHit || !found || r.x < tmax || f2u(r.w) < hit.prim) {
                            found = true; tmax = r.x;
                            hit.t = r.x; hit.u = r.y; hit.v = r.z; hit.prim = f2u(r.w);
                            ws.ray[1][lane].w = r.x;
                        }
                        ws.best[lane] = ~0ull;
                    }
                    __syncwarp();
                }
            }
            // ---- next group / completion
            if (active) {
                bool done = false;
                if (AnyHit && found) done = true;
                else if (tm == 0u && (G.y & 0xff000000u) == 0u) {
                    if (sp == 0) done = true;
                    else { sp--; G = sp < kStackFast ? s_stack[sp * kBlock + threadIdx.x] : slow[(sp - kStackFast) < kStackSlow ? (sp - kStackFast) : kStackSlow - 1]; }
                }
                if (done) { sink(ray, found, hit); active = false; tm = 0u; }
            }
            uint32_t a = __ballot_sync(0xffffffffu, active);
            if (a == 0u || (!exhausted && (uint32_t) __popc(a) < refill)) break;
        }
    }
}

template <bool AnyHit, bool Shapes = false>
__global__ void __launch_bounds__(kBlock, kTraceBlocksPerSM)
k_trace(Bvh bvh, const uint32_t *count_ptr, uint32_t count_max, const float4 *__restrict__ o_mint,
        const float4 *__restrict__ d_maxt, float4 *__restrict__ hits, uint32_t *__restrict__ occluded,
        uint32_t *work, uint32_t refill, const uint8_t *__restrict__ prim_class) {
    uint32_t count = count_ptr ? min(*count_ptr, count_max) : count_max;
    trace_persistent<AnyHit, Shapes>(bvh, count, o_mint, d_maxt, work, refill,
        [=](uint32_t i, bool found, const Hit &h) {
            if (AnyHit) occluded[i] = found ? 1u : 0u;
            else {
                // wavefront launches tag the hit with its shade class in the two top bits of the primitive id
                // (kPrimMask strips them again in make_si); the public intersect entry points pass no table
                uint32_t p = h.prim;
                if (prim_class != nullptr && found) p |= (uint32_t) prim_class[p] << 30;
                hits[i] = make_float4(h.t, h.u, h.v, u2f(p));
            }
        });
}
// brute force = Scene::ray_intersect_naive (kdtree.h:2304-2329) with the same triangle test
__global__ void k_trace_naive(Bvh bvh, uint32_t count, const float4 *o_mint, const float4 *d_maxt, float4 *hits) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float4 o = o_mint[i], d = d_maxt[i];
    RayWoop rw = woop_setup(xyz(o), xyz(d));
    Hit h; h.t = INFINITY; h.u = h.v = 0.f; h.prim = 0xffffffffu;
    float tmax = d.w;
    for (uint32_t k = 0; k < bvh.n_tris; ++k) {
        float4 a0 = bvh.tris[3 * k], a1 = bvh.tris[3 * k + 1], a2 = bvh.tris[3 * k + 2];
        float t, u, v;
        bool ok = f2u(a1.w) != 0u ? shape_intersect(bvh.shapes[f2u(a2.w)], xyz(o), xyz(d), o.w, tmax, t, u, v)
                                  : woop_intersect(rw, xyz(o), xyz(a0), xyz(a1), xyz(a2), o.w, tmax, t, u, v);
        if (ok) { tmax = t; h.t = t; h.u = u; h.v = v; h.prim = f2u(a0.w); }
    }
    hits[i] = make_float4(h.t, h.u, h.v, u2f(h.prim));
}

// ---------------- wavefront stages ---------------------------------------------------------------
// counters layout (uint32): [0..1023] live paths entering depth d (index d), [1024..2047] shadow rays of depth d
// stats (uint64): [0] closest rays, [1] shadow rays, [2] segments
template <typename M>
__global__ void k_raygen(SensorDev sn, RenderBatch b, PathQueue q, uint32_t *counters) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) counters[1] = b.n_paths;
    if (i < b.n_paths) raygen_body<M>(sn, b, q, i);
}

__device__ __forceinline__ uint32_t warp_alloc(uint32_t *counter, bool want) {
    unsigned mask = __ballot_sync(0xffffffffu, want);
    if (mask == 0) return 0;
    int lane = threadIdx.x & 31, leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, (uint32_t) __popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + (uint32_t) __popc(mask & ((1u << lane) - 1u));
}

// Two queue appends of a whole block (next-bounce queue, shadow queue) with ONE global atomic per queue and block: per-warp
// ballots, a block prefix in shared memory, then base + warp offset + lane rank.  (One atomic per warp on a single hot
// counter serialises: block aggregation took the shade stage of a C3 step from 12.3 to 10.9 ms in k_classify alone.)
// Every thread of the block must call it.
__device__ __forceinline__ void block_alloc2(uint32_t *counter_a, bool want_a, uint32_t &slot_a, uint32_t *counter_b, bool want_b, uint32_t &slot_b) {
    __shared__ uint32_t s_cnt[kBlock / 32][2], s_base[2];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, lt = (1u << lane) - 1u;
    unsigned ma = __ballot_sync(0xffffffffu, want_a), mb = __ballot_sync(0xffffffffu, want_b);
    if (lane == 0) { s_cnt[warp][0] = (uint32_t) __popc(ma); s_cnt[warp][1] = (uint32_t) __popc(mb); }
    __syncthreads();
    if (threadIdx.x < 2) {
        uint32_t total = 0;
        for (int w = 0; w < kBlock / 32; ++w) { uint32_t n = s_cnt[w][threadIdx.x]; s_cnt[w][threadIdx.x] = total; total += n; }
        s_base[threadIdx.x] = total ? atomicAdd(threadIdx.x == 0 ? counter_a : counter_b, total) : 0u;
    }
    __syncthreads();
    slot_a = s_base[0] + s_cnt[warp][0] + (uint32_t) __popc(ma & lt);
    slot_b = s_base[1] + s_cnt[warp][1] + (uint32_t) __popc(mb & lt);
}

// ---- material binning -------------------------------------------------------------------------------------------------
// After a bounce the queue is incoherent: neighbouring entries hit different materials, so a warp of the shade kernel
// would run the diffuse, rough-conductor, dielectric and environment branches one after the other.  k_classify appends
// every queue index to the list of its class (warp-aggregated atomics); the shade kernel then walks the four lists
// back to back, each padded to a warp boundary, so that a warp sees one class only.
constexpr int kShadeClasses = 8;          // 0 = miss (environment); hits: 1 + 2 * (BSDF class - 1) + emitter pick (see k_classify)
struct ShadeBins {
    uint32_t *lists;                      // kShadeClasses x pool queue indices
    uint32_t *counts;                     // kShadeClasses counters of the current depth
    uint32_t pool;
};
__global__ void __launch_bounds__(kBlock)
k_classify(const uint32_t *count_ptr, const float4 *__restrict__ hit, ShadeBins bins, const uint4 *__restrict__ rng, int rr_draw, int n_emitters) {
    // block-aggregated append: one global atomic per class and block (instead of one per class and warp)
    __shared__ uint32_t s_warp[kBlock / 32][kShadeClasses], s_base[kShadeClasses];
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t count = *count_ptr;
    if (blockIdx.x * blockDim.x >= count) return;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    int cls = -1;
    if (i < count) {
        float4 h = hit[i];
        cls = h.x != INFINITY ? (int) (f2u(h.w) >> 30) : 0;       // class tag written by k_trace
        if (cls > 0) {
            // with two emitters every path picks one from its own sample (scene.cpp:141-160): the area light's triangle sample and
            // the environment map's hierarchical warp would run one after the other in a mixed warp.  The pick is the next draw of
            // the path's stream (after the Russian-roulette draw where there is one), replayed here.
            int pick = 0;
            if (rng != nullptr && n_emitters == 2) {
                Pcg32 r = rng_load(rng[i]);
                if (rr_draw) (void) r.next_f32();
                pick = r.next_f32() * 2.f >= 1.f ? 1 : 0;
            }
            cls = 1 + 2 * (cls - 1) + pick;
        }
    }
    uint32_t rank = 0;
#pragma unroll
    for (int c = 0; c < kShadeClasses; ++c) {
        unsigned m = __ballot_sync(0xffffffffu, cls == c);
        if (cls == c) rank = (uint32_t) __popc(m & ((1u << lane) - 1u));
        if (lane == 0) s_warp[warp][c] = (uint32_t) __popc(m);
    }
    __syncthreads();
    if (threadIdx.x < kShadeClasses) {
        uint32_t total = 0;
        for (int w = 0; w < kBlock / 32; ++w) { uint32_t n = s_warp[w][threadIdx.x]; s_warp[w][threadIdx.x] = total; total += n; }
        s_base[threadIdx.x] = total ? atomicAdd(&bins.counts[threadIdx.x], total) : 0u;
    }
    __syncthreads();
    if (cls >= 0) bins.lists[(size_t) cls * bins.pool + s_base[cls] + s_warp[warp][cls] + rank] = i;
}
// thread -> queue index through the bins; returns false for padding threads
__device__ __forceinline__ bool shade_index(const ShadeBins &bins, uint32_t t, uint32_t &i) {
    if (bins.lists == nullptr) { i = t; return t < bins.counts[0]; }      // identity: counts[0] = live paths
    uint32_t start = 0;
#pragma unroll
    for (int c = 0; c < kShadeClasses; ++c) {
        uint32_t n = bins.counts[c], padded = (n + 31u) & ~31u;
        if (t < start + padded) {
            uint32_t k = t - start;
            if (k >= n) return false;
            i = bins.lists[(size_t) c * bins.pool + k];
            return true;
        }
        start += padded;
    }
    return false;
}

// M = Rgb (3 channels) or Spectral (4 wavelengths; the spectral instance needs more registers, so it is not held to
// the rgb kernel's occupancy target)
template <typename M>
#ifndef MTSB_SPECTRAL_BLOCKS
#define MTSB_SPECTRAL_BLOCKS 5      // measured on C3 spectral: 4 -> 22.1 ms, 5 -> 21.4 ms, 6 -> 22.8 ms (spills)
#endif
__global__ void __launch_bounds__(kBlock, M::spectral ? MTSB_SPECTRAL_BLOCKS : kShadeBlocksPerSM)
k_shade(SceneDev sc, RenderBatch b, PathQueue qin, PathQueue qout, ShadowQueue sq, uint32_t *counters,
        unsigned long long *stats, uint32_t depth, ShadeBins bins) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x, i = 0;
    uint32_t count = counters[depth];
    if (t == 0) { atomicAdd(&stats[0], (unsigned long long) count); atomicAdd(&stats[2], (unsigned long long) count); }
    // the grid is sized for the batch; at depth d only the live entries (padded per class) have work: whole blocks leave
    if (blockIdx.x * blockDim.x >= count + 32u * kShadeClasses) return;
    bool live = shade_index(bins, t, i);
    ShadeOut out; out.extend = false; out.shadow = false;
    if (live) out = shade_body<M>(sc, qin, b, i);
    uint32_t slot, sslot;
    block_alloc2(&counters[depth + 1], out.extend, slot, &counters[1024 + depth], out.shadow, sslot);
    if (out.extend) {
        qout.o_mint[slot] = out.o_mint; qout.d_maxt[slot] = out.d_maxt; qout.thr_eta[slot] = out.thr_eta;
        qout.misc[slot] = out.misc; qout.rng[slot] = out.rng;
        if (M::spectral) qout.wav_eta[slot] = out.wav_eta;
    }
    if (out.shadow) { sq.o_mint[sslot] = out.s_o_mint; sq.d_maxt[sslot] = out.s_d_maxt; sq.contrib_id[sslot] = out.s_contrib_id; }
}

// adjoint replay of k_shade: identical control flow, plus the texel-gradient scatter of texture_adjoint() and, with
// Vtx, the vertex-position adjoint of adjoint.h (which carries the two previous path vertices through the queue)
struct DevAdd { __device__ __forceinline__ void operator()(float *dst, float v) const { if (v != 0.f) atomicAdd(dst, v); } };
template <bool Vtx, typename M>
__global__ void __launch_bounds__(kBlock)
k_shade_grad(SceneDev sc, RenderBatch b, PathQueue qin, PathQueue qout, ShadowQueue sq, uint32_t *counters,
             unsigned long long *stats, uint32_t depth, GradCtx g) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t count = counters[depth];
    if (i == 0) { atomicAdd(&stats[0], (unsigned long long) count); atomicAdd(&stats[2], (unsigned long long) count); }
    if (blockIdx.x * blockDim.x >= count) return;          // whole block beyond the live entries
    bool live = i < count;
    ShadeOut out; out.extend = false; out.shadow = false;
    if (live) { ReplayGrad<DevAdd, Vtx> rg; rg.g = g; out = shade_body<M>(sc, qin, b, i, rg); }
    uint32_t slot, sslot;
    block_alloc2(&counters[depth + 1], out.extend, slot, &counters[1024 + depth], out.shadow, sslot);
    if (out.extend) {
        qout.o_mint[slot] = out.o_mint; qout.d_maxt[slot] = out.d_maxt; qout.thr_eta[slot] = out.thr_eta;
        qout.misc[slot] = out.misc; qout.rng[slot] = out.rng;
        if (Vtx) { qout.prev1[slot] = out.prev1; qout.prev2[slot] = out.prev2; }
    }
    if (out.shadow) { sq.o_mint[sslot] = out.s_o_mint; sq.d_maxt[sslot] = out.s_d_maxt; sq.contrib_id[sslot] = out.s_contrib_id;
                      if (Vtx) sq.nee_rec[sslot] = out.s_nee_rec; }
}
// vertex adjoint of the emitter samples of one depth, after k_shadow has told which of them were unoccluded
__global__ void __launch_bounds__(kBlock)
k_nee_grad(SceneDev sc, RenderBatch b, PathQueue qin, ShadowQueue sq, const uint32_t *counters, uint32_t depth, uint32_t count_max, GradCtx g) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t count = min(counters[1024 + depth], count_max);
    if (j < count && sq.occluded[j] == 0u) nee_adjoint(sc, g, qin, b, sq.nee_rec[j], sq.contrib_id[j], DevAdd());
}
// vertex-normal adjoint (adjoint.h): per face of the differentiated mesh
__global__ void k_vnormal_accum(SceneDev sc, GradCtx g, uint32_t face_begin, uint32_t face_count, float *acc) {
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < face_count) vnormal_accum_body(sc, g, face_begin + f, acc, DevAdd());
}
__global__ void k_vnormal_adjoint(SceneDev sc, GradCtx g, uint32_t face_begin, uint32_t face_count, const float *acc) {
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < face_count) vnormal_adjoint_body(sc, g, acc, face_begin + f, DevAdd());
}
// boundary term of the vertex adjoint (adjoint.h): projected edge lengths, two primary rays per edge sample, scatter
__global__ void k_edge_len(SceneDev sc, EdgeCtx e, float3 cam_o, float *len) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < e.n_edges) len[k] = edge_len_body(sc, e, cam_o, k);
}
__global__ void k_edge_raygen(SceneDev sc, SensorDev sn, EdgeCtx e, const float *total_len, RenderBatch b, PathQueue q, uint32_t *counters) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j == 0) counters[1] = b.n_paths;
    e.total_len = *total_len;
    if (2 * j < b.n_paths) edge_raygen_body(sc, sn, e, b, q, j);
}
__global__ void k_edge_scatter(SceneDev sc, EdgeCtx e, const float *total_len, GradCtx g, const float4 *result, uint32_t n_samples) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    e.total_len = *total_len;
    if (j < n_samples) edge_scatter_body(sc, e, g, result, j, DevAdd());
}

// ---- gradient pass under a reconstruction filter wider than a pixel (path.h film_gather) ----------------------------
// weight channel of the film over the samples of `b` (positions only: the first two draws of every path's stream)
__global__ void k_film_weights(SensorDev sn, RenderBatch b, float *w_sum) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= b.n_paths) return;
    float fx, fy; path_film_pos(sn, b, i, fx, fy);
    film_put(sn, fx, fy, make_float4(0.f, 0.f, 0.f, 0.f), 1.f, [&](uint32_t pix, float4, float w) { atomicAdd(&w_sum[pix], w); });
}
__global__ void k_film_weights_invert(float *w_sum, uint32_t n, float spp_total) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { float w = w_sum[i]; w_sum[i] = w != 0.f ? spp_total / w : 0.f; }
}
// per path of `b`: dL/d(sample radiance) gathered over the sample's footprint
__global__ void k_path_adjoint(SensorDev sn, RenderBatch b, GradCtx g, float4 *path_dl) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= b.n_paths) return;
    float fx, fy; path_film_pos(sn, b, i, fx, fy);
    path_dl[i] = mk4(film_gather(g.filt, g.dL_dimage, g.inv_w, fx, fy) * g.scale, 0.f);
}

// any-hit traversal of the shadow queue fused with the MIS contribution accumulate (path.cpp:170-171)
template <bool Shapes = false>
__global__ void __launch_bounds__(kBlock, kTraceBlocksPerSM)
k_shadow(Bvh bvh, ShadowQueue sq, const uint32_t *counters, unsigned long long *stats, uint32_t depth, uint32_t count_max,
         float4 *result, uint32_t *work, uint32_t refill) {
    uint32_t count = min(counters[1024 + depth], count_max);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&stats[1], (unsigned long long) count);
    trace_persistent<true, Shapes>(bvh, count, sq.o_mint, sq.d_maxt, work, refill,
        [=](uint32_t i, bool found, const Hit &) {
            if (sq.occluded != nullptr) sq.occluded[i] = found ? 1u : 0u;
            if (found) return;
            float4 c = sq.contrib_id[i];
            uint32_t pid = f2u(c.w);            // one shadow ray per path and depth: no concurrent writer
            float4 r = result[pid];
            r.x += c.x; r.y += c.y; r.z += c.z;
            result[pid] = r;
        });
}

__device__ __forceinline__ void film_atomic_add(const SensorDev &sn, uint32_t pix, float4 v, float w) {
#if __CUDA_ARCH__ >= 900
    atomicAdd(&sn.film_xyza[pix], v);
#else
    atomicAdd(&sn.film_xyza[pix].x, v.x); atomicAdd(&sn.film_xyza[pix].y, v.y);
    atomicAdd(&sn.film_xyza[pix].z, v.z); atomicAdd(&sn.film_xyza[pix].w, v.w);
#endif
    atomicAdd(&sn.film_w[pix], w);
}

// second moments of the per-sample XYZ value, weighted like every other film channel: the `m2_<nested>.X/Y/Z` AOVs of the
// `moment` integrator (src/integrators/moment.cpp:55-90); v = filter weight * xyz, x = xyz
__device__ __forceinline__ void film_atomic_add_m2(const SensorDev &sn, uint32_t pix, float4 v, float4 x) {
    atomicAdd(&sn.film_m2[pix], make_float4(v.x * x.x, v.y * x.y, v.z * x.z, 0.f));
}

// film splat at batch end with warp-aggregated atomics (one atomic set per warp when the warp's samples share a pixel)
template <typename M>
__global__ void __launch_bounds__(kBlock)
k_splat(SensorDev sn, RenderBatch b) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool live = i < b.n_paths;
    float4 xyza = make_float4(0.f, 0.f, 0.f, 0.f); float wgt = 0.f; float fx = -10.f, fy = -10.f;
    if (live) { xyza = result_to_xyza<M>(b.result[i]); wgt = 1.f; path_film_pos(sn, b, i, fx, fy); }
    if (live && !film_sample_valid(xyza, wgt)) { live = false; xyza = make_float4(0.f, 0.f, 0.f, 0.f); wgt = 0.f; }   // imageblock.cpp:84-108
    if (sn.radius <= 1.f) {
        int lx = (int) ceilf(fx - 1.f), ly = (int) ceilf(fy - 1.f);
        bool ok = live && lx >= 0 && ly >= 0 && lx < (int) sn.width && ly < (int) sn.height;
        uint32_t pix = ok ? (uint32_t) (ly * (int) sn.width + lx) : 0xffffffffu;
        if (!ok) { xyza = make_float4(0.f, 0.f, 0.f, 0.f); wgt = 0.f; }
        uint32_t p0 = __shfl_sync(0xffffffffu, pix, 0);
        const bool m2 = sn.film_m2 != nullptr;
        float4 sq = make_float4(xyza.x * xyza.x, xyza.y * xyza.y, xyza.z * xyza.z, 0.f);
        if (__all_sync(0xffffffffu, pix == p0)) {
            for (int o = 16; o > 0; o >>= 1) {
                xyza.x += __shfl_xor_sync(0xffffffffu, xyza.x, o); xyza.y += __shfl_xor_sync(0xffffffffu, xyza.y, o);
                xyza.z += __shfl_xor_sync(0xffffffffu, xyza.z, o); xyza.w += __shfl_xor_sync(0xffffffffu, xyza.w, o);
                wgt += __shfl_xor_sync(0xffffffffu, wgt, o);
                if (m2) { sq.x += __shfl_xor_sync(0xffffffffu, sq.x, o); sq.y += __shfl_xor_sync(0xffffffffu, sq.y, o); sq.z += __shfl_xor_sync(0xffffffffu, sq.z, o); }
            }
            if ((threadIdx.x & 31) == 0 && pix != 0xffffffffu) { film_atomic_add(sn, pix, xyza, wgt); if (m2) atomicAdd(&sn.film_m2[pix], sq); }
        } else if (ok) {
            film_atomic_add(sn, pix, xyza, wgt);
            if (m2) atomicAdd(&sn.film_m2[pix], sq);
        }
    } else if (live) {
        film_put(sn, fx, fy, xyza, wgt, [&](uint32_t pix, float4 v, float w) { film_atomic_add(sn, pix, v, w); if (sn.film_m2 != nullptr) film_atomic_add_m2(sn, pix, v, xyza); });
    }
}

// ImageBlock::put parity hook
__global__ void k_film_put(SensorDev sn, uint32_t n, const float *pos2, const float *value5) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v = make_float4(value5[5 * i], value5[5 * i + 1], value5[5 * i + 2], value5[5 * i + 3]);
    film_put(sn, pos2[2 * i], pos2[2 * i + 1], v, value5[5 * i + 4], [&](uint32_t pix, float4 vv, float w) { film_atomic_add(sn, pix, vv, w); });
}

// film export / import / develop (src/films/hdrfilm.cpp:249-320; Bitmap::convert XYZAW -> RGB)
__global__ void k_film_pack(SensorDev sn, float *xyzaw) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sn.width * sn.height) return;
    float4 v = sn.film_xyza[i];
    xyzaw[5 * i] = v.x; xyzaw[5 * i + 1] = v.y; xyzaw[5 * i + 2] = v.z; xyzaw[5 * i + 3] = v.w; xyzaw[5 * i + 4] = sn.film_w[i];
}
__global__ void k_film_pack_m2(SensorDev sn, float *m2) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sn.width * sn.height) return;
    float4 v = sn.film_m2[i];
    m2[3 * i] = v.x; m2[3 * i + 1] = v.y; m2[3 * i + 2] = v.z;
}
__global__ void k_film_unpack_add(SensorDev sn, const float *xyzaw, int replace) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sn.width * sn.height) return;
    float4 v = replace ? make_float4(0.f, 0.f, 0.f, 0.f) : sn.film_xyza[i];
    float w = replace ? 0.f : sn.film_w[i];
    v.x += xyzaw[5 * i]; v.y += xyzaw[5 * i + 1]; v.z += xyzaw[5 * i + 2]; v.w += xyzaw[5 * i + 3]; w += xyzaw[5 * i + 4];
    sn.film_xyza[i] = v; sn.film_w[i] = w;
}
__global__ void k_film_develop(SensorDev sn, float *rgb) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sn.width * sn.height) return;
    float4 v = sn.film_xyza[i]; float w = sn.film_w[i];
    float inv = w != 0.f ? 1.f / w : 0.f;   // Bitmap::convert: weight-normalise (0 weight -> 0)
    float X = v.x * inv, Y = v.y * inv, Z = v.z * inv;
    rgb[3 * i + 0] = 3.240479f * X - 1.537150f * Y - 0.498535f * Z;
    rgb[3 * i + 1] = -0.969256f * X + 1.875991f * Y + 0.041556f * Z;
    rgb[3 * i + 2] = 0.055648f * X - 0.204043f * Y + 1.057311f * Z;
}

// boundary term seen from x1 (adjoint.h): 3D edge lengths, far-ray generation from the primary hits, scatter
__global__ void k_edge_len3(SceneDev sc, Edge2Ctx e, float *len) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < e.n_edges) len[k] = edge_len3_body(sc, e, k);
}
// every live queue entry of depth e.depth continues 1:1 as the far ray of its edge sample (a zero-length ray where there is none)
__global__ void k_edge2_raygen(SceneDev sc, Edge2Ctx e, const float *total_len, RenderBatch b, PathQueue qin, PathQueue qout, uint32_t *counters) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t count = counters[e.depth];
    if (j == 0) counters[e.depth + 1] = count;
    e.total_len = *total_len;
    if (j < count) edge2_raygen_body(sc, e, b, qin, qout, j);
}
__global__ void k_edge2_scatter(SceneDev sc, Edge2Ctx e, GradCtx g, RenderBatch b) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < b.n_paths) edge2_scatter_body(sc, e, g, b, j, DevAdd());
}

// ---------------- parity hooks -------------------------------------------------------------------
__global__ void k_pcg32(uint32_t n, const uint64_t *initstate, const uint64_t *initseq, uint32_t n_draws, float *out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Pcg32 r; r.seed(initstate[i], initseq[i]);
    for (uint32_t k = 0; k < n_draws; ++k) out[(size_t) i * n_draws + k] = r.next_f32();
}
__global__ void k_sampler_seed(uint32_t n, uint64_t seed, uint64_t *state, uint64_t *inc) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Pcg32 r = sampler_seed(seed, i); state[i] = r.state; inc[i] = r.inc;
}
__global__ void k_sensor_ray(SensorDev sn, uint32_t n, const float *pos2, float4 *o, float4 *d) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    sensor_sample_ray(sn, pos2[2 * i], pos2[2 * i + 1], o[i], d[i]);
}
HD SI local_si(float3 wi, float u, float v) {
    SI si; si.p = mk3(0.f); si.n = mk3(0.f, 0.f, 1.f); si.sh.s = mk3(1.f, 0.f, 0.f); si.sh.t = mk3(0.f, 1.f, 0.f); si.sh.n = mk3(0.f, 0.f, 1.f);
    si.u = u; si.v = v; si.wi = wi; si.t = 1.f; si.prim = 0; si.mesh = 0; si.valid = true; return si;
}
__global__ void k_bsdf_sample(SceneDev sc, uint32_t bsdf_id, uint32_t n, const float *wi, const float *uv, const float *s1, const float *s2, float *out9) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    SI si = local_si(ld3(wi, i), uv[2 * i], uv[2 * i + 1]);
    BsdfSample bs = bsdf_sample(sc, sc.bsdfs[bsdf_id], si, s1[i], s2[2 * i], s2[2 * i + 1]);
    float *o = out9 + 9 * (size_t) i;
    o[0] = bs.wo.x; o[1] = bs.wo.y; o[2] = bs.wo.z; o[3] = bs.pdf; o[4] = bs.eta; o[5] = bs.weight.x; o[6] = bs.weight.y; o[7] = bs.weight.z; o[8] = bs.delta ? 1.f : 0.f;
}
__global__ void k_bsdf_eval_pdf(SceneDev sc, uint32_t bsdf_id, uint32_t n, const float *wi, const float *uv, const float *wo, float *eval3, float *pdf) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    SI si = local_si(ld3(wi, i), uv[2 * i], uv[2 * i + 1]);
    float3 val; float p;
    bsdf_eval_pdf(sc, sc.bsdfs[bsdf_id], si, ld3(wo, i), val, p);
    eval3[3 * i] = val.x; eval3[3 * i + 1] = val.y; eval3[3 * i + 2] = val.z; pdf[i] = p;
}
__global__ void k_sample_emitter(SceneDev sc, uint32_t n, const float *ref_p, const float *s2, float *out11) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    DirSampleT<RgbX> ds = sample_emitter_direction<RgbX>(sc, ld3(ref_p, i), s2[2 * i], s2[2 * i + 1]);     // RgbX: incl. analytic emitter shapes
    float *o = out11 + 11 * (size_t) i;
    o[0] = ds.d.x; o[1] = ds.d.y; o[2] = ds.d.z; o[3] = ds.dist; o[4] = ds.pdf; o[5] = ds.value.x; o[6] = ds.value.y; o[7] = ds.value.z;
    o[8] = ds.n.x; o[9] = ds.n.y; o[10] = ds.n.z;
}

} // namespace mtsb
