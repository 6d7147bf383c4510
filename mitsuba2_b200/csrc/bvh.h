// bvh.h -- GPU-built 8-wide compressed BVH: per-thread build bodies, node encoding, and the
// traversal core.  Replaces the SAH kd-tree of the reference (include/mitsuba/render/kdtree.h:1710
// build, :2080-2175 traversal) -- semantics kept (closest hit in [mint,maxt] / any hit), algorithm new:
//   Morton codes (63 bit) -> radix sort -> Karras LBVH -> bottom-up fit -> greedy surface-area collapse
//   to 8 children -> 80-byte quantised nodes with octant-ordered child slots (after Ylitie et al. 2017)
//   -> leaf triangles re-laid out contiguously per node (48 B, watertight test after Woop et al. 2013).
// All bodies are HD so tests/hostsim can execute them sequentially on the CPU.
#pragma once
#include "shapes.h"

namespace mtsb {

template <typename T> HD T atomic_add(T *p, T v) {
#ifdef __CUDA_ARCH__
    return atomicAdd(p, v);
#else
    T o = *p; *p = o + v; return o;
#endif
}
HD void mem_fence() {
#ifdef __CUDA_ARCH__
    __threadfence();
#endif
}
HD int clz32(uint32_t x) {
#ifdef __CUDA_ARCH__
    return __clz((int) x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}
HD int clz64(uint64_t x) {
#ifdef __CUDA_ARCH__
    return __clzll((long long) x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}
HD int popc32(uint32_t x) {
#ifdef __CUDA_ARCH__
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}

constexpr int kMaxLeafTris = 3;
constexpr uint32_t kAnalyticTag = 0xffffffffu;   // faces[prim].y of an analytic shape primitive

struct BvhBuild {
    // inputs
    const float *pos;        // 3*V, all meshes concatenated
    const uint4 *faces;      // (i0,i1,i2,mesh_id) with global vertex indices; index = global prim id
                             // analytic shape: (shape index, kAnalyticTag, type, mesh_id)
    const ShapeDev *shapes;  // analytic shape table (may be null when there is none)
    uint32_t n;              // primitives (triangles + analytic shapes)
    // scratch
    float4 *tlo, *thi;       // per-triangle bounds (input order)
    uint64_t *keys; uint32_t *order;     // morton keys / sorted triangle order
    int *left, *right;       // per internal node (n-1)
    int *parent;             // per node id (2n-1); id < n-1 internal, else leaf (id-(n-1))
    uint32_t *rfirst, *rlast;// per internal node: covered sorted range
    float4 *nlo, *nhi;       // per node id bounds
    uint32_t *flags;         // per internal node arrival counter
    // outputs
    uint4 *nodes;            // 5 x uint4 per wide node
    float4 *tris;            // 3 x float4 per triangle slot
    int *wide_bin;           // per wide node: binary node id it collapses
    uint32_t *counters;      // [0] wide nodes allocated, [1] triangle slots allocated
    float3 scene_lo, scene_hi;
};

HD void tri_bounds_body(const BvhBuild &b, uint32_t i, float3 &lo, float3 &hi) {
    uint4 f = b.faces[i];
    if (f.y == kAnalyticTag) {
        shape_bounds(b.shapes[f.x], lo, hi);
        b.tlo[i] = mk4(lo, 0.f); b.thi[i] = mk4(hi, 0.f);
        return;
    }
    float3 p0 = mk3(b.pos[3 * f.x], b.pos[3 * f.x + 1], b.pos[3 * f.x + 2]);
    float3 p1 = mk3(b.pos[3 * f.y], b.pos[3 * f.y + 1], b.pos[3 * f.y + 2]);
    float3 p2 = mk3(b.pos[3 * f.z], b.pos[3 * f.z + 1], b.pos[3 * f.z + 2]);
    lo = fmin3(p0, fmin3(p1, p2)); hi = fmax3(p0, fmax3(p1, p2));
    b.tlo[i] = mk4(lo, 0.f); b.thi[i] = mk4(hi, 0.f);
}

HD uint64_t expand21(uint32_t v) {   // spread 21 bits to every third bit
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffULL;
    x = (x | x << 16) & 0x1f0000ff0000ffULL;
    x = (x | x << 8) & 0x100f00f00f00f00fULL;
    x = (x | x << 4) & 0x10c30c30c30c30c3ULL;
    x = (x | x << 2) & 0x1249249249249249ULL;
    return x;
}
HD void morton_body(const BvhBuild &b, uint32_t i) {
    float3 c = (xyz(b.tlo[i]) + xyz(b.thi[i])) * 0.5f;
    float3 ext = fmax3(b.scene_hi - b.scene_lo, mk3(1e-30f));
    float3 q = (c - b.scene_lo) / ext * 2097152.f;
    uint32_t x = (uint32_t) fminf(fmaxf(q.x, 0.f), 2097151.f);
    uint32_t y = (uint32_t) fminf(fmaxf(q.y, 0.f), 2097151.f);
    uint32_t z = (uint32_t) fminf(fmaxf(q.z, 0.f), 2097151.f);
    b.keys[i] = (expand21(x) << 2) | (expand21(y) << 1) | expand21(z);
    b.order[i] = i;
}

HD int lbvh_delta(const BvhBuild &b, int i, int j) {
    if (j < 0 || j >= (int) b.n) return -1;
    uint64_t a = b.keys[i], c = b.keys[j];
    if (a == c) return 64 + clz32((uint32_t) i ^ (uint32_t) j);
    return clz64(a ^ c);
}
// Karras 2012: one thread per internal node i in [0, n-2]
HD void lbvh_node_body(const BvhBuild &b, int i) {
    int n = (int) b.n;
    int d = (lbvh_delta(b, i, i + 1) - lbvh_delta(b, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = lbvh_delta(b, i, i - d);
    int lmax = 2;
    while (lbvh_delta(b, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lbvh_delta(b, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = lbvh_delta(b, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) / 2;
        if (lbvh_delta(b, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + (d < 0 ? -1 : 0);
    int lo = i < j ? i : j, hi = i < j ? j : i;
    int lc = (lo == gamma) ? (n - 1) + gamma : gamma;
    int rc = (hi == gamma + 1) ? (n - 1) + gamma + 1 : gamma + 1;
    b.left[i] = lc; b.right[i] = rc;
    b.parent[lc] = i; b.parent[rc] = i;
    b.rfirst[i] = (uint32_t) lo; b.rlast[i] = (uint32_t) hi;
    if (i == 0) b.parent[0] = -1;
}
// bottom-up fit: one thread per sorted leaf j
HD void lbvh_fit_body(const BvhBuild &b, int j) {
    int n = (int) b.n;
    uint32_t t = b.order[j];
    int id = (n - 1) + j;
    b.nlo[id] = b.tlo[t]; b.nhi[id] = b.thi[t];
    if (n == 1) return;
    int p = b.parent[id];
    while (p >= 0) {
        mem_fence();
        uint32_t old = atomic_add(&b.flags[p], 1u);
        if (old == 0) return;
        mem_fence();
        int l = b.left[p], r = b.right[p];
#ifdef __CUDA_ARCH__
        float4 llo = __ldcg(&b.nlo[l]), lhi = __ldcg(&b.nhi[l]), rlo = __ldcg(&b.nlo[r]), rhi = __ldcg(&b.nhi[r]);
#else
        float4 llo = b.nlo[l], lhi = b.nhi[l], rlo = b.nlo[r], rhi = b.nhi[r];
#endif
        b.nlo[p] = mk4(fmin3(xyz(llo), xyz(rlo)), 0.f);
        b.nhi[p] = mk4(fmax3(xyz(lhi), xyz(rhi)), 0.f);
        p = b.parent[p];
    }
}

HD float half_area(float3 lo, float3 hi) {
    float3 e = hi - lo;
    return e.x * e.y + e.y * e.z + e.z * e.x;
}

// ------------------------------------------------------------------------------------------------
// PLOC: parallel locally-ordered clustering (Meister & Bittner 2018) over the Morton-sorted triangles.  Bottom-up
// agglomeration: every cluster looks at its 2R neighbours in the current (Morton-ordered) cluster array and picks the
// one whose union box has the smallest area; mutual nearest neighbours merge.  Gives a markedly better tree than the
// plain Morton-split LBVH (fewer node visits per ray) at a build cost of a few passes over a shrinking array.
// Node numbering is the one the collapse expects: internal nodes [0, n-1) with the ROOT = 0 (ids are handed out
// downwards from n-2, the last merge creates the root), leaves n-1+position.  After the last merge the leaves are
// renumbered in depth-first order so that every subtree again covers a contiguous range [rfirst, rlast] of `order`.
// Scratch use in PLOC mode: flags[id] = triangles below internal node id; parent[c] = 2*parent + (c is right child).
// ------------------------------------------------------------------------------------------------
#ifndef MTSB_PLOC_RADIUS
#define MTSB_PLOC_RADIUS 6
#endif
constexpr int kPlocRadius = MTSB_PLOC_RADIUS;      // neighbours examined on each side

HD void ploc_init_body(const BvhBuild &b, int *cid, uint32_t i) {       // cluster i = leaf i (sorted position)
    int leaf = (int) b.n - 1 + (int) i;
    uint32_t prim = b.order[i];
    b.nlo[leaf] = b.tlo[prim]; b.nhi[leaf] = b.thi[prim];
    cid[i] = leaf;
}
HD void ploc_nearest_body(const BvhBuild &b, const int *cid, uint32_t m, int *nn, uint32_t i) {
    float3 lo = xyz(b.nlo[cid[i]]), hi = xyz(b.nhi[cid[i]]);
    int j0 = (int) i - kPlocRadius, j1 = (int) i + kPlocRadius;
    if (j0 < 0) j0 = 0;
    if (j1 > (int) m - 1) j1 = (int) m - 1;
    float best = INFINITY; int bj = -1;
    for (int j = j0; j <= j1; ++j) {
        if (j == (int) i) continue;
        float a = half_area(fmin3(lo, xyz(b.nlo[cid[j]])), fmax3(hi, xyz(b.nhi[cid[j]])));
        if (a < best) { best = a; bj = j; }          // ties: lowest index (deterministic)
    }
    nn[i] = bj;
}
// returns 1 if cluster i survives this round (as itself or as the merged node), 0 if it was absorbed by its partner
HD uint32_t ploc_merge_body(const BvhBuild &b, const int *cid, const int *nn, int *cid_out, uint32_t i) {
    int j = nn[i];
    if (j < 0 || nn[j] != (int) i) { cid_out[i] = cid[i]; return 1u; }
    if ((int) i > j) { cid_out[i] = -1; return 0u; }
    int l = cid[i], r = cid[j], n = (int) b.n;
    int id = n - 2 - (int) atomic_add(&b.counters[0], 1u);
    b.left[id] = l; b.right[id] = r;
    b.parent[l] = 2 * id; b.parent[r] = 2 * id + 1;
    b.nlo[id] = mk4(fmin3(xyz(b.nlo[l]), xyz(b.nlo[r])), 0.f);
    b.nhi[id] = mk4(fmax3(xyz(b.nhi[l]), xyz(b.nhi[r])), 0.f);
    b.flags[id] = (l >= n - 1 ? 1u : b.flags[l]) + (r >= n - 1 ? 1u : b.flags[r]);
    cid_out[i] = id;
    return 1u;
}
// depth-first position of the leftmost leaf below `node`: sum of the left siblings' sizes on the way up (root = 0)
HD uint32_t ploc_first_leaf(const BvhBuild &b, int node) {
    int n = (int) b.n; uint32_t pos = 0;
    while (node != 0) {
        int pc = b.parent[node], p = pc >> 1;
        if (pc & 1) { int l = b.left[p]; pos += l >= n - 1 ? 1u : b.flags[l]; }
        node = p;
    }
    return pos;
}
HD void ploc_finalize_internal_body(const BvhBuild &b, int id) {
    uint32_t first = ploc_first_leaf(b, id);
    b.rfirst[id] = first; b.rlast[id] = first + b.flags[id] - 1u;
}
// leaf at sorted position i moves to depth-first position pos; `leaf_new[i]` = its new id (children are patched afterwards)
HD void ploc_finalize_leaf_body(const BvhBuild &b, uint32_t *order_out, int *leaf_new, uint32_t i) {
    int n = (int) b.n;
    uint32_t pos = ploc_first_leaf(b, n - 1 + (int) i);
    order_out[pos] = b.order[i];
    leaf_new[i] = n - 1 + (int) pos;
}
// second half: patch the parent's child pointer and write the leaf bounds at the new id (nothing reads the old ones now)
HD void ploc_patch_leaf_body(const BvhBuild &b, const int *leaf_new, uint32_t i) {
    int n = (int) b.n, old_id = n - 1 + (int) i, new_id = leaf_new[i];
    int pc = b.parent[old_id], p = pc >> 1;
    if (pc & 1) b.right[p] = new_id; else b.left[p] = new_id;
    (void) old_id;
}
HD void ploc_leaf_bounds_body(const BvhBuild &b, const uint32_t *order_out, uint32_t pos) {
    int leaf = (int) b.n - 1 + (int) pos;
    uint32_t prim = order_out[pos];
    b.nlo[leaf] = b.tlo[prim]; b.nhi[leaf] = b.thi[prim];
}
HD uint32_t bin_count(const BvhBuild &b, int id) {
    int n = (int) b.n;
    return id >= n - 1 ? 1u : b.rlast[id] - b.rfirst[id] + 1u;
}
HD uint32_t bin_first(const BvhBuild &b, int id) {
    int n = (int) b.n;
    return id >= n - 1 ? (uint32_t) (id - (n - 1)) : b.rfirst[id];
}

// Quantisation frame of a wide node with box [lo, hi]: per axis the power-of-two cell size 2^ex (sc) that maps the extent to 8 bits
HD void quant_frame(float3 lo, float3 hi, int ex[3], float sc[3]) {
    float ext[3] = { hi.x - lo.x, hi.y - lo.y, hi.z - lo.z };
    for (int a = 0; a < 3; ++a) {
        int e = (int) ceilf(log2f(fmaxf(ext[a], 1e-30f) / 255.f));
        if (e < -126) e = -126; if (e > 126) e = 126;
        while (ceilf(ext[a] / exp2f((float) e)) > 254.f) e++;   // one unit of slack for the safety pass below
        ex[a] = e; sc[a] = exp2f((float) e);
    }
}
// child box [cl, chh] -> quantised planes of slot s, conservative under the exact arithmetic the traversal uses
HD void quant_child(const float cl[3], const float chh[3], const float plo[3], const float sc[3], int s, uint8_t qlo[3][8], uint8_t qhi[3][8]) {
    for (int a = 0; a < 3; ++a) {
        float fl = floorf((cl[a] - plo[a]) / sc[a]), fh = ceilf((chh[a] - plo[a]) / sc[a]);
        fl = fminf(fmaxf(fl, 0.f), 255.f); fh = fminf(fmaxf(fh, 0.f), 255.f);
        while (fl > 0.f && plo[a] + fl * sc[a] > cl[a]) fl -= 1.f;
        while (fh < 255.f && plo[a] + fh * sc[a] < chh[a]) fh += 1.f;
        qlo[a][s] = (uint8_t) fl; qhi[a][s] = (uint8_t) fh;
    }
}
HD uint32_t pack4u8(const uint8_t *p) { return (uint32_t) p[0] | ((uint32_t) p[1] << 8) | ((uint32_t) p[2] << 16) | ((uint32_t) p[3] << 24); }
// words 0 and 2..4 of a node (word 1 = child base, triangle base, valid24 is the topology)
HD void write_node_planes(uint4 *o, float3 lo, const int ex[3], uint32_t imask, const uint8_t qlo[3][8], const uint8_t qhi[3][8]) {
    uint32_t ew = (uint32_t) (ex[0] + 127) | ((uint32_t) (ex[1] + 127) << 8) | ((uint32_t) (ex[2] + 127) << 16) | (imask << 24);
    o[0] = make_uint4(f2u(lo.x), f2u(lo.y), f2u(lo.z), ew);
    o[2] = make_uint4(pack4u8(qlo[0]), pack4u8(qlo[0] + 4), pack4u8(qlo[1]), pack4u8(qlo[1] + 4));
    o[3] = make_uint4(pack4u8(qlo[2]), pack4u8(qlo[2] + 4), pack4u8(qhi[0]), pack4u8(qhi[0] + 4));
    o[4] = make_uint4(pack4u8(qhi[1]), pack4u8(qhi[1] + 4), pack4u8(qhi[2]), pack4u8(qhi[2] + 4));
}

// Collapse: one thread per wide node w of the current level.
HD void collapse_body(const BvhBuild &b, uint32_t w) {
    int n = (int) b.n;
    int root = b.wide_bin[w];
    int ch[8]; int nch = 0;
    if (root >= n - 1 || bin_count(b, root) <= 1) { ch[nch++] = root; }
    else { ch[nch++] = b.left[root]; ch[nch++] = b.right[root]; }
    // phase 1: open children that are too big to be leaves (largest area first); phase 2: use spare
    // slots to split multi-triangle leaves.
    for (int phase = 0; phase < 2; ++phase) {
        uint32_t limit = phase == 0 ? (uint32_t) kMaxLeafTris : 1u;
        while (nch < 8) {
            int best = -1; float best_a = -1.f;
            for (int k = 0; k < nch; ++k) {
                if (ch[k] >= n - 1 || bin_count(b, ch[k]) <= limit) continue;
                float a = half_area(xyz(b.nlo[ch[k]]), xyz(b.nhi[ch[k]]));
                if (a > best_a) { best_a = a; best = k; }
            }
            if (best < 0) break;
            int id = ch[best];
            ch[best] = b.left[id]; ch[nch++] = b.right[id];
        }
    }
    // classify
    bool internal[8]; uint32_t n_int = 0, n_tri = 0;
    for (int k = 0; k < nch; ++k) {
        internal[k] = ch[k] < n - 1 && bin_count(b, ch[k]) > (uint32_t) kMaxLeafTris;
        if (internal[k]) n_int++; else n_tri += bin_count(b, ch[k]);
    }
    uint32_t child_base = n_int ? atomic_add(&b.counters[0], n_int) : 0u;
    uint32_t tri_base = n_tri ? atomic_add(&b.counters[1], n_tri) : 0u;
    float3 lo = xyz(b.nlo[root]), hi = xyz(b.nhi[root]);
    float3 nc = (lo + hi) * 0.5f;
    // octant slot assignment (greedy auction)
    int slot_of[8]; bool slot_used[8]; bool done[8];
    for (int k = 0; k < 8; ++k) { slot_used[k] = false; done[k] = false; slot_of[k] = -1; }
    for (int it = 0; it < nch; ++it) {
        float bc = -FLT_MAX; int bk = -1, bs = -1;
        for (int k = 0; k < nch; ++k) {
            if (done[k]) continue;
            float3 c = (xyz(b.nlo[ch[k]]) + xyz(b.nhi[ch[k]])) * 0.5f - nc;
            for (int s = 0; s < 8; ++s) {
                if (slot_used[s]) continue;
                float cost = ((s & 1) ? c.x : -c.x) + ((s & 2) ? c.y : -c.y) + ((s & 4) ? c.z : -c.z);
                if (cost > bc) { bc = cost; bk = k; bs = s; }
            }
        }
        done[bk] = true; slot_used[bs] = true; slot_of[bk] = bs;
    }
    // quantisation frame
    int ex[3]; float sc[3];
    quant_frame(lo, hi, ex, sc);
    uint8_t qlo[3][8], qhi[3][8];
    for (int s = 0; s < 8; ++s) for (int a = 0; a < 3; ++a) { qlo[a][s] = 255; qhi[a][s] = 0; }   // empty slot: inverted box, never hit
    uint32_t imask = 0, valid24 = 0;
    // internal children take consecutive indices in slot order; leaf triangles are stored in slot order too: candidate bit
    // 3*slot + k (k < kMaxLeafTris) of `valid24` <-> triangle tri_base + popc(valid24 below that bit)
    int kid_of_slot[8];
    for (int s = 0; s < 8; ++s) kid_of_slot[s] = -1;
    for (int k = 0; k < nch; ++k) kid_of_slot[slot_of[k]] = k;
    uint32_t int_rank = 0, tri_off = 0;
    float plo[3] = { lo.x, lo.y, lo.z };
    for (int s = 0; s < 8; ++s) {
        int k = kid_of_slot[s];
        if (k < 0) continue;
        float4 cl4 = b.nlo[ch[k]], ch4 = b.nhi[ch[k]];
        float cl[3] = { cl4.x, cl4.y, cl4.z }, chh[3] = { ch4.x, ch4.y, ch4.z };
        quant_child(cl, chh, plo, sc, s, qlo, qhi);
        if (internal[k]) {
            imask |= 1u << s;
            b.wide_bin[child_base + int_rank] = ch[k];
            int_rank++;
        } else {
            uint32_t cnt = bin_count(b, ch[k]), first = bin_first(b, ch[k]);
            valid24 |= ((1u << cnt) - 1u) << (3 * s);
            for (uint32_t q = 0; q < cnt; ++q) {
                uint32_t prim = b.order[first + q];
                uint4 f = b.faces[prim];
                uint32_t o = 3 * (tri_base + tri_off + q);
                if (f.y == kAnalyticTag) {      // analytic leaf: type tag in a1.w (0 for triangles), shape index in a2.w
                    b.tris[o + 0] = make_float4(0.f, 0.f, 0.f, u2f(prim));
                    b.tris[o + 1] = make_float4(0.f, 0.f, 0.f, u2f(f.z));
                    b.tris[o + 2] = make_float4(0.f, 0.f, 0.f, u2f(f.x));
                    continue;
                }
                b.tris[o + 0] = make_float4(b.pos[3 * f.x], b.pos[3 * f.x + 1], b.pos[3 * f.x + 2], u2f(prim));
                b.tris[o + 1] = make_float4(b.pos[3 * f.y], b.pos[3 * f.y + 1], b.pos[3 * f.y + 2], 0.f);
                b.tris[o + 2] = make_float4(b.pos[3 * f.z], b.pos[3 * f.z + 1], b.pos[3 * f.z + 2], 0.f);
            }
            tri_off += cnt;
        }
    }
    uint4 *o = b.nodes + 5 * (size_t) w;
    write_node_planes(o, lo, ex, imask, qlo, qhi);
    o[1] = make_uint4(child_base, tri_base, valid24, 0u);
}

// ------------------------------------------------------------------------------------------------
// Refit: the vertices moved, the topology of the last build is kept (Mesh::parameters_changed, mesh.cpp:794-804, asks the
// reference for a kd-tree rebuild; an optimisation loop moves vertices a little every step).  Leaf slots are re-read from the
// position array, then the wide nodes are re-boxed and re-quantised level by level from the deepest one up: the collapse
// numbers the nodes level by level, so [begin, end) of every level is known from the build.  Slot assignment, child order and
// the triangle lists do not change.
// ------------------------------------------------------------------------------------------------
struct BvhRefit {
    uint4 *nodes; float4 *tris;
    const float *pos; const uint4 *faces; const ShapeDev *shapes;
    float4 *box_lo, *box_hi;          // per wide node: its box after the refit (children read them one level later)
};
HD void refit_tri_body(const BvhRefit &r, uint32_t t) {
    float4 *tp = r.tris + 3 * (size_t) t;
    if (f2u(tp[1].w) != 0u) return;                         // analytic shapes do not move
    uint4 f = r.faces[f2u(tp[0].w)];
    tp[0] = make_float4(r.pos[3 * f.x], r.pos[3 * f.x + 1], r.pos[3 * f.x + 2], tp[0].w);
    tp[1] = make_float4(r.pos[3 * f.y], r.pos[3 * f.y + 1], r.pos[3 * f.y + 2], 0.f);
    tp[2] = make_float4(r.pos[3 * f.z], r.pos[3 * f.z + 1], r.pos[3 * f.z + 2], 0.f);
}
HD void refit_node_body(const BvhRefit &r, uint32_t w) {
    uint4 *o = r.nodes + 5 * (size_t) w;
    uint4 n0 = o[0], n1 = o[1];
    uint32_t imask = n0.w >> 24, valid24 = n1.z;
    float3 clo[8], chi[8]; uint32_t present = 0, int_rank = 0, tri_off = 0;
    float3 lo = mk3(INFINITY), hi = mk3(-INFINITY);
    for (int s = 0; s < 8; ++s) {
        if ((imask >> s) & 1u) {
            uint32_t c = n1.x + int_rank++;
            clo[s] = xyz(r.box_lo[c]); chi[s] = xyz(r.box_hi[c]);
        } else {
            uint32_t cnt = (uint32_t) popc32((valid24 >> (3 * s)) & 7u);
            if (cnt == 0) continue;
            float3 l = mk3(INFINITY), h = mk3(-INFINITY);
            for (uint32_t q = 0; q < cnt; ++q) {
                const float4 *tp = r.tris + 3 * (size_t) (n1.y + tri_off + q);
                float4 a0 = tp[0], a1 = tp[1], a2 = tp[2];
                if (f2u(a1.w) != 0u) { float3 sl, sh; shape_bounds(r.shapes[f2u(a2.w)], sl, sh); l = fmin3(l, sl); h = fmax3(h, sh); }
                else { l = fmin3(l, fmin3(xyz(a0), fmin3(xyz(a1), xyz(a2)))); h = fmax3(h, fmax3(xyz(a0), fmax3(xyz(a1), xyz(a2)))); }
            }
            tri_off += cnt;
            clo[s] = l; chi[s] = h;
        }
        present |= 1u << s;
        lo = fmin3(lo, clo[s]); hi = fmax3(hi, chi[s]);
    }
    r.box_lo[w] = mk4(lo, 0.f); r.box_hi[w] = mk4(hi, 0.f);
    int ex[3]; float sc[3];
    quant_frame(lo, hi, ex, sc);
    uint8_t qlo[3][8], qhi[3][8];
    for (int s = 0; s < 8; ++s) for (int a = 0; a < 3; ++a) { qlo[a][s] = 255; qhi[a][s] = 0; }
    float plo[3] = { lo.x, lo.y, lo.z };
    for (int s = 0; s < 8; ++s) {
        if (!((present >> s) & 1u)) continue;
        float cl[3] = { clo[s].x, clo[s].y, clo[s].z }, chh[3] = { chi[s].x, chi[s].y, chi[s].z };
        quant_child(cl, chh, plo, sc, s, qlo, qhi);
    }
    write_node_planes(o, lo, ex, imask, qlo, qhi);
}

// ------------------------------------------------------------------------------------------------
// Traversal
// ------------------------------------------------------------------------------------------------
struct Bvh {
    const uint4 *nodes;
    const float4 *tris;
    uint32_t n_tris;
    const ShapeDev *shapes;   // analytic leaf primitives (null: triangles only)
};

// Woop et al. 2013 watertight test.  The per-ray axis permutation + shear is folded into three row vectors so the
// per-triangle work is branch-free FMAs: rx = e_kx - Sx e_kz, ry = e_ky - Sy e_kz, rz = Sz e_kz (kx/ky swapped when
// d[kz] < 0 to keep the winding).  Zero coefficients multiply exactly, so shared vertices still map to identical
// sheared coordinates and shared edges stay watertight.
struct RayWoop { float3 rx, ry, rz; };
HD float3 axis3(int k, float s) { return mk3(k == 0 ? s : 0.f, k == 1 ? s : 0.f, k == 2 ? s : 0.f); }
HD RayWoop woop_setup(float3 o, float3 d) {
    RayWoop r; (void) o;
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int kz = (ax > ay) ? (ax > az ? 0 : 2) : (ay > az ? 1 : 2);
    int kx = kz == 2 ? 0 : kz + 1, ky = kx == 2 ? 0 : kx + 1;
    float dz = kz == 0 ? d.x : (kz == 1 ? d.y : d.z);
    if (dz < 0.f) { int tmp = kx; kx = ky; ky = tmp; }
    float dx = kx == 0 ? d.x : (kx == 1 ? d.y : d.z), dy = ky == 0 ? d.x : (ky == 1 ? d.y : d.z);
    float Sx = dx / dz, Sy = dy / dz, Sz = 1.f / dz;
    r.rx = axis3(kx, 1.f) + axis3(kz, -Sx);
    r.ry = axis3(ky, 1.f) + axis3(kz, -Sy);
    r.rz = axis3(kz, Sz);
    return r;
}
HD float dot3f(float3 a, float3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
// returns true on hit in [mint, maxt]; (bu,bv) = barycentrics of v1, v2 (Mitsuba's si.uv convention, mesh.cpp:404-407)
HD bool woop_intersect(const RayWoop &r, float3 o, float3 v0, float3 v1, float3 v2, float mint, float maxt,
                       float &t, float &bu, float &bv) {
    float3 A = v0 - o, B = v1 - o, C = v2 - o;
    float Ax = dot3f(r.rx, A), Ay = dot3f(r.ry, A);
    float Bx = dot3f(r.rx, B), By = dot3f(r.ry, B);
    float Cx = dot3f(r.rx, C), Cy = dot3f(r.ry, C);
    float U = Cx * By - Cy * Bx, V = Ax * Cy - Ay * Cx, W = Bx * Ay - By * Ax;
    if (U == 0.f || V == 0.f || W == 0.f) {
        double CxBy = (double) Cx * (double) By, CyBx = (double) Cy * (double) Bx;
        U = (float) (CxBy - CyBx);
        double AxCy = (double) Ax * (double) Cy, AyCx = (double) Ay * (double) Cx;
        V = (float) (AxCy - AyCx);
        double BxAy = (double) Bx * (double) Ay, ByAx = (double) By * (double) Ax;
        W = (float) (BxAy - ByAx);
    }
    if ((U < 0.f || V < 0.f || W < 0.f) && (U > 0.f || V > 0.f || W > 0.f)) return false;
    float det = U + V + W;
    if (det == 0.f) return false;
    float T = U * dot3f(r.rz, A) + V * dot3f(r.rz, B) + W * dot3f(r.rz, C);
    float inv = 1.f / det;
    float tt = T * inv;
    if (!(tt >= mint && tt <= maxt)) return false;
    t = tt; bu = V * inv; bv = W * inv;
    return true;
}

struct Hit { float t, u, v; uint32_t prim; };

// Stack policy: fixed small array supplied by the caller (shared memory on the device) + local overflow.
template <int NFast, int NSlow> struct TravStack {
    uint2 *fast; int stride;       // fast[i*stride]
    uint2 slow[NSlow];
    int sp;
    HD void push(uint2 v) { if (sp < NFast) fast[sp * stride] = v; else if (sp - NFast < NSlow) slow[sp - NFast] = v; sp++; }
    HD uint2 pop() { sp--; return sp < NFast ? fast[sp * stride] : slow[(sp - NFast) < NSlow ? (sp - NFast) : NSlow - 1]; }
    HD bool empty() const { return sp == 0; }
};

HD uint32_t byte_of(uint32_t w, int i) { return (w >> (8 * i)) & 0xffu; }

// Node words: n0 = (origin xyz, ex | ey<<8 | ez<<16 | imask<<24), n1 = (first child, first triangle, valid24, -),
// n2..n4 = quantised planes lo x,y,z / hi x,y,z, 8 bytes each.  valid24: bit 3*slot+k set <-> leaf child `slot` holds a k-th
// triangle; triangles of a node are stored in that bit order, so candidate bit b is triangle n1.y + popc(valid24 below b).
// Hit mask word of a node test: bits 24..31 inner children in near-to-far priority (bit = 24 + (slot ^ octinv)), bits
// 0..23 candidate triangles (the valid24 bits of the leaf children whose box is hit).
HD uint32_t perm_xor8(uint32_t x, uint32_t k) {       // bit i of x -> bit i ^ k, 8 bits
    if (k & 1u) x = ((x & 0x55u) << 1) | ((x >> 1) & 0x55u);
    if (k & 2u) x = ((x & 0x33u) << 2) | ((x >> 2) & 0x33u);
    if (k & 4u) x = ((x & 0x0fu) << 4) | ((x >> 4) & 0x0fu);
    return x;
}
HD uint32_t spread3x8(uint32_t x) {                   // bit s of x -> bits 3s, 3s+1, 3s+2
    x = (x | (x << 8)) & 0x00f00fu;
    x = (x | (x << 4)) & 0x0c30c3u;
    x = (x | (x << 2)) & 0x249249u;
    return x * 7u;
}
HD uint32_t hitmask_from_slots(uint32_t h8, uint32_t imask, uint32_t valid24, uint32_t octinv) {
    return (perm_xor8(h8 & imask, octinv) << 24) | (spread3x8(h8) & valid24);
}
// Reference (host / test) formulation: integer -> float conversion per plane.
HD uint32_t node_hitmask_ref(const uint4 n0, const uint4 n1, const uint4 n2, const uint4 n3, const uint4 n4,
                             float3 o, float3 idir, uint32_t octinv, float tmin, float tmax) {
    float3 p = mk3(u2f(n0.x), u2f(n0.y), u2f(n0.z));
    float sx = u2f(byte_of(n0.w, 0) << 23), sy = u2f(byte_of(n0.w, 1) << 23), sz = u2f(byte_of(n0.w, 2) << 23);
    float3 adir = mk3(sx * idir.x, sy * idir.y, sz * idir.z);
    float3 org = mk3((p.x - o.x) * idir.x, (p.y - o.y) * idir.y, (p.z - o.z) * idir.z);
    uint32_t h8 = 0;
    bool nx = idir.x < 0.f, ny = idir.y < 0.f, nz = idir.z < 0.f;
    for (int half = 0; half < 2; ++half) {
        uint32_t lox = half ? n2.y : n2.x, loy = half ? n2.w : n2.z, loz = half ? n3.y : n3.x;
        uint32_t hix = half ? n3.w : n3.z, hiy = half ? n4.y : n4.x, hiz = half ? n4.w : n4.z;
        uint32_t ax = nx ? hix : lox, bx = nx ? lox : hix;
        uint32_t ay = ny ? hiy : loy, by = ny ? loy : hiy;
        uint32_t az = nz ? hiz : loz, bz = nz ? loz : hiz;
        for (int j = 0; j < 4; ++j) {
            float t0x = fmaf((float) byte_of(ax, j), adir.x, org.x), t1x = fmaf((float) byte_of(bx, j), adir.x, org.x);
            float t0y = fmaf((float) byte_of(ay, j), adir.y, org.y), t1y = fmaf((float) byte_of(by, j), adir.y, org.y);
            float t0z = fmaf((float) byte_of(az, j), adir.z, org.z), t1z = fmaf((float) byte_of(bz, j), adir.z, org.z);
            float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));
            float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
            if (tn <= tf * 1.0000004f) h8 |= 1u << (half * 4 + j);
        }
    }
    return hitmask_from_slots(h8, n0.w >> 24, n1.z, octinv);
}

#ifdef __CUDACC__
// Device formulation: no I2F and no branches.  A quantised byte q is turned into the float 1 + q*2^-15 with one PRMT
// (byte placed at mantissa bits 8..15 of 1.0f); the plane distance is then one FFMA: t = f*A + B with A = adir*2^15 and
// B = org - A.  B carries a rounding error <= |adir|*2^-9 (1/512 of a grid cell); the near/far offsets are padded by
// |adir|*2^-7 so the slab test stays conservative.  The outcome of child j is the sign bit of tf - tn (one FADD on the FMA
// pipe), shifted into the slot mask by one funnel shift -- the ALU pipe is this kernel's limiter (ncu: 64 % of its peak),
// so compares / selects / per-child bit arithmetic were moved off it.  Returns the 8-bit mask of the child slots hit.
__device__ __forceinline__ uint32_t node_slots_hit(const uint4 n0, const uint4 n2, const uint4 n3, const uint4 n4,
                                                   float3 o, float3 idir, float tmin, float tmax) {
    float sx = __uint_as_float((n0.w & 0xffu) << 23), sy = __uint_as_float((n0.w & 0xff00u) << 15), sz = __uint_as_float((n0.w & 0xff0000u) << 7);
    float ax_ = sx * idir.x, ay_ = sy * idir.y, az_ = sz * idir.z;
    float Ax = ax_ * 32768.f, Ay = ay_ * 32768.f, Az = az_ * 32768.f;
    float ox = (__uint_as_float(n0.x) - o.x) * idir.x - Ax, oy = (__uint_as_float(n0.y) - o.y) * idir.y - Ay, oz = (__uint_as_float(n0.z) - o.z) * idir.z - Az;
    float px = fabsf(ax_) * 0.0078125f, py = fabsf(ay_) * 0.0078125f, pz = fabsf(az_) * 0.0078125f;
    float Bnx = ox - px, Bny = oy - py, Bnz = oz - pz, Bfx = ox + px, Bfy = oy + py, Bfz = oz + pz;
    bool nx = idir.x < 0.f, ny = idir.y < 0.f, nz = idir.z < 0.f;
    uint32_t miss = 0;
    const uint32_t one = 0x3f800000u;
#pragma unroll
    for (int half = 1; half >= 0; --half) {
        uint32_t lox = half ? n2.y : n2.x, loy = half ? n2.w : n2.z, loz = half ? n3.y : n3.x;
        uint32_t hix = half ? n3.w : n3.z, hiy = half ? n4.y : n4.x, hiz = half ? n4.w : n4.z;
        uint32_t ax = nx ? hix : lox, bx = nx ? lox : hix;
        uint32_t ay = ny ? hiy : loy, by = ny ? loy : hiy;
        uint32_t az = nz ? hiz : loz, bz = nz ? loz : hiz;
#pragma unroll
        for (int j = 3; j >= 0; --j) {      // slot 7 first: every funnel shift moves the earlier outcomes one bit up
            const uint32_t sel = 0x3240u + ((uint32_t) j << 4);      // result bytes: 0x00, q_j, 0x80, 0x3f
            float t0x = fmaf(__uint_as_float(__byte_perm(one, ax, sel)), Ax, Bnx);
            float t0y = fmaf(__uint_as_float(__byte_perm(one, ay, sel)), Ay, Bny);
            float t0z = fmaf(__uint_as_float(__byte_perm(one, az, sel)), Az, Bnz);
            float t1x = fmaf(__uint_as_float(__byte_perm(one, bx, sel)), Ax, Bfx);
            float t1y = fmaf(__uint_as_float(__byte_perm(one, by, sel)), Ay, Bfy);
            float t1z = fmaf(__uint_as_float(__byte_perm(one, bz, sel)), Az, Bfz);
            float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));
            float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
            miss = __funnelshift_l(__float_as_uint(tf - tn), miss, 1);      // (miss << 1) | sign(tf - tn)
        }
    }
    return ~miss & 0xffu;
}
#endif
HD uint32_t node_hitmask(const uint4 n0, const uint4 n1, const uint4 n2, const uint4 n3, const uint4 n4,
                         float3 o, float3 idir, uint32_t octinv, float tmin, float tmax) {
#ifdef __CUDA_ARCH__
    return hitmask_from_slots(node_slots_hit(n0, n2, n3, n4, o, idir, tmin, tmax), n0.w >> 24, n1.z, octinv);
#else
    return node_hitmask_ref(n0, n1, n2, n3, n4, o, idir, octinv, tmin, tmax);
#endif
}

HD uint4 ldg4(const uint4 *p) {
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}
HD float4 ldg4f(const float4 *p) {
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}

#ifdef MTSB_TRAV_STATS
static unsigned long long g_trav_nodes = 0, g_trav_tris = 0;
#endif
#if defined(MTSB_TRAV_STATS) && !defined(__CUDA_ARCH__)
#define TRAV_NODE() (g_trav_nodes++)
#define TRAV_TRI() (g_trav_tris++)
#else
#define TRAV_NODE()
#define TRAV_TRI()
#endif

// Traversal state of one ray (scalars only, so it lives in registers; the stack is a separate object because its
// overflow array is dynamically indexed and would drag the whole struct into local memory).  `trav_step` advances the
// ray by one node visit (plus the triangles that visit selected), so a kernel can interleave rays on a lane
// (persistent threads with dynamic fetch) while the host simulator just loops.
struct Trav {
    float3 o, d, idir; RayWoop rw;
    float mint, tmax;
    uint32_t octinv;
    uint2 G;            // current node group: x = first child index, y = hits<<24 | imask
    Hit hit; bool found;
};

template <typename Stack>
HD void trav_init(Trav &tr, Stack &st, float3 o, float3 d, float mint, float maxt) {
    const float big = 1e30f;
    tr.o = o; tr.d = d;
    tr.idir = mk3(fabsf(d.x) > 1e-30f ? 1.f / d.x : copysignf(big, d.x),
                  fabsf(d.y) > 1e-30f ? 1.f / d.y : copysignf(big, d.y),
                  fabsf(d.z) > 1e-30f ? 1.f / d.z : copysignf(big, d.z));
    uint32_t oct = (d.x < 0.f ? 1u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 4u : 0u);
    tr.octinv = 7u ^ oct;
    tr.rw = woop_setup(o, d);
    tr.mint = mint; tr.tmax = maxt;
    tr.found = false;
    tr.hit.t = INFINITY; tr.hit.u = 0.f; tr.hit.v = 0.f; tr.hit.prim = 0xffffffffu;
    tr.G = make_uint2(0u, 0x80000000u);
    st.sp = 0;
}

// returns true when the ray is finished (closest: stack exhausted; any-hit: first hit or stack exhausted)
template <bool AnyHit, typename Stack>
HD bool trav_step(const Bvh &bvh, Trav &tr, Stack &st) {
    uint2 G = tr.G;
    if (G.y & 0xff000000u) {
        uint32_t bit = 31u - (uint32_t) clz32(G.y);
        G.y &= ~(1u << bit);
        if (G.y & 0xff000000u) st.push(G);
        uint32_t slot = (bit - 24u) ^ tr.octinv;
        uint32_t rel = (uint32_t) popc32(G.y & ((1u << slot) - 1u) & 0xffu);
        const uint4 *np = bvh.nodes + 5 * (size_t) (G.x + rel);
        TRAV_NODE();
        uint4 n0 = ldg4(np), n1 = ldg4(np + 1), n2 = ldg4(np + 2), n3 = ldg4(np + 3), n4 = ldg4(np + 4);
        uint32_t hm = node_hitmask(n0, n1, n2, n3, n4, tr.o, tr.idir, tr.octinv, tr.mint, tr.tmax);
        G = make_uint2(n1.x, (hm & 0xff000000u) | (n0.w >> 24));
        uint32_t tb = n1.y, tm = hm & 0x00ffffffu;
        while (tm) {
            uint32_t b = 31u - (uint32_t) clz32(tm);
            tm &= ~(1u << b);
            const float4 *tp = bvh.tris + 3 * (size_t) (tb + (uint32_t) popc32(n1.z & ((1u << b) - 1u)));
            float4 a0 = ldg4f(tp), a1 = ldg4f(tp + 1), a2 = ldg4f(tp + 2);
            float t, u, v;
            TRAV_TRI();
            bool ok = f2u(a1.w) != 0u ? shape_intersect(bvh.shapes[f2u(a2.w)], tr.o, tr.d, tr.mint, tr.tmax, t, u, v)
                                      : woop_intersect(tr.rw, tr.o, xyz(a0), xyz(a1), xyz(a2), tr.mint, tr.tmax, t, u, v);
            if (ok) {
                tr.found = true;
                tr.hit.t = t;
                if (AnyHit) return true;
                tr.tmax = t;
                tr.hit.u = u; tr.hit.v = v; tr.hit.prim = f2u(a0.w);
            }
        }
    }
    if ((G.y & 0xff000000u) == 0) {
        if (st.empty()) return true;
        G = st.pop();
    }
    tr.G = G;
    return false;
}

template <bool AnyHit, typename Stack>
HD bool bvh_trace(const Bvh &bvh, float3 o, float3 d, float mint, float maxt, Stack &st, Hit &hit) {
    Trav tr;
    trav_init(tr, st, o, d, mint, maxt);
    if (bvh.n_tris != 0)
        while (!trav_step<AnyHit>(bvh, tr, st)) { }
    hit = tr.hit;
    return tr.found;
}

} // namespace mtsb
