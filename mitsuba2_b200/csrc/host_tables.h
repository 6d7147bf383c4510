// host_tables.h -- host-side construction of sampling tables uploaded once per scene.
#pragma once
#include <cmath>
#include <vector>
#include <vector>
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <cstring>
#include "shapes.h"

namespace mtsb {

// DiscreteDistribution::update (include/mitsuba/core/distr_1d.h:49-88) over Mesh::face_area (mesh.h:108-116)
struct AreaTable { std::vector<float> cdf; float sum = 0, normalization = 0; uint32_t valid_lo = 0, valid_hi = 0; };
inline AreaTable build_area_table(const float *pos, const uint32_t *faces, uint32_t n_faces) {
    AreaTable t; t.cdf.resize(n_faces);
    double sum = 0.0; uint32_t lo = (uint32_t) -1, hi = (uint32_t) -1;
    for (uint32_t i = 0; i < n_faces; ++i) {
        const float *p0 = pos + 3 * faces[3 * i], *p1 = pos + 3 * faces[3 * i + 1], *p2 = pos + 3 * faces[3 * i + 2];
        float ax = p1[0] - p0[0], ay = p1[1] - p0[1], az = p1[2] - p0[2];
        float bx = p2[0] - p0[0], by = p2[1] - p0[1], bz = p2[2] - p0[2];
        float cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
        float area = 0.5f * std::sqrt(cx * cx + cy * cy + cz * cz);
        sum += (double) area;
        t.cdf[i] = (float) sum;
        if (area > 0.f) { if (lo == (uint32_t) -1) lo = i; hi = i; }
    }
    if (lo == (uint32_t) -1) throw std::runtime_error("DiscreteDistribution: no probability mass found!");
    t.sum = (float) sum; t.normalization = (float) (1.0 / sum); t.valid_lo = lo; t.valid_hi = hi;
    return t;
}

inline uint32_t log2i_ceil(uint32_t v) { uint32_t r = 0; while ((1u << r) < v) ++r; return r; }
inline uint32_t h2d_index_host(uint32_t x, uint32_t y, uint32_t width) {
    return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + ((y & ~1u) * width);
}
// Hierarchical2D<Float,0> constructor (include/mitsuba/core/distr_2d.h:206-310), normalize = true
struct H2DTable { std::vector<std::vector<float>> levels; std::vector<uint32_t> widths; };
inline H2DTable build_h2d(const float *data, uint32_t w, uint32_t h) {
    H2DTable t;
    uint32_t npx = w - 1, npy = h - 1;
    uint32_t max_level = log2i_ceil(npx > npy ? npx : npy);
    t.levels.emplace_back((size_t) w * h, 0.f); t.widths.push_back(w);
    uint32_t lx = npx, ly = npy;
    for (int level = (int) max_level; level >= 0; --level) {
        lx += lx & 1u; ly += ly & 1u;
        t.levels.emplace_back((size_t) lx * ly, 0.f); t.widths.push_back(lx);
        lx >>= 1; ly >>= 1;
    }
    double sum = 0.0;
    const float *in = data;
    for (uint32_t y = 0; y < npy; ++y) {
        for (uint32_t x = 0; x < npx; ++x) {
            float avg = (in[0] + in[1] + in[w] + in[w + 1]) * .25f;
            sum += (double) avg;
            t.levels[1][h2d_index_host(x, y, t.widths[1])] = avg;
            ++in;
        }
        ++in;
    }
    float scale = (float) ((double) (npx * npy) / sum);
    for (size_t i = 0; i < t.levels[0].size(); ++i) t.levels[0][i] = data[i] * scale;
    for (size_t i = 0; i < t.levels[1].size(); ++i) t.levels[1][i] *= scale;
    lx = npx; ly = npy;
    for (uint32_t level = 2; level <= max_level + 1; ++level) {
        const std::vector<float> &l0 = t.levels[level - 1]; std::vector<float> &l1 = t.levels[level];
        lx = (lx + 1) >> 1; ly = (ly + 1) >> 1;
        for (uint32_t y = 0; y < ly; ++y)
            for (uint32_t x = 0; x < lx; ++x) {
                const float *d0 = l0.data() + h2d_index_host(x * 2, y * 2, t.widths[level - 1]);
                l1[h2d_index_host(x, y, t.widths[level])] = d0[0] + d0[1] + d0[2] + d0[3];
            }
    }
    return t;
}
// EnvironmentMapEmitter constructor (src/emitters/envmap.cpp:87-111): luminance * sin(theta_row)
inline std::vector<float> env_importance(const float *rgba, uint32_t w, uint32_t h) {
    std::vector<float> lum((size_t) w * h);
    for (uint32_t y = 0; y < h; ++y) {
        float sin_theta = std::sin((float) y / (float) (h - 1) * 3.14159265358979323846f);
        for (uint32_t x = 0; x < w; ++x) {
            const float *p = rgba + 4 * ((size_t) y * w + x);
            float l = p[0] * 0.212671f + p[1] * 0.715160f + p[2] * 0.072169f;
            lum[(size_t) y * w + x] = l * sin_theta;
        }
    }
    return lum;
}

// row-major 4x4 inverse (general, via cofactors)
inline bool invert4x4(const float *m, float *inv) {
    double a[16], o[16];
    for (int i = 0; i < 16; ++i) a[i] = m[i];
    o[0] = a[5]*a[10]*a[15] - a[5]*a[11]*a[14] - a[9]*a[6]*a[15] + a[9]*a[7]*a[14] + a[13]*a[6]*a[11] - a[13]*a[7]*a[10];
    o[4] = -a[4]*a[10]*a[15] + a[4]*a[11]*a[14] + a[8]*a[6]*a[15] - a[8]*a[7]*a[14] - a[12]*a[6]*a[11] + a[12]*a[7]*a[10];
    o[8] = a[4]*a[9]*a[15] - a[4]*a[11]*a[13] - a[8]*a[5]*a[15] + a[8]*a[7]*a[13] + a[12]*a[5]*a[11] - a[12]*a[7]*a[9];
    o[12] = -a[4]*a[9]*a[14] + a[4]*a[10]*a[13] + a[8]*a[5]*a[14] - a[8]*a[6]*a[13] - a[12]*a[5]*a[10] + a[12]*a[6]*a[9];
    o[1] = -a[1]*a[10]*a[15] + a[1]*a[11]*a[14] + a[9]*a[2]*a[15] - a[9]*a[3]*a[14] - a[13]*a[2]*a[11] + a[13]*a[3]*a[10];
    o[5] = a[0]*a[10]*a[15] - a[0]*a[11]*a[14] - a[8]*a[2]*a[15] + a[8]*a[3]*a[14] + a[12]*a[2]*a[11] - a[12]*a[3]*a[10];
    o[9] = -a[0]*a[9]*a[15] + a[0]*a[11]*a[13] + a[8]*a[1]*a[15] - a[8]*a[3]*a[13] - a[12]*a[1]*a[11] + a[12]*a[3]*a[9];
    o[13] = a[0]*a[9]*a[14] - a[0]*a[10]*a[13] - a[8]*a[1]*a[14] + a[8]*a[2]*a[13] + a[12]*a[1]*a[10] - a[12]*a[2]*a[9];
    o[2] = a[1]*a[6]*a[15] - a[1]*a[7]*a[14] - a[5]*a[2]*a[15] + a[5]*a[3]*a[14] + a[13]*a[2]*a[7] - a[13]*a[3]*a[6];
    o[6] = -a[0]*a[6]*a[15] + a[0]*a[7]*a[14] + a[4]*a[2]*a[15] - a[4]*a[3]*a[14] - a[12]*a[2]*a[7] + a[12]*a[3]*a[6];
    o[10] = a[0]*a[5]*a[15] - a[0]*a[7]*a[13] - a[4]*a[1]*a[15] + a[4]*a[3]*a[13] + a[12]*a[1]*a[7] - a[12]*a[3]*a[5];
    o[14] = -a[0]*a[5]*a[14] + a[0]*a[6]*a[13] + a[4]*a[1]*a[14] - a[4]*a[2]*a[13] - a[12]*a[1]*a[6] + a[12]*a[2]*a[5];
    o[3] = -a[1]*a[6]*a[11] + a[1]*a[7]*a[10] + a[5]*a[2]*a[11] - a[5]*a[3]*a[10] - a[9]*a[2]*a[7] + a[9]*a[3]*a[6];
    o[7] = a[0]*a[6]*a[11] - a[0]*a[7]*a[10] - a[4]*a[2]*a[11] + a[4]*a[3]*a[10] + a[8]*a[2]*a[7] - a[8]*a[3]*a[6];
    o[11] = -a[0]*a[5]*a[11] + a[0]*a[7]*a[9] + a[4]*a[1]*a[11] - a[4]*a[3]*a[9] - a[8]*a[1]*a[7] + a[8]*a[3]*a[5];
    o[15] = a[0]*a[5]*a[10] - a[0]*a[6]*a[9] - a[4]*a[1]*a[10] + a[4]*a[2]*a[9] + a[8]*a[1]*a[6] - a[8]*a[2]*a[5];
    double det = a[0]*o[0] + a[1]*o[4] + a[2]*o[8] + a[3]*o[12];
    if (det == 0.0) return false;
    for (int i = 0; i < 16; ++i) inv[i] = (float) (o[i] / det);
    return true;
}

// The analytic shape plugins' constructor + update(): sphere.cpp:100-135, cylinder.cpp:84-127, disk.cpp:76-98,
// rectangle.cpp:78-95.  `tw` = the plugin's object-to-world matrix (row-major 4x4) with center / radius / p0 / p1 already
// folded in.  Sphere and cylinder decompose it into scale, rotation and translation (enoki transform_decompose: polar
// decomposition by ten scaled Newton steps Q <- (g Q + Q^-T / g) / 2) and rebuild it without shear.
inline bool make_shape_dev(int type, int flip_normals, const float *tw, ShapeDev &out, const char **err) {
    memset(&out, 0, sizeof(out));
    out.type = type; out.flip = flip_normals ? 1 : 0;
    double A[3][3], T[3] = { tw[3], tw[7], tw[11] };
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) A[i][j] = tw[4 * i + j];
    auto inv_t3 = [](const double M[3][3], double R[3][3]) {      // inverse transpose
        double c00 = M[1][1] * M[2][2] - M[1][2] * M[2][1], c01 = M[1][2] * M[2][0] - M[1][0] * M[2][2], c02 = M[1][0] * M[2][1] - M[1][1] * M[2][0];
        double det = M[0][0] * c00 + M[0][1] * c01 + M[0][2] * c02;
        if (det == 0.0) return false;
        double id = 1.0 / det;
        R[0][0] = c00 * id; R[0][1] = c01 * id; R[0][2] = c02 * id;
        R[1][0] = (M[0][2] * M[2][1] - M[0][1] * M[2][2]) * id; R[1][1] = (M[0][0] * M[2][2] - M[0][2] * M[2][0]) * id; R[1][2] = (M[0][1] * M[2][0] - M[0][0] * M[2][1]) * id;
        R[2][0] = (M[0][1] * M[1][2] - M[0][2] * M[1][1]) * id; R[2][1] = (M[0][2] * M[1][0] - M[0][0] * M[1][2]) * id; R[2][2] = (M[0][0] * M[1][1] - M[0][1] * M[1][0]) * id;
        return true;
    };
    float m[16] = { 0 };
    m[15] = 1.f;
    if (type == kShapeSphere || type == kShapeCylinder) {
        double Q[3][3], Qi[3][3];
        memcpy(Q, A, sizeof(Q));
        for (int it = 0; it < 10; ++it) {
            if (!inv_t3(Q, Qi)) { *err = "shape to_world is singular"; return false; }
            double fq = 0, fi = 0;
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { fq += Q[i][j] * Q[i][j]; fi += Qi[i][j] * Qi[i][j]; }
            double gamma = std::sqrt(std::sqrt(fi) / std::sqrt(fq));
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Q[i][j] = Q[i][j] * (gamma * 0.5) + Qi[i][j] * (0.5 / gamma);
        }
        double S[3][3];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { S[i][j] = 0; for (int k = 0; k < 3; ++k) S[i][j] += Q[k][i] * A[k][j]; }
        double detq = Q[0][0] * (Q[1][1] * Q[2][2] - Q[1][2] * Q[2][1]) - Q[0][1] * (Q[1][0] * Q[2][2] - Q[1][2] * Q[2][0]) + Q[0][2] * (Q[1][0] * Q[2][1] - Q[1][1] * Q[2][0]);
        if (detq < 0) for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { Q[i][j] = -Q[i][j]; S[i][j] = -S[i][j]; }
        double radius = S[0][0], length = S[2][2];
        if (radius <= 0) { radius = std::fabs(radius); out.flip = !out.flip; }
        out.radius = (float) radius; out.length = (float) length;
        out.center = make_float3((float) T[0], (float) T[1], (float) T[2]);
        double sc = type == kShapeSphere ? (double) out.radius : 1.0;
        for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) m[4 * i + j] = (float) (Q[i][j] * sc); m[4 * i + 3] = (float) T[i]; }
        out.inv_area = type == kShapeSphere ? 1.f / (4.f * 3.14159265358979323846f * out.radius * out.radius)
                                            : 1.f / (2.f * 3.14159265358979323846f * out.radius * out.length);
    } else if (type == kShapeDisk || type == kShapeRectangle) {
        for (int i = 0; i < 16; ++i) m[i] = tw[i];
        if (flip_normals) for (int i = 0; i < 4; ++i) m[4 * i + 2] = -m[4 * i + 2];      // to_world * scale(1, 1, -1)
        out.flip = 0;
    } else { *err = "unknown shape type"; return false; }
    memcpy(out.to_world, m, sizeof(m));
    if (!invert4x4(out.to_world, out.to_object)) { *err = "shape to_world is singular"; return false; }
    if (type == kShapeDisk || type == kShapeRectangle) {
        const float *w = out.to_world, *o = out.to_object;
        float k = type == kShapeRectangle ? 2.f : 1.f;
        float3 du = make_float3(w[0] * k, w[4] * k, w[8] * k), dv = make_float3(w[1] * k, w[5] * k, w[9] * k);
        float3 n = make_float3(o[8], o[9], o[10]);               // Transform * Normal = inverse transpose
        float nl = std::sqrt(n.x * n.x + n.y * n.y + n.z * n.z);
        out.n = make_float3(n.x / nl, n.y / nl, n.z / nl);
        out.dp_du = du; out.dp_dv = dv;
        if (type == kShapeRectangle) {
            float3 c = make_float3(du.y * dv.z - du.z * dv.y, du.z * dv.x - du.x * dv.z, du.x * dv.y - du.y * dv.x);
            out.inv_area = 1.f / std::sqrt(c.x * c.x + c.y * c.y + c.z * c.z);
        } else {
            float lu = std::sqrt(du.x * du.x + du.y * du.y + du.z * du.z), lv = std::sqrt(dv.x * dv.x + dv.y * dv.y + dv.z * dv.z);
            float st = (du.x * dv.x + du.y * dv.y + du.z * dv.z) / (lu * lv);      // dot(frame.t, frame.s)
            float h = std::sqrt(lv * lv - (lv * st) * (lv * st));
            out.inv_area = 1.f / (3.14159265358979323846f * lu * h);
        }
    }
    return true;
}

} // namespace mtsb

// quad::gauss_legendre (include/mitsuba/core/quad.h:26, src/libcore/quad.cpp): nodes and weights of the n-point rule on
// [-1, 1] -- Newton iteration on the Legendre polynomial from the Chebyshev guess, in double.
inline void gauss_legendre_rule(int n, std::vector<float> &nodes, std::vector<float> &weights) {
    nodes.assign(n, 0.f); weights.assign(n, 0.f);
    const double pi = 3.14159265358979323846;
    for (int i = 0; i < (n + 1) / 2; ++i) {
        double x = -std::cos((2 * i + 1) / (double) (2 * n) * pi), dp = 1.0;
        for (int it = 0; it < 100; ++it) {
            double p0 = 1.0, p1 = x;
            for (int k = 2; k <= n; ++k) { double p2 = ((2 * k - 1) * x * p1 - (k - 1) * p0) / k; p0 = p1; p1 = p2; }
            dp = n * (x * p1 - p0) / (x * x - 1.0);
            double step = p1 / dp;
            x -= step;
            if (std::fabs(step) <= 4e-16 * std::fabs(x)) break;
        }
        double p0 = 1.0, p1 = x;
        for (int k = 2; k <= n; ++k) { double p2 = ((2 * k - 1) * x * p1 - (k - 1) * p0) / k; p0 = p1; p1 = p2; }
        dp = n * (x * p1 - p0) / (x * x - 1.0);
        double w = 2.0 / ((1.0 - x * x) * dp * dp);
        nodes[i] = (float) x; nodes[n - 1 - i] = (float) -x; weights[i] = weights[n - 1 - i] = (float) w;
    }
    if (n % 2 == 1) nodes[n / 2] = 0.f;
}

// `irregular` spectrum = nodes then values; appended here: a uniform bucket index over [node_0, node_last] so that the device
// finds the interval of a wavelength in ~2 steps instead of a 6-step binary search (same interval, same result).
// Layout after the 2 w floats: nb start indices (as floats), then the scale nb / (node_last - node_0).  Bucket j stores the
// last node strictly below the lower bound of bucket j - 1 (one bucket of slack against rounding of the bucket number).
constexpr uint32_t kIrregularBuckets = 64;
inline std::vector<float> irregular_with_buckets(const float *data, uint32_t w) {
    std::vector<float> out(data, data + 2 * (size_t) w);
    const uint32_t nb = kIrregularBuckets;
    double lo = data[0], width = ((double) data[w - 1] - lo) / nb;
    for (uint32_t j = 0; j < nb; ++j) {
        double bound = lo + width * (j > 0 ? j - 1 : 0);
        uint32_t idx = 0;
        while (idx + 2 < w && (double) data[idx + 1] < bound) ++idx;
        out.push_back((float) idx);
    }
    out.push_back((float) (nb / ((double) data[w - 1] - lo)));
    return out;
}

