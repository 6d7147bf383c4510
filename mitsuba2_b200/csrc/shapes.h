// shapes.h -- analytic shapes as BVH leaf primitives: sphere, disk, cylinder, rectangle
// (src/shapes/sphere.cpp, disk.cpp, cylinder.cpp, rectangle.cpp).  One primitive per shape; the leaf slot carries the
// primitive id, the type tag and the index into the shape table, the table the plugin's state after its update().
#pragma once
#include "vec.h"

namespace mtsb {

enum : int { kShapeSphere = 1, kShapeDisk = 2, kShapeCylinder = 3, kShapeRectangle = 4 };

struct ShapeDev {
    int type;
    int flip;                      // sphere / cylinder: flip_normals (disk / rectangle fold it into to_world, disk.cpp:77-79)
    float to_world[16], to_object[16];   // after update(): sphere = uniform scale r (sphere.cpp:112-135), cylinder = rigid (cylinder.cpp:103-127)
    float3 center;                 // sphere
    float radius, length;          // sphere / cylinder radius, cylinder length
    float3 n, dp_du, dp_dv;        // disk: frame normal; rectangle: frame (to_world * (2,0,0), to_world * (0,2,0), normal)
    float inv_area;
};

// math::solve_quadratic (include/mitsuba/core/math.h:371-413) in double precision, as the shapes call it
HD bool solve_quadratic_d(double a, double b, double c, double &x0, double &x1) {
    bool linear = a == 0.0;
    bool active = !linear || b > 0.0;
    x0 = x1 = -c / b;
    double discrim = b * b - 4.0 * a * c;
    active = active && (linear || discrim >= 0.0);
    if (active && !linear) {
        double sq = sqrt(discrim);
        double temp = -0.5 * (b + copysign(sq, b));
        double x0p = temp / a, x1p = c / temp;
        x0 = fmin(x0p, x1p); x1 = fmax(x0p, x1p);
    }
    return active;
}

// Shape::ray_intersect of the four plugins.  (u, v) = the two cached floats (disk / rectangle: local x, y); sphere and
// cylinder cache nothing.  sphere.cpp:276-300, cylinder.cpp:248-287, disk.cpp:144-165, rectangle.cpp:142-164.
HD bool shape_intersect(const ShapeDev &sh, float3 o, float3 d, float mint, float maxt, float &t, float &u, float &v) {
    u = v = 0.f;
    if (sh.type == kShapeSphere) {
        double ox = (double) o.x - (double) sh.center.x, oy = (double) o.y - (double) sh.center.y, oz = (double) o.z - (double) sh.center.z;
        double dx = d.x, dy = d.y, dz = d.z;
        double A = dx * dx + dy * dy + dz * dz, B = 2.0 * (ox * dx + oy * dy + oz * dz);
        double C = ox * ox + oy * oy + oz * oz - (double) sh.radius * (double) sh.radius;
        double nt, ft;
        bool found = solve_quadratic_d(A, B, C, nt, ft);
        bool out_bounds = !(nt <= (double) maxt && ft >= (double) mint);
        bool in_bounds = nt < (double) mint && ft > (double) maxt;
        if (!found || out_bounds || in_bounds) return false;
        t = (float) (nt < (double) mint ? ft : nt);
        return true;
    }
    float3 lo = xform_point_affine(sh.to_object, o), ld = xform_vec(sh.to_object, d);
    if (sh.type == kShapeCylinder) {
        double ox = lo.x, oy = lo.y, oz = lo.z, dx = ld.x, dy = ld.y, dz = ld.z;
        double radius = sh.radius, length = sh.length;
        double A = dx * dx + dy * dy, B = 2.0 * (dx * ox + dy * oy), C = ox * ox + oy * oy - radius * radius;
        double nt, ft;
        bool found = solve_quadratic_d(A, B, C, nt, ft);
        bool out_bounds = !(nt <= (double) maxt && ft >= (double) mint);
        double zn = oz + dz * nt, zf = oz + dz * ft;
        bool in_bounds = nt < (double) mint && ft > (double) maxt;
        bool near_ok = zn >= 0 && zn <= length && nt >= (double) mint, far_ok = zf >= 0 && zf <= length && ft <= (double) maxt;
        if (!found || out_bounds || in_bounds || !(near_ok || far_ok)) return false;
        t = (float) (near_ok ? nt : ft);
        return true;
    }
    // disk / rectangle: the plane z = 0 of object space
    float tt = -lo.z * (1.f / ld.z);
    float lx = fmaf(ld.x, tt, lo.x), ly = fmaf(ld.y, tt, lo.y);
    bool inside = sh.type == kShapeDisk ? (lx * lx + ly * ly <= 1.f) : (fabsf(lx) <= 1.f && fabsf(ly) <= 1.f);
    if (!(tt >= mint && tt <= maxt && inside)) return false;
    t = tt; u = lx; v = ly;
    return true;
}

// Shape::bbox (sphere.cpp:137-142, cylinder.cpp:129-141, disk.cpp:103-110, rectangle.cpp:97-104)
HD void shape_bounds(const ShapeDev &sh, float3 &lo, float3 &hi) {
    if (sh.type == kShapeSphere) { lo = sh.center - mk3(sh.radius); hi = sh.center + mk3(sh.radius); return; }
    if (sh.type == kShapeCylinder) {
        float3 x1 = xform_vec(sh.to_world, mk3(sh.radius, 0.f, 0.f)), x2 = xform_vec(sh.to_world, mk3(0.f, sh.radius, 0.f));
        float3 x = mk3(sqrtf(x1.x * x1.x + x2.x * x2.x), sqrtf(x1.y * x1.y + x2.y * x2.y), sqrtf(x1.z * x1.z + x2.z * x2.z));
        float3 p0 = xform_point_affine(sh.to_world, mk3(0.f)), p1 = xform_point_affine(sh.to_world, mk3(0.f, 0.f, sh.length));
        lo = fmin3(p0 - x, p1 - x); hi = fmax3(p0 + x, p1 + x);
        return;
    }
    lo = mk3(INFINITY); hi = mk3(-INFINITY);
    for (int k = 0; k < 4; ++k) {
        float3 p = xform_point_affine(sh.to_world, mk3((k & 1) ? 1.f : -1.f, (k & 2) ? 1.f : -1.f, 0.f));
        lo = fmin3(lo, p); hi = fmax3(hi, p);
    }
}

// warp::square_to_uniform_disk_concentric (include/mitsuba/core/warp.h:95-126)
HD void square_to_uniform_disk_concentric(float sx, float sy, float &px, float &py) {
    float x = fmaf(2.f, sx, -1.f), y = fmaf(2.f, sy, -1.f);
    bool is_zero = x == 0.f && y == 0.f, quadrant_1_or_3 = fabsf(x) < fabsf(y);
    float r = quadrant_1_or_3 ? y : x, rp = quadrant_1_or_3 ? x : y;
    float phi = 0.25f * kPi * rp / r;
    if (quadrant_1_or_3) phi = 0.5f * kPi - phi;
    if (is_zero) phi = 0.f;
    float s, c; sincosf(phi, &s, &c);
    px = r * c; py = r * s;
}

} // namespace mtsb
