// label.cuh - ground-truth collision labels for one pose, float64, in numpy's operation order with FMA
// contraction made impossible by explicit round-to-nearest intrinsics.
#pragma once
#include <math.h>

#include "common.cuh"

namespace nfopp {

struct CheckerParams {
    int kind;          // 0 boundary only, 1 circle, 2 rectangle, 3 grid
    int has_bounds;
    double bounds[4];  // x0,x1,y0,y1
    double radius;
    double radius_sq;  // smallest float64 s with sqrt(s) >= radius (circle_threshold): sqrt(s) < radius  <=>  s < radius_sq
    double box[4];
    int rows, cols;
    double ox, oy, cell;
};

// smallest float64 s >= 0 with sqrt(s) >= r (IEEE correctly rounded sqrt on the host = __dsqrt_rn on the device); 0 for r <= 0
// (sqrt(s) < r is then never true), NaN for a NaN radius
inline double circle_threshold(double r) {
    if (!(r == r)) return r;
    if (r <= 0.0) return 0.0;
    double s = r * r;
    if (!(s < 1.0e308)) return s;                       // r so large that r * r overflows: every finite distance is inside
    while (s > 0.0 && sqrt(s) >= r) s = nextafter(s, -1.0);
    while (sqrt(s) < r) s = nextafter(s, 1.0e308 * 10.0);
    return s;
}

// collision_checker/collision_checker.py:12-19 (strict inequalities)
__device__ __forceinline__ bool label_boundary(const CheckerParams &c, double x, double y) {
    if (!c.has_bounds) return false;
    return (x > c.bounds[1]) | (x < c.bounds[0]) | (y > c.bounds[3]) | (y < c.bounds[2]);
}

// circle_collision_checker.py:11-14: any_m sqrt(dx*dx + dy*dy) < r.  The correctly rounded square root is monotone, so
// sqrt(s) < r is the same predicate as s < thr with thr = the smallest float64 whose square root rounds to >= r (found on
// the host, circle_threshold): no square root per pose-obstacle pair, the same labels bit for bit (NaN / inf included:
// both comparisons are false).
__device__ __forceinline__ bool label_circle(const double *__restrict__ obs, int m, double thr, double x, double y) {
    bool hit = false;
    for (int i = 0; i < m; ++i) {
        double dx = __dsub_rn(x, obs[2 * i]), dy = __dsub_rn(y, obs[2 * i + 1]);
        double s = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
        hit |= s < thr;
    }
    return hit;
}

// rectangle_collision_checker.py:11-26 with position2.py:91-94 (inv)
__device__ __forceinline__ bool label_rectangle(const double *__restrict__ obs, int m, const double *box, double x,
                                                double y, double th) {
    double c = cos(th), s = sin(th);
    double ix = __dsub_rn(__dmul_rn(-x, c), __dmul_rn(y, s));
    double iy = __dsub_rn(__dmul_rn(x, s), __dmul_rn(y, c));
    double ci = cos(-th), si = sin(-th);
    bool hit = false;
    for (int i = 0; i < m; ++i) {
        double ox = obs[2 * i], oy = obs[2 * i + 1];
        double x1 = __dadd_rn(__dsub_rn(__dmul_rn(ox, ci), __dmul_rn(oy, si)), ix);
        double y1 = __dadd_rn(__dadd_rn(__dmul_rn(ox, si), __dmul_rn(oy, ci)), iy);
        hit |= (x1 > box[0]) & (x1 < box[1]) & (y1 > box[2]) & (y1 < box[3]);
    }
    return hit;
}

// notebooks/onf_planner_image_map.ipynb:65-80: half-cell offset, C truncation, last row/col rejected
__device__ __forceinline__ bool label_grid(const uint8_t *__restrict__ grid, const CheckerParams &c, double x,
                                           double y) {
    double half = __ddiv_rn(c.cell, 2.0);
    double fx = __ddiv_rn(__dsub_rn(__dsub_rn(x, c.ox), half), c.cell);
    double fy = __ddiv_rn(__dsub_rn(__dsub_rn(y, c.oy), half), c.cell);
    int ix = __double2int_rz(fx), iy = __double2int_rz(fy);   // saturating; out-of-range => invalid either way
    bool valid = (ix >= 0) & (iy >= 0) & (iy < c.rows - 1) & (ix < c.cols - 1) & (fx == fx) & (fy == fy);
    if (!valid) return true;
    return grid[(size_t)iy * c.cols + ix] > 0;
}

// one pose against one environment; obs may point to shared memory
__device__ __forceinline__ bool label_pose(const CheckerParams &c, const double *__restrict__ obs, int m,
                                           const uint8_t *__restrict__ grid, double x, double y, double th) {
    bool hit = false;
    if (c.kind == 1) hit = label_circle(obs, m, c.radius_sq, x, y);
    else if (c.kind == 2) hit = label_rectangle(obs, m, c.box, x, y, th);
    else if (c.kind == 3) hit = label_grid(grid, c, x, y);
    return hit | label_boundary(c, x, y);
}

}  // namespace nfopp
