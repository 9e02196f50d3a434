// postprocess.cu - PathPostprocessor.process (ros/path_postprocessor.py:13-71) for a batch of paths on the device
// (what the ROS adapter does to every path before publishing it).  One thread per path - the work per path
// is a 100-point banded solve and a few hundred spline evaluations, the batch supplies the parallelism:
//   1. _filter_trajectory (:40-49)      walking back from the goal, interior nodes closer than `minimal_distance` (float32)
//                                       to the last kept node are dropped
//   2. parametrisation (:28-38)         float32 chord lengths + 1e-6, float32 running sum normalised in float64; the total
//                                       is numpy's PAIRWISE float32 sum; point_count = int(float32(total) / step)
//   3. unfold_angles (utils/math.py:38-43) in float32 (python-sign modulo, two sequential `where`s, running sum), final add
//                                       in float64, stored back as float32
//   4. interp1d(kind="quadratic")       = make_interp_spline(k=2), not-a-knot knots, collocation matrix in band storage,
//                                       elimination without pivoting, de Boor evaluation at linspace(0, 1, point_count)
//   5. _find_minimal_filter_index (:58-71) the leading poses up to a direction flip within the first 6 segments are dropped
//                                       (the first pose always is)
// oracle/postprocess.py is the CPU restatement (pinned against the live reference, tests/golden/postprocess.npz).
#include "common.cuh"
#include "spline.cuh"

namespace nfopp {

struct PostLayout {
    int n;
    size_t off_pts, off_keep, off_dist, off_param, off_knots, off_band, off_coef, per_path;
};

__host__ __device__ inline size_t pp_al(size_t x) { return (x + 255) / 256 * 256; }

__host__ __device__ inline PostLayout post_layout(int n) {
    PostLayout l{};
    l.n = n;
    size_t o = 0;
    l.off_pts = o; o = pp_al(o + (size_t)n * 3 * 4);
    l.off_keep = o; o = pp_al(o + (size_t)n * 4);
    l.off_dist = o; o = pp_al(o + (size_t)n * 4);
    l.off_param = o; o = pp_al(o + (size_t)n * 8);
    l.off_knots = o; o = pp_al(o + (size_t)(n + 3) * 8);
    l.off_band = o; o = pp_al(o + (size_t)n * 5 * 8);
    l.off_coef = o; o = pp_al(o + (size_t)n * 3 * 8);
    l.per_path = o;
    return l;
}

struct PostArgs {
    const float *paths;      // [P, n, 3]
    int P, n, max_points;
    double minimal_distance, distance_step;
    PostLayout l;
    unsigned char *ws;
    double *out;             // [P, max_points, 3]
    int32_t *count;          // [P]
};

// numpy's pairwise float32 sum (umath loops: < 8 sequential; <= 128 eight running sums; else split at a multiple of 8)
__device__ float pairwise_sum_f32(const float *a, int n) {
    if (n < 8) {
        float r = 0.f;
        for (int i = 0; i < n; ++i) r = __fadd_rn(r, a[i]);
        return r;
    }
    if (n <= 128) {
        float r[8];
        for (int j = 0; j < 8; ++j) r[j] = a[j];
        int i = 8;
        for (; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] = __fadd_rn(r[j], a[i + j]);
        float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                              __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
        for (; i < n; ++i) res = __fadd_rn(res, a[i]);
        return res;
    }
    int n2 = n / 2;
    n2 -= n2 % 8;
    return __fadd_rn(pairwise_sum_f32(a, n2), pairwise_sum_f32(a + n2, n - n2));
}

// float32 `(a + pi) % (2 pi) - pi` with numpy's remainder (sign of the divisor)
__device__ __forceinline__ float np_wrap_f32(float a) {
    float m = fmodf(__fadd_rn(a, NFOPP_PI_F), NFOPP_2PI_F);
    if (m != 0.f) {
        if (m < 0.f) m = __fadd_rn(m, NFOPP_2PI_F);
    } else {
        m = 0.f;
    }
    return __fsub_rn(m, NFOPP_PI_F);
}

__device__ __forceinline__ double np_wrap_f64(double a) {
    const double pi = 3.141592653589793, two_pi = 6.283185307179586;
    double m = fmod(a + pi, two_pi);
    if (m != 0.0 && m < 0.0) m += two_pi;
    return m - pi;
}

__global__ void __launch_bounds__(32) postprocess_kernel(const PostArgs a) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.P) return;
    const int n = a.n;
    const float *path = a.paths + (size_t)p * n * 3;
    double *out = a.out + (size_t)p * a.max_points * 3;
    if (n < 3) {                                                     // :14-15: returned unchanged
        for (int i = 0; i < n * 3; ++i) out[i] = (double)path[i];
        a.count[p] = n;
        return;
    }
    unsigned char *base = a.ws + (size_t)p * a.l.per_path;
    float *pts = reinterpret_cast<float *>(base + a.l.off_pts);
    int32_t *keep = reinterpret_cast<int32_t *>(base + a.l.off_keep);
    float *dist = reinterpret_cast<float *>(base + a.l.off_dist);
    double *param = reinterpret_cast<double *>(base + a.l.off_param);
    double *t = reinterpret_cast<double *>(base + a.l.off_knots);
    double *band = reinterpret_cast<double *>(base + a.l.off_band);      // [m][5]: columns i-2 .. i+2 of row i
    double *coef = reinterpret_cast<double *>(base + a.l.off_coef);      // [m][3]
    // ---- 1. distance filter, from the goal backwards ----
    const float min_d = (float)a.minimal_distance;
    int m = 0, prev = n - 1;
    keep[m++] = n - 1;
    for (int i = n - 2; i >= 1; --i) {
        const float dx = __fsub_rn(path[3 * prev], path[3 * i]), dy = __fsub_rn(path[3 * prev + 1], path[3 * i + 1]);
        if (__fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy))) > min_d) {
            keep[m++] = i;
            prev = i;
        }
    }
    keep[m++] = 0;
    if (m < 3) {                                                     // scipy cannot fit a quadratic through two nodes
        a.count[p] = -1;
        return;
    }
    for (int i = 0; i < m; ++i) {
        const int src = keep[m - 1 - i];
        pts[3 * i] = path[3 * src]; pts[3 * i + 1] = path[3 * src + 1]; pts[3 * i + 2] = path[3 * src + 2];
    }
    // ---- 2. chord-length parametrisation, number of output poses ----
    float cum = 0.f;
    param[0] = 0.0;
    for (int i = 1; i < m; ++i) {
        const float dx = __fsub_rn(pts[3 * i], pts[3 * i - 3]), dy = __fsub_rn(pts[3 * i + 1], pts[3 * i - 2]);
        dist[i - 1] = __fadd_rn(__fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy))), 1e-6f);
        cum = __fadd_rn(cum, dist[i - 1]);
        param[i] = (double)cum;
    }
    const double last = param[m - 1];
    for (int i = 0; i < m; ++i) param[i] = __ddiv_rn(param[i], last);
    const int count = (int)__fdiv_rn(pairwise_sum_f32(dist, m - 1), (float)a.distance_step);
    if (count > a.max_points) {                                      // the caller's buffer is too small: report how many
        a.count[p] = -count;
        return;
    }
    if (count <= 1) {                                                // linspace(0, 1, <=1) minus the first pose: nothing left
        a.count[p] = 0;
        return;
    }
    // ---- 3. headings unfolded ----
    {
        float w_prev = np_wrap_f32(pts[2]);
        const float w0 = w_prev;
        float run = 0.f;
        pts[2] = w0;
        for (int i = 1; i < m; ++i) {
            const float w = np_wrap_f32(pts[3 * i + 2]);
            float d = __fsub_rn(w, w_prev);
            if (d > NFOPP_PI_F) d = __fsub_rn(d, NFOPP_2PI_F);
            if (d < -NFOPP_PI_F) d = __fadd_rn(d, NFOPP_2PI_F);
            run = __fadd_rn(run, d);
            pts[3 * i + 2] = (float)__dadd_rn((double)w0, (double)run);
            w_prev = w;
        }
    }
    // ---- 4. quadratic interpolating spline through (param, pts) ----
    for (int i = 0; i < m; ++i)
        for (int c = 0; c < 3; ++c) coef[3 * i + c] = (double)pts[3 * i + c];
    quad_spline_fit<3>(m, param, t, band, coef);
    const double step = __ddiv_rn(1.0, (double)(count - 1));           // numpy linspace: arange * step, last = stop
    int j = 2;
    for (int q = 0; q < count; ++q) {
        double y[3];
        quad_spline_eval<3>(t, m, coef, q == count - 1 ? 1.0 : __dmul_rn((double)q, step), j, y);
        for (int c = 0; c < 3; ++c) out[3 * q + c] = y[c];
    }
    // ---- 5. direction filter at the head of the path ----
    int first = 1;
    {
        auto forward = [&](int i) {
            const double dx = out[3 * (i + 1)] - out[3 * i], dy = out[3 * (i + 1) + 1] - out[3 * i + 1];
            const double ma = out[3 * i + 2] + np_wrap_f64(out[3 * (i + 1) + 2] - out[3 * i + 2]) / 2;
            return cos(ma) * dx + sin(ma) * dy > 0;
        };
        const bool d0 = forward(0);
        for (int i = 1; i < 6 && i < count - 1; ++i)
            if (forward(i) != d0) { first = i; break; }
    }
    for (int i = 0; i < (count - first) * 3; ++i) out[i] = out[i + 3 * first];
    for (int i = (count - first) * 3; i < count * 3; ++i) out[i] = 0.0;      // rows past the count stay zero
    a.count[p] = count - first;
}

}  // namespace nfopp

using namespace nfopp;

extern "C" int nfopp_postprocess_workspace_bytes(int P, int n, size_t *bytes) {
    NFOPP_REQUIRE(bytes && P > 0 && n > 0, "postprocess_workspace_bytes: bad arguments");
    *bytes = (size_t)P * post_layout(n).per_path;
    return NFOPP_OK;
}

extern "C" int nfopp_postprocess_paths(const float *paths_dev, int P, int n, double minimal_distance, double distance_step,
                                       int max_points, void *workspace_dev, size_t workspace_bytes, double *out_dev,
                                       int32_t *count_dev, void *stream) {
    NFOPP_REQUIRE(paths_dev && out_dev && count_dev && workspace_dev && P > 0 && n > 0, "postprocess_paths: bad arguments");
    NFOPP_REQUIRE(distance_step > 0.0 && minimal_distance >= 0.0, "postprocess_paths: distance_step=%g minimal_distance=%g",
                  distance_step, minimal_distance);
    NFOPP_REQUIRE(max_points >= n, "postprocess_paths: max_points=%d < %d path nodes", max_points, n);
    PostArgs a{};
    a.paths = paths_dev; a.P = P; a.n = n; a.max_points = max_points;
    a.minimal_distance = minimal_distance; a.distance_step = distance_step;
    a.l = post_layout(n);
    NFOPP_REQUIRE(workspace_bytes >= (size_t)P * a.l.per_path, "postprocess_paths: workspace of %zu bytes, %zu needed",
                  workspace_bytes, (size_t)P * a.l.per_path);
    a.ws = static_cast<unsigned char *>(workspace_dev);
    a.out = out_dev; a.count = count_dev;
    postprocess_kernel<<<(P + 31) / 32, 32, 0, (cudaStream_t)stream>>>(a);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}
