// label.cu - K1 standalone labelling kernels (C ABI: nfopp_label_*).  HBM stream: 8*pose_dim bytes in,
// 1 byte out per pose (16-byte vector loads, 4-byte label stores); obstacle sets are staged in shared memory once per CTA.
#include "label.cuh"

namespace nfopp {

constexpr int LABEL_THREADS = 256;
constexpr int LABEL_OBS_SMEM = 1024;   // obstacle points staged per chunk (16 KB)

// One thread labels LABEL_V = 4 consecutive poses: the 4 * pose_dim doubles arrive as 16-byte loads (ld.global.v2.f64, a warp
// reads one contiguous 3 KB / 2 KB run) and the 4 labels leave as one 32-bit store.  A problem whose pose / label rows are not
// 16- / 4-byte aligned (odd n * pose_dim) and the last n % 4 poses take scalar accesses; the arithmetic is the same.
constexpr int LABEL_V = 4;

__global__ void __launch_bounds__(LABEL_THREADS)
label_kernel(CheckerParams c, const double *__restrict__ poses, int n, int pose_dim,
             const double *__restrict__ obstacles, const int32_t *__restrict__ obstacle_count, int env_stride_p,
             int m_max, const uint8_t *__restrict__ grid, uint8_t *__restrict__ labels) {
    __shared__ double s_obs[2 * LABEL_OBS_SMEM];
    const int p = blockIdx.y;
    const int pe = env_stride_p ? p : 0;
    const double *obs = obstacles ? obstacles + (size_t)pe * m_max * 2 : nullptr;
    const int m = obstacles ? (obstacle_count ? obstacle_count[pe] : m_max) : 0;
    const uint8_t *g = grid ? grid + (size_t)pe * c.rows * c.cols : nullptr;
    const double *pp = poses + (size_t)p * n * pose_dim;
    uint8_t *out = labels + (size_t)p * n;
    const bool vec = ((reinterpret_cast<uintptr_t>(pp) & 15) == 0) && ((reinterpret_cast<uintptr_t>(out) & 3) == 0);
    // each CTA walks contiguous slabs of LABEL_THREADS * 4 poses; obstacles are consumed chunk by chunk from smem
    for (int base = blockIdx.x * LABEL_THREADS * LABEL_V; base < n; base += gridDim.x * LABEL_THREADS * LABEL_V) {
        const int i0 = base + threadIdx.x * LABEL_V;
        const bool full = vec && i0 + LABEL_V <= n;
        double x[LABEL_V], y[LABEL_V], th[LABEL_V];
        if (full) {
            const double2 *q = reinterpret_cast<const double2 *>(pp + (size_t)i0 * pose_dim);
            if (pose_dim == 3) {
                const double2 a = q[0], b = q[1], cc2 = q[2], d = q[3], e = q[4], f = q[5];
                x[0] = a.x; y[0] = a.y; th[0] = b.x; x[1] = b.y; y[1] = cc2.x; th[1] = cc2.y;
                x[2] = d.x; y[2] = d.y; th[2] = e.x; x[3] = e.y; y[3] = f.x; th[3] = f.y;
            } else {
#pragma unroll
                for (int k = 0; k < LABEL_V; ++k) { const double2 a = q[k]; x[k] = a.x; y[k] = a.y; th[k] = 0; }
            }
        } else {
#pragma unroll
            for (int k = 0; k < LABEL_V; ++k) {
                const int i = i0 + k;
                x[k] = y[k] = th[k] = 0;
                if (i < n) {
                    x[k] = pp[(size_t)i * pose_dim];
                    y[k] = pp[(size_t)i * pose_dim + 1];
                    if (pose_dim == 3) th[k] = pp[(size_t)i * pose_dim + 2];
                }
            }
        }
        bool hit[LABEL_V] = {false, false, false, false};
        if (c.kind == 1 || c.kind == 2) {
            CheckerParams cc = c;
            cc.has_bounds = 0;
            for (int m0 = 0; m0 < m; m0 += LABEL_OBS_SMEM) {
                int mc = min(LABEL_OBS_SMEM, m - m0);
                __syncthreads();
                for (int k = threadIdx.x; k < 2 * mc; k += LABEL_THREADS) s_obs[k] = obs[2 * m0 + k];
                __syncthreads();
#pragma unroll
                for (int k = 0; k < LABEL_V; ++k) hit[k] |= label_pose(cc, s_obs, mc, nullptr, x[k], y[k], th[k]);
            }
#pragma unroll
            for (int k = 0; k < LABEL_V; ++k) hit[k] |= label_boundary(c, x[k], y[k]);
        } else {
#pragma unroll
            for (int k = 0; k < LABEL_V; ++k) hit[k] = label_pose(c, nullptr, 0, g, x[k], y[k], th[k]);
        }
        if (full) {
            *reinterpret_cast<uchar4 *>(out + i0) = make_uchar4(hit[0], hit[1], hit[2], hit[3]);
        } else {
#pragma unroll
            for (int k = 0; k < LABEL_V; ++k)
                if (i0 + k < n) out[i0 + k] = hit[k] ? 1 : 0;
        }
    }
}

static int launch_label(const CheckerParams &c, const double *poses, int P, int n, int pose_dim,
                        const double *obstacles, const int32_t *count, int P_env, int m_max, const uint8_t *grid,
                        uint8_t *labels, void *stream) {
    NFOPP_REQUIRE(poses && labels && P > 0 && n >= 0, "label: bad arguments");
    NFOPP_REQUIRE(pose_dim == 2 || pose_dim == 3, "label: pose_dim must be 2 or 3");
    NFOPP_REQUIRE(P_env == 1 || P_env == P, "label: P_env must be 1 or P");
    if (n == 0) return NFOPP_OK;
    int bx = (n + LABEL_THREADS * LABEL_V - 1) / (LABEL_THREADS * LABEL_V);
    int cap = max(1, (148 * 8 + P - 1) / P);      // ~8 resident CTAs per SM over the whole grid
    bx = min(bx, cap);
    dim3 grid_dim(bx, P);
    label_kernel<<<grid_dim, LABEL_THREADS, 0, (cudaStream_t)stream>>>(c, poses, n, pose_dim, obstacles, count,
                                                                      P_env == P && P > 1 ? 1 : 0, m_max, grid,
                                                                      labels);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

static void fill_bounds(CheckerParams &c, const double *bounds) {
    c.has_bounds = bounds != nullptr;
    for (int i = 0; i < 4; ++i) c.bounds[i] = bounds ? bounds[i] : 0.0;
}

}  // namespace nfopp

using namespace nfopp;

extern "C" int nfopp_label_boundary(const double *poses, int P, int n, int pose_dim, const double *bounds,
                                    uint8_t *labels, void *stream) {
    CheckerParams c{};
    c.kind = 0;
    fill_bounds(c, bounds);
    return launch_label(c, poses, P, n, pose_dim, nullptr, nullptr, 1, 0, nullptr, labels, stream);
}

extern "C" int nfopp_label_circle(const double *poses, int P, int n, int pose_dim, const double *obstacles,
                                  const int32_t *count, int P_env, int m_max, double radius, const double *bounds,
                                  uint8_t *labels, void *stream) {
    CheckerParams c{};
    c.kind = 1;
    c.radius = radius;
    c.radius_sq = circle_threshold(radius);
    fill_bounds(c, bounds);
    NFOPP_REQUIRE(obstacles || m_max == 0, "label_circle: obstacles missing");
    return launch_label(c, poses, P, n, pose_dim, m_max ? obstacles : nullptr, count, P_env, m_max, nullptr, labels,
                        stream);
}

extern "C" int nfopp_label_rectangle(const double *poses, int P, int n, const double *obstacles,
                                     const int32_t *count, int P_env, int m_max, const double *box,
                                     const double *bounds, uint8_t *labels, void *stream) {
    CheckerParams c{};
    c.kind = 2;
    NFOPP_REQUIRE(box, "label_rectangle: box missing");
    for (int i = 0; i < 4; ++i) c.box[i] = box[i];
    fill_bounds(c, bounds);
    NFOPP_REQUIRE(obstacles || m_max == 0, "label_rectangle: obstacles missing");
    return launch_label(c, poses, P, n, 3, m_max ? obstacles : nullptr, count, P_env, m_max, nullptr, labels, stream);
}

extern "C" int nfopp_label_grid(const double *poses, int P, int n, int pose_dim, const uint8_t *grid, int P_env,
                                int rows, int cols, double ox, double oy, double cell, const double *bounds,
                                uint8_t *labels, void *stream) {
    CheckerParams c{};
    c.kind = 3;
    c.rows = rows;
    c.cols = cols;
    c.ox = ox;
    c.oy = oy;
    c.cell = cell;
    fill_bounds(c, bounds);
    NFOPP_REQUIRE(grid && rows > 1 && cols > 1, "label_grid: grid missing");
    return launch_label(c, poses, P, n, pose_dim, nullptr, nullptr, P_env, 0, grid, labels, stream);
}
