// astar.cu - batched A* trajectory initialiser: AstarTrajectoryInitializer.initialize_trajectory
// (astar/astar_trajectory_initializer.py:15-46) for all P problems of an engine at once.
//
//   1. astar_label_kernel   the checker-labelled cell grid: cell centres at heading 3 pi / 4 through the engine's own
//                           collision checker (K1 arithmetic, float64, bit-exact labels)            (:29-39)
//   2. astar_search_kernel  JPS.search with jps=False (astar/jps.py:40-130, 282-322): plain A* on the 8-connected
//                           grid, sqrt(2) diagonals, Euclidean heuristic, start key 0, heap ordered by
//                           (key, (row, col)), strict-improvement relaxation; one thread per problem with a binary
//                           heap in global memory - the total order on (key, cell) makes the expansion sequence, and
//                           so the returned cell path, identical to the reference's heapq run.  Goal unreachable ->
//                           path to the last expanded state (jps.py:321-322)
//   3. astar_spline_kernel  cell centres (float32) + start + goal -> reparametrize_path (utils/math.py:57-65): float32
//                           chord lengths + 1e-6, scipy interp1d(kind="quadratic") = make_interp_spline(k=2) with
//                           not-a-knot knots written out (collocation matrix, banded elimination - the matrix is
//                           totally positive, no pivoting needed - de Boor evaluation), float64; x, y of the T
//                           interior points go to the trajectory buffer as float32.
// oracle/astar.py is the CPU restatement (pinned bit-for-bit against the live reference and scipy).
#include "engine.cuh"
#include "spline.cuh"

namespace nfopp {

struct AstarLayout {
    int X, Y, cells, cap, nmax;
    size_t off_matrix, off_cost, off_parent, off_closed, off_hkey, off_hloc, off_path, off_param, off_pts, off_knots,
        off_band, off_coef, per_problem;
};

__host__ __device__ inline size_t al256(size_t x) { return (x + 255) / 256 * 256; }

__host__ __device__ inline AstarLayout astar_layout(int X, int Y) {
    AstarLayout l{};
    l.X = X; l.Y = Y; l.cells = X * Y; l.cap = 8 * l.cells + 8; l.nmax = l.cells + 2;
    size_t o = 0;
    l.off_matrix = o; o = al256(o + (size_t)l.cells);
    l.off_cost = o; o = al256(o + (size_t)l.cells * 8);
    l.off_parent = o; o = al256(o + (size_t)l.cells * 4);
    l.off_closed = o; o = al256(o + (size_t)l.cells);
    l.off_hkey = o; o = al256(o + (size_t)l.cap * 8);
    l.off_hloc = o; o = al256(o + (size_t)l.cap * 4);
    l.off_path = o; o = al256(o + (size_t)l.cells * 4);
    l.off_param = o; o = al256(o + (size_t)l.nmax * 8);
    l.off_pts = o; o = al256(o + (size_t)l.nmax * 8);
    l.off_knots = o; o = al256(o + (size_t)(l.nmax + 3) * 8);
    l.off_band = o; o = al256(o + (size_t)l.nmax * 5 * 8);
    l.off_coef = o; o = al256(o + (size_t)l.nmax * 2 * 8);
    l.per_problem = o;
    return l;
}

struct AstarArgs {
    EngineDev e;
    AstarLayout l;
    double resolution, ox, oy;               // cell size, origin (x_min, y_min) of the cell grid
    const int32_t *start_cell, *goal_cell;   // [P, 2] (row = y cell, col = x cell)
    unsigned char *ws;
    int32_t *path_len;                       // [P] cells on the A* path (or null)
};

// ---- 1. labelled cell grid ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) astar_label_kernel(const AstarArgs a) {
    const EngineDev &e = a.e;
    const int p = blockIdx.y;
    const int pe = e.env_shared ? 0 : p;
    const int m = (e.checker.kind == 1 || e.checker.kind == 2) ? (e.obstacle_n ? e.obstacle_n[pe] : e.m_max) : 0;
    const double *obs = e.obstacles ? e.obstacles + (size_t)pe * e.m_max * 2 : nullptr;
    const uint8_t *grid = e.grid ? e.grid + (size_t)pe * e.checker.rows * e.checker.cols : nullptr;
    uint8_t *matrix = a.ws + (size_t)p * a.l.per_problem + a.l.off_matrix;
    const double half = __ddiv_rn(a.resolution, 2.0), th = __ddiv_rn(__dmul_rn(3.0, 3.141592653589793), 4.0);
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < a.l.cells; c += gridDim.x * blockDim.x) {
        const int iy = c / a.l.X, ix = c - iy * a.l.X;
        const double x = __dadd_rn(__dadd_rn(__dmul_rn((double)ix, a.resolution), half), a.ox);
        const double y = __dadd_rn(__dadd_rn(__dmul_rn((double)iy, a.resolution), half), a.oy);
        matrix[c] = label_pose(e.checker, obs, m, grid, x, y, th) ? 1 : 0;
    }
}

// ---- 2. A* ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool heap_less(double ka, int la, double kb, int lb) { return ka < kb || (ka == kb && la < lb); }

__global__ void __launch_bounds__(32) astar_search_kernel(const AstarArgs a) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.e.P) return;
    const AstarLayout &l = a.l;
    unsigned char *base = a.ws + (size_t)p * l.per_problem;
    const uint8_t *matrix = base + l.off_matrix;
    double *cost = reinterpret_cast<double *>(base + l.off_cost);
    int32_t *parent = reinterpret_cast<int32_t *>(base + l.off_parent);
    uint8_t *closed = base + l.off_closed;
    double *hkey = reinterpret_cast<double *>(base + l.off_hkey);
    int32_t *hloc = reinterpret_cast<int32_t *>(base + l.off_hloc);
    int32_t *path = reinterpret_cast<int32_t *>(base + l.off_path);
    const int X = l.X, Y = l.Y;
    const int sr = a.start_cell[2 * p], sc = a.start_cell[2 * p + 1], gr = a.goal_cell[2 * p], gc = a.goal_cell[2 * p + 1];
    const int start = sr * X + sc, goal = gr * X + gc;
    const double inf = __longlong_as_double(0x7ff0000000000000LL), diag_cost = __dsqrt_rn(2.0);
    for (int c = 0; c < l.cells; ++c) { cost[c] = inf; parent[c] = -1; closed[c] = 0; }
    int hn = 0;
    auto push = [&](double k, int loc) {
        int i = hn++;
        while (i > 0) {
            const int up = (i - 1) >> 1;
            if (!heap_less(k, loc, hkey[up], hloc[up])) break;
            hkey[i] = hkey[up]; hloc[i] = hloc[up];
            i = up;
        }
        hkey[i] = k; hloc[i] = loc;
    };
    auto pop = [&]() {
        const int top = hloc[0];
        --hn;
        if (hn > 0) {
            const double k = hkey[hn];
            const int loc = hloc[hn];
            int i = 0;
            for (;;) {
                int ch = 2 * i + 1;
                if (ch >= hn) break;
                if (ch + 1 < hn && heap_less(hkey[ch + 1], hloc[ch + 1], hkey[ch], hloc[ch])) ++ch;
                if (!heap_less(hkey[ch], hloc[ch], k, loc)) break;
                hkey[i] = hkey[ch]; hloc[i] = hloc[ch];
                i = ch;
            }
            hkey[i] = k; hloc[i] = loc;
        }
        return top;
    };
    int last = start, found = -1;
    if (sr >= 0 && sr < Y && sc >= 0 && sc < X && gr >= 0 && gr < Y && gc >= 0 && gc < X) {
        cost[start] = 0.0;
        push(0.0, start);                                          // start.key = 0 (jps.py:285)
        while (hn > 0) {
            const int loc = pop();
            if (closed[loc]) continue;
            last = loc;
            if (loc == goal) { found = loc; break; }
            closed[loc] = 1;
            const int r = loc / X, c = loc - r * X;
            const double base_cost = cost[loc];
#pragma unroll
            for (int nb = 0; nb < 8; ++nb) {
                const int dr = nb < 4 ? (nb == 0 ? -1 : nb == 2 ? 1 : 0) : (nb < 6 ? -1 : 1);
                const int dc = nb < 4 ? (nb == 1 ? -1 : nb == 3 ? 1 : 0) : ((nb & 1) ? 1 : -1);
                const int rr = r + dr, cc = c + dc;
                if (rr < 0 || rr >= Y || cc < 0 || cc >= X) continue;
                const int nl = rr * X + cc;
                if ((matrix[nl] && nl != goal) || closed[nl]) continue;   // matrix[goal] = False (:40)
                const double nc = __dadd_rn(base_cost, nb < 4 ? 1.0 : diag_cost);
                if (cost[nl] > nc) {                               // strict improvement (jps.py:96)
                    cost[nl] = nc;
                    parent[nl] = loc;
                    const double dr2 = (double)(rr - gr), dc2 = (double)(cc - gc);
                    const double h = __dsqrt_rn(__dadd_rn(__dmul_rn(dr2, dr2), __dmul_rn(dc2, dc2)));
                    if (hn < l.cap) push(__dadd_rn(h, nc), nl);    // (unimproved re-pushes are exact duplicates: no effect)
                }
            }
        }
    }
    const int node0 = found >= 0 ? found : last;
    int n = 0;
    for (int node = node0; node >= 0; node = parent[node]) ++n;
    int i = n;
    for (int node = node0; node >= 0; node = parent[node]) path[--i] = node;
    if (a.path_len) a.path_len[p] = n;
    reinterpret_cast<int32_t *>(closed)[0] = n;                    // closed[] is dead: hand the length to the spline kernel
}

// ---- 3. reparametrize_path: quadratic interpolating spline through the path, T+2 equally spaced parameters ----------
__global__ void __launch_bounds__(32) astar_spline_kernel(const AstarArgs a) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const EngineDev &e = a.e;
    if (p >= e.P) return;
    const AstarLayout &l = a.l;
    unsigned char *base = a.ws + (size_t)p * l.per_problem;
    const int32_t *path = reinterpret_cast<const int32_t *>(base + l.off_path);
    const int n_cells = reinterpret_cast<const int32_t *>(base + l.off_closed)[0];
    double *param = reinterpret_cast<double *>(base + l.off_param);
    float *pts = reinterpret_cast<float *>(base + l.off_pts);
    double *t = reinterpret_cast<double *>(base + l.off_knots);
    double *band = reinterpret_cast<double *>(base + l.off_band);      // [n][5]: columns i-2 .. i+2 of row i
    double *coef = reinterpret_cast<double *>(base + l.off_coef);      // [n][2]
    const int n = n_cells + 2, D = e.D, T = e.T;
    const double half = __ddiv_rn(a.resolution, 2.0);
    // points: start, cell centres rounded to float32 (:43-46), goal
    pts[0] = e.start[(size_t)p * D]; pts[1] = e.start[(size_t)p * D + 1];
    for (int i = 0; i < n_cells; ++i) {
        const int r = path[i] / l.X, c = path[i] - r * l.X;
        pts[2 * (i + 1)] = (float)__dadd_rn(__dadd_rn(__dmul_rn((double)c, a.resolution), half), a.ox);
        pts[2 * (i + 1) + 1] = (float)__dadd_rn(__dadd_rn(__dmul_rn((double)r, a.resolution), half), a.oy);
    }
    pts[2 * (n - 1)] = e.goal[(size_t)p * D]; pts[2 * (n - 1) + 1] = e.goal[(size_t)p * D + 1];
    // float32 chord lengths + 1e-6, float32 running sum, normalised in float64 (utils/math.py:58-61)
    float cum = 0.f;
    param[0] = 0.0;
    for (int i = 1; i < n; ++i) {
        const float dx = __fsub_rn(pts[2 * i], pts[2 * i - 2]), dy = __fsub_rn(pts[2 * i + 1], pts[2 * i - 1]);
        const float dist = __fadd_rn(__fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy))), 1e-6f);
        cum = __fadd_rn(cum, dist);
        param[i] = (double)cum;
    }
    const double total = param[n - 1];
    for (int i = 0; i < n; ++i) param[i] = __ddiv_rn(param[i], total);
    for (int i = 0; i < n; ++i) {
        coef[2 * i] = (double)pts[2 * i];
        coef[2 * i + 1] = (double)pts[2 * i + 1];
    }
    quad_spline_fit<2>(n, param, t, band, coef);
    // evaluation at linspace(0, 1, T+2)[1:-1] (numpy: i * step)
    const double step = __ddiv_rn(1.0, (double)(T + 1));
    int j = 2;
    for (int q = 1; q <= T; ++q) {
        double y[2];
        quad_spline_eval<2>(t, n, coef, __dmul_rn((double)q, step), j, y);
        float *dst = e.traj + ((size_t)p * T + (q - 1)) * D;
        dst[0] = (float)y[0];
        dst[1] = (float)y[1];
    }
}

// ---- host side ---------------------------------------------------------------------------------------------------
size_t astar_workspace_bytes(int P, int X, int Y) { return (size_t)P * astar_layout(X, Y).per_problem; }

int launch_astar_init(const EngineDev &e, int X, int Y, double resolution, const double *origin_xy, const int32_t *start_cell,
                      const int32_t *goal_cell, unsigned char *ws, size_t ws_bytes, int32_t *path_len, cudaStream_t s) {
    NFOPP_REQUIRE(X >= 1 && Y >= 1 && (long long)X * Y <= (1 << 24), "astar: grid %d x %d cells", X, Y);
    NFOPP_REQUIRE(resolution > 0 && (origin_xy || e.checker.has_bounds), "astar: needs a positive resolution and an origin (checker boundaries)");
    NFOPP_REQUIRE(start_cell && goal_cell && ws && ws_bytes >= astar_workspace_bytes(e.P, X, Y), "astar: workspace too small");
    AstarArgs a{};
    a.e = e; a.l = astar_layout(X, Y); a.resolution = resolution;
    a.ox = origin_xy ? origin_xy[0] : e.checker.bounds[0];
    a.oy = origin_xy ? origin_xy[1] : e.checker.bounds[2];
    a.start_cell = start_cell; a.goal_cell = goal_cell; a.ws = ws; a.path_len = path_len;
    dim3 grid((a.l.cells + 255) / 256 < 64 ? (a.l.cells + 255) / 256 : 64, e.P);
    astar_label_kernel<<<grid, 256, 0, s>>>(a);
    astar_search_kernel<<<(e.P + 31) / 32, 32, 0, s>>>(a);
    astar_spline_kernel<<<(e.P + 31) / 32, 32, 0, s>>>(a);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

}  // namespace nfopp
