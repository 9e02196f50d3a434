// traj.cu - K3: the trajectory half of a planner step, one CTA per problem, everything in shared memory.
//   traj_pre : random collision sample points on the current path (+ interpolated collision multipliers)
//   traj_post: analytic gradient of every loss term (CHOMP distance stencil, non-holonomic Lagrangian +
//              quadratic term, forward-only direction term, boundary term, collision term through the ONF
//              input gradients), inverse-Hessian preconditioning, Adam, multiplier ascent, and (every
//              reparametrize_trajectory_freq steps) arc-length re-parametrisation.
//   tick     : advances the device-side step counters.
// The reference does all of this with ~1900 autograd ops per step; formulas follow
// SURVEY.md appendix A.2-A.4 and cite the reference lines they restate.
#include "engine.cuh"

namespace nfopp {

constexpr int TRAJ_THREADS = 128;

// nerf_opt_planner.py:113-117,158 (holonomic) / constrained_nerf_opt_planner.py:78-83
__global__ void __launch_bounds__(TRAJ_THREADS) traj_pre_kernel(const EngineDev e) {
    const StepParams sp = *e.sp;
    const int p = blockIdx.x, T = e.T, D = e.D, n = T - 1;
    if (e.frozen[p]) {                               // finished problem: no path samples -> the ONF launch skips it
        if (threadIdx.x == 0) e.path_n[p] = 0;
        return;
    }
    const float *X = e.traj + (size_t)p * T * D;
    for (int i = threadIdx.x; i < n; i += TRAJ_THREADS) {
        float t;
        if (e.rng_mode == 1) {
            t = e.var_t_path[(size_t)p * n + i];
        } else {
            uint4 r = Philox(e.seed)(i, RNG_T_PATH, sp.step_count, e.p_base + p);
            t = e.constrained ? (float)(r.x >> 8) * (1.0f / 16777216.0f)      // torch.rand float32: 24 bits
                              : (float)u53(r.x, r.y);                           // np.random.rand -> float32
        }
        e.path_t[(size_t)p * n + i] = t;
        float *q = e.path_q + ((size_t)p * n + i) * D;
        if (e.constrained) {
            for (int d = 0; d < D; ++d) {
                float delta = __fsub_rn(X[i * D + d], X[(i + 1) * D + d]);
                if (d == 2) delta = wrap_angle_f(delta);
                q[d] = __fadd_rn(X[(i + 1) * D + d], __fmul_rn(t, delta));
            }
            const float *mu = e.mu + (size_t)p * T;
            e.path_mu[(size_t)p * n + i] = __fadd_rn(__fmul_rn(mu[i + 1], __fsub_rn(1.f, t)), __fmul_rn(mu[i], t));
        } else {
            const float omt = __fsub_rn(1.f, t);
            for (int d = 0; d < D; ++d)
                q[d] = __fadd_rn(__fmul_rn(X[(i + 1) * D + d], omt), __fmul_rn(X[i * D + d], t));
        }
    }
    if (threadIdx.x == 0) e.path_n[p] = n;
}

__device__ __forceinline__ float relu_f(float v) { return v > 0.f ? v : 0.f; }

struct TrajSmem {
    float *full;     // [(T+2)*D]
    float *hi, *lo;  // per segment: gradient contribution to node j+1 / node j, [(T+1)*D] each
    float *g;        // raw gradient per trajectory point [T][4] (x, y, theta | 0, pad): one 16-byte read per point in the mat-vec
    float *gm;       // g_mu [T]
    float *cseg;     // non-holonomic residual per segment [T+1]
    float *wd;       // wrapped dtheta per segment / segment length / cdf [T+2]
    float *aux;      // [T+2]
    float *red;      // [4]
};

__device__ TrajSmem carve_traj(float *base, int T, int D) {
    TrajSmem s;
    float *p = base;
    s.full = p; p += (T + 2) * D;
    s.hi = p; p += (T + 1) * D;
    s.lo = p; p += (T + 1) * D;
    p += (4 - ((p - base) & 3)) & 3;             // 16-byte aligned (base is)
    s.g = p; p += T * 4;
    s.gm = p; p += T;
    s.cseg = p; p += T + 1;
    s.wd = p; p += T + 2;
    s.aux = p; p += T + 2;
    s.red = p;
    return s;
}

static size_t traj_smem_bytes(int T, int D) {
    return (size_t)((T + 2) * D + 2 * (T + 1) * D + 3 + T * 4 + T + (T + 1) + 2 * (T + 2) + 4) * sizeof(float);
}

// Writes the updated path / multipliers held in shared memory (s.full, s.gm = mu, s.cseg = lambda) back to
// global memory, optionally through arc-length re-parametrisation.
__device__ void store_path(const EngineDev &e, TrajSmem &s, int p, bool do_reparam) {
    const int tid = threadIdx.x, T = e.T, D = e.D;
    float *X = e.traj + (size_t)p * T * D;
    float *lam = e.lam + (size_t)p * (T + 1), *mu = e.mu + (size_t)p * T;
    if (do_reparam) {
        // ---- nerf_opt_planner.py:224-248 / constrained:132-171 -------------------------------------
        for (int j = tid; j <= T; j += TRAJ_THREADS) {
            float acc = 0.f;
            const int dn = e.constrained ? 2 : D;              // constrained:45-47 uses x,y only
            for (int d = 0; d < dn; ++d) {
                float df = s.full[(j + 1) * D + d] - s.full[j * D + d];
                acc += df * df;
            }
            s.wd[j + 1] = sqrtf(acc);
        }
        __syncthreads();
        if (tid == 0) {
            float tot = 0.f;
            for (int j = 1; j <= T + 1; ++j) tot += s.wd[j];
            float acc = 0.f;
            s.wd[0] = 0.f;
            for (int j = 1; j <= T + 1; ++j) { acc += s.wd[j] / tot; s.wd[j] = acc; }   // cdf
        }
        __syncthreads();
        const float *cdf = s.wd;
        // node-centred constraint multipliers (constrained:161-164): [l0, (l_j+l_{j+1})/2 ..., l_T]
        if (e.constrained) {
            for (int j = tid; j <= T + 1; j += TRAJ_THREADS)
                s.aux[j] = j == 0 ? s.cseg[0] : (j == T + 1 ? s.cseg[T] : (s.cseg[j - 1] + s.cseg[j]) * 0.5f);
        }
        __syncthreads();
        for (int k = tid; k < T; k += TRAJ_THREADS) {
            const float u = __ldg(e.lin_s + k);
            int lo_i = 0, hi_i = T + 2;                        // torch.searchsorted (left): first cdf[idx] >= u
            while (lo_i < hi_i) {
                int mid = (lo_i + hi_i) >> 1;
                if (cdf[mid] < u) lo_i = mid + 1; else hi_i = mid;
            }
            const int above = min(lo_i, T + 1), below = max(lo_i - 1, 0);
            float den = cdf[above] - cdf[below];
            if (den < 1e-5f) den = 1e-5f;
            const float tau = (u - cdf[below]) / den;
            const float *A = s.full + below * D, *B = s.full + above * D;
            float *o = s.hi + k * D;                            // s.hi reused as the new trajectory
            o[0] = __fadd_rn(__fmul_rn(1.f - tau, A[0]), __fmul_rn(tau, B[0]));
            o[1] = __fadd_rn(__fmul_rn(1.f - tau, A[1]), __fmul_rn(tau, B[1]));
            if (e.constrained) {
                o[2] = __fadd_rn(A[2], __fmul_rn(tau, wrap_angle_f(B[2] - A[2])));
                const float mb = below == 0 || below == T + 1 ? 0.f : s.gm[below - 1];   // zero-padded mu
                const float ma = above == 0 || above == T + 1 ? 0.f : s.gm[above - 1];
                s.lo[k] = __fadd_rn(__fmul_rn(1.f - tau, mb), __fmul_rn(tau, ma));       // new mu
                s.lo[T + k] = __fadd_rn(__fmul_rn(1.f - tau, s.aux[below]), __fmul_rn(tau, s.aux[above]));
            }
        }
        __syncthreads();
        for (int i = tid; i < T * D; i += TRAJ_THREADS) X[i] = s.hi[i];
        if (e.constrained) {
            for (int k = tid; k < T; k += TRAJ_THREADS) mu[k] = s.lo[k];
            const float *cn = s.lo + T;                        // constrained:168-171 back to segments
            for (int j = tid; j <= T; j += TRAJ_THREADS)
                lam[j] = j == 0 ? cn[0] : (j == T ? cn[T - 1] : (cn[j - 1] + cn[j]) * 0.5f);
        }
    } else {
        for (int i = tid; i < T * D; i += TRAJ_THREADS) X[i] = s.full[D + i];
        if (e.constrained) {
            for (int k = tid; k < T; k += TRAJ_THREADS) mu[k] = s.gm[k];
            for (int j = tid; j <= T; j += TRAJ_THREADS) lam[j] = s.cseg[j];
        }
    }
}


__global__ void __launch_bounds__(TRAJ_THREADS) traj_post_kernel(const EngineDev e) {
    extern __shared__ __align__(16) float smem_f[];
    const StepParams sp = *e.sp;
    const int p = blockIdx.x, tid = threadIdx.x, T = e.T, D = e.D, n = T - 1;
    if (e.frozen[p]) return;                         // finished problem: trajectory, moments and multipliers stay as they are
    TrajSmem s = carve_traj(smem_f, T, D);
    float *X = e.traj + (size_t)p * T * D;
    const float *gq = e.path_gq + (size_t)p * n * D;
    const float *pz = e.path_z + (size_t)p * n;
    const float *pt = e.path_t + (size_t)p * n;
    float *lam = e.lam + (size_t)p * (T + 1), *mu = e.mu + (size_t)p * T;

    // full trajectory = start | trajectory | goal  (nerf_opt_planner.py:73-74)
    for (int i = tid; i < (T + 2) * D; i += TRAJ_THREADS) {
        int node = i / D, d = i - node * D;
        s.full[i] = node == 0 ? e.start[(size_t)p * D + d]
                              : (node == T + 1 ? e.goal[(size_t)p * D + d] : X[(node - 1) * D + d]);
    }
    __syncthreads();
    float ang_sum = 0.f;
    if (e.constrained) {
        // constrained:123-125: s = sum(wrap(dtheta)).detach() - theta_goal + theta_start
        for (int j = tid; j <= T; j += TRAJ_THREADS)
            s.wd[j] = wrap_angle_f(__fsub_rn(s.full[(j + 1) * 3 + 2], s.full[j * 3 + 2]));
        __syncthreads();
        if (tid == 0) {
            float acc = 0.f;
            for (int j = 0; j <= T; ++j) acc += s.wd[j];
            s.red[0] = acc - s.full[(T + 1) * 3 + 2] + s.full[2];
        }
        __syncthreads();
        ang_sum = s.red[0];
    }
    // ---- per-segment terms -------------------------------------------------------------------------
    for (int j = tid; j <= T; j += TRAJ_THREADS) {
        const float *a = s.full + j * D, *b = s.full + (j + 1) * D;
        const float dx = b[0] - a[0], dy = b[1] - a[1];
        float hx = 2.f * dx, hy = 2.f * dy, lx = -hx, ly = -hy, hth = 0.f, lth = 0.f;   // distance term
        if (e.constrained) {
            const float aw = e.angle_weight;
            float dth = b[2] - a[2];
            if (j == T) dth += ang_sum;                      // constrained:126
            dth *= aw;                                       // :127
            hth = 2.f * aw * dth;
            lth = -hth;
            // non-holonomic residual (:102-109) with multiplier + quadratic weight (:94-96)
            const float abar = a[2] + wrap_angle_f(b[2] - a[2]) * 0.5f;
            float sn, cs;
            sincosf(abar, &sn, &cs);
            const float c = dx * sn - dy * cs;
            s.cseg[j] = c;
            const float kap = lam[j] + 2.f * e.constraint_weight * c;
            hx += kap * sn; lx -= kap * sn;
            hy -= kap * cs; ly += kap * cs;
            const float dcda = kap * (dx * cs + dy * sn) * 0.5f;
            hth += dcda; lth += dcda;
            if (e.direction_weight != 0.f) {                 // :111-118,93,98 (wrap(a_j - a_{j+1}))
                const float abar2 = a[2] + wrap_angle_f(a[2] - b[2]) * 0.5f;
                float sn2, cs2;
                sincosf(abar2, &sn2, &cs2);
                const float dd = -(cs2 * dx + sn2 * dy);
                const float rho = 2.f * e.direction_weight * relu_f(dd);
                hx -= rho * cs2; lx += rho * cs2;
                hy -= rho * sn2; ly += rho * sn2;
                const float ddda = rho * (sn2 * dx - cs2 * dy);
                lth += 1.5f * ddda;
                hth -= 0.5f * ddda;
            }
            s.hi[j * 3 + 2] = hth; s.lo[j * 3 + 2] = lth;
        }
        s.hi[j * D] = hx; s.hi[j * D + 1] = hy;
        s.lo[j * D] = lx; s.lo[j * D + 1] = ly;
    }
    __syncthreads();
    // ---- per-point gradient ------------------------------------------------------------------------
    for (int i = tid; i < T; i += TRAJ_THREADS) {
        float g[3];
        for (int d = 0; d < D; ++d) g[d] = s.hi[i * D + d] + s.lo[(i + 1) * D + d];
        if (e.constrained) {                                  // boundary term nerf_opt_planner.py:171-176
            const float x = s.full[(i + 1) * D], y = s.full[(i + 1) * D + 1];
            g[0] += e.boundary_weight * 2.f * (relu_f(x - e.border_f[1]) - relu_f(-x + e.border_f[0]));
            g[1] += e.boundary_weight * 2.f * (relu_f(y - e.border_f[3]) - relu_f(-y + e.border_f[2]));
        }
        float gmu = 0.f;
        if (i < n) {                                          // q_i depends on X[i] with weight t_i
            const float t = pt[i];
            for (int d = 0; d < D; ++d) g[d] += t * gq[i * D + d];
            gmu += t * tanhf(pz[i]);
        }
        if (i >= 1) {                                         // and on X[i] via q_{i-1} with weight 1-t
            const float omt = 1.f - pt[i - 1];
            for (int d = 0; d < D; ++d) g[d] += omt * gq[(i - 1) * D + d];
            gmu += omt * tanhf(pz[i - 1]);
        }
        *reinterpret_cast<float4 *>(s.g + 4 * i) = make_float4(g[0], g[1], D == 3 ? g[2] : 0.f, 0.f);
        s.gm[i] = gmu;
    }
    __syncthreads();
    // ---- grad <- inv_hessian @ grad (nerf_opt_planner.py:151), Adam (:154) --------------------------
    for (int i = tid; i < T; i += TRAJ_THREADS) {
        float acc[3] = {0.f, 0.f, 0.f};
        // (the matrix is L2-resident: ten rows' loads in flight at a time - one after the other the kernel was bound by 100
        // dependent L2 latencies per thread; the accumulation order is unchanged)
        int j = 0;
        for (; j + 10 <= T; j += 10) {
            float h[10];
#pragma unroll
            for (int u = 0; u < 10; ++u) h[u] = __ldg(e.inv_hessian_t + (size_t)(j + u) * T + i);   // = inv_hessian[i][j + u], coalesced
#pragma unroll
            for (int u = 0; u < 10; ++u) {
                const float4 g4 = *reinterpret_cast<const float4 *>(s.g + 4 * (j + u));
                acc[0] = fmaf(h[u], g4.x, acc[0]); acc[1] = fmaf(h[u], g4.y, acc[1]); acc[2] = fmaf(h[u], g4.z, acc[2]);
            }
        }
        for (; j < T; ++j) {
            const float h = __ldg(e.inv_hessian_t + (size_t)j * T + i);
            const float4 g4 = *reinterpret_cast<const float4 *>(s.g + 4 * j);
            acc[0] = fmaf(h, g4.x, acc[0]); acc[1] = fmaf(h, g4.y, acc[1]); acc[2] = fmaf(h, g4.z, acc[2]);
        }
        for (int d = 0; d < D; ++d) {
            const size_t o = (size_t)p * T * D + i * D + d;
            float m = e.traj_m[o], v = e.traj_v[o], w = X[i * D + d];
            const float gr = acc[d];
            m = fmaf(e.traj_w1, gr - m, m);
            v = fmaf(e.traj_w2 * gr, gr, v * e.traj_beta2);
            const float denom = sqrtf(v) / sp.traj_bc2_sqrt + e.traj_eps;
            w = fmaf(-sp.traj_step_size, m / denom, w);
            e.traj_m[o] = m; e.traj_v[o] = v;
            s.full[(i + 1) * D + d] = w;
            if (!(w == w) || fabsf(w) > 3.0e38f) atomicOr(&e.status[p], STATUS_NONFINITE);
        }
        if (e.constrained) {                                   // constrained:68-73
            float m2 = mu[i] + e.mu_lr * s.gm[i];
            s.gm[i] = m2 > 0.f ? m2 : 0.f;
        }
    }
    if (e.constrained)                                         // constrained:66 (ascent on the residual)
        for (int j = tid; j <= T; j += TRAJ_THREADS) s.cseg[j] = lam[j] + e.multipliers_lr * s.cseg[j];
    __syncthreads();
    // now: s.full = updated path, s.gm = updated mu, s.cseg = updated lambda
    store_path(e, s, p, sp.do_reparam != 0);
}

// update_goal_point / update_start_point (nerf_opt_planner.py:202-218, constrained:178-194) and tests:
// re-parametrise the stored trajectory (and multipliers) of every problem without an optimisation step.
__global__ void __launch_bounds__(TRAJ_THREADS) reparam_kernel(const EngineDev e) {
    extern __shared__ __align__(16) float smem_f[];
    const int p = blockIdx.x, tid = threadIdx.x, T = e.T, D = e.D;
    TrajSmem s = carve_traj(smem_f, T, D);
    const float *X = e.traj + (size_t)p * T * D;
    for (int i = tid; i < (T + 2) * D; i += TRAJ_THREADS) {
        int node = i / D, d = i - node * D;
        s.full[i] = node == 0 ? e.start[(size_t)p * D + d]
                              : (node == T + 1 ? e.goal[(size_t)p * D + d] : X[(node - 1) * D + d]);
    }
    for (int k = tid; k < T; k += TRAJ_THREADS) s.gm[k] = e.mu[(size_t)p * T + k];
    for (int j = tid; j <= T; j += TRAJ_THREADS) s.cseg[j] = e.lam[(size_t)p * (T + 1) + j];
    __syncthreads();
    store_path(e, s, p, true);
}

// last kernel of a step: advance counters, precompute next step's flags and Adam bias corrections
__global__ void tick_kernel(const EngineDev e, int trained_this_step_unknown) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    StepParams sp = *e.sp;
    if (sp.do_train) { sp.onf_step += 1; sp.has_prev = 1; }
    sp.traj_step += 1;
    sp.step_count += 1;
    sp.do_train = (sp.step_count % e.train_freq) == 0;
    sp.do_reparam = (sp.step_count % e.reparam_freq) == 0;
    const double ko = sp.onf_step + 1, kt = sp.traj_step + 1;
    sp.onf_step_size = (float)(e.onf_lr_d / (1.0 - pow(e.onf_b1_d, ko)));
    sp.onf_bc2_sqrt = (float)sqrt(1.0 - pow(e.onf_b2_d, ko));
    sp.traj_step_size = (float)(e.traj_lr_d / (1.0 - pow(e.traj_b1_d, kt)));
    sp.traj_bc2_sqrt = (float)sqrt(1.0 - pow(e.traj_b2_d, kt));
    *e.sp = sp;
}

// warm-up iterations only advance the ONF optimiser
__global__ void tick_warm_kernel(const EngineDev e) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    StepParams sp = *e.sp;
    sp.onf_step += 1;
    const double ko = sp.onf_step + 1;
    sp.onf_step_size = (float)(e.onf_lr_d / (1.0 - pow(e.onf_b1_d, ko)));
    sp.onf_bc2_sqrt = (float)sqrt(1.0 - pow(e.onf_b2_d, ko));
    *e.sp = sp;
}

__global__ void paths_kernel(const EngineDev e, float *__restrict__ out) {
    const int p = blockIdx.x, T = e.T, D = e.D;
    for (int i = threadIdx.x; i < (T + 2) * D; i += blockDim.x) {
        int node = i / D, d = i - node * D;
        out[(size_t)p * (T + 2) * D + i] = node == 0 ? e.start[(size_t)p * D + d]
                                           : (node == T + 1 ? e.goal[(size_t)p * D + d]
                                                            : e.traj[((size_t)p * T + node - 1) * D + d]);
    }
}

int launch_traj_pre(const EngineDev &e, cudaStream_t s) {
    traj_pre_kernel<<<e.P, TRAJ_THREADS, 0, s>>>(e);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}
int traj_prepare(const EngineDev &e) {
    size_t smem = traj_smem_bytes(e.T, e.D);
    NFOPP_REQUIRE(smem <= 227 * 1024, "trajectory_length=%d exceeds shared memory", e.T);
    if (smem > 48 * 1024)
    {
        NFOPP_CHECK_CUDA(cudaFuncSetAttribute(traj_post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        NFOPP_CHECK_CUDA(cudaFuncSetAttribute(reparam_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    return NFOPP_OK;
}
int launch_traj_post(const EngineDev &e, cudaStream_t s) {
    traj_post_kernel<<<e.P, TRAJ_THREADS, traj_smem_bytes(e.T, e.D), s>>>(e);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}
int launch_reparam(const EngineDev &e, cudaStream_t s) {
    reparam_kernel<<<e.P, TRAJ_THREADS, traj_smem_bytes(e.T, e.D), s>>>(e);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}
int launch_tick(const EngineDev &e, bool warm, cudaStream_t s) {
    if (warm) tick_warm_kernel<<<1, 32, 0, s>>>(e);
    else tick_kernel<<<1, 32, 0, s>>>(e, 0);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}
int launch_paths(const EngineDev &e, float *out, cudaStream_t s) {
    paths_kernel<<<e.P, 128, 0, s>>>(e, out);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

// ---- path evaluation: what scripts/run_bench_mr.py:119-126 asks bench-mr's evaluatePath for every
// check_collision_frequence iterations (BenchmarkAdapter.cpp:201-208: interpolate the path, validity-check every
// interpolated pose, path length).  bench-mr's interpolation is in a missing submodule; ours is `k` equally spaced poses
// per segment (x, y linear, heading along the wrapped difference), float64, the arithmetic oracle/evaluate.py restates.
// Each pose is labelled by the engine's own checker (K1).  One CTA per problem.
constexpr int EVAL_THREADS = 128;

__device__ __forceinline__ double wrap_angle_d(double a) {             // numpy: (a + pi) % (2 pi) - pi, python-sign modulo
    const double two_pi = 6.283185307179586, pi = 3.141592653589793;
    double r = fmod(__dadd_rn(a, pi), two_pi);
    if (r != 0.0 && r < 0.0) r = __dadd_rn(r, two_pi);
    return __dsub_rn(r, pi);
}

__global__ void __launch_bounds__(EVAL_THREADS) evaluate_paths_kernel(const EngineDev e, int k, uint8_t *__restrict__ collides,
                                                                        float *__restrict__ length, int32_t *__restrict__ n_hit) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *s_obs = reinterpret_cast<double *>(smem_raw);
    __shared__ int s_hits;
    __shared__ double s_len[EVAL_THREADS / 32];
    const int p = blockIdx.x, tid = threadIdx.x, T = e.T, D = e.D;
    const int pe = e.env_shared ? 0 : p;
    const int m = (e.checker.kind == 1 || e.checker.kind == 2) ? (e.obstacle_n ? e.obstacle_n[pe] : e.m_max) : 0;
    const bool staged = e.m_max <= OBS_SMEM_POINTS;
    const double *obs = staged ? s_obs : e.obstacles + (size_t)pe * e.m_max * 2;
    if (staged)
        for (int i = tid; i < 2 * m; i += EVAL_THREADS) s_obs[i] = e.obstacles[(size_t)pe * e.m_max * 2 + i];
    if (tid == 0) s_hits = 0;
    __syncthreads();
    const uint8_t *grid = e.grid ? e.grid + (size_t)pe * e.checker.rows * e.checker.cols : nullptr;
    auto node = [&](int j, int d) -> double {
        return (double)(j == 0 ? e.start[(size_t)p * D + d] : (j == T + 1 ? e.goal[(size_t)p * D + d] : e.traj[((size_t)p * T + j - 1) * D + d]));
    };
    const int segs = T + 1, poses = segs * k + 1;
    int hits = 0;
    for (int i = tid; i < poses; i += EVAL_THREADS) {
        const int j = i < segs * k ? i / k : segs - 1, r = i - j * k;          // the last pose is the goal itself
        const double sfrac = __ddiv_rn((double)r, (double)k);
        const double x0 = node(j, 0), y0 = node(j, 1), x1 = node(j + 1, 0), y1 = node(j + 1, 1);
        const double x = __dadd_rn(x0, __dmul_rn(sfrac, __dsub_rn(x1, x0)));
        const double y = __dadd_rn(y0, __dmul_rn(sfrac, __dsub_rn(y1, y0)));
        double th = 0.0;
        if (D == 3) {
            const double t0 = node(j, 2), t1 = node(j + 1, 2);
            th = __dadd_rn(t0, __dmul_rn(sfrac, wrap_angle_d(__dsub_rn(t1, t0))));
        }
        hits += label_pose(e.checker, obs, m, grid, x, y, th) ? 1 : 0;
    }
    double len = 0.0;                                                          // sum of segment lengths (x, y)
    for (int j = tid; j < segs; j += EVAL_THREADS) {
        const double dx = __dsub_rn(node(j + 1, 0), node(j, 0)), dy = __dsub_rn(node(j + 1, 1), node(j, 1));
        len += __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    }
    for (int o = 16; o > 0; o >>= 1) {
        hits += __shfl_xor_sync(0xffffffffu, hits, o);
        len += __shfl_xor_sync(0xffffffffu, len, o);
    }
    if ((tid & 31) == 0) {
        atomicAdd(&s_hits, hits);
        s_len[tid >> 5] = len;
    }
    __syncthreads();
    if (tid == 0) {
        double tot = 0.0;
        for (int w = 0; w < EVAL_THREADS / 32; ++w) tot += s_len[w];
        collides[p] = s_hits > 0;
        length[p] = (float)tot;
        if (n_hit) n_hit[p] = s_hits;
    }
}

int launch_evaluate_paths(const EngineDev &e, int k, uint8_t *collides, float *length, int32_t *n_hit, cudaStream_t s) {
    NFOPP_REQUIRE(k >= 1 && k <= 4096, "evaluate_paths: samples_per_segment=%d", k);
    const size_t smem = (size_t)2 * obs_staged_points(e.m_max) * sizeof(double);
    evaluate_paths_kernel<<<e.P, EVAL_THREADS, smem, s>>>(e, k, collides, length, n_hit);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

// ---- path-quality statistics: the numbers BenchmarkAdapter::evaluate (BenchmarkAdapter.cpp:68-103) files
// per path - length, max / normalised curvature, angle over length, OMPL smoothness, cusps, clearing - on the T+2 nodes
// of every problem's path (x, y; float64).  computeCusps follows BenchmarkAdapter.cpp:105-135 line by line; the other
// metrics live in the missing bench-mr / OMPL submodules and are restated from their published definitions
// (oracle/path_stats.py is the CPU restatement; parity unpinned for those).  One warp per problem.
constexpr int STATS_PER_PROBLEM = 8;   // length, max curvature, normalised curvature, aol, smoothness, cusps, min / mean clearing

__device__ __forceinline__ double norm_angle_d(double a) {   // PlannerUtils::normalizeAngle: into [-pi, pi)
    const double two_pi = 6.283185307179586, pi = 3.141592653589793;
    a = fmod(a + pi, two_pi);
    if (a < 0.0) a += two_pi;
    return a - pi;
}

__device__ __forceinline__ double warp_sum_d(double v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max_d(double v) {
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_min_d(double v) {
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// distance from a pose to the nearest obstacle of the engine's environment (>= 0 outside, 0 in collision)
__device__ double clearing_distance(const EngineDev &e, int pe, int m, double x, double y, double th, int window) {
    const CheckerParams &c = e.checker;
    double best = 1e30;
    if (c.kind == 1 || c.kind == 2) {
        const double *obs = e.obstacles + (size_t)pe * e.m_max * 2;
        const double ct = cos(th), st = sin(th);
        for (int i = 0; i < m; ++i) {
            const double dx = obs[2 * i] - x, dy = obs[2 * i + 1] - y;
            double dist;
            if (c.kind == 1) {
                dist = sqrt(dx * dx + dy * dy) - c.radius;
            } else {                                              // obstacle point in the robot frame vs the box
                const double rx = ct * dx + st * dy, ry = -st * dx + ct * dy;
                const double ex = fmax(fmax(c.box[0] - rx, rx - c.box[1]), 0.0), ey = fmax(fmax(c.box[2] - ry, ry - c.box[3]), 0.0);
                dist = sqrt(ex * ex + ey * ey);
            }
            best = fmin(best, dist);
        }
    } else if (c.kind == 3) {
        const uint8_t *grid = e.grid + (size_t)pe * c.rows * c.cols;
        const int cx = (int)floor((x - c.ox) / c.cell), cy = (int)floor((y - c.oy) / c.cell);
        for (int iy = max(cy - window, 0); iy <= min(cy + window, c.rows - 1); ++iy)
            for (int ix = max(cx - window, 0); ix <= min(cx + window, c.cols - 1); ++ix)
                if (grid[(size_t)iy * c.cols + ix] > 0) {
                    // nearest point of the occupied cell's square [ox + ix cell, +cell] x [oy + iy cell, +cell]
                    const double x0 = c.ox + ix * c.cell, y0 = c.oy + iy * c.cell;
                    const double ex = fmax(fmax(x0 - x, x - (x0 + c.cell)), 0.0), ey = fmax(fmax(y0 - y, y - (y0 + c.cell)), 0.0);
                    best = fmin(best, sqrt(ex * ex + ey * ey));
                }
        best = fmin(best, (double)window * c.cell);               // nothing inside the search window
    }
    if (c.has_bounds) best = fmin(best, fmin(fmin(x - c.bounds[0], c.bounds[1] - x), fmin(y - c.bounds[2], c.bounds[3] - y)));
    return fmax(best, 0.0);
}

__global__ void __launch_bounds__(128) path_statistics_kernel(const EngineDev e, double cusp_threshold, int window, float *__restrict__ out) {
    const int p = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (p >= e.P) return;
    const int T = e.T, D = e.D, n = T + 2;
    const int pe = e.env_shared ? 0 : p;
    const int m = (e.checker.kind == 1 || e.checker.kind == 2) ? (e.obstacle_n ? e.obstacle_n[pe] : e.m_max) : 0;
    auto node = [&](int j, int d) -> double {
        return (double)(j == 0 ? e.start[(size_t)p * D + d] : (j == T + 1 ? e.goal[(size_t)p * D + d] : e.traj[((size_t)p * T + j - 1) * D + d]));
    };
    double len = 0.0, kmax = 0.0, knorm = 0.0, smooth = 0.0, cmin = 1e30, csum = 0.0;
    for (int j = lane; j < n; j += 32) {
        const double x = node(j, 0), y = node(j, 1);
        if (j + 1 < n) {
            const double dx = node(j + 1, 0) - x, dy = node(j + 1, 1) - y;
            len += sqrt(dx * dx + dy * dy);
        }
        if (j >= 1 && j + 1 < n) {
            const double ax = x - node(j - 1, 0), ay = y - node(j - 1, 1), bx = node(j + 1, 0) - x, by = node(j + 1, 1) - y;
            const double a = sqrt(ax * ax + ay * ay), b = sqrt(bx * bx + by * by);
            const double cx = ax + bx, cy = ay + by, c = sqrt(cx * cx + cy * cy);
            if (a > 1e-12 && b > 1e-12 && c > 1e-12) {            // Menger curvature of (p_{j-1}, p_j, p_{j+1})
                const double kj = 2.0 * fabs(ax * by - ay * bx) / (a * b * c);
                kmax = fmax(kmax, kj);
                knorm += kj * 0.5 * (a + b);
            }
            if (a > 0.0 && b > 0.0) {                             // ompl::geometric::PathGeometric::smoothness
                const double acv = (a * a + b * b - c * c) / (2.0 * a * b);
                if (acv > -1.0 && acv < 1.0) {
                    const double ang = 3.141592653589793 - acos(acv);
                    const double kk = 2.0 * ang / (a + b);
                    smooth += kk * kk;
                }
            }
        }
        const double cd = clearing_distance(e, pe, m, x, y, D == 3 ? node(j, 2) : 0.0, window);
        cmin = fmin(cmin, cd);
        csum += cd;
    }
    len = warp_sum_d(len); kmax = warp_max_d(kmax); knorm = warp_sum_d(knorm); smooth = warp_sum_d(smooth);
    cmin = warp_min_d(cmin); csum = warp_sum_d(csum);
    if (lane == 0) {
        // BenchmarkAdapter::computeCusps (BenchmarkAdapter.cpp:105-135) + the summed heading change for angle-over-length
        int cusps = 0, prev = 0, cur = 0, next = 0;
        double turn = 0.0;
        auto dist = [&](int i, int j) { const double dx = node(i, 0) - node(j, 0), dy = node(i, 1) - node(j, 1); return sqrt(dx * dx + dy * dy); };
        while (next < n) {
            if (dist(prev, cur) <= 0.0) { ++cur; ++next; if (cur >= n) break; }
            else if (dist(cur, next) <= 0.0) { ++next; }
            else {
                const double yaw_prev = atan2(node(cur, 1) - node(prev, 1), node(cur, 0) - node(prev, 0));
                const double yaw_next = atan2(node(next, 1) - node(cur, 1), node(next, 0) - node(cur, 0));
                const double change = fabs(norm_angle_d(yaw_next - yaw_prev));
                if (change > cusp_threshold) ++cusps;
                turn += change;
                prev = cur; cur = next; ++next;
            }
        }
        float *o = out + (size_t)p * STATS_PER_PROBLEM;
        o[0] = (float)len; o[1] = (float)kmax; o[2] = (float)knorm; o[3] = (float)(len > 0.0 ? turn / len : 0.0);
        o[4] = (float)smooth; o[5] = (float)cusps; o[6] = (float)cmin; o[7] = (float)(csum / n);
    }
}

int launch_path_statistics(const EngineDev &e, double cusp_threshold, int window, float *out, cudaStream_t s) {
    NFOPP_REQUIRE(window >= 0 && window <= 4096, "path_statistics: clearing window=%d cells", window);
    path_statistics_kernel<<<(e.P + 3) / 4, 128, 0, s>>>(e, cusp_threshold, window, out);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

}  // namespace nfopp
