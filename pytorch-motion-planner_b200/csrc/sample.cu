// sample.cu - K1 inside the engine: near-path / uniform pose sampling, replay-buffer resampling with numpy's
// legacy weighted choice, and ground-truth labelling of the training set.  One CTA per problem.
// float64 where the reference is float64 (numpy), float32 where it is float32 (torch); explicit
// round-to-nearest intrinsics keep nvcc from contracting a*b+c where numpy/torch round twice.
#include "engine.cuh"

namespace nfopp {

// ---- numpy pairwise summation (numpy/core/src/umath/loops_utils.h.src, DOUBLE_pairwise_sum) ---------
__device__ double np_pairwise_sum(const double *a, int n) {
    if (n < 8) {
        double r = 0.;
        for (int i = 0; i < n; ++i) r = __dadd_rn(r, a[i]);
        return r;
    }
    if (n <= 128) {
        double r[8];
        for (int j = 0; j < 8; ++j) r[j] = a[j];
        int i;
        for (i = 8; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], a[i + j]);
        double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                               __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
        for (; i < n; ++i) res = __dadd_rn(res, a[i]);
        return res;
    }
    int n2 = n / 2;
    n2 -= n2 % 8;
    return __dadd_rn(np_pairwise_sum(a, n2), np_pairwise_sum(a + n2, n - n2));
}

// first index with cdf[idx] > u  (numpy searchsorted side='right')
__device__ __forceinline__ int search_right(const double *cdf, int n, double u) {
    int lo = 0, hi = n;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (u < cdf[mid]) hi = mid; else lo = mid + 1;
    }
    return lo;
}

struct UniformSource {
    const double *injected;   // this problem's injected stream or null
    int n_injected;
    uint64_t seed;
    uint32_t problem, step;
    __device__ __forceinline__ double get(int i, int *exhausted) const {
        if (injected) {
            if (i < n_injected) return injected[i];
            *exhausted = 1;
            return 0.5;
        }
        double u0, u1;
        rng_uniform2(seed, problem, step, RNG_CHOICE, (uint32_t)(i >> 1), &u0, &u1);
        return (i & 1) ? u1 : u0;
    }
};

struct ChoiceScratch {
    double *p, *cdf, *u;     // [n], [n], [size]
    int *newv, *found, *mark;  // [size], [size], [n]
    int *ctl;                // [4 + 32]: n_uniq, consumed, exhausted, tmp | per-warp counts of the compaction
};

// numpy.random.RandomState.choice(n, size, replace, p) (mtrand.pyx, legacy); block-cooperative.
// Result in sc.found[0..size); returns uniforms consumed (negative when the injected stream ran out).
__device__ int choice_block(const ChoiceScratch &sc, int n, int size, bool replace, const UniformSource &src) {
    const int tid = threadIdx.x, nt = blockDim.x;
    if (tid == 0) { sc.ctl[0] = 0; sc.ctl[1] = 0; sc.ctl[2] = 0; }
    __syncthreads();
    for (int guard = 0; guard < 4096; ++guard) {
        const int n_uniq = sc.ctl[0], consumed = sc.ctl[1];
        if (n_uniq >= size) break;
        const int k = size - n_uniq;
        if (!replace)
            for (int i = tid; i < n_uniq; i += nt) sc.p[sc.found[i]] = 0.0;
        for (int i = tid; i < k; i += nt) sc.u[i] = src.get(consumed + i, &sc.ctl[2]);
        for (int i = tid; i < n; i += nt) sc.mark[i] = 0x7fffffff;
        __syncthreads();
        if (tid == 0) {                      // np.cumsum: strictly sequential float64 adds (the order is the contract);
            double acc = 0.0;                // operands are fetched 8 at a time so only the adds sit on the dependent chain
            int i = 0;
            for (; i + 8 <= n; i += 8) {
                double v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = sc.p[i + j];
#pragma unroll
                for (int j = 0; j < 8; ++j) { acc = __dadd_rn(acc, v[j]); v[j] = acc; }
#pragma unroll
                for (int j = 0; j < 8; ++j) sc.cdf[i + j] = v[j];
            }
            for (; i < n; ++i) { acc = __dadd_rn(acc, sc.p[i]); sc.cdf[i] = acc; }
        }
        __syncthreads();
        const double last = sc.cdf[n - 1];
        __syncthreads();
        for (int i = tid; i < n; i += nt) sc.cdf[i] = __ddiv_rn(sc.cdf[i], last);
        __syncthreads();
        for (int i = tid; i < k; i += nt) {
            int v = min(search_right(sc.cdf, n, sc.u[i]), n - 1);
            sc.newv[i] = v;
            if (!replace) atomicMin(&sc.mark[v], i);
        }
        __syncthreads();
        // np.unique(return_index) + sort: first occurrences in draw order = an order-preserving compaction of the
        // draws whose mark is their own index (all draws when sampling with replacement): ballot + prefix over warps
        {
            int *wcount = &sc.ctl[4];                       // [nt / 32] (carve_choice reserves 4 + 32 ints)
            const int lane = tid & 31, wid = tid >> 5, nw = (nt + 31) >> 5;
            int base = n_uniq;
            for (int i0 = 0; i0 < k; i0 += nt) {            // k <= size; one round when size <= blockDim
                const int i = i0 + tid;
                const bool keep = i < k && (replace || sc.mark[sc.newv[i]] == i);
                const unsigned bal = __ballot_sync(0xffffffffu, keep);
                if (lane == 0) wcount[wid] = __popc(bal);
                __syncthreads();
                int off = base;
                for (int w = 0; w < wid; ++w) off += wcount[w];
                if (keep) sc.found[off + __popc(bal & ((1u << lane) - 1u))] = sc.newv[i];
                for (int w = 0; w < nw; ++w) base += wcount[w];
                __syncthreads();
            }
            if (tid == 0) {
                sc.ctl[0] = base;
                sc.ctl[1] = consumed + k;
            }
        }
        __syncthreads();
    }
    return sc.ctl[2] ? -sc.ctl[1] : sc.ctl[1];
}

__host__ __device__ static size_t choice_smem_bytes(int n, int size) {
    return (size_t)(2 * n + size) * sizeof(double) + (size_t)(2 * size + n + 4 + 32) * sizeof(int);
}

__device__ ChoiceScratch carve_choice(unsigned char *base, int n, int size) {
    ChoiceScratch sc;
    double *d = reinterpret_cast<double *>(base);
    sc.p = d; d += n;
    sc.cdf = d; d += n;
    sc.u = d; d += size;
    int *ip = reinterpret_cast<int *>(d);
    sc.newv = ip; ip += size;
    sc.found = ip; ip += size;
    sc.mark = ip; ip += n;
    sc.ctl = ip;
    return sc;
}

__global__ void __launch_bounds__(128) choice_kernel(const double *__restrict__ p, int n, int size, int replace,
                                                      const double *__restrict__ u, int n_u, int32_t *idx,
                                                      int32_t *used) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ChoiceScratch sc = carve_choice(smem_raw, n, size);
    const int pr = blockIdx.x;
    for (int i = threadIdx.x; i < n; i += blockDim.x) sc.p[i] = p[(size_t)pr * n + i];
    __syncthreads();
    UniformSource src{u + (size_t)pr * n_u, n_u, 0, 0, 0};
    int c = choice_block(sc, n, size, replace != 0, src);
    for (int i = threadIdx.x; i < size; i += blockDim.x) idx[(size_t)pr * size + i] = sc.found[i];
    if (threadIdx.x == 0) used[pr] = c;
}

// ---- engine kernels ------------------------------------------------------------------------------
constexpr int SAMPLE_THREADS = 128;

// nerf_opt_planner.py:101-120,135-141 / constrained:57-61,173-176: coarse + fine + uniform samples, candidates
__global__ void __launch_bounds__(SAMPLE_THREADS) sample_pre_kernel(const EngineDev e) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *s_prev = reinterpret_cast<float *>(smem_raw);      // [T*D]
    const StepParams sp = *e.sp;
    if (e.frozen[blockIdx.x]) {                      // finished problem: no candidates -> the ONF launches skip it
        if (threadIdx.x == 0) e.cand_n[blockIdx.x] = 0;
        return;
    }
    if (!sp.do_train) return;
    const int p = blockIdx.x, tid = threadIdx.x, T = e.T, D = e.D, n = T - 1;
    const bool inj = e.rng_mode == 1;
    // nerf_opt_planner.py:79-82: sample around the PREVIOUS step's trajectory, then remember the current one
    const float *src = (sp.has_prev ? e.prev_traj : e.traj) + (size_t)p * T * D;
    for (int i = tid; i < T * D; i += SAMPLE_THREADS) s_prev[i] = src[i];
    __syncthreads();
    for (int i = tid; i < T * D; i += SAMPLE_THREADS) e.prev_traj[(size_t)p * T * D + i] = e.traj[(size_t)p * T * D + i];

    const int nb = e.buf_n[p];
    const int nc = nb + n;
    const int new_n = nc < e.C ? nc : e.C;
    double *tr = e.train_pos + (size_t)p * e.NTR * D;
    double *cp = e.cand_pos + (size_t)p * e.NC * D;
    double *ca = e.cand_age + (size_t)p * e.NC;
    float *cx = e.cand_x + (size_t)p * e.NC * D;
    for (int i = tid; i < nb; i += SAMPLE_THREADS) {
        for (int d = 0; d < D; ++d) {
            double v = e.buf_pos[((size_t)p * e.C + i) * D + d];
            cp[i * D + d] = v;
            cx[i * D + d] = (float)v;
        }
        ca[i] = e.buf_age[(size_t)p * e.C + i];
    }
    for (int i = tid; i < n; i += SAMPLE_THREADS) {
        // nerf_opt_planner.py:113-117 (float32): traj[1:]*(1-t) + traj[:-1]*t
        float t;
        double zc[3], zf[3];
        if (inj) {
            t = (float)e.var_t_sample[(size_t)p * n + i];
            zc[0] = e.var_coarse_xy[((size_t)p * n + i) * 2]; zc[1] = e.var_coarse_xy[((size_t)p * n + i) * 2 + 1];
            zf[0] = e.var_fine_xy[((size_t)p * n + i) * 2]; zf[1] = e.var_fine_xy[((size_t)p * n + i) * 2 + 1];
            zc[2] = D == 3 ? e.var_coarse_th[(size_t)p * n + i] : 0.0;
            zf[2] = D == 3 ? e.var_fine_th[(size_t)p * n + i] : 0.0;
        } else {
            double u0, u1, w0, w1;
            rng_uniform2(e.seed, e.p_base + p, sp.step_count, RNG_T_SAMPLE, i, &u0, &u1);
            t = (float)u0;
            rng_normal2(e.seed, e.p_base + p, sp.step_count, RNG_COARSE, 2 * i, &zc[0], &zc[1]);
            rng_normal2(e.seed, e.p_base + p, sp.step_count, RNG_COARSE, 2 * i + 1, &zc[2], &w0);
            rng_normal2(e.seed, e.p_base + p, sp.step_count, RNG_FINE, 2 * i, &zf[0], &zf[1]);
            rng_normal2(e.seed, e.p_base + p, sp.step_count, RNG_FINE, 2 * i + 1, &zf[2], &w1);
        }
        const float omt = __fsub_rn(1.f, t);
        for (int d = 0; d < D; ++d) {
            float base = __fadd_rn(__fmul_rn(s_prev[(i + 1) * D + d], omt), __fmul_rn(s_prev[i * D + d], t));
            double off_c = d < 2 ? e.coarse_offset : e.angle_offset;
            double off_f = d < 2 ? e.fine_offset : e.angle_offset;
            double c = __dadd_rn((double)base, __dmul_rn(zc[d], off_c));
            double f = __dadd_rn((double)base, __dmul_rn(zf[d], off_f));
            if (e.constrained) {      // constrained:57-61 assigns into a float32 array (rounds)
                c = (double)__double2float_rn(c);
                f = (double)__double2float_rn(f);
            }
            tr[i * D + d] = c;
            cp[(nb + i) * D + d] = f;
            cx[(nb + i) * D + d] = (float)f;
        }
        ca[nb + i] = 0.0;
    }
    // uniform field points (nerf_opt_planner.py:135-141; constrained:173-176) behind coarse and buffer
    for (int i = tid; i < e.NF; i += SAMPLE_THREADS) {
        double u[3];
        if (inj) {
            u[0] = e.var_field_xy[((size_t)p * e.NF + i) * 2];
            u[1] = e.var_field_xy[((size_t)p * e.NF + i) * 2 + 1];
            u[2] = D == 3 ? e.var_field_th[(size_t)p * e.NF + i] : 0.0;
        } else {
            double w;
            rng_uniform2(e.seed, e.p_base + p, sp.step_count, RNG_FIELD, 2 * i, &u[0], &u[1]);
            rng_uniform2(e.seed, e.p_base + p, sp.step_count, RNG_FIELD, 2 * i + 1, &u[2], &w);
        }
        double *o = tr + (size_t)(n + new_n + i) * D;
        o[0] = __dadd_rn(e.border[0], __dmul_rn(u[0], __dsub_rn(e.border[1], e.border[0])));
        o[1] = __dadd_rn(e.border[2], __dmul_rn(u[1], __dsub_rn(e.border[3], e.border[2])));
        if (D == 3) o[2] = __dmul_rn(__dmul_rn(u[2], 2.0), 3.141592653589793);
    }
    if (tid == 0) {
        e.cand_n[p] = nc;
        e.train_n[p] = n + new_n + e.NF;
    }
}

// warm-up iteration input (nerf_opt_planner.py:197-200): NW uniform field points as the whole training set
__global__ void __launch_bounds__(SAMPLE_THREADS) warm_sample_kernel(const EngineDev e, int iteration) {
    const int p = blockIdx.x, D = e.D;
    const bool inj = e.rng_mode == 1;
    double *tr = e.train_pos + (size_t)p * e.NTR * D;
    for (int i = threadIdx.x; i < e.NW; i += SAMPLE_THREADS) {
        double u[3];
        if (inj) {
            u[0] = e.var_warm_xy[((size_t)p * e.NW + i) * 2];
            u[1] = e.var_warm_xy[((size_t)p * e.NW + i) * 2 + 1];
            u[2] = D == 3 ? e.var_warm_th[(size_t)p * e.NW + i] : 0.0;
        } else {
            double w;
            rng_uniform2(e.seed, e.p_base + p, iteration, RNG_WARM, 2 * i, &u[0], &u[1]);
            rng_uniform2(e.seed, e.p_base + p, iteration, RNG_WARM, 2 * i + 1, &u[2], &w);
        }
        double *o = tr + (size_t)i * D;
        o[0] = __dadd_rn(e.border[0], __dmul_rn(u[0], __dsub_rn(e.border[1], e.border[0])));
        o[1] = __dadd_rn(e.border[2], __dmul_rn(u[1], __dsub_rn(e.border[3], e.border[2])));
        if (D == 3) o[2] = __dmul_rn(__dmul_rn(u[2], 2.0), 3.141592653589793);
    }
    if (threadIdx.x == 0) e.train_n[p] = e.NW;
}

constexpr int RESAMPLE_THREADS = 128;

// nerf_opt_planner.py:122-133 (resample the replay buffer) + :86,93-96 (ground-truth labels of the train set)
// resample == 0: labelling only (warm-up iterations)
__global__ void __launch_bounds__(RESAMPLE_THREADS, 8) resample_label_kernel(const EngineDev e, int resample) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const StepParams sp = *e.sp;
    if (resample && e.frozen[blockIdx.x]) {          // finished problem: empty training set -> no ONF update
        if (threadIdx.x == 0) e.train_n[blockIdx.x] = 0;
        return;
    }
    if (resample && !sp.do_train) return;
    const int p = blockIdx.x, tid = threadIdx.x, D = e.D, n = e.T - 1, C = e.C;
    ChoiceScratch sc = carve_choice(smem_raw, e.NC, C);
    double *s_obs = reinterpret_cast<double *>(smem_raw + ((choice_smem_bytes(e.NC, C) + 15) / 16) * 16);
    double *tr = e.train_pos + (size_t)p * e.NTR * D;
    if (resample) {
        const int nc = e.cand_n[p];
        const double *cp = e.cand_pos + (size_t)p * e.NC * D;
        const double *ca = e.cand_age + (size_t)p * e.NC;
        double *bp = e.buf_pos + (size_t)p * C * D;
        double *ba = e.buf_age + (size_t)p * C;
        int new_n;
        if (nc < C) {                                    // :128-129 keep all, ages untouched
            new_n = nc;
            for (int i = tid; i < nc; i += RESAMPLE_THREADS) {
                for (int d = 0; d < D; ++d) bp[i * D + d] = cp[i * D + d];
                ba[i] = ca[i];
            }
            if (tid == 0) e.choice_used[p] = 0;
        } else {
            new_n = C;
            const float *cz = e.cand_z + (size_t)p * e.NC;
            for (int i = tid; i < nc; i += RESAMPLE_THREADS) {   // :125-126
                double sg = (double)sigmoid_f(cz[i]);
                sc.p[i] = __dadd_rn(__dmul_rn(sg, exp(__dmul_rn(-ca[i], 0.03))), 1e-6);
            }
            __syncthreads();
            if (tid == 0) sc.u[0] = np_pairwise_sum(sc.p, nc);   // :127 np.sum
            __syncthreads();
            const double tot = sc.u[0];
            __syncthreads();
            int big = 0;
            for (int i = tid; i < nc; i += RESAMPLE_THREADS) {
                double w = __ddiv_rn(sc.p[i], tot);
                sc.p[i] = w;
                big += w > 1e-6;
            }
            // block-wide count of w > 1e-6 (:130)
            for (int o = 16; o > 0; o >>= 1) big += __shfl_xor_sync(0xffffffffu, big, o);
            if ((tid & 31) == 0) sc.mark[tid >> 5] = big;
            __syncthreads();
            int total_big = 0;
            for (int w = 0; w < RESAMPLE_THREADS / 32; ++w) total_big += sc.mark[w];
            __syncthreads();
            const bool replace = total_big < C;
            UniformSource src{e.rng_mode == 1 ? e.var_choice_u + (size_t)p * e.max_choice_u : nullptr, e.max_choice_u,
                              e.seed, (uint32_t)(e.p_base + p), (uint32_t)sp.step_count};
            int used = choice_block(sc, nc, C, replace, src);             // :131
            if (tid == 0) {
                e.choice_used[p] = used;
                if (used < 0) atomicOr(&e.status[p], STATUS_CHOICE_EXHAUSTED);
            }
            for (int j = tid; j < C; j += RESAMPLE_THREADS) {              // :132-133
                int idx = sc.found[j];
                e.choice_idx[(size_t)p * C + j] = idx;
                for (int d = 0; d < D; ++d) bp[j * D + d] = cp[idx * D + d];
                ba[j] = __dadd_rn(ca[idx], 1.0);
            }
        }
        __syncthreads();
        if (tid == 0) e.buf_n[p] = new_n;
        for (int j = tid; j < new_n; j += RESAMPLE_THREADS)                // :109-111 middle block = buffer
            for (int d = 0; d < D; ++d) tr[(size_t)(n + j) * D + d] = bp[j * D + d];
        __syncthreads();
    }
    // ---- labels (collision_checker/*.py) + float32 copy for the ONF (:99) ----
    const int pe = e.env_shared ? 0 : p;
    const int m = (e.checker.kind == 1 || e.checker.kind == 2) ? (e.obstacle_n ? e.obstacle_n[pe] : e.m_max) : 0;
    const bool staged = e.m_max <= OBS_SMEM_POINTS;
    const double *obs = staged ? s_obs : e.obstacles + (size_t)pe * e.m_max * 2;
    if (staged)
        for (int k = tid; k < 2 * m; k += RESAMPLE_THREADS) s_obs[k] = e.obstacles[(size_t)pe * e.m_max * 2 + k];
    __syncthreads();
    const uint8_t *grid = e.grid ? e.grid + (size_t)pe * e.checker.rows * e.checker.cols : nullptr;
    const int nt = e.train_n[p];
    for (int i = tid; i < nt; i += RESAMPLE_THREADS) {
        double x = tr[(size_t)i * D], y = tr[(size_t)i * D + 1], th = D == 3 ? tr[(size_t)i * D + 2] : 0.0;
        bool hit = label_pose(e.checker, obs, m, grid, x, y, th);
        e.train_y[(size_t)p * e.NTR + i] = hit ? 1.f : 0.f;
        for (int d = 0; d < D; ++d) e.train_x[((size_t)p * e.NTR + i) * D + d] = (float)tr[(size_t)i * D + d];
    }
}

// ---- host launchers --------------------------------------------------------------------------------
int launch_sample_pre(const EngineDev &e, cudaStream_t s) {
    sample_pre_kernel<<<e.P, SAMPLE_THREADS, (size_t)e.T * e.D * sizeof(float), s>>>(e);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}
int launch_warm_sample(const EngineDev &e, int iteration, cudaStream_t s) {
    warm_sample_kernel<<<e.P, SAMPLE_THREADS, 0, s>>>(e, iteration);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}
size_t resample_smem_bytes(const EngineDev &e) {
    return ((choice_smem_bytes(e.NC, e.C) + 15) / 16) * 16 + (size_t)2 * obs_staged_points(e.m_max) * sizeof(double);
}
int resample_prepare(const EngineDev &e) {
    size_t smem = resample_smem_bytes(e);
    NFOPP_REQUIRE(smem <= 227 * 1024, "resample: %zu bytes of shared memory", smem);
    if (smem > 48 * 1024)
        NFOPP_CHECK_CUDA(cudaFuncSetAttribute(resample_label_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem));
    return NFOPP_OK;
}
int launch_resample_label(const EngineDev &e, int resample, cudaStream_t s) {
    resample_label_kernel<<<e.P, RESAMPLE_THREADS, resample_smem_bytes(e), s>>>(e, resample);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

}  // namespace nfopp

using namespace nfopp;

extern "C" int nfopp_choice(const double *p_dev, int P, int n, int size, int replace, const double *u_dev, int n_u,
                            int32_t *idx_dev, int32_t *used_dev, void *stream) {
    NFOPP_REQUIRE(p_dev && u_dev && idx_dev && used_dev && P > 0 && n > 0 && size > 0, "choice: bad arguments");
    NFOPP_REQUIRE(replace || size <= n, "choice: cannot take %d of %d without replacement", size, n);
    size_t smem = choice_smem_bytes(n, size);
    NFOPP_REQUIRE(smem <= 200 * 1024, "choice: n=%d too large", n);
    if (smem > 48 * 1024)
        NFOPP_CHECK_CUDA(cudaFuncSetAttribute(choice_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    choice_kernel<<<P, 128, smem, (cudaStream_t)stream>>>(p_dev, n, size, replace, u_dev, n_u, idx_dev, used_dev);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}
