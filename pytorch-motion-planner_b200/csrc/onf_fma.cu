// onf_fma.cu - K2, CUDA-core path: fused ONF encode -> MLP -> logit, with recompute-free backward
// (dL/dz -> per-tile weight-gradient partials and input gradients) in ONE kernel per call, and the fused
// gradient-reduce + Adam kernel.  fp32 FMA, no tensor cores: this is the narrow-layer fallback of the design
// (north_star: "warp-level FMA/shuffle code for narrow layers") and the on-device fp32 reference the
// tcgen05 path (onf_tc.cu) is validated against.
//
// Work decomposition: grid = (tiles, P); one CTA = one tile of TP points of one problem.  Inside a warp a
// lane owns one point (lane % TP) and a register strip of JT outputs, so activations come from shared
// memory as conflict-free 128-bit loads (row stride/4 is odd) and weights as warp-uniform 128-bit global
// loads served by L1/L2 (each weight is used by exactly one warp per tile).
#include "onf.cuh"

namespace nfopp {

constexpr int ONF_THREADS = 512;
constexpr int ONF_WARPS = ONF_THREADS / 32;

__host__ __device__ inline int smem_stride(int k) {   // multiple of 4 floats, (stride/4) odd
    int s = (k + 3) / 4 * 4;
    if (((s / 4) & 1) == 0) s += 4;
    return s;
}

template <int TP>
struct OnfSmem {
    float *e, *de, *ge, *h[NFOPP_MAX_LAYERS], *gA, *gB, *u, *th, *z, *gbar, *part;
    int S_e, S_h;
    __device__ OnfSmem(float *base, const OnfDims &d, bool bwd) {
        S_e = smem_stride(d.lay.f_in);
        S_h = smem_stride(d.H);
        float *p = base;
        e = p; p += TP * S_e;
        for (int l = 0; l < d.L; ++l) { h[l] = p; p += TP * S_h; }
        u = p; p += TP * 2;
        th = p; p += TP;
        z = p; p += TP;
        gbar = p; p += TP;
        part = p; p += 2 * ONF_WARPS * TP;
        if (bwd) {
            de = p; p += TP * S_e;
            ge = p; p += TP * S_e;
            gA = p; p += TP * S_h;
            gB = p; p += TP * S_h;
        } else {
            de = ge = gA = gB = nullptr;
        }
    }
};

static size_t onf_smem_bytes(const OnfDims &d, int TP, bool bwd) {
    size_t S_e = smem_stride(d.lay.f_in), S_h = smem_stride(d.H);
    size_t f = TP * S_e + (size_t)d.L * TP * S_h + TP * 5 + 2 * ONF_WARPS * TP;
    if (bwd) f += 2 * TP * S_e + 2 * TP * S_h;
    return f * sizeof(float);
}

__device__ __forceinline__ float dot4(const float4 &a, const float4 &b, float acc) {
    acc = fmaf(a.x, b.x, acc);
    acc = fmaf(a.y, b.y, acc);
    acc = fmaf(a.z, b.z, acc);
    return fmaf(a.w, b.w, acc);
}

// out[pt][j] = act( b[j] + sum_k in[pt][k] * W[j][k] )
template <int TP, int JT>
__device__ __forceinline__ void fwd_layer(const float *__restrict__ in, int S_in, int K, const float *__restrict__ W,
                                          const float *__restrict__ b, float *__restrict__ out, int S_out, int H) {
    constexpr int SUB = 32 / TP;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pt = lane % TP, sub = lane / TP;
    const int nchunks = (H + JT - 1) / JT;
    for (int chunk = warp * SUB + sub; chunk < nchunks; chunk += ONF_WARPS * SUB) {
        const int j0 = chunk * JT;
        float acc[JT];
        const float *wrow[JT];
#pragma unroll
        for (int i = 0; i < JT; ++i) {
            int j = min(j0 + i, H - 1);
            acc[i] = __ldg(b + j);
            wrow[i] = W + (size_t)j * K;
        }
        const float *arow = in + pt * S_in;
#pragma unroll 2
        for (int k = 0; k < K; k += 4) {
            float4 a = *reinterpret_cast<const float4 *>(arow + k);
#pragma unroll
            for (int i = 0; i < JT; ++i) {
                float4 w = __ldg(reinterpret_cast<const float4 *>(wrow[i] + k));
                acc[i] = dot4(a, w, acc[i]);
            }
        }
#pragma unroll
        for (int i = 0; i < JT; ++i)
            if (j0 + i < H) out[pt * S_out + j0 + i] = fmaxf(acc[i], 0.f);
    }
}

// gin[pt][k] (op)= sum_j gout[pt][j] * W[j][k];  mask != null: gin = relu'(mask) * sum ; else gin += sum
template <int TP>
__device__ __forceinline__ void dgrad_layer(const float *__restrict__ gout, int S_g, int H, const float *__restrict__ W,
                                            int K, float *__restrict__ gin, int S_in,
                                            const float *__restrict__ mask) {
    constexpr int SUB = 32 / TP;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pt = lane % TP, sub = lane / TP;
    const int K4 = K / 4, nchunks = (K4 + 1) / 2;
    for (int chunk = warp * SUB + sub; chunk < nchunks; chunk += ONF_WARPS * SUB) {
        const int k0 = chunk * 8;
        const bool two = (2 * chunk + 1) < K4;
        const int k1 = two ? k0 + 4 : k0;
        float4 a0 = make_float4(0, 0, 0, 0), a1 = a0;
        const float *grow = gout + pt * S_g;
#pragma unroll 2
        for (int j = 0; j < H; j += 4) {
            float4 g = *reinterpret_cast<const float4 *>(grow + j);
            const float gv[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                const float *wr = W + (size_t)(j + jj) * K;
                float4 w0 = __ldg(reinterpret_cast<const float4 *>(wr + k0));
                float4 w1 = __ldg(reinterpret_cast<const float4 *>(wr + k1));
                a0.x = fmaf(gv[jj], w0.x, a0.x); a0.y = fmaf(gv[jj], w0.y, a0.y);
                a0.z = fmaf(gv[jj], w0.z, a0.z); a0.w = fmaf(gv[jj], w0.w, a0.w);
                a1.x = fmaf(gv[jj], w1.x, a1.x); a1.y = fmaf(gv[jj], w1.y, a1.y);
                a1.z = fmaf(gv[jj], w1.z, a1.z); a1.w = fmaf(gv[jj], w1.w, a1.w);
            }
        }
        float *dst = gin + pt * S_in;
        if (mask) {
            const float *mrow = mask + pt * S_in;
            float4 m0 = *reinterpret_cast<const float4 *>(mrow + k0);
            a0.x = m0.x > 0.f ? a0.x : 0.f; a0.y = m0.y > 0.f ? a0.y : 0.f;
            a0.z = m0.z > 0.f ? a0.z : 0.f; a0.w = m0.w > 0.f ? a0.w : 0.f;
            *reinterpret_cast<float4 *>(dst + k0) = a0;
            if (two) {
                float4 m1 = *reinterpret_cast<const float4 *>(mrow + k1);
                a1.x = m1.x > 0.f ? a1.x : 0.f; a1.y = m1.y > 0.f ? a1.y : 0.f;
                a1.z = m1.z > 0.f ? a1.z : 0.f; a1.w = m1.w > 0.f ? a1.w : 0.f;
                *reinterpret_cast<float4 *>(dst + k1) = a1;
            }
        } else {
            float4 o0 = *reinterpret_cast<float4 *>(dst + k0);
            o0.x += a0.x; o0.y += a0.y; o0.z += a0.z; o0.w += a0.w;
            *reinterpret_cast<float4 *>(dst + k0) = o0;
            if (two) {
                float4 o1 = *reinterpret_cast<float4 *>(dst + k1);
                o1.x += a1.x; o1.y += a1.y; o1.z += a1.z; o1.w += a1.w;
                *reinterpret_cast<float4 *>(dst + k1) = o1;
            }
        }
    }
}

// dW[j][k] = sum_n gout[n][j] * in[n][k]  -> global partial (float4 stores);  db[j] = sum_n gout[n][j]
template <int TP, int JT>
__device__ __forceinline__ void wgrad_layer(const float *__restrict__ gout, int S_g, int H, const float *__restrict__ in,
                                            int S_in, int K, float *__restrict__ dW, float *__restrict__ db) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int K4 = K / 4, nchunks = (H + JT - 1) / JT;
    for (int chunk = warp; chunk < nchunks; chunk += ONF_WARPS) {
        const int j0 = chunk * JT;
        for (int kc = lane; kc < K4; kc += 32) {
            float4 acc[JT];
#pragma unroll
            for (int i = 0; i < JT; ++i) acc[i] = make_float4(0, 0, 0, 0);
#pragma unroll 2
            for (int n = 0; n < TP; ++n) {
                float4 a = *reinterpret_cast<const float4 *>(in + n * S_in + 4 * kc);
                const float *grow = gout + n * S_g + j0;
#pragma unroll
                for (int i = 0; i < JT; ++i) {
                    float g = grow[min(i, H - 1 - j0)];
                    acc[i].x = fmaf(g, a.x, acc[i].x); acc[i].y = fmaf(g, a.y, acc[i].y);
                    acc[i].z = fmaf(g, a.z, acc[i].z); acc[i].w = fmaf(g, a.w, acc[i].w);
                }
            }
#pragma unroll
            for (int i = 0; i < JT; ++i)
                if (j0 + i < H) *reinterpret_cast<float4 *>(dW + (size_t)(j0 + i) * K + 4 * kc) = acc[i];
        }
    }
    for (int j = threadIdx.x; j < H; j += ONF_THREADS) {
        float s = 0.f;
        for (int n = 0; n < TP; ++n) s += gout[n * S_g + j];
        db[j] = s;
    }
}

template <int TP, int JT>
__global__ void __launch_bounds__(ONF_THREADS, 1) onf_tile_kernel(const OnfArgs a) {
    extern __shared__ __align__(16) float smem_f[];
    const OnfDims &d = a.d;
    const nfopp_onf_layout &L = d.lay;
    OnfSmem<TP> s(smem_f, d, a.want_bwd != 0);
    const int p = blockIdx.y, tile = blockIdx.x, tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    if (a.sp && !a.sp->do_train) return;                     // engine: not a field-training step
    const int cnt = a.count ? a.count[p] : a.n;
    const int n0 = tile * TP;
    if (n0 >= cnt) return;                                   // whole tile beyond this problem's points
    const float *W = a.weights + (size_t)p * L.n_params;
    const int F = L.f_pos, A = L.n_angle, FIN = L.f_in, H = d.H, D = L.in_dim;

    // ---- phase 0: inputs (onf_model.py:37-38) ----
    if (tid < TP) {
        int n = n0 + tid;
        float x0 = 0.f, x1 = 0.f, x2 = 0.f;
        if (n < cnt) {
            const float *xp = a.x + ((size_t)p * a.n + n) * a.x_stride;
            x0 = xp[0]; x1 = xp[1];
            if (D == 3) x2 = xp[2];
        }
        s.u[2 * tid] = (x0 - d.mean) / d.sigma;
        s.u[2 * tid + 1] = (x1 - d.mean) / d.sigma;
        s.th[tid] = x2;
    }
    __syncthreads();
    // ---- phase 1: Fourier / angle encoding (onf_model.py:39-46, angle_encoder.py:15-18) ----
    {
        constexpr int SUB = 32 / TP;
        const int pt = lane % TP, sub = lane / TP;
        const float u0 = s.u[2 * pt], u1 = s.u[2 * pt + 1], th = s.th[pt];
        for (int f = warp * SUB + sub; f < FIN; f += ONF_WARPS * SUB) {
            float arg;
            bool use_cosine;
            if (f < F) {
                arg = fmaf(u1, __ldg(W + L.off_enc_w + 2 * f + 1),
                           fmaf(u0, __ldg(W + L.off_enc_w + 2 * f), __ldg(W + L.off_enc_b + f)));
                use_cosine = d.use_cos && f >= 100;
            } else {
                int k = f - F;
                arg = (th + __ldg(W + L.off_ang_b + k)) * __ldg(W + L.off_ang_f + k);
                use_cosine = k >= A / 2;
            }
            float sn, cs;
            sincosf(arg, &sn, &cs);
            s.e[pt * s.S_e + f] = use_cosine ? cs : sn;
            if (a.want_bwd) s.de[pt * s.S_e + f] = use_cosine ? -sn : cs;
        }
    }
    __syncthreads();
    // ---- phase 2: hidden layers (onf_model.py:18-23,47) ----
    for (int l = 0; l < d.L; ++l) {
        const float *in = l == 0 ? s.e : s.h[l - 1];
        fwd_layer<TP, JT>(in, l == 0 ? s.S_e : s.S_h, l == 0 ? FIN : H, W + L.off_w[l], W + L.off_b[l], s.h[l], s.S_h, H);
        __syncthreads();
    }
    // ---- phase 3: logit with skip connection of the encoding (onf_model.py:24-26,48-49) ----
    {
        constexpr int SUB = 32 / TP;
        const int pt = lane % TP, sub = lane / TP, grp = warp * SUB + sub;
        const float *hl = s.h[d.L - 1] + pt * s.S_h, *er = s.e + pt * s.S_e;
        const float *wo = W + L.off_w_out;
        float acc = 0.f;
        const int c_h = H / 4, c_all = (H + FIN) / 4;
        for (int c = grp; c < c_all; c += ONF_WARPS * SUB) {
            float4 w = __ldg(reinterpret_cast<const float4 *>(wo + 4 * c));
            float4 v = c < c_h ? *reinterpret_cast<const float4 *>(hl + 4 * c)
                               : *reinterpret_cast<const float4 *>(er + 4 * (c - c_h));
            acc = dot4(v, w, acc);
        }
        s.part[grp * TP + pt] = acc;
    }
    __syncthreads();
    if (tid < TP) {
        constexpr int SUB = 32 / TP;
        float z = __ldg(W + L.off_b_out);
        for (int g = 0; g < ONF_WARPS * SUB; ++g) z += s.part[g * TP + tid];
        s.z[tid] = z;
        const int n = n0 + tid;
        const bool valid = n < cnt;
        if (valid && a.z_out) a.z_out[(size_t)p * a.n + n] = z;
        if (a.want_bwd) {
            // ---- dL/dz ----
            float g = 0.f;
            if (valid) {
                const size_t gi = (size_t)p * a.n + n;
                if (a.gbar_mode == 0) g = a.gz[gi];
                else if (a.gbar_mode == 1) g = (sigmoid_f(z) - a.aux[gi]) / (float)cnt;
                else {
                    float zb = z * a.beta;
                    float ez = expf(zb);
                    g = a.wc * (zb > 20.f ? 1.f : ez / (ez + 1.f));
                    if (a.gbar_mode == 3) {
                        float t = tanhf(z);
                        g += a.aux[gi] * (1.f - t * t);
                    }
                }
            }
            s.gbar[tid] = g;
        }
    }
    if (!a.want_bwd) return;
    __syncthreads();
    float *gp = a.gpart ? a.gpart + ((size_t)p * a.tiles + tile) * L.n_params : nullptr;
    // ---- backward through the output layer ----
    {
        const float *wo = W + L.off_w_out;
        const float *hl = s.h[d.L - 1];
        for (int i = tid; i < TP * H; i += ONF_THREADS) {
            int pt = i / H, j = i - pt * H;
            s.gA[pt * s.S_h + j] = hl[pt * s.S_h + j] > 0.f ? s.gbar[pt] * __ldg(wo + j) : 0.f;
        }
        for (int i = tid; i < TP * FIN; i += ONF_THREADS) {
            int pt = i / FIN, k = i - pt * FIN;
            s.ge[pt * s.S_e + k] = s.gbar[pt] * __ldg(wo + H + k);
        }
        if (gp) {
            for (int k = tid; k < H + FIN; k += ONF_THREADS) {
                const float *col = k < H ? hl + k : s.e + (k - H);
                const int S = k < H ? s.S_h : s.S_e;
                float acc = 0.f;
                for (int n = 0; n < TP; ++n) acc = fmaf(s.gbar[n], col[n * S], acc);
                gp[L.off_w_out + k] = acc;
            }
            if (tid == 0) {
                float acc = 0.f;
                for (int n = 0; n < TP; ++n) acc += s.gbar[n];
                *reinterpret_cast<float4 *>(gp + L.off_b_out) = make_float4(acc, 0.f, 0.f, 0.f);
            }
        }
    }
    __syncthreads();
    // ---- backward through the hidden layers ----
    float *gcur = s.gA, *gnext = s.gB;
    for (int l = d.L - 1; l >= 0; --l) {
        const float *in = l == 0 ? s.e : s.h[l - 1];
        const int S_in = l == 0 ? s.S_e : s.S_h, K = l == 0 ? FIN : H;
        if (gp) wgrad_layer<TP, JT>(gcur, s.S_h, H, in, S_in, K, gp + L.off_w[l], gp + L.off_b[l]);
        if (l > 0) {
            dgrad_layer<TP>(gcur, s.S_h, H, W + L.off_w[l], K, gnext, s.S_h, s.h[l - 1]);
        } else {
            dgrad_layer<TP>(gcur, s.S_h, H, W + L.off_w[0], K, s.ge, s.S_e, nullptr);
        }
        __syncthreads();
        float *t = gcur; gcur = gnext; gnext = t;
    }
    // ---- backward through the encoding: ge <- ge * d(enc)/d(arg) ----
    for (int i = tid; i < TP * FIN; i += ONF_THREADS) {
        int pt = i / FIN, k = i - pt * FIN;
        s.ge[pt * s.S_e + k] *= s.de[pt * s.S_e + k];
    }
    __syncthreads();
    if (gp) {
        for (int f = tid; f < F; f += ONF_THREADS) {
            float g0 = 0.f, g1 = 0.f, gb = 0.f;
            for (int n = 0; n < TP; ++n) {
                float g = s.ge[n * s.S_e + f];
                g0 = fmaf(g, s.u[2 * n], g0);
                g1 = fmaf(g, s.u[2 * n + 1], g1);
                gb += g;
            }
            gp[L.off_enc_w + 2 * f] = g0;
            gp[L.off_enc_w + 2 * f + 1] = g1;
            gp[L.off_enc_b + f] = d.bias ? gb : 0.f;
        }
        for (int k = tid; k < A; k += ONF_THREADS) {
            const float fk = __ldg(W + L.off_ang_f + k), bk = __ldg(W + L.off_ang_b + k);
            float gb = 0.f, gf = 0.f;
            for (int n = 0; n < TP; ++n) {
                float g = s.ge[n * s.S_e + F + k];
                gb = fmaf(g, fk, gb);
                gf = fmaf(g, s.th[n] + bk, gf);       // _frequencies IS trained
            }
            gp[L.off_ang_b + k] = gb;
            gp[L.off_ang_f + k] = gf;
        }
    }
    if (a.gx_out) {
        // one warp per (point, coordinate): g_x = (sum_f g_p[f] * We[f][c]) / sigma ; g_theta = sum_k g_a[k] f_k
        for (int item = warp; item < TP * D; item += ONF_WARPS) {
            int pt = item / D, c = item - pt * D;
            float acc = 0.f;
            if (c < 2) {
                for (int f = lane; f < F; f += 32) acc = fmaf(s.ge[pt * s.S_e + f], __ldg(W + L.off_enc_w + 2 * f + c), acc);
            } else {
                for (int k = lane; k < A; k += 32) acc = fmaf(s.ge[pt * s.S_e + F + k], __ldg(W + L.off_ang_f + k), acc);
            }
            acc = warp_sum(acc);
            int n = n0 + pt;
            if (lane == 0 && n < cnt) a.gx_out[((size_t)p * a.n + n) * a.x_stride + c] = c < 2 ? acc / d.sigma : acc;
        }
        // pose columns the ONF does not read (theta without angle encoding) get a zero gradient
        for (int i = tid; i < TP * (a.x_stride - D); i += ONF_THREADS) {
            int pt = i / (a.x_stride - D), c = D + i % (a.x_stride - D), n = n0 + pt;
            if (n < cnt) a.gx_out[((size_t)p * a.n + n) * a.x_stride + c] = 0.f;
        }
    }
}

// ---- fused gradient reduce (+ Adam) --------------------------------------------------------------
__global__ void __launch_bounds__(256) onf_adam_kernel(const AdamArgs a) {
    const int p = blockIdx.y;
    const int i4 = blockIdx.x * blockDim.x + threadIdx.x;
    if (i4 * 4 >= a.n_params) return;
    float step_size = a.step_size, bc2_sqrt = a.bc2_sqrt;
    if (a.sp) {
        if (!a.warm && !a.sp->do_train) return;
        step_size = a.sp->onf_step_size;
        bc2_sqrt = a.sp->onf_bc2_sqrt;
    }
    const int cnt = a.count ? a.count[p] : a.n;
    if (a.sp && !a.grad_out && cnt <= 0) return;     // engine: a finished (frozen) problem has an empty training set - no update
    const int tiles = min(a.tiles, (cnt + a.tp - 1) / a.tp);
    const float *gp = a.gpart + (size_t)p * a.tiles * a.n_params + 4 * i4;
    float4 g = make_float4(0, 0, 0, 0);
    for (int t = 0; t < tiles; ++t) {
        float4 q = __ldg(reinterpret_cast<const float4 *>(gp + (size_t)t * a.n_params));
        g.x += q.x; g.y += q.y; g.z += q.z; g.w += q.w;
    }
    const size_t o = (size_t)p * a.n_params + 4 * i4;
    if (a.grad_out) {
        *reinterpret_cast<float4 *>(a.grad_out + o) = g;
        return;
    }
    float4 w = *reinterpret_cast<float4 *>(a.w + o), m = *reinterpret_cast<float4 *>(a.m + o),
           v = *reinterpret_cast<float4 *>(a.v + o);
    float gv[4] = {g.x, g.y, g.z, g.w}, wv[4] = {w.x, w.y, w.z, w.w}, mv[4] = {m.x, m.y, m.z, m.w},
          vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        // torch/optim/adam.py _single_tensor_adam: lerp_, mul_/addcmul_, sqrt/div/add_, addcdiv_
        mv[c] = fmaf(a.w1, gv[c] - mv[c], mv[c]);
        vv[c] = fmaf(a.w2 * gv[c], gv[c], vv[c] * a.beta2);
        float denom = sqrtf(vv[c]) / bc2_sqrt + a.eps;
        wv[c] = fmaf(-step_size, mv[c] / denom, wv[c]);
    }
    *reinterpret_cast<float4 *>(a.w + o) = make_float4(wv[0], wv[1], wv[2], wv[3]);
    *reinterpret_cast<float4 *>(a.m + o) = make_float4(mv[0], mv[1], mv[2], mv[3]);
    *reinterpret_cast<float4 *>(a.v + o) = make_float4(vv[0], vv[1], vv[2], vv[3]);
}

// ---- host side -----------------------------------------------------------------------------------
static int pick_tp(const OnfDims &d, bool bwd) {
    if (onf_smem_bytes(d, 32, bwd) <= 227 * 1024) return 32;
    if (onf_smem_bytes(d, 16, bwd) <= 227 * 1024) return 16;
    return 0;
}

int onf_fma_tile_points(const OnfDims &d) { return pick_tp(d, true); }

template <int TP, int JT>
static int launch_tile(const OnfArgs &a, int P, bool bwd, bool prepare_only, cudaStream_t stream) {
    auto kern = onf_tile_kernel<TP, JT>;
    if (prepare_only) {
        NFOPP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)onf_smem_bytes(a.d, TP, true)));
        return NFOPP_OK;
    }
    size_t smem = onf_smem_bytes(a.d, TP, bwd);
    dim3 grid(a.tiles, P);
    kern<<<grid, ONF_THREADS, smem, stream>>>(a);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

template <int TP>
static int dispatch_jt(const OnfArgs &a, int P, bool bwd, bool prep, cudaStream_t stream) {
    constexpr int SUB = 32 / TP;
    int per = (a.d.H + ONF_WARPS * SUB - 1) / (ONF_WARPS * SUB);
    if (per <= 4) return launch_tile<TP, 4>(a, P, bwd, prep, stream);
    if (per <= 7) return launch_tile<TP, 7>(a, P, bwd, prep, stream);
    if (per <= 8) return launch_tile<TP, 8>(a, P, bwd, prep, stream);
    return launch_tile<TP, 16>(a, P, bwd, prep, stream);
}

int onf_fma_prepare(const OnfDims &d) {
    int tp = pick_tp(d, true);
    NFOPP_REQUIRE(tp != 0, "onf(fma): hidden=%d layers=%d does not fit shared memory; use the tensor path", d.H, d.L);
    OnfArgs a{};
    a.d = d;
    return tp == 32 ? dispatch_jt<32>(a, 0, true, true, nullptr) : dispatch_jt<16>(a, 0, true, true, nullptr);
}

// Tile size is a function of the network shape only (fwd and bwd calls must agree on `tiles`).
int onf_fma_launch(OnfArgs a, int P, cudaStream_t stream) {
    int tp = pick_tp(a.d, true);
    NFOPP_REQUIRE(tp != 0, "onf(fma): hidden=%d layers=%d does not fit shared memory; use the tensor path", a.d.H,
                  a.d.L);
    a.tiles = (a.n + tp - 1) / tp;
    if (a.x_stride == 0) a.x_stride = a.d.lay.in_dim;
    if (a.n == 0 || P == 0) return NFOPP_OK;
    return tp == 32 ? dispatch_jt<32>(a, P, a.want_bwd != 0, false, stream)
                    : dispatch_jt<16>(a, P, a.want_bwd != 0, false, stream);
}

int onf_adam_launch(const AdamArgs &a, int P, cudaStream_t stream) {
    int n4 = a.n_params / 4;
    dim3 grid((n4 + 255) / 256, P);
    onf_adam_kernel<<<grid, 256, 0, stream>>>(a);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

}  // namespace nfopp
