// onf_tc.cu - K2, tensor-core path: the fused ONF kernel on tcgen05 (sm_100a).
//
// One CTA = one planning problem (its own weights); the problem's points are walked in tiles of <= 64.  Every
// dense contraction of the forward pass (onf_model.py:33-50), of the input-gradient pass and of the
// weight-gradient pass (what autograd does for nerf_opt_planner.py:85-90 / constrained...:85-100) is a
// tcgen05.mma with the HIDDEN index on the UMMA M axis (128 TMEM lanes) and the POINT index on the N axis:
//
//   fwd1   D1[j,n]  = sum_k W1p[j,k] E[n,k]      A = W1p  (K-major)   B = E    (K-major)    K = K1
//   fwd2   D2[i,n]  = sum_j W2p[i,j] H1[j,n]     A = W2p  (K-major)   B = H1   (MN-major)   K = KH
//   wgrad2 dW2[i,j]+= sum_n gH2[i,n] H1[j,n]     A = gH2  (K-major)   B = H1   (K-major)    K = points
//   dgrad2 gD1[j,n] = sum_i W2p[i,j] gH2[i,n]    A = W2p  (MN-major)  B = gH2  (MN-major)   K = KH
//   dgrad1 gE[k,n]  = sum_j W1p[j,k] gH1[j,n]    A = W1p  (MN-major)  B = gH1  (MN-major)   K = KH  (2 M blocks)
//   wgrad1 dW1[j,k]+= sum_n gH1[j,n] E[n,k]      A = gH1  (K-major)   B = E^T  (K-major)    K = points
//
// Precision: every fp32 operand is split into bf16 hi + bf16 lo and each product is three kind::f16 MMAs
// (hi*hi + lo*hi + hi*lo) accumulated in fp32 in TMEM: ~3e-6 relative (umma_probe.cu), far inside the 1e-3
// bar; single-pass bf16 (4e-3) or tf32 (5e-4 per operand) would not be.
//
// Row H of the layer-1 operand W1p holds the skip-connection weights w_out[H:], so the skip term of the logit
// (onf_model.py:48-49) falls out of fwd1 as D1[H,n], its input gradient out of dgrad1 (gH1 row H := dL/dz) and
// its weight gradient out of wgrad1 (row H of dW1) - no extra pass.  Biases, ReLU, the logit reduction, dL/dz,
// the Fourier/angle encoding and its backward run on the CUDA cores between the MMAs, one TMEM lane (hidden
// unit or feature) per thread, so every bias / per-unit reduction is thread-local.
//
// The weight-gradient accumulators (dW1: K1 columns, dW2: KH columns, fp32) stay in TMEM across all point tiles
// of the problem; the epilogue streams them out once - either as a plain gradient or straight through Adam
// (torch.optim.Adam, planner_factory.py:68), which also rewrites the bf16 hi/lo operand copy of the weights.
// Nothing like the FMA path's per-tile gradient partials ever reaches HBM.
#include "onf.cuh"
#include "tc05.cuh"

namespace nfopp {
using namespace tc05;

constexpr int TC_THREADS = 512;
constexpr int TC_TP = 64;                        // points per tile = max UMMA N used for point-indexed outputs
constexpr int TC_REGION = 32768;                 // E ring / gH2 / E^T scratch (two 16 KB halves: hi, lo)
constexpr uint32_t TM_DW1 = 0, TM_DW2 = 224, TM_WORK = 336;   // TMEM column map (512 columns allocated)

struct TcDims {
    int F, A, FIN, H, K1, KH, MB;                // MB = M blocks of dgrad1 = feature groups of 128
    int S_colW;                                  // column-octet stride of the packed weights = KH * 16 bytes
    int w1_half, w2_half, h1_half;               // bytes of one hi (or lo) buffer
    int off_w1, off_w2, off_h1, off_r, off_misc; // shared-memory byte offsets
    int smem_bytes;
    int packed_bytes;                            // per problem, global
};

__host__ __device__ inline int ru16(int x) { return (x + 15) / 16 * 16; }

__host__ __device__ inline bool tc_make_dims(const OnfDims &d, TcDims *out) {
    TcDims t{};
    t.F = d.lay.f_pos; t.A = d.lay.n_angle; t.FIN = d.lay.f_in; t.H = d.H;
    t.K1 = ru16(t.FIN); t.KH = ru16(t.H + 1); t.MB = (t.K1 + 127) / 128;
    t.S_colW = t.KH * 16;
    t.w1_half = (t.K1 / 8) * t.S_colW;
    t.w2_half = (t.KH / 8) * t.S_colW;
    t.h1_half = (t.KH / 8) * 1024;
    t.off_w1 = 0;
    t.off_w2 = 2 * t.w1_half;
    t.off_h1 = t.off_w2 + 2 * t.w2_half;
    t.off_r = t.off_h1 + 2 * t.h1_half;
    t.off_misc = t.off_r + TC_REGION;
    int misc_floats = 3 * t.K1 + (t.A + 3) / 4 * 4 + 3 * t.H + 4 /*b3 + pad*/ + 5 * TC_TP + 4 * TC_TP + 4 * TC_TP * 3;
    t.smem_bytes = t.off_misc + misc_floats * 4 + 64;
    t.packed_bytes = 2 * t.w1_half + 2 * t.w2_half;
    *out = t;
    // the epilogue stages the whole gradient blob (+ 20 KB of per-unit partial sums) over the dead operand buffers
    const int stage_bytes = ((d.lay.n_params + 31) / 32 * 32 + 10 * 512) * 4;
    return d.L == 2 && t.KH <= 112 && t.K1 <= 224 && t.H >= 8 && t.smem_bytes <= 227 * 1024 && stage_bytes <= t.off_misc;
}

struct OnfTcArgs {
    OnfDims d;
    TcDims t;
    const float *weights;          // [P, n_params] fp32 blob (encoder, biases, w_out[:H], b_out are read from here)
    const unsigned char *packed;   // [P, packed_bytes] bf16 hi/lo operand copy of W1p / W2p
    const float *x;
    const int32_t *count;
    const StepParams *sp;
    int n, x_stride;
    int want_bwd;                  // 0: logits only
    int gbar_mode;
    const float *gz, *aux;
    float wc, beta;
    float *z_out, *gx_out;
    float *grad_out;               // [P, n_params] plain gradient (API), or null
    // fused Adam (engine): updates w/m/v and the packed copy in place
    float *adam_w, *adam_m, *adam_v;
    unsigned char *packed_out;
    float w1, beta2, w2, eps;
    int warm;
};

struct TcPackArgs {
    OnfDims d;
    TcDims t;
    const float *weights;
    unsigned char *packed;
};

// ---- fp32 blob -> packed bf16 hi/lo operands (API entry points and engine start; the training kernel keeps
// the copy current afterwards) -----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) onf_tc_pack_kernel(const TcPackArgs a) {
    const int p = blockIdx.y;
    const TcDims &t = a.t;
    const nfopp_onf_layout &L = a.d.lay;
    const float *W = a.weights + (size_t)p * L.n_params;
    unsigned char *out = a.packed + (size_t)p * t.packed_bytes;
    const int rows = t.KH;
    const int n1 = (t.K1 / 8) * rows, n2 = (t.KH / 8) * rows;
    for (int item = blockIdx.x * blockDim.x + threadIdx.x; item < n1 + n2; item += gridDim.x * blockDim.x) {
        const bool second = item >= n1;
        const int it = second ? item - n1 : item;
        const int oct = it / rows, r = it - oct * rows;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int c = oct * 8 + e;
            float val = 0.f;
            if (!second) {
                if (c < t.FIN) {
                    if (r < t.H) val = W[L.off_w[0] + r * t.FIN + c];
                    else if (r == t.H) val = W[L.off_w_out + t.H + c];
                }
            } else if (r < t.H && c < t.H) {
                val = W[L.off_w[1] + r * t.H + c];
            }
            v[e] = val;
        }
        uint4 hi, lo;
        split8(v, hi, lo);
        unsigned char *base = out + (second ? 2 * t.w1_half : 0);
        const int half = second ? t.w2_half : t.w1_half;
        const int off = oct * t.S_colW + (r / 8) * 128 + (r % 8) * 16;
        *reinterpret_cast<uint4 *>(base + off) = hi;
        *reinterpret_cast<uint4 *>(base + half + off) = lo;
    }
}

// ---- device helpers ------------------------------------------------------------------------------------
// column sums over the 32 lanes of a warp for 16 columns: every lane returns sum_lanes v[lane >> 1]  (16 shuffles)
__device__ __forceinline__ float colsum16(float (&v)[16], int lane) {
#pragma unroll
    for (int o = 16; o >= 2; o >>= 1) {
        const bool up = (lane & o) != 0;
        const int h = o >> 1;
#pragma unroll
        for (int c = 0; c < h; ++c) {
            const float keep = up ? v[c + h] : v[c];
            const float send = up ? v[c] : v[c + h];
            v[c] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
}

struct TcSmem {
    unsigned char *w1h, *w1l, *w2h, *w2l, *h1h, *h1l, *rh, *rl;
    // per input feature f < K1: arg = x1 * pb + u0 * pa + pc  (x1 = u1 for Fourier features, theta for heading features;
    // pc carries the bias and +pi/2 for cosine features, so every feature is sin(arg) with derivative cos(arg))
    float *pa, *pb, *pc, *ang_b, *b1, *b2, *w3h, *b3;
    float *u0, *u1, *th, *skip, *gbar, *zp, *gx;
    uint64_t *bars;
    uint32_t *tmem_slot;
    __device__ TcSmem(unsigned char *base, const TcDims &t) {
        w1h = base + t.off_w1; w1l = w1h + t.w1_half;
        w2h = base + t.off_w2; w2l = w2h + t.w2_half;
        h1h = base + t.off_h1; h1l = h1h + t.h1_half;
        rh = base + t.off_r; rl = rh + TC_REGION / 2;
        float *f = reinterpret_cast<float *>(base + t.off_misc);
        pa = f; f += t.K1;
        pb = f; f += t.K1;
        pc = f; f += t.K1;
        ang_b = f; f += (t.A + 3) / 4 * 4;
        b1 = f; f += t.H;
        b2 = f; f += t.H;
        w3h = f; f += t.H;
        b3 = f; f += 4;
        u0 = f; f += TC_TP;
        u1 = f; f += TC_TP;
        th = f; f += TC_TP;
        skip = f; f += TC_TP;
        gbar = f; f += TC_TP;
        zp = f; f += 4 * TC_TP;
        gx = f; f += 4 * TC_TP * 3;
        bars = reinterpret_cast<uint64_t *>(f);
        tmem_slot = reinterpret_cast<uint32_t *>(bars + 4);
    }
};

// sin and cos of the feature argument (onf_model.py:39-46, angle_encoder.py:15-18 with the phase folded into pc)
__device__ __forceinline__ float feature_arg(float pa, float pb, float pc, float u0, float x1) {
    return fmaf(x1, pb, fmaf(u0, pa, pc));
}
__device__ __forceinline__ float reduce_2pi(float x) {
    // Cody-Waite reduction to [-pi, pi]; the MUFU sin/cos that follow have abs error ~5e-7 there
    const float k = rintf(x * 0.15915494309189535f);
    float r = fmaf(k, -6.2831854820251465f, x);
    return fmaf(k, 1.7484555e-7f, r);
}

// store 16 consecutive columns (col0 .. col0+15) of row `row` into a [row/8][col/8][row%8][col%8] operand buffer
__device__ __forceinline__ void store_row16(unsigned char *hi_buf, unsigned char *lo_buf, int row, int col0,
                                            const float *v, int S_row) {
    uint4 h0, l0, h1, l1;
    split8(v, h0, l0);
    split8(v + 8, h1, l1);
    const int off = (row >> 3) * S_row + (col0 >> 3) * 128 + (row & 7) * 16;
    *reinterpret_cast<uint4 *>(hi_buf + off) = h0;
    *reinterpret_cast<uint4 *>(hi_buf + off + 128) = h1;
    *reinterpret_cast<uint4 *>(lo_buf + off) = l0;
    *reinterpret_cast<uint4 *>(lo_buf + off + 128) = l1;
}

// three-pass split product: D (+)= Ahi*Bhi + Alo*Bhi + Ahi*Blo
__device__ __forceinline__ void mma3(uint32_t tmem_d, uint64_t ah, uint64_t al, uint64_t bh, uint64_t bl, uint32_t idesc,
                                     bool accumulate, uint32_t issue) {
    mma_bf16(tmem_d, ah, bh, idesc, accumulate ? 1u : 0u, issue);
    mma_bf16(tmem_d, al, bh, idesc, 1u, issue);
    mma_bf16(tmem_d, ah, bl, idesc, 1u, issue);
}

// Thread layout: 16 warps.  warp = 4*cq + q: q = TMEM lane quarter (rows 32q..32q+31), cq = column quarter of the
// point tile (columns 16cq..16cq+15).  A thread therefore owns one hidden unit / feature (`row`) and 16 points.
__global__ void __launch_bounds__(TC_THREADS, 1) onf_tc_kernel(const OnfTcArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const OnfDims &d = a.d;
    const TcDims &t = a.t;
    const nfopp_onf_layout &L = d.lay;
    const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q = warp & 3, cq = warp >> 2;
    const int row = q * 32 + lane;
    if (a.sp && !a.warm && !a.sp->do_train) return;    // engine: not a field-training step
    const int cnt = a.count ? a.count[p] : a.n;
    if (cnt <= 0) return;
    const bool want_bwd = a.want_bwd != 0;
    const bool want_wgrad = want_bwd && (a.grad_out != nullptr || a.adam_w != nullptr);
    const bool want_gx = want_bwd && a.gx_out != nullptr;
    const int F = t.F, FIN = t.FIN, H = t.H, K1 = t.K1, KH = t.KH;

    TcSmem s(smem_raw, t);
    const float *W = a.weights + (size_t)p * L.n_params;
    uint64_t *bar_w = s.bars, *bar_mma = s.bars + 1, *bar_aux = s.bars + 2;

    // ---- one-time setup: barriers, TMEM, packed weights by bulk copy (TMA), small fp32 parameters ----
    if (tid == 0) {
        mbar_init(bar_w, 1);
        mbar_init(bar_mma, 1);
        mbar_init(bar_aux, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(s.tmem_slot, 512);
    __syncthreads();
    if (tid == 0) {
        const unsigned char *src = a.packed + (size_t)p * t.packed_bytes;
        mbar_expect_tx(bar_w, (uint32_t)t.packed_bytes);
        bulk_g2s(s.w1h, src, 2 * t.w1_half, bar_w);
        bulk_g2s(s.w2h, src + 2 * t.w1_half, 2 * t.w2_half, bar_w);
    }
    for (int f = tid; f < K1; f += TC_THREADS) {
        float pa = 0.f, pb = 0.f, pc = 0.f;
        if (f < F) {
            pa = W[L.off_enc_w + 2 * f];
            pb = W[L.off_enc_w + 2 * f + 1];
            pc = W[L.off_enc_b + f] + ((d.use_cos && f >= 100) ? 1.57079632679489662f : 0.f);
        } else if (f < FIN) {
            const int k = f - F;
            const float b = W[L.off_ang_b + k], fr = W[L.off_ang_f + k];
            s.ang_b[k] = b;
            pb = fr;
            pc = fmaf(b, fr, k >= t.A / 2 ? 1.57079632679489662f : 0.f);
        }
        s.pa[f] = pa; s.pb[f] = pb; s.pc[f] = pc;
    }
    for (int i = tid; i < H; i += TC_THREADS) {
        s.b1[i] = W[L.off_b[0] + i];
        s.b2[i] = W[L.off_b[1] + i];
        s.w3h[i] = W[L.off_w_out + i];
    }
    if (tid == 0) s.b3[0] = W[L.off_b_out];
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *s.tmem_slot;
    const uint32_t tm_lane = tmem + ((uint32_t)(q * 32) << 16);
    if (warp == 0) mbar_wait(bar_w, 0);    // the MMA-issuing warp observes the bulk copies

    uint32_t ph_mma = 0, ph_aux = 0;
    bool pending_aux = false;
    // thread-local gradient accumulators (hidden unit `row`; features row and 128+row; point tid)
    float acc_db1 = 0.f, acc_db2 = 0.f, acc_dw3 = 0.f, acc_db3 = 0.f;
    float fa0 = 0.f, fb0 = 0.f, fc0 = 0.f, fa1 = 0.f, fb1 = 0.f, fc1 = 0.f;
    const float my_b1 = row < H ? s.b1[row] : 0.f, my_b2 = row < H ? s.b2[row] : 0.f, my_w3 = row < H ? s.w3h[row] : 0.f;

    const int tiles = (cnt + TC_TP - 1) / TC_TP;
    for (int tile = 0; tile < tiles; ++tile) {
        const int n0 = tile * TC_TP;
        const int nv = min(TC_TP, cnt - n0);
        const int Nt = ru16(nv);
        const int c0 = 16 * cq;                        // this warp's 16 columns of the tile
        const bool cols_on = c0 < Nt;
        // ---- P0: inputs (onf_model.py:37-38) ----
        if (tid < TC_TP) {
            float x0 = 0.f, x1 = 0.f, x2 = 0.f;
            if (tid < nv) {
                const float *xp = a.x + ((size_t)p * a.n + n0 + tid) * a.x_stride;
                x0 = xp[0]; x1 = xp[1];
                if (L.in_dim == 3) x2 = xp[2];
            }
            s.u0[tid] = (x0 - d.mean) / d.sigma;
            s.u1[tid] = (x1 - d.mean) / d.sigma;
            s.th[tid] = x2;
        }
        __syncthreads();
        // ---- P1: encoding (CUDA cores) feeding fwd1 in groups of 128 features ----
        {
            const int n = tid & 63, oc = tid >> 6;       // point, feature octet (0..7) within a 64-feature half group
            const float u0 = s.u0[n], u1 = s.u1[n], th = s.th[n];
#pragma unroll 1
            for (int g = 0; g < t.MB; ++g) {
                uint4 ehi[2], elo[2];
#pragma unroll
                for (int it = 0; it < 2; ++it) {
                    const int f0 = 128 * g + 8 * (oc + 8 * it);
                    float e[8];
                    if (f0 < K1 && n < Nt) {
                        float pa[8], pb[8], pc[8];
                        *reinterpret_cast<float4 *>(pa) = *reinterpret_cast<const float4 *>(s.pa + f0);
                        *reinterpret_cast<float4 *>(pa + 4) = *reinterpret_cast<const float4 *>(s.pa + f0 + 4);
                        *reinterpret_cast<float4 *>(pb) = *reinterpret_cast<const float4 *>(s.pb + f0);
                        *reinterpret_cast<float4 *>(pb + 4) = *reinterpret_cast<const float4 *>(s.pb + f0 + 4);
                        *reinterpret_cast<float4 *>(pc) = *reinterpret_cast<const float4 *>(s.pc + f0);
                        *reinterpret_cast<float4 *>(pc + 4) = *reinterpret_cast<const float4 *>(s.pc + f0 + 4);
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            e[i] = __sinf(reduce_2pi(feature_arg(pa[i], pb[i], pc[i], u0, f0 + i < F ? u1 : th)));
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; ++i) e[i] = 0.f;
                    }
                    split8(e, ehi[it], elo[it]);
                }
                if (g > 0 || pending_aux) {          // the scratch region is still being read by earlier MMAs
                    mbar_wait(bar_aux, ph_aux);
                    ph_aux ^= 1;
                    pending_aux = false;
                }
#pragma unroll
                for (int it = 0; it < 2; ++it) {
                    const int fo = oc + 8 * it;
                    if (128 * g + 8 * fo < K1 && n < Nt) {
                        *reinterpret_cast<uint4 *>(s.rh + fo * 1024 + n * 16) = ehi[it];
                        *reinterpret_cast<uint4 *>(s.rl + fo * 1024 + n * 16) = elo[it];
                    }
                }
                fence_proxy_async();
                fence_before_sync();
                __syncthreads();
                if (warp == 0) {                 // warp-uniform: descriptor arithmetic stays on the uniform datapath
                    fence_after_sync();
                    const uint32_t lead = elect_one();
                    const uint32_t idesc = instr_desc_bf16(128, Nt, 0, 0);
                    const int ks_n = min(8, (K1 - 128 * g) / 16);
                    for (int ks = 0; ks < ks_n; ++ks) {
                        const uint32_t aoff = (uint32_t)(16 * g + 2 * ks) * t.S_colW, boff = (uint32_t)(2 * ks) * 1024;
                        mma3(tmem + TM_WORK, desc_kmajor(smem_u32(s.w1h) + aoff, 128, t.S_colW),
                             desc_kmajor(smem_u32(s.w1l) + aoff, 128, t.S_colW), desc_kmajor(smem_u32(s.rh) + boff, 128, 1024),
                             desc_kmajor(smem_u32(s.rl) + boff, 128, 1024), idesc, g > 0 || ks > 0, lead);
                    }
                    if (g + 1 < t.MB) mma_commit(bar_aux, lead);
                    else mma_commit(bar_mma, lead);
                }
            }
        }
        // ---- P2: epilogue 1: bias + ReLU -> H1 operand; row H = skip term of the logit ----
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        fence_after_sync();
        if (cols_on) {
            float v[16];
            tmem_ld16(tm_lane + TM_WORK + c0, v);
            if (row < H) {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i] + my_b1, 0.f);
            } else {
                if (row == H) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) s.skip[c0 + i] = v[i];
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = 0.f;
            }
            if (row < KH) store_row16(s.h1h, s.h1l, row, c0, v, 1024);
        }
        fence_proxy_async();
        fence_before_sync();
        __syncthreads();
        // ---- P3: fwd2 ----
        if (warp == 0) {                 // warp-uniform: descriptor arithmetic stays on the uniform datapath
            fence_after_sync();
            const uint32_t lead = elect_one();
            const uint32_t idesc = instr_desc_bf16(128, Nt, 0, 1);
            for (int ks = 0; ks < KH / 16; ++ks) {
                const uint32_t aoff = (uint32_t)(2 * ks) * t.S_colW, boff = (uint32_t)(2 * ks) * 1024;
                mma3(tmem + TM_WORK + 64, desc_kmajor(smem_u32(s.w2h) + aoff, 128, t.S_colW),
                     desc_kmajor(smem_u32(s.w2l) + aoff, 128, t.S_colW), desc_mnmajor(smem_u32(s.h1h) + boff, 1024, 128),
                     desc_mnmajor(smem_u32(s.h1l) + boff, 1024, 128), idesc, ks > 0, lead);
            }
            mma_commit(bar_mma, lead);
        }
        // ---- P4: epilogue 2: logit, dL/dz, gH2 operand ----
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        fence_after_sync();
        float r2[16];
        if (cols_on) {
            tmem_ld16(tm_lane + TM_WORK + 64 + c0, r2);
#pragma unroll
            for (int i = 0; i < 16; ++i) r2[i] = row < H ? fmaxf(r2[i] + my_b2, 0.f) : 0.f;
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) r2[i] = 0.f;
        }
        {
            float pz[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) pz[i] = my_w3 * r2[i];
            const float zs = colsum16(pz, lane);
            if ((lane & 1) == 0) s.zp[q * TC_TP + c0 + (lane >> 1)] = zs;
        }
        __syncthreads();
        if (tid < TC_TP) {
            const int n = tid;
            const float z = s.zp[n] + s.zp[TC_TP + n] + s.zp[2 * TC_TP + n] + s.zp[3 * TC_TP + n] + s.skip[n] + s.b3[0];
            const bool valid = n < nv;
            const size_t gi = (size_t)p * a.n + n0 + n;
            if (valid && a.z_out) a.z_out[gi] = z;
            float g = 0.f;
            if (valid && want_bwd) {                        // dL/dz
                if (a.gbar_mode == 0) g = a.gz[gi];
                else if (a.gbar_mode == 1) g = (sigmoid_f(z) - a.aux[gi]) / (float)cnt;
                else {
                    const float zb = z * a.beta;
                    const float ez = expf(zb);
                    g = a.wc * (zb > 20.f ? 1.f : ez / (ez + 1.f));
                    if (a.gbar_mode == 3) {
                        const float th = tanhf(z);
                        g += a.aux[gi] * (1.f - th * th);
                    }
                }
            }
            s.gbar[n] = g;
            acc_db3 += g;
        }
        __syncthreads();
        if (!want_bwd) continue;
        if (cols_on) {
            float v[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float gb = s.gbar[c0 + i];
                acc_dw3 = fmaf(gb, r2[i], acc_dw3);
                const float g = r2[i] > 0.f ? gb * my_w3 : 0.f;
                acc_db2 += g;
                v[i] = g;
            }
            if (row < KH) store_row16(s.rh, s.rl, row, c0, v, 1024);
        }
        fence_proxy_async();
        fence_before_sync();
        __syncthreads();
        // ---- P5: wgrad2 (persistent accumulator) + dgrad2 ----
        if (warp == 0) {                 // warp-uniform: descriptor arithmetic stays on the uniform datapath
            fence_after_sync();
            const uint32_t lead = elect_one();
            if (want_wgrad) {
                const uint32_t idesc = instr_desc_bf16(128, KH, 0, 0);
                for (int ks = 0; ks < Nt / 16; ++ks) {
                    const uint32_t off = (uint32_t)(2 * ks) * 128;
                    mma3(tmem + TM_DW2, desc_kmajor(smem_u32(s.rh) + off, 1024, 128), desc_kmajor(smem_u32(s.rl) + off, 1024, 128),
                         desc_kmajor(smem_u32(s.h1h) + off, 1024, 128), desc_kmajor(smem_u32(s.h1l) + off, 1024, 128), idesc,
                         tile > 0 || ks > 0, lead);
                }
            }
            const uint32_t idesc = instr_desc_bf16(128, Nt, 1, 1);
            for (int ks = 0; ks < KH / 16; ++ks) {
                const uint32_t aoff = (uint32_t)(2 * ks) * 128, boff = (uint32_t)(2 * ks) * 1024;
                mma3(tmem + TM_WORK, desc_mnmajor(smem_u32(s.w2h) + aoff, 128, t.S_colW),
                     desc_mnmajor(smem_u32(s.w2l) + aoff, 128, t.S_colW), desc_mnmajor(smem_u32(s.rh) + boff, 1024, 128),
                     desc_mnmajor(smem_u32(s.rl) + boff, 1024, 128), idesc, ks > 0, lead);
            }
            mma_commit(bar_mma, lead);
        }
        // ---- P6: epilogue 3: ReLU mask of layer 1 -> gH1 operand (in place over H1); row H := dL/dz ----
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        fence_after_sync();
        if (cols_on) {
            float v[16];
            tmem_ld16(tm_lane + TM_WORK + c0, v);
            if (row < H) {
                const int off = (row >> 3) * 1024 + (c0 >> 3) * 128 + (row & 7) * 16;
                const uint4 m0 = *reinterpret_cast<const uint4 *>(s.h1h + off);
                const uint4 m1 = *reinterpret_cast<const uint4 *>(s.h1h + off + 128);
                const uint32_t mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const uint32_t bits = (mw[i >> 1] >> ((i & 1) * 16)) & 0x7FFFu;
                    v[i] = bits != 0 ? v[i] : 0.f;
                    acc_db1 += v[i];
                }
            } else if (row == H) {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = s.gbar[c0 + i];
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = 0.f;
            }
            if (row < KH) store_row16(s.h1h, s.h1l, row, c0, v, 1024);
        }
        fence_proxy_async();
        fence_before_sync();
        __syncthreads();
        // ---- P7: dgrad1: gE[k,n] for both feature blocks ----
        if (warp == 0) {                 // warp-uniform: descriptor arithmetic stays on the uniform datapath
            fence_after_sync();
            const uint32_t lead = elect_one();
            const uint32_t idesc = instr_desc_bf16(128, Nt, 1, 1);
            for (int mb = 0; mb < t.MB; ++mb)
                for (int ks = 0; ks < KH / 16; ++ks) {
                    const uint32_t aoff = (uint32_t)(16 * mb) * t.S_colW + (uint32_t)(2 * ks) * 128, boff = (uint32_t)(2 * ks) * 1024;
                    mma3(tmem + TM_WORK + 64 * mb, desc_mnmajor(smem_u32(s.w1h) + aoff, 128, t.S_colW),
                         desc_mnmajor(smem_u32(s.w1l) + aoff, 128, t.S_colW), desc_mnmajor(smem_u32(s.h1h) + boff, 1024, 128),
                         desc_mnmajor(smem_u32(s.h1l) + boff, 1024, 128), idesc, ks > 0, lead);
                }
            mma_commit(bar_mma, lead);
        }
        // ---- P8/P9: epilogue 4: encoding backward (+ input gradient), E^T operand, wgrad1 per feature block ----
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        fence_after_sync();
#pragma unroll 1
        for (int mb = 0; mb < t.MB; ++mb) {
            const int f = 128 * mb + row;
            const int pf = min(128, K1 - 128 * mb);              // features of this block that are MMA rows
            const bool f_on = row < pf;                          // rows beyond hold garbage
            const bool fourier = f < F;
            const float ppa = f_on ? s.pa[f] : 0.f, ppb = f_on ? s.pb[f] : 0.f, ppc = f_on ? s.pc[f] : 0.f;
            const float ew0 = fourier ? ppa : 0.f, ew1 = fourier ? ppb : 0.f, eaf = (!fourier && f < FIN) ? ppb : 0.f;
            float ta = 0.f, tb = 0.f, tc = 0.f;                  // sum gp*u0, sum gp*x1, sum gp
            float gq[16];
            uint4 ehi[2], elo[2];
            {
                float ev[16];
                if (cols_on) {
                    float v[16], xu[16], x1[16];
                    tmem_ld16(tm_lane + TM_WORK + 64 * mb + c0, v);
                    const float *x1src = fourier ? s.u1 : s.th;
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4) {
                        *reinterpret_cast<float4 *>(xu + 4 * i4) = *reinterpret_cast<const float4 *>(s.u0 + c0 + 4 * i4);
                        *reinterpret_cast<float4 *>(x1 + 4 * i4) = *reinterpret_cast<const float4 *>(x1src + c0 + 4 * i4);
                    }
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float r = reduce_2pi(feature_arg(ppa, ppb, ppc, xu[i], x1[i]));
                        const float gp = f_on ? v[i] * __cosf(r) : 0.f;
                        ev[i] = __sinf(r);
                        gq[i] = gp;
                        ta = fmaf(gp, xu[i], ta);
                        tb = fmaf(gp, x1[i], tb);
                        tc += gp;
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) { ev[i] = 0.f; gq[i] = 0.f; }
                }
                split8(ev, ehi[0], elo[0]);
                split8(ev + 8, ehi[1], elo[1]);
            }
            if (!fourier) {                                      // heading features: d ang_b = f_k * sum gp,
                const float b = f < FIN ? s.ang_b[f - F] : 0.f;  // d ang_f = sum gp * (theta + b_k)  (trained)
                tb = fmaf(b, tc, tb);
                ta = eaf * tc;
            }
            if (mb == 0) { fa0 += ta; fb0 += tb; fc0 += tc; }
            else { fa1 += ta; fb1 += tb; fc1 += tc; }
            if (want_gx) {
                // gx[n] = (sum_f gp[f,n] * We[f,c]) / sigma ; g_theta[n] = sum_k gp[k,n] * f_k   (reduce over lanes)
                float cv[16];
                float *dst = s.gx + ((size_t)q * TC_TP + c0 + (lane >> 1)) * 3;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float wgt = c == 0 ? ew0 : c == 1 ? ew1 : eaf;
#pragma unroll
                    for (int i = 0; i < 16; ++i) cv[i] = gq[i] * wgt;
                    const float sum = colsum16(cv, lane);
                    if ((lane & 1) == 0) dst[c] = mb == 0 ? sum : dst[c] + sum;
                }
            }
            if (want_wgrad) {
                if (mb > 0) {                       // wgrad1 of the previous block still reads the scratch region
                    mbar_wait(bar_aux, ph_aux);
                    ph_aux ^= 1;
                }
                if (row < pf && cols_on) {
#pragma unroll
                    for (int o = 0; o < 2; ++o) {
                        const int off = ((c0 >> 3) + o) * 2048 + (row >> 3) * 128 + (row & 7) * 16;
                        *reinterpret_cast<uint4 *>(s.rh + off) = ehi[o];
                        *reinterpret_cast<uint4 *>(s.rl + off) = elo[o];
                    }
                }
                fence_proxy_async();
                fence_before_sync();
                __syncthreads();
                if (warp == 0) {                 // warp-uniform: descriptor arithmetic stays on the uniform datapath
                    fence_after_sync();
                    const uint32_t lead = elect_one();
                    const uint32_t idesc = instr_desc_bf16(128, pf, 0, 0);
                    for (int ks = 0; ks < Nt / 16; ++ks) {
                        const uint32_t aoff = (uint32_t)(2 * ks) * 128, boff = (uint32_t)(2 * ks) * 2048;
                        mma3(tmem + TM_DW1 + 128 * mb, desc_kmajor(smem_u32(s.h1h) + aoff, 1024, 128),
                             desc_kmajor(smem_u32(s.h1l) + aoff, 1024, 128), desc_kmajor(smem_u32(s.rh) + boff, 128, 2048),
                             desc_kmajor(smem_u32(s.rl) + boff, 128, 2048), idesc, tile > 0 || ks > 0, lead);
                    }
                    mma_commit(bar_aux, lead);
                }
                if (mb + 1 >= t.MB) pending_aux = true;
            }
        }
        if (want_gx) {
            __syncthreads();
            if (tid < nv) {
                const int n = tid;
                float g0 = 0.f, g1 = 0.f, g2 = 0.f;
#pragma unroll
                for (int qq = 0; qq < 4; ++qq) {
                    const float *src = s.gx + ((size_t)qq * TC_TP + n) * 3;
                    g0 += src[0]; g1 += src[1]; g2 += src[2];
                }
                float *out = a.gx_out + ((size_t)p * a.n + n0 + n) * a.x_stride;
                out[0] = g0 / d.sigma;
                out[1] = g1 / d.sigma;
                if (L.in_dim == 3) out[2] = g2;
                for (int c = L.in_dim; c < a.x_stride; ++c) out[c] = 0.f;   // pose columns the ONF does not read
            }
        }
        fence_before_sync();
        __syncthreads();           // TMEM work columns and s.* scratch are reused by the next tile
    }

    // ---- weight gradients: TMEM accumulators -> shared memory in blob order -> coalesced Adam / gradient store ----
    if (want_wgrad) {
        if (pending_aux) {
            mbar_wait(bar_aux, ph_aux);
            ph_aux ^= 1;
        }
        fence_after_sync();
        // every MMA has completed: the operand buffers are dead, shared memory [0, n_params) floats now holds the
        // gradient of the whole parameter blob (same element order as the fp32 weights)
        float *G = reinterpret_cast<float *>(smem_raw);
        float *red = G + ((L.n_params + 31) & ~31);                    // [10][4][128] partial per-unit sums
        const int np4 = L.n_params / 4;
        for (int i = tid; i < np4; i += TC_THREADS) reinterpret_cast<float4 *>(G)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        {
            float *r0 = red + cq * 128 + row;
            r0[0 * 512] = acc_db1; r0[1 * 512] = acc_db2; r0[2 * 512] = acc_dw3;
            r0[3 * 512] = fa0; r0[4 * 512] = fb0; r0[5 * 512] = fc0;
            r0[6 * 512] = fa1; r0[7 * 512] = fb1; r0[8 * 512] = fc1;
            if (tid < TC_TP) red[9 * 512 + tid] = acc_db3;
        }
        __syncthreads();
        // dW1: rows j < H -> mlp.0.weight[j,:], row H -> mlp2.0.weight[0, H:];  dW2: rows i < H -> mlp.2.weight[i,:]
        for (int ch = cq; ch < K1 / 16; ch += 4) {
            float g[16];
            tmem_ld16(tm_lane + TM_DW1 + 16 * ch, g);
            if (row <= H) {
                float *dst = G + (row < H ? L.off_w[0] + row * FIN : L.off_w_out + H) + 16 * ch;
#pragma unroll
                for (int c4 = 0; c4 < 4; ++c4)
                    if (16 * ch + 4 * c4 < FIN)
                        *reinterpret_cast<float4 *>(dst + 4 * c4) = make_float4(g[4 * c4], g[4 * c4 + 1], g[4 * c4 + 2], g[4 * c4 + 3]);
            }
        }
        for (int ch = cq; ch < KH / 16; ch += 4) {
            float g[16];
            tmem_ld16(tm_lane + TM_DW2 + 16 * ch, g);
            if (row < H) {
                float *dst = G + L.off_w[1] + row * H + 16 * ch;
#pragma unroll
                for (int c4 = 0; c4 < 4; ++c4)
                    if (16 * ch + 4 * c4 < H)
                        *reinterpret_cast<float4 *>(dst + 4 * c4) = make_float4(g[4 * c4], g[4 * c4 + 1], g[4 * c4 + 2], g[4 * c4 + 3]);
            }
        }
        if (tid < 128) {
            auto tot = [&](int qn) {
                const float *r = red + qn * 512 + tid;
                return (r[0] + r[128]) + (r[256] + r[384]);
            };
            if (tid < H) {
                G[L.off_b[0] + tid] = tot(0);
                G[L.off_b[1] + tid] = tot(1);
                G[L.off_w_out + tid] = tot(2);
            }
            for (int mb = 0; mb < t.MB; ++mb) {
                const int f = 128 * mb + tid;
                const float ga = tot(3 + 3 * mb), gb = tot(4 + 3 * mb), gc = tot(5 + 3 * mb);
                if (f < F) {
                    G[L.off_enc_w + 2 * f] = ga;
                    G[L.off_enc_w + 2 * f + 1] = gb;
                    G[L.off_enc_b + f] = d.bias ? gc : 0.f;
                } else if (f < FIN) {
                    G[L.off_ang_b + f - F] = ga;
                    G[L.off_ang_f + f - F] = gb;
                }
            }
            if (tid == 0) {
                float g3 = 0.f;
                for (int i = 0; i < TC_TP; ++i) g3 += red[9 * 512 + i];
                G[L.off_b_out] = g3;
            }
        }
        __syncthreads();
        const size_t po = (size_t)p * L.n_params;
        if (a.adam_w == nullptr) {
            float4 *dst = reinterpret_cast<float4 *>(a.grad_out + po);
            for (int i = tid; i < np4; i += TC_THREADS) dst[i] = reinterpret_cast<const float4 *>(G)[i];
        } else {
            // torch.optim.Adam (_single_tensor_adam) over the whole blob, fully coalesced; the new weights replace the
            // gradient in shared memory and are re-split into the bf16 operand copy below
            const float step_size = a.sp->onf_step_size, inv_bc2 = 1.f / a.sp->onf_bc2_sqrt;
            float4 *w4 = reinterpret_cast<float4 *>(a.adam_w + po), *m4 = reinterpret_cast<float4 *>(a.adam_m + po),
                   *v4 = reinterpret_cast<float4 *>(a.adam_v + po);
            constexpr int U = 4;
            for (int i0 = tid; i0 < np4; i0 += U * TC_THREADS) {
                float4 wv[U], mv[U], vv[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int i = i0 + u * TC_THREADS;
                    if (i < np4) { wv[u] = w4[i]; mv[u] = m4[i]; vv[u] = v4[i]; }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int i = i0 + u * TC_THREADS;
                    if (i >= np4) continue;
                    const float4 g4 = reinterpret_cast<const float4 *>(G)[i];
                    float gg[4] = {g4.x, g4.y, g4.z, g4.w}, ww[4] = {wv[u].x, wv[u].y, wv[u].z, wv[u].w},
                          mm[4] = {mv[u].x, mv[u].y, mv[u].z, mv[u].w}, v2[4] = {vv[u].x, vv[u].y, vv[u].z, vv[u].w};
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        mm[c] = fmaf(a.w1, gg[c] - mm[c], mm[c]);
                        v2[c] = fmaf(a.w2 * gg[c], gg[c], v2[c] * a.beta2);
                        float sq;
                        asm("sqrt.approx.f32 %0, %1;" : "=f"(sq) : "f"(v2[c]));
                        const float denom = fmaf(sq, inv_bc2, a.eps);
                        ww[c] = fmaf(-step_size, __fdividef(mm[c], denom), ww[c]);
                    }
                    w4[i] = make_float4(ww[0], ww[1], ww[2], ww[3]);
                    m4[i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
                    v4[i] = make_float4(v2[0], v2[1], v2[2], v2[3]);
                    reinterpret_cast<float4 *>(G)[i] = make_float4(ww[0], ww[1], ww[2], ww[3]);
                }
            }
            __syncthreads();
            // refresh the packed operand copy: item = (matrix, column octet, row) -> 16 B of hi + 16 B of lo
            unsigned char *pk = a.packed_out + (size_t)p * t.packed_bytes;
            const int rows = KH, n1 = (K1 / 8) * rows, n2 = (KH / 8) * rows;
            for (int item = tid; item < n1 + n2; item += TC_THREADS) {
                const bool second = item >= n1;
                const int it = second ? item - n1 : item;
                const int oct = it / rows, r = it - oct * rows;
                const int kdim = second ? H : FIN;
                float v[8];
                const float *src = nullptr;
                if (!second) {
                    if (r < H) src = G + L.off_w[0] + r * FIN;
                    else if (r == H) src = G + L.off_w_out + H;
                } else if (r < H) {
                    src = G + L.off_w[1] + r * H;
                }
#pragma unroll
                for (int h4 = 0; h4 < 2; ++h4) {
                    const int c = oct * 8 + 4 * h4;
                    float4 x4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (src && c < kdim) x4 = *reinterpret_cast<const float4 *>(src + c);
                    v[4 * h4] = x4.x; v[4 * h4 + 1] = x4.y; v[4 * h4 + 2] = x4.z; v[4 * h4 + 3] = x4.w;
                }
                uint4 hi, lo;
                split8(v, hi, lo);
                unsigned char *base = pk + (second ? 2 * t.w1_half : 0);
                const int half = second ? t.w2_half : t.w1_half;
                const int off = oct * t.S_colW + (r >> 3) * 128 + (r & 7) * 16;
                *reinterpret_cast<uint4 *>(base + off) = hi;
                *reinterpret_cast<uint4 *>(base + half + off) = lo;
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}


// ---- host side ---------------------------------------------------------------------------------------------
bool onf_tc_supported(const OnfDims &d) {
    TcDims t;
    return tc_make_dims(d, &t);
}

size_t onf_tc_packed_bytes(const OnfDims &d) {
    TcDims t;
    return tc_make_dims(d, &t) ? (size_t)t.packed_bytes : 0;
}

int onf_tc_prepare(const OnfDims &d) {
    TcDims t;
    NFOPP_REQUIRE(tc_make_dims(d, &t), "onf(tensor): needs 2 hidden layers, hidden <= 111, f_in <= 224 (hidden=%d layers=%d)", d.H,
                  d.L);
    NFOPP_CHECK_CUDA(cudaFuncSetAttribute(onf_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, t.smem_bytes));
    return NFOPP_OK;
}

int onf_tc_pack(const OnfDims &d, const float *weights, unsigned char *packed, int P, cudaStream_t stream) {
    TcPackArgs a{};
    a.d = d;
    NFOPP_REQUIRE(tc_make_dims(d, &a.t), "onf(tensor): unsupported network shape");
    a.weights = weights; a.packed = packed;
    if (P == 0) return NFOPP_OK;
    const int items = (a.t.K1 / 8 + a.t.KH / 8) * a.t.KH;
    dim3 grid((items + 255) / 256, P);
    onf_tc_pack_kernel<<<grid, 256, 0, stream>>>(a);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

int onf_tc_launch(const OnfArgs &o, const OnfTcExtra &x, int P, cudaStream_t stream) {
    OnfTcArgs a{};
    a.d = o.d;
    NFOPP_REQUIRE(tc_make_dims(o.d, &a.t), "onf(tensor): unsupported network shape");
    a.weights = o.weights; a.packed = x.packed; a.x = o.x; a.count = o.count; a.sp = o.sp;
    a.n = o.n; a.x_stride = o.x_stride ? o.x_stride : o.d.lay.in_dim;
    a.want_bwd = o.want_bwd; a.gbar_mode = o.gbar_mode; a.gz = o.gz; a.aux = o.aux; a.wc = o.wc; a.beta = o.beta;
    a.z_out = o.z_out; a.gx_out = o.gx_out; a.grad_out = x.grad_out;
    a.adam_w = x.adam_w; a.adam_m = x.adam_m; a.adam_v = x.adam_v; a.packed_out = x.packed_out;
    a.w1 = x.w1; a.beta2 = x.beta2; a.w2 = x.w2; a.eps = x.eps; a.warm = x.warm;
    NFOPP_REQUIRE(a.packed, "onf(tensor): packed operand copy missing");
    NFOPP_REQUIRE(!a.adam_w || (a.sp && a.packed_out), "onf(tensor): fused Adam needs step parameters and the packed copy");
    if (a.n == 0 || P == 0) return NFOPP_OK;
    onf_tc_kernel<<<P, TC_THREADS, a.t.smem_bytes, stream>>>(a);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

}  // namespace nfopp
