// onf_tc.cu - K2, tensor-core path: the fused ONF kernel on tcgen05 (sm_100a).
//
// One CTA = one planning problem (its own weights); the problem's points are walked in tiles of <= 32 by two groups
// of 8 warps (alternating tiles) while a 17th warp issues every MMA, so one group's epilogue overlaps the other's MMAs.  Every
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
#include <stdlib.h>

#include "onf.cuh"
#include "tc05.cuh"

namespace nfopp {
using namespace tc05;

// optional phase trace (build with -DNFOPP_TC_TRACE; scripts/tc_trace.py): clock64 stamps of the group leaders of two CTAs
#ifdef NFOPP_TC_TRACE
__device__ long long g_tc_trace[3][2][2][96];   // [launch kind: fwd | train | path][first / last CTA][group][slot]
#define TC_TRACE(slot)                                                                                      \
    do {                                                                                                    \
        if ((tid & (TC_GT - 1)) == 0 && tid < TC_WORKERS && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1) && (slot) < 96) \
            g_tc_trace[!a.want_bwd ? 0 : ((a.grad_out || a.adam_w) ? 1 : 2)][blockIdx.x == 0 ? 0 : 1][tid / TC_GT][(slot)] = clock64();                           \
    } while (0)
#else
#define TC_TRACE(slot) do { } while (0)
#endif

constexpr int TC_WORKERS = 512;                  // 2 groups x 8 epilogue warps
constexpr int TC_THREADS = TC_WORKERS + 32;      // + 1 MMA-issue warp
constexpr int TC_GT = 256;                       // threads per group
constexpr int TC_TP = 32;                        // points per tile (one group works on one tile at a time)
constexpr int TC_RHALF = 8192;                   // scratch region half (hi or lo): E ring / gH2 / E^T for 32 points
constexpr uint32_t TM_DW1 = 0, TM_DW2 = 224, TM_WORK = 336;   // TMEM columns; group g works in [336 + 64 g, +64)

struct TcDims {
    int F, A, FIN, H, K1, KH, MB;                // MB = M blocks of dgrad1 = feature groups of 128
    int S_colW;                                  // column-octet stride of the packed weights = KH * 16 bytes
    int w1_half, w2_half, h1_half;               // bytes of one hi (or lo) buffer
    int off_w1, off_w2, off_grp, grp_bytes, off_misc;   // shared-memory byte offsets
    int smem_bytes;
    int packed_bytes;                            // per problem, global
};

__host__ __device__ inline int ru16(int x) { return (x + 15) / 16 * 16; }

constexpr int TC_GRP_FLOATS = 8 * TC_TP + 4 * TC_TP + 4 * TC_TP * 3;     // u0,u1,th (x2: next tile prefetched),skip,gbar | zp | gx

__host__ __device__ inline bool tc_make_dims(const OnfDims &d, TcDims *out) {
    TcDims t{};
    t.F = d.lay.f_pos; t.A = d.lay.n_angle; t.FIN = d.lay.f_in; t.H = d.H;
    t.K1 = ru16(t.FIN); t.KH = ru16(t.H + 1); t.MB = (t.K1 + 127) / 128;
    t.S_colW = t.KH * 16;
    t.w1_half = (t.K1 / 8) * t.S_colW;
    t.w2_half = (t.KH / 8) * t.S_colW;
    t.h1_half = (t.KH / 8) * (TC_TP / 8) * 128;
    t.off_w1 = 0;
    t.off_w2 = 2 * t.w1_half;
    t.off_grp = t.off_w2 + 2 * t.w2_half;
    t.grp_bytes = 2 * t.h1_half + 2 * TC_RHALF;
    t.off_misc = t.off_grp + 2 * t.grp_bytes;
    int misc_floats = 3 * t.K1 + (t.A + 3) / 4 * 4 + 3 * t.H + 4 /*b3 + pad*/ + 2 * TC_GRP_FLOATS;
    t.smem_bytes = t.off_misc + misc_floats * 4 + 128;
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

// shared-memory views: weights and per-feature parameters are shared by the CTA, everything else is per group
struct TcShared {
    unsigned char *w1h, *w1l, *w2h, *w2l;
    // per input feature f < K1: arg = x1 * pb + u0 * pa + pc  (x1 = u1 for Fourier features, theta for heading features;
    // pc carries the bias and +pi/2 for cosine features, so every feature is sin(arg) with derivative cos(arg))
    float *pa, *pb, *pc, *ang_b, *b1, *b2, *w3h, *b3;
    float *grp_f;                 // 2 x TC_GRP_FLOATS
    uint64_t *bars;               // [0] weights, then per group: ready, done, aux
    uint32_t *tmem_slot;
    __device__ TcShared(unsigned char *base, const TcDims &t) {
        w1h = base + t.off_w1; w1l = w1h + t.w1_half;
        w2h = base + t.off_w2; w2l = w2h + t.w2_half;
        float *f = reinterpret_cast<float *>(base + t.off_misc);
        pa = f; f += t.K1;
        pb = f; f += t.K1;
        pc = f; f += t.K1;
        ang_b = f; f += (t.A + 3) / 4 * 4;
        b1 = f; f += t.H;
        b2 = f; f += t.H;
        w3h = f; f += t.H;
        b3 = f; f += 4;
        grp_f = f; f += 2 * TC_GRP_FLOATS;
        bars = reinterpret_cast<uint64_t *>(f);
        tmem_slot = reinterpret_cast<uint32_t *>(bars + 8);
    }
};
struct TcGroup {
    unsigned char *h1h, *h1l, *rh, *rl;      // H1 / gH1 operand, scratch region (E ring / gH2 / E^T)
    float *in0, *skip, *gbar, *zp, *gx;            // in0: [2][u0 | u1 | th] input buffers (this tile / next tile)
    uint64_t *ready, *done, *aux;
    __device__ TcGroup(unsigned char *base, const TcDims &t, const TcShared &s, int g) {
        h1h = base + t.off_grp + g * t.grp_bytes; h1l = h1h + t.h1_half;
        rh = h1l + t.h1_half; rl = rh + TC_RHALF;
        float *f = s.grp_f + g * TC_GRP_FLOATS;
        in0 = f; f += 6 * TC_TP;
        skip = f; f += TC_TP;
        gbar = f; f += TC_TP;
        zp = f; f += 4 * TC_TP;
        gx = f;
        ready = s.bars + 1 + 3 * g; done = ready + 1; aux = ready + 2;
    }
};

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

// K loop of one GEMM: `ksteps` x (three-pass split product), descriptors advanced by constant strides
__device__ __forceinline__ void mma_kloop(uint32_t tmem_d, uint64_t ah, uint64_t al, uint64_t bh, uint64_t bl, uint32_t a_step,
                                          uint32_t b_step, int ksteps, uint32_t idesc, bool accumulate, uint32_t issue) {
    const uint64_t da = a_step >> 4, db = b_step >> 4;
#pragma unroll 1
    for (int ks = 0; ks < ksteps; ++ks) {
        mma3(tmem_d, ah, al, bh, bl, idesc, accumulate || ks > 0, issue);
        ah += da; al += da; bh += db; bl += db;
    }
}

__device__ __forceinline__ void group_sync(int g) { asm volatile("bar.sync %0, %1;" ::"r"(1 + g), "r"(TC_GT) : "memory"); }
// a worker thread has finished writing MMA operands (generic proxy) / reading TMEM: publish to the MMA warp
__device__ __forceinline__ void signal_ready(uint64_t *bar) {
    fence_proxy_async();
    fence_before_sync();
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(bar);     // one arrival per warp (8 per group), not 256 smem atomics
}
__device__ __forceinline__ void wait_done(uint64_t *bar, uint32_t &phase) {
    mbar_wait(bar, phase);
    phase ^= 1;
    fence_after_sync();
}

// S_row of the point-indexed operand buffers (H1, gH1, gH2): 4 point octets x 128 B
constexpr uint32_t S_PT = (TC_TP / 8) * 128;

// ---- the MMA-issue warp: serves both groups, one tile phase at a time, in whatever order they become ready ----
__device__ __forceinline__ void tc_issue_warp(const OnfTcArgs &a, const TcShared &s, unsigned char *smem_raw, uint32_t tmem,
                                              int cnt, bool want_bwd, bool want_wgrad) {
    const TcDims &t = a.t;
    const int K1 = t.K1, KH = t.KH, MB = t.MB;
    const int n_tiles = (cnt + TC_TP - 1) / TC_TP;
    const int n_phases = !want_bwd ? MB + 1 : (want_wgrad ? 2 * MB + 3 : MB + 3);
    const uint32_t lead = elect_one();
    const uint32_t w1h = smem_u32(s.w1h), w1l = smem_u32(s.w1l), w2h = smem_u32(s.w2h), w2l = smem_u32(s.w2l);
    mbar_wait(s.bars, 0);                                     // packed weights have landed (bulk copies)
    int tile[2] = {0, 1}, phase[2] = {0, 0};
    uint32_t par[2] = {0, 0};
    bool dw2_started = false, dw1_started[2] = {false, false};
    const long long t_start = clock64();
    while (tile[0] < n_tiles || tile[1] < n_tiles) {
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            if (tile[g] >= n_tiles) continue;
            TcGroup G(smem_raw, t, s, g);
            if (!mbar_try_wait(G.ready, par[g])) continue;
            par[g] ^= 1;
            fence_after_sync();
            const int Nt = ru16(min(TC_TP, cnt - tile[g] * TC_TP));
            const uint32_t tw = tmem + TM_WORK + 64 * g;
            const uint32_t h1h = smem_u32(G.h1h), h1l = smem_u32(G.h1l), rh = smem_u32(G.rh), rl = smem_u32(G.rl);
            const int ph = phase[g];
            if (ph < MB) {
                // fwd1, feature group ph: D1[j,n] (+)= W1p[j,k] E[n,k]
                const uint32_t aoff = (uint32_t)(16 * ph) * t.S_colW;
                mma_kloop(tw, desc_kmajor(w1h + aoff, 128, t.S_colW), desc_kmajor(w1l + aoff, 128, t.S_colW),
                          desc_kmajor(rh, 128, S_PT), desc_kmajor(rl, 128, S_PT), 2 * t.S_colW, 2 * S_PT,
                          min(8, (K1 - 128 * ph) / 16), instr_desc_bf16(128, Nt, 0, 0), ph > 0, lead);
                mma_commit(ph + 1 < MB ? G.aux : G.done, lead);
            } else if (ph == MB) {
                // fwd2: D2[i,n] = W2p[i,j] H1[j,n]
                mma_kloop(tw + 32, desc_kmajor(w2h, 128, t.S_colW), desc_kmajor(w2l, 128, t.S_colW),
                          desc_mnmajor(h1h, S_PT, 128), desc_mnmajor(h1l, S_PT, 128), 2 * t.S_colW, 2 * S_PT, KH / 16,
                          instr_desc_bf16(128, Nt, 0, 1), false, lead);
                mma_commit(G.done, lead);
            } else if (ph == MB + 1) {
                // wgrad2 (persistent accumulator): dW2[i,j] += gH2[i,n] H1[j,n];  dgrad2: gD1[j,n] = W2p[i,j] gH2[i,n]
                if (want_wgrad) {
                    mma_kloop(tmem + TM_DW2, desc_kmajor(rh, S_PT, 128), desc_kmajor(rl, S_PT, 128), desc_kmajor(h1h, S_PT, 128),
                              desc_kmajor(h1l, S_PT, 128), 256, 256, Nt / 16, instr_desc_bf16(128, KH, 0, 0), dw2_started, lead);
                    dw2_started = true;
                }
                mma_kloop(tw, desc_mnmajor(w2h, 128, t.S_colW), desc_mnmajor(w2l, 128, t.S_colW), desc_mnmajor(rh, S_PT, 128),
                          desc_mnmajor(rl, S_PT, 128), 256, 2 * S_PT, KH / 16, instr_desc_bf16(128, Nt, 1, 1), false, lead);
                mma_commit(G.done, lead);
            } else if (ph == MB + 2) {
                // dgrad1: gE[k,n] = W1p[j,k] gH1[j,n] for all feature blocks
                for (int mb = 0; mb < MB; ++mb) {
                    const uint32_t aoff = (uint32_t)(16 * mb) * t.S_colW;
                    mma_kloop(tw + 32 * mb, desc_mnmajor(w1h + aoff, 128, t.S_colW), desc_mnmajor(w1l + aoff, 128, t.S_colW),
                              desc_mnmajor(h1h, S_PT, 128), desc_mnmajor(h1l, S_PT, 128), 256, 2 * S_PT, KH / 16,
                              instr_desc_bf16(128, Nt, 1, 1), false, lead);
                }
                mma_commit(G.done, lead);
            } else {
                // wgrad1, feature block mb: dW1[j, 128 mb + f] += gH1[j,n] E[n,f]
                const int mb = ph - (MB + 3);
                mma_kloop(tmem + TM_DW1 + 128 * mb, desc_kmajor(h1h, S_PT, 128), desc_kmajor(h1l, S_PT, 128),
                          desc_kmajor(rh, 128, 2048), desc_kmajor(rl, 128, 2048), 256, 4096, Nt / 16,
                          instr_desc_bf16(128, min(128, K1 - 128 * mb), 0, 0), dw1_started[mb], lead);
                dw1_started[mb] = true;
                mma_commit(G.aux, lead);
            }
            if (++phase[g] == n_phases) { phase[g] = 0; tile[g] += 2; }
        }
        if (clock64() - t_start > 20000000000LL) __trap();    // protocol bug: fail loudly instead of hanging
    }
}

// Thread layout: 17 warps.  Warps 0..15 are workers in two groups of 8 (group g = warp / 8 walks the tiles g, g+2, ...
// of 32 points each); inside a group warp lw = 4*cq + q: q = TMEM lane quarter (rows 32q..32q+31), cq = column half
// of the tile (columns 16cq..16cq+15), so a thread owns one hidden unit / feature (`row`) and 16 points.  Warp 16
// issues every tcgen05.mma: while one group runs an epilogue on the CUDA cores the other group's MMAs execute.
__global__ void __launch_bounds__(TC_THREADS, 1) onf_tc_kernel(const OnfTcArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const OnfDims &d = a.d;
    const TcDims &t = a.t;
    const nfopp_onf_layout &L = d.lay;
    const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (a.sp && !a.warm && !a.sp->do_train) return;    // engine: not a field-training step
    const int cnt = a.count ? a.count[p] : a.n;
    if (cnt <= 0) return;
    const bool want_bwd = a.want_bwd != 0;
    const bool want_wgrad = want_bwd && (a.grad_out != nullptr || a.adam_w != nullptr);
    const bool want_gx = want_bwd && a.gx_out != nullptr;
    const int F = t.F, FIN = t.FIN, H = t.H, K1 = t.K1, KH = t.KH;
    const bool is_issuer = warp == TC_WORKERS / 32;
    TC_TRACE(0);

    TcShared s(smem_raw, t);
    const float *W = a.weights + (size_t)p * L.n_params;

    // ---- one-time setup: barriers, TMEM, packed weights by bulk copy (TMA), per-feature / per-unit fp32 parameters ----
    if (tid == 0) {
        mbar_init(s.bars, 1);
        for (int g = 0; g < 2; ++g) {
            mbar_init(s.bars + 1 + 3 * g, TC_GT / 32); // ready: one arrival per worker warp of the group
            mbar_init(s.bars + 2 + 3 * g, 1);          // done: one tcgen05.commit
            mbar_init(s.bars + 3 + 3 * g, 1);          // aux: one tcgen05.commit
        }
        fence_mbar_init();
    }
    if (is_issuer) tmem_alloc(s.tmem_slot, 512);
    __syncthreads();
    if (is_issuer && lane == 0) {
        const unsigned char *src = a.packed + (size_t)p * t.packed_bytes;
        mbar_expect_tx(s.bars, (uint32_t)t.packed_bytes);
        bulk_g2s(s.w1h, src, 2 * t.w1_half, s.bars);
        bulk_g2s(s.w2h, src + 2 * t.w1_half, 2 * t.w2_half, s.bars);
        if (a.adam_w != nullptr && want_bwd) {
            // the Adam epilogue streams this problem's w / m / v once: start pulling them into L2 now, under the compute
            const size_t po = (size_t)p * L.n_params;
            const uint32_t bytes = (uint32_t)(L.n_params * sizeof(float)) & ~15u;
            bulk_prefetch_l2(a.adam_w + po, bytes);
            bulk_prefetch_l2(a.adam_m + po, bytes);
            bulk_prefetch_l2(a.adam_v + po, bytes);
        }
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
    TC_TRACE(1);

    // worker identity (the issuer warp computes harmless values)
    const int g = (warp >> 3) & 1, gt = tid & (TC_GT - 1), lw = warp & 7;
    const int q = lw & 3, cq = lw >> 2;
    const int row = q * 32 + lane;
    const uint32_t tm_lane = tmem + ((uint32_t)(q * 32) << 16);
    // thread-local gradient accumulators (hidden unit `row`; features row and 128+row; point gt)
    float acc_db1 = 0.f, acc_db2 = 0.f, acc_dw3 = 0.f, acc_db3 = 0.f;
    float fa0 = 0.f, fb0 = 0.f, fc0 = 0.f, fa1 = 0.f, fb1 = 0.f, fc1 = 0.f;

    if (is_issuer) {
        tc_issue_warp(a, s, smem_raw, tmem, cnt, want_bwd, want_wgrad);
    } else {
        TcGroup G(smem_raw, t, s, g);
        const uint32_t tw = tm_lane + TM_WORK + 64 * g;
        uint32_t ph_done = 0, ph_aux = 0;
        bool pending_aux = false;
        const float my_b1 = row < H ? s.b1[row] : 0.f, my_b2 = row < H ? s.b2[row] : 0.f, my_w3 = row < H ? s.w3h[row] : 0.f;
        const int n_tiles = (cnt + TC_TP - 1) / TC_TP;
        const int c0 = 16 * cq;                            // this warp's 16 columns of the tile
        if (gt < TC_TP) {                                  // inputs of the group's first tile
            float x0 = 0.f, x1 = 0.f, x2 = 0.f;
            if (g * TC_TP + gt < cnt) {
                const float *xp = a.x + ((size_t)p * a.n + g * TC_TP + gt) * a.x_stride;
                x0 = xp[0]; x1 = xp[1];
                if (L.in_dim == 3) x2 = xp[2];
            }
            G.in0[gt] = (x0 - d.mean) / d.sigma;
            G.in0[TC_TP + gt] = (x1 - d.mean) / d.sigma;
            G.in0[2 * TC_TP + gt] = x2;
        }
        group_sync(g);
        for (int tile = g; tile < n_tiles; tile += 2) {
            const int n0 = tile * TC_TP;
            const int nv = min(TC_TP, cnt - n0);
            const int Nt = ru16(nv);
            const bool cols_on = c0 < Nt;
            const int tsl = 2 + 16 * (tile >> 1);
            TC_TRACE(tsl);
            // ---- P0: this tile's inputs were staged by the previous tile (or before the loop); threads gt < 32 now
            // fetch the next tile's inputs (onf_model.py:37-38) and this tile's label / upstream gradient, consumed later ----
            float *cu0 = G.in0 + ((tile >> 1) & 1) * 3 * TC_TP, *cu1 = cu0 + TC_TP, *cth = cu1 + TC_TP;
            float nx0 = 0.f, nx1 = 0.f, nx2 = 0.f, my_aux = 0.f;
            if (gt < TC_TP) {
                const int nn = n0 + 2 * TC_TP + gt;
                if (nn < cnt) {
                    const float *xp = a.x + ((size_t)p * a.n + nn) * a.x_stride;
                    nx0 = xp[0]; nx1 = xp[1];
                    if (L.in_dim == 3) nx2 = xp[2];
                }
                if (gt < nv && want_bwd) {
                    const size_t gi = (size_t)p * a.n + n0 + gt;
                    if (a.gbar_mode == 0) my_aux = a.gz[gi];
                    else if (a.gbar_mode != 2) my_aux = a.aux[gi];
                }
            }
            TC_TRACE(tsl + 1);
            // ---- P1: encoding (CUDA cores) feeding fwd1 in groups of 128 features ----
            {
                const int n = gt & 31, oc = gt >> 5;         // point, feature octet (0..7) within a 64-feature half group
                const float u0 = cu0[n], u1 = cu1[n], th = cth[n];
#pragma unroll 1
                for (int fg = 0; fg < t.MB; ++fg) {
                    uint4 ehi[2], elo[2];
#pragma unroll
                    for (int it = 0; it < 2; ++it) {
                        const int f0 = 128 * fg + 8 * (oc + 8 * it);
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
                    if (fg > 0 || pending_aux) {          // the scratch region is still being read by earlier MMAs
                        mbar_wait(G.aux, ph_aux);
                        ph_aux ^= 1;
                        pending_aux = false;
                    }
#pragma unroll
                    for (int it = 0; it < 2; ++it) {
                        const int fo = oc + 8 * it;
                        if (128 * fg + 8 * fo < K1 && n < Nt) {
                            *reinterpret_cast<uint4 *>(G.rh + fo * S_PT + n * 16) = ehi[it];
                            *reinterpret_cast<uint4 *>(G.rl + fo * S_PT + n * 16) = elo[it];
                        }
                    }
                    signal_ready(G.ready);                // -> fwd1(fg)
                    TC_TRACE(tsl + 2 + fg);
                }
            }
            // ---- P2: epilogue 1: bias + ReLU -> H1 operand; row H = skip term of the logit ----
            wait_done(G.done, ph_done);
            TC_TRACE(tsl + 4);
            if (gt < TC_TP) {                             // next tile's inputs -> the other input buffer
                float *nu = G.in0 + (((tile >> 1) & 1) ^ 1) * 3 * TC_TP;
                nu[gt] = (nx0 - d.mean) / d.sigma;
                nu[TC_TP + gt] = (nx1 - d.mean) / d.sigma;
                nu[2 * TC_TP + gt] = nx2;
            }
            if (cols_on) {
                float v[16];
                tmem_ld16(tw + c0, v);
                if (row < H) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i] + my_b1, 0.f);
                } else {
                    if (row == H) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) G.skip[c0 + i] = v[i];
                    }
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = 0.f;
                }
                if (row < KH) store_row16(G.h1h, G.h1l, row, c0, v, S_PT);
            }
            signal_ready(G.ready);                        // -> fwd2
            TC_TRACE(tsl + 5);
            // ---- P4: epilogue 2: logit, dL/dz, gH2 operand ----
            wait_done(G.done, ph_done);
            TC_TRACE(tsl + 6);
            float r2[16];
            if (cols_on) {
                tmem_ld16(tw + 32 + c0, r2);
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
                if ((lane & 1) == 0) G.zp[q * TC_TP + c0 + (lane >> 1)] = zs;
            }
            group_sync(g);
            TC_TRACE(tsl + 7);
            if (gt < TC_TP) {
                const int n = gt;
                const float z = G.zp[n] + G.zp[TC_TP + n] + G.zp[2 * TC_TP + n] + G.zp[3 * TC_TP + n] + G.skip[n] + s.b3[0];
                const bool valid = n < nv;
                const size_t gi = (size_t)p * a.n + n0 + n;
                if (valid && a.z_out) a.z_out[gi] = z;
                float gb = 0.f;
                if (valid && want_bwd) {                        // dL/dz
                    if (a.gbar_mode == 0) gb = my_aux;
                    else if (a.gbar_mode == 1) gb = (sigmoid_f(z) - my_aux) / (float)cnt;
                    else {
                        const float zb = z * a.beta;
                        const float ez = expf(zb);
                        gb = a.wc * (zb > 20.f ? 1.f : ez / (ez + 1.f));
                        if (a.gbar_mode == 3) {
                            const float th = tanhf(z);
                            gb += my_aux * (1.f - th * th);
                        }
                    }
                }
                G.gbar[n] = gb;
                acc_db3 += gb;
            }
            group_sync(g);
            TC_TRACE(tsl + 8);
            if (!want_bwd) continue;
            if (cols_on) {
                float v[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float gb = G.gbar[c0 + i];
                    acc_dw3 = fmaf(gb, r2[i], acc_dw3);
                    const float gg = r2[i] > 0.f ? gb * my_w3 : 0.f;
                    acc_db2 += gg;
                    v[i] = gg;
                }
                if (row < KH) store_row16(G.rh, G.rl, row, c0, v, S_PT);
            }
            signal_ready(G.ready);                        // -> wgrad2 + dgrad2
            TC_TRACE(tsl + 9);
            // ---- P6: epilogue 3: ReLU mask of layer 1 -> gH1 operand (in place over H1); row H := dL/dz ----
            wait_done(G.done, ph_done);
            TC_TRACE(tsl + 10);
            if (cols_on) {
                float v[16];
                tmem_ld16(tw + c0, v);
                if (row < H) {
                    const int off = (row >> 3) * S_PT + (c0 >> 3) * 128 + (row & 7) * 16;
                    const uint4 m0 = *reinterpret_cast<const uint4 *>(G.h1h + off);
                    const uint4 m1 = *reinterpret_cast<const uint4 *>(G.h1h + off + 128);
                    const uint32_t mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const uint32_t bits = (mw[i >> 1] >> ((i & 1) * 16)) & 0x7FFFu;
                        v[i] = bits != 0 ? v[i] : 0.f;
                        acc_db1 += v[i];
                    }
                } else if (row == H) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = G.gbar[c0 + i];
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = 0.f;
                }
                if (row < KH) store_row16(G.h1h, G.h1l, row, c0, v, S_PT);
            }
            signal_ready(G.ready);                        // -> dgrad1
            TC_TRACE(tsl + 11);
            // ---- P8/P9: epilogue 4: encoding backward (+ input gradient), E^T operand, wgrad1 per feature block ----
            wait_done(G.done, ph_done);
            TC_TRACE(tsl + 12);
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
                        tmem_ld16(tw + 32 * mb + c0, v);
                        const float *x1src = fourier ? cu1 : cth;
#pragma unroll
                        for (int i4 = 0; i4 < 4; ++i4) {
                            *reinterpret_cast<float4 *>(xu + 4 * i4) = *reinterpret_cast<const float4 *>(cu0 + c0 + 4 * i4);
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
                    float *dst = G.gx + ((size_t)q * TC_TP + c0 + (lane >> 1)) * 3;
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
                        mbar_wait(G.aux, ph_aux);
                        ph_aux ^= 1;
                    }
                    if (f_on && cols_on) {
#pragma unroll
                        for (int o = 0; o < 2; ++o) {
                            const int off = ((c0 >> 3) + o) * 2048 + (row >> 3) * 128 + (row & 7) * 16;
                            *reinterpret_cast<uint4 *>(G.rh + off) = ehi[o];
                            *reinterpret_cast<uint4 *>(G.rl + off) = elo[o];
                        }
                    }
                    signal_ready(G.ready);                // -> wgrad1(mb)
                    TC_TRACE(tsl + 13 + mb);
                    if (mb + 1 >= t.MB) pending_aux = true;
                }
            }
            if (want_gx) {
                group_sync(g);
                if (gt < nv) {
                    const int n = gt;
                    float g0 = 0.f, g1 = 0.f, g2 = 0.f;
#pragma unroll
                    for (int qq = 0; qq < 4; ++qq) {
                        const float *src = G.gx + ((size_t)qq * TC_TP + n) * 3;
                        g0 += src[0]; g1 += src[1]; g2 += src[2];
                    }
                    float *out = a.gx_out + ((size_t)p * a.n + n0 + n) * a.x_stride;
                    out[0] = g0 / d.sigma;
                    out[1] = g1 / d.sigma;
                    if (L.in_dim == 3) out[2] = g2;
                    for (int c = L.in_dim; c < a.x_stride; ++c) out[c] = 0.f;   // pose columns the ONF does not read
                }
            }
            group_sync(g);             // zp, gx, skip, gbar are rewritten by the group's next tile
            TC_TRACE(tsl + 15);
        }
        if (pending_aux) {             // the last wgrad1 of this group
            mbar_wait(G.aux, ph_aux);
            ph_aux ^= 1;
        }
    }
    TC_TRACE(80);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    TC_TRACE(81);

    // ---- weight gradients: TMEM accumulators -> shared memory in blob order -> coalesced Adam / gradient store ----
    if (want_wgrad) {
        // every MMA has completed: the operand buffers are dead, shared memory [0, n_params) floats now holds the
        // gradient of the whole parameter blob (same element order as the fp32 weights)
        float *G = reinterpret_cast<float *>(smem_raw);
        float *red = G + ((L.n_params + 31) & ~31);                    // [10][4][128] partial per-unit sums
        const int np4 = L.n_params / 4;
        for (int i = tid; i < np4; i += TC_THREADS) reinterpret_cast<float4 *>(G)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        const int cq4 = warp >> 2;                                    // drain: 16 worker warps = 4 lane quarters x 4 column slots
        if (!is_issuer) {
            float *r0 = red + cq4 * 128 + row;
            r0[0 * 512] = acc_db1; r0[1 * 512] = acc_db2; r0[2 * 512] = acc_dw3;
            r0[3 * 512] = fa0; r0[4 * 512] = fb0; r0[5 * 512] = fc0;
            r0[6 * 512] = fa1; r0[7 * 512] = fb1; r0[8 * 512] = fc1;
            if (gt < TC_TP) red[9 * 512 + g * TC_TP + gt] = acc_db3;
        }
        __syncthreads();
        // dW1: rows j < H -> mlp.0.weight[j,:], row H -> mlp2.0.weight[0, H:];  dW2: rows i < H -> mlp.2.weight[i,:]
        for (int ch = cq4; ch < (is_issuer ? 0 : K1 / 16); ch += 4) {
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
        for (int ch = cq4; ch < (is_issuer ? 0 : KH / 16); ch += 4) {
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
                for (int i = 0; i < 2 * TC_TP; ++i) g3 += red[9 * 512 + i];
                G[L.off_b_out] = g3;
            }
        }
        __syncthreads();
        TC_TRACE(82);
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
            TC_TRACE(83);
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
    TC_TRACE(84);
    if (is_issuer) tmem_dealloc(tmem, 512);
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

int onf_tc_variant(const OnfDims &d, int n_wgrad) {
    static const bool legacy = [] { const char *e = getenv("NFOPP_TC_LEGACY"); return e && e[0] == '1'; }();
    (void)n_wgrad;
    return (!legacy && onf_tcs_supported(d)) ? 1 : 0;
}

int onf_tc_prepare(const OnfDims &d) {
    if (onf_tcs_supported(d)) {
        const int rc = onf_tcs_prepare(d);
        if (rc) return rc;
    }
    TcDims t;
    NFOPP_REQUIRE(tc_make_dims(d, &t), "onf(tensor): needs 2 hidden layers, hidden <= 111, f_in <= 224 (hidden=%d layers=%d)", d.H,
                  d.L);
    NFOPP_CHECK_CUDA(cudaFuncSetAttribute(onf_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, t.smem_bytes));
    return NFOPP_OK;
}

int onf_tc_pack(const OnfDims &d, const float *weights, unsigned char *packed, int P, cudaStream_t stream, int variant) {
    if (variant == 1) return onf_tcs_pack(d, weights, packed, P, stream);
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
    if (x.variant == 1) return onf_tcs_launch(o, x, P, stream);
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

#ifdef NFOPP_TC_TRACE
extern "C" int nfopp_debug_tc_trace(long long *host_out) {
    return (int)cudaMemcpyFromSymbol(host_out, nfopp::g_tc_trace, sizeof(nfopp::g_tc_trace));
}
#endif
