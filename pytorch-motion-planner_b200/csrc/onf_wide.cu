// onf_wide.cu - K2 for WIDE / DEEP fields and LARGE point sets (BASELINE configs 4 and 5: hidden 64..512, up to
// 8 hidden layers, 2^8..2^24 points per weight set): every dense contraction of ONF.forward (onf_model.py:33-50)
// and of its autograd backward is a grouped GEMM on tcgen05 with the POINT index on the UMMA M axis (128 TMEM
// lanes = 128 points) and the layer width on N (<= 256 per tile):
//
//   fwd   l : H_l [pt, out] = relu(sum_in  H_{l-1}[pt, in] W_l[out, in] + b_l)   A = H_{l-1} (K-major)  B = W_l (K-major)
//   dgrad l : G_{l-1}[pt,in] = 1[H_{l-1}>0] sum_out G_l[pt,out] W_l[out,in]      A = G_l     (K-major)  B = W_l (MN-major)
//   wgrad l : dW_l[out, in] = sum_pt G_l[pt, out] H_{l-1}[pt, in]                A = G_l     (MN-major) B = H_{l-1} (MN-major)
//
// (the encoder gradients, the output-layer gradient and every bias gradient are wgrad jobs too: B = [u0,u1,theta,1],
// B = dL/dz, B = ones).  Weights do not fit shared memory at these widths, so they are STREAMED: a producer warp
// issues bulk async copies (TMA) of K-chunks of the packed weights and of the activation tiles into a 4-stage ring,
// one warp issues the MMAs, four warps run the epilogue out of a double-buffered TMEM accumulator (bias + ReLU +
// bf16 hi/lo split -> the next layer's operand, or ReLU mask, or fp32 gradient).  Kernels are persistent (one CTA
// per SM walks a list of tiles).
//
// Precision: as in onf_tc.cu every fp32 operand is a bf16 hi + bf16 lo pair and every product three kind::f16
// MMAs accumulated in fp32 (~3e-6 relative).
//
// Operand storage in global memory = the shared-memory layout (SWIZZLE_NONE core blocks, tc05.cuh), so a stage is
// a handful of contiguous bulk copies:
//   activations [P][n_pad/64 tiles][hi|lo][C/8 column octets][64 points][8 columns] bf16      (tile = C*256 bytes)
//   Wf_l (fwd)  [n-tile of 256 out][hi|lo][in/8 octets][out rows][8 in]                       K-chunk contiguous
//   Wd_l (dgrad)[n-tile of 256 in ][hi|lo][out/8 octets][in/8 octets][8 out][8 in]            K-chunk contiguous
#include <stdlib.h>

#include "onf.cuh"
#include "tc05.cuh"

namespace nfopp {
using namespace tc05;

namespace {

constexpr int WG_THREADS = 192;          // warp 0: producer, warp 1: MMA issue + TMEM owner, warps 2..5: epilogue
constexpr int WG_KC = 32;                // K per stage (fwd / dgrad)
constexpr int WG_STAGES = 4;
constexpr int WG_NT = 256;               // N per tile
constexpr int WG_A_STAGE = 128 * WG_KC * 4;       // 128 points x KC x (hi + lo)
constexpr int WG_B_STAGE = WG_NT * WG_KC * 4;
constexpr int WG_STAGE = WG_A_STAGE + WG_B_STAGE;
constexpr int WG_SMEM = WG_STAGES * WG_STAGE + 256 + 2 * WG_NT * 4;   // + double-buffered per-tile bias / skip weights

constexpr int WW_STAGES = 2;             // wgrad: one stage = one 64-point tile of A (128 columns) and B (<= 256 columns)
constexpr int WW_A_STAGE = 128 * 64 * 4;
constexpr int WW_B_STAGE = WG_NT * 64 * 4;
constexpr int WW_STAGE = WW_A_STAGE + WW_B_STAGE;
constexpr int WW_ONES = 2048;            // [2 octets][64 points][8] bf16 1.0
constexpr int WW_SMEM = WW_STAGES * WW_STAGE + WW_ONES + 256;

__host__ __device__ inline int ru16i(int x) { return (x + 15) / 16 * 16; }

struct WideDims {
    int F, A, FIN, H, L, K1, Hp;
    int kin[NFOPP_MAX_LAYERS];            // padded input width of layer l
    int kin_real[NFOPP_MAX_LAYERS];
    size_t wf_off[NFOPP_MAX_LAYERS], wd_off[NFOPP_MAX_LAYERS];
    size_t packed_bytes;
};

inline bool make_wide_dims(const OnfDims &d, WideDims *out) {
    WideDims w{};
    w.F = d.lay.f_pos; w.A = d.lay.n_angle; w.FIN = d.lay.f_in; w.H = d.H; w.L = d.L;
    w.K1 = ru16i(w.FIN); w.Hp = ru16i(w.H);
    size_t off = 0;
    for (int l = 0; l < w.L; ++l) {
        w.kin[l] = l == 0 ? w.K1 : w.Hp;
        w.kin_real[l] = l == 0 ? w.FIN : w.H;
        const size_t bytes = (size_t)w.Hp * w.kin[l] * 4;
        w.wf_off[l] = off; off += bytes;
        w.wd_off[l] = off; off += bytes;
    }
    w.packed_bytes = off;
    *out = w;
    return w.Hp <= 1024 && w.L >= 1 && w.L <= NFOPP_MAX_LAYERS && (w.H % 4) == 0;
}

// sin of the feature argument, the same arithmetic as onf_tc.cu (phase folded into pc; cos = sin(. + pi/2))
__device__ __forceinline__ float w_reduce_2pi(float x) {
    const float k = rintf(x * 0.15915494309189535f);
    float r = fmaf(k, -6.2831854820251465f, x);
    return fmaf(k, 1.7484555e-7f, r);
}
struct FeatParams { float pa[8], pb[8], pc[8]; };
// parameters of features 8*co .. 8*co+7 of problem blob W
__device__ __forceinline__ void load_feat_params(const float *W, const OnfDims &d, int co, FeatParams &fp) {
    const nfopp_onf_layout &L = d.lay;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int f = co * 8 + i;
        float pa = 0.f, pb = 0.f, pc = 0.f;
        if (f < L.f_pos) {
            pa = __ldg(W + L.off_enc_w + 2 * f);
            pb = __ldg(W + L.off_enc_w + 2 * f + 1);
            pc = __ldg(W + L.off_enc_b + f) + ((d.use_cos && f >= 100) ? 1.57079632679489662f : 0.f);
        } else if (f < L.f_in) {
            const int k = f - L.f_pos;
            const float b = __ldg(W + L.off_ang_b + k), fr = __ldg(W + L.off_ang_f + k);
            pb = fr;
            pc = fmaf(b, fr, k >= L.n_angle / 2 ? 1.57079632679489662f : 0.f);
        }
        fp.pa[i] = pa; fp.pb[i] = pb; fp.pc[i] = pc;
    }
}

__device__ __forceinline__ void unpack8(const uint4 &hi, const uint4 &lo, float *v) {
    const uint32_t h[4] = {hi.x, hi.y, hi.z, hi.w}, l[4] = {lo.x, lo.y, lo.z, lo.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        v[2 * i] = __uint_as_float(h[i] << 16) + __uint_as_float(l[i] << 16);
        v[2 * i + 1] = __uint_as_float(h[i] & 0xFFFF0000u) + __uint_as_float(l[i] & 0xFFFF0000u);
    }
}

// ---- fp32 blob -> packed operand copies ------------------------------------------------------------------------
struct WidePackArgs {
    OnfDims d;
    WideDims w;
    const float *weights;
    unsigned char *packed;
    // fused Adam (engine): adam_m != null -> weights/m/v are updated from `grad` first (adam_w aliases weights)
    float *adam_w, *adam_m, *adam_v;
    const float *grad;
    const StepParams *sp;
    const int32_t *count;        // fused Adam: problems with count <= 0 (finished / frozen) are left untouched
    int warm;
    float w1, beta2, w2, eps;
};

__global__ void __launch_bounds__(256) wide_pack_kernel(const WidePackArgs a) {
    const int p = blockIdx.y;
    if (a.adam_m && a.count && a.count[p] <= 0) return;
    const WideDims &w = a.w;
    const float *W = a.weights + (size_t)p * a.d.lay.n_params;
    unsigned char *out = a.packed + (size_t)p * ((w.packed_bytes + 255) / 256 * 256);
    const int gid = blockIdx.x * blockDim.x + threadIdx.x, gstride = gridDim.x * blockDim.x;
    float step_size = 0.f, bc2_sqrt = 1.f;
    if (a.adam_m) {
        if (!a.warm && !a.sp->do_train) return;
        step_size = a.sp->onf_step_size;
        bc2_sqrt = a.sp->onf_bc2_sqrt;
        // everything that is not a hidden-layer matrix (encoder, biases, output layer): plain Adam
        const nfopp_onf_layout &L = a.d.lay;
        const size_t po = (size_t)p * L.n_params;
        for (int i4 = gid; i4 < L.n_params / 4; i4 += gstride) {
            const int e = 4 * i4;
            bool matrix = false;
            for (int l = 0; l < w.L; ++l) matrix |= e >= L.off_w[l] && e < L.off_b[l];
            if (matrix) continue;
            const float4 g = *reinterpret_cast<const float4 *>(a.grad + po + e);
            float4 x = *reinterpret_cast<float4 *>(a.adam_w + po + e), m = *reinterpret_cast<float4 *>(a.adam_m + po + e),
                   vv = *reinterpret_cast<float4 *>(a.adam_v + po + e);
            float xs[4] = {x.x, x.y, x.z, x.w}, gs[4] = {g.x, g.y, g.z, g.w}, ms[4] = {m.x, m.y, m.z, m.w}, vs[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                ms[c] = fmaf(a.w1, gs[c] - ms[c], ms[c]);
                vs[c] = fmaf(a.w2 * gs[c], gs[c], vs[c] * a.beta2);
                const float denom = sqrtf(vs[c]) / bc2_sqrt + a.eps;
                xs[c] = fmaf(-step_size, ms[c] / denom, xs[c]);
            }
            *reinterpret_cast<float4 *>(a.adam_w + po + e) = make_float4(xs[0], xs[1], xs[2], xs[3]);
            *reinterpret_cast<float4 *>(a.adam_m + po + e) = make_float4(ms[0], ms[1], ms[2], ms[3]);
            *reinterpret_cast<float4 *>(a.adam_v + po + e) = make_float4(vs[0], vs[1], vs[2], vs[3]);
        }
    }
    // hidden-layer matrices in tiles of 32 out-rows x 8 in-octets: phase 1 walks a row (in-octet fastest: coalesced fp32
    // traffic for Adam), phase 2 walks a column (out-row fastest: the operand copies are contiguous along the out index)
    __shared__ uint4 t_hi[8][33], t_lo[8][33];
    const int tid = threadIdx.x;
    for (int l = 0; l < w.L; ++l) {
        const int K = w.kin[l], kr = w.kin_real[l], KO = K / 8;
        const int ot_n = (w.Hp + 31) / 32, it_n = (KO + 7) / 8;
        for (int tile = blockIdx.x; tile < ot_n * it_n; tile += gridDim.x) {
            const int o0 = (tile / it_n) * 32, io0 = (tile % it_n) * 8;
            {
                const int io = io0 + (tid & 7), o = o0 + (tid >> 3);
                float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (o < w.H && io < KO) {
#pragma unroll
                    for (int h4 = 0; h4 < 2; ++h4) {
                        const int i = io * 8 + 4 * h4;
                        if (i >= kr) continue;
                        const size_t e = (size_t)a.d.lay.off_w[l] + (size_t)o * kr + i;
                        float4 x = *reinterpret_cast<const float4 *>(W + e);
                        if (a.adam_m) {
                            // torch.optim.Adam on these four weights, then the operand copies from the NEW values
                            const size_t ge = (size_t)p * a.d.lay.n_params + e;
                            const float4 g = *reinterpret_cast<const float4 *>(a.grad + ge);
                            float4 m = *reinterpret_cast<float4 *>(a.adam_m + ge), vv = *reinterpret_cast<float4 *>(a.adam_v + ge);
                            float xs[4] = {x.x, x.y, x.z, x.w}, gs[4] = {g.x, g.y, g.z, g.w}, ms[4] = {m.x, m.y, m.z, m.w},
                                  vs[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
                            for (int c = 0; c < 4; ++c) {
                                ms[c] = fmaf(a.w1, gs[c] - ms[c], ms[c]);
                                vs[c] = fmaf(a.w2 * gs[c], gs[c], vs[c] * a.beta2);
                                const float denom = sqrtf(vs[c]) / bc2_sqrt + a.eps;
                                xs[c] = fmaf(-step_size, ms[c] / denom, xs[c]);
                            }
                            x = make_float4(xs[0], xs[1], xs[2], xs[3]);
                            *reinterpret_cast<float4 *>(a.adam_w + ge) = x;
                            *reinterpret_cast<float4 *>(a.adam_m + ge) = make_float4(ms[0], ms[1], ms[2], ms[3]);
                            *reinterpret_cast<float4 *>(a.adam_v + ge) = make_float4(vs[0], vs[1], vs[2], vs[3]);
                        }
                        v[4 * h4] = x.x; v[4 * h4 + 1] = x.y; v[4 * h4 + 2] = x.z; v[4 * h4 + 3] = x.w;
                    }
                }
                uint4 hi, lo;
                split8(v, hi, lo);
                t_hi[tid & 7][tid >> 3] = hi;
                t_lo[tid & 7][tid >> 3] = lo;
            }
            __syncthreads();
            {
                const int io = io0 + (tid >> 5), o = o0 + (tid & 31);
                if (o < w.Hp && io < KO) {
                    const uint4 hi = t_hi[tid >> 5][tid & 31], lo = t_lo[tid >> 5][tid & 31];
                    {   // Wf: n-tile over out
                        const int nt = o / WG_NT, ow = min(WG_NT, w.Hp - nt * WG_NT), oo = o - nt * WG_NT;
                        unsigned char *b = out + w.wf_off[l] + (size_t)nt * WG_NT * K * 4;
                        const size_t off = (size_t)io * ow * 16 + (size_t)oo * 16;
                        *reinterpret_cast<uint4 *>(b + off) = hi;
                        *reinterpret_cast<uint4 *>(b + (size_t)K * ow * 2 + off) = lo;
                    }
                    {   // Wd: n-tile over in
                        const int i0 = io * 8, nt = i0 / WG_NT, iw = min(WG_NT, K - nt * WG_NT), ii = i0 - nt * WG_NT;
                        unsigned char *b = out + w.wd_off[l] + (size_t)nt * WG_NT * w.Hp * 4;
                        const size_t off = (size_t)(o / 8) * iw * 16 + (size_t)(ii / 8) * 128 + (o % 8) * 16;
                        *reinterpret_cast<uint4 *>(b + off) = hi;
                        *reinterpret_cast<uint4 *>(b + (size_t)w.Hp * iw * 2 + off) = lo;
                    }
                }
            }
            __syncthreads();
        }
    }
}

// ---- encoding: x -> E tiles (+ U16 = [u0, u1, theta, 1] for the encoder gradients) -------------------------------
struct WideEncArgs {
    OnfDims d;
    const float *weights, *x;
    const int32_t *count;
    int n, x_stride, T64, K1;
    unsigned char *E, *U16;      // U16 may be null
};

__global__ void __launch_bounds__(256) wide_enc_kernel(const WideEncArgs a) {
    const int p = blockIdx.y, t = blockIdx.x, tid = threadIdx.x;
    const nfopp_onf_layout &L = a.d.lay;
    const float *W = a.weights + (size_t)p * L.n_params;
    const int cnt = a.count ? a.count[p] : a.n;
    __shared__ float su0[64], su1[64], sth[64];
    if (tid < 64) {
        const int idx = t * 64 + tid;
        float x0 = 0.f, x1 = 0.f, x2 = 0.f;
        if (idx < cnt) {
            const float *xp = a.x + ((size_t)p * a.n + idx) * a.x_stride;
            x0 = xp[0]; x1 = xp[1];
            if (L.in_dim == 3) x2 = xp[2];
        }
        su0[tid] = (x0 - a.d.mean) / a.d.sigma;
        su1[tid] = (x1 - a.d.mean) / a.d.sigma;
        sth[tid] = x2;
    }
    __syncthreads();
    const size_t tile = (size_t)p * a.T64 + t;
    unsigned char *Et = a.E + tile * a.K1 * 256;
    const int nco = a.K1 / 8;
    for (int item = tid; item < nco * 64; item += 256) {
        const int co = item >> 6, pt = item & 63;
        const bool valid = t * 64 + pt < cnt;
        FeatParams fp;
        load_feat_params(W, a.d, co, fp);
        const float u0 = su0[pt], u1 = su1[pt], th = sth[pt];
        float e[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int f = co * 8 + i;
            const float arg = fmaf(f < L.f_pos ? u1 : th, fp.pb[i], fmaf(u0, fp.pa[i], fp.pc[i]));
            e[i] = (valid && f < L.f_in) ? __sinf(w_reduce_2pi(arg)) : 0.f;
        }
        uint4 hi, lo;
        split8(e, hi, lo);
        *reinterpret_cast<uint4 *>(Et + co * 1024 + pt * 16) = hi;
        *reinterpret_cast<uint4 *>(Et + a.K1 * 128 + co * 1024 + pt * 16) = lo;
    }
    if (a.U16 && tid < 128) {
        const int co = tid >> 6, pt = tid & 63;
        const bool valid = t * 64 + pt < cnt;
        float e[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (co == 0 && valid) { e[0] = su0[pt]; e[1] = su1[pt]; e[2] = sth[pt]; e[3] = 1.f; }
        uint4 hi, lo;
        split8(e, hi, lo);
        unsigned char *Ut = a.U16 + tile * 16 * 256;
        *reinterpret_cast<uint4 *>(Ut + co * 1024 + pt * 16) = hi;
        *reinterpret_cast<uint4 *>(Ut + 16 * 128 + co * 1024 + pt * 16) = lo;
    }
}

// ---- fwd / dgrad GEMM ------------------------------------------------------------------------------------------
struct WideGemmArgs {
    const unsigned char *A;       // activations in: [P][T64][CA*256]
    int CA;
    const unsigned char *Wpk;     // packed weights of this layer (problem 0); problems are w_stride bytes apart
    size_t w_stride;
    int K;                        // reduction length (multiple of 16)
    int Nfull;                    // output width (multiple of 16) = column count of Out
    unsigned char *Out;
    const float *weights;         // fp32 blobs [P][n_params]
    int n_params;
    int bias_off, n_valid;        // fwd: bias offset in the blob, real output width
    const unsigned char *Mask;    // dgrad: H_{l-1} tiles (ReLU mask) or null
    const float *gbar;            // dgrad of layer 0: + gbar[pt] * W[skip_off + col], col < skip_n   (skip connection)
    int skip_off, skip_n;
    int P, T64;
};

__device__ __forceinline__ void wmma3(uint32_t tmem_d, uint64_t ah, uint64_t al, uint64_t bh, uint64_t bl, uint32_t idesc,
                                      bool accumulate, uint32_t issue) {
    mma_bf16(tmem_d, ah, bh, idesc, accumulate ? 1u : 0u, issue);
    mma_bf16(tmem_d, al, bh, idesc, 1u, issue);
    mma_bf16(tmem_d, ah, bl, idesc, 1u, issue);
}

// MODE 0: forward layer, 1: input-gradient layer
template <int MODE>
__global__ void __launch_bounds__(WG_THREADS, 1) wide_gemm_kernel(const WideGemmArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + WG_STAGES * WG_STAGE);
    uint64_t *full = bars, *empty = bars + WG_STAGES, *tfull = bars + 2 * WG_STAGES, *tempty = tfull + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);
    float *svec = reinterpret_cast<float *>(smem + WG_STAGES * WG_STAGE + 256);    // [2][WG_NT] bias (fwd) / skip weights (dgrad)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int NTL = (a.Nfull + WG_NT - 1) / WG_NT, RB = a.T64 / 2;
    const int items = a.P * RB * NTL;
    const int nstages = (a.K + WG_KC - 1) / WG_KC;
    if (threadIdx.x == 0) {
        for (int s = 0; s < WG_STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(tfull + b, 1); mbar_init(tempty + b, 4); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ---------------- producer: bulk copies of the A (activation) and B (weight) K-chunks ----------------
        uint32_t it = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x) {
            const int nt = item % NTL, rb = (item / NTL) % RB, p = item / (NTL * RB);
            const int NTw = min(WG_NT, a.Nfull - nt * WG_NT);
            const unsigned char *At = a.A + ((size_t)p * a.T64 + 2 * rb) * a.CA * 256;
            const unsigned char *Wp = a.Wpk + (size_t)p * a.w_stride + (size_t)nt * WG_NT * a.K * 4;
            for (int st = 0; st < nstages; ++st, ++it) {
                const int s = it % WG_STAGES;
                mbar_wait(empty + s, ((it / WG_STAGES) & 1) ^ 1);
                const int k0 = st * WG_KC, kc = min(WG_KC, a.K - k0), nj = kc / 8;
                unsigned char *As = smem + s * WG_STAGE, *Bs = As + WG_A_STAGE;
                if (lane == 0) mbar_expect_tx(full + s, (uint32_t)(128 * kc * 4 + NTw * kc * 4));
                __syncwarp();
                if (lane < 4 * nj) {
                    const int j = lane % nj, pl = (lane / nj) & 1, t = lane / (2 * nj);
                    bulk_g2s(As + pl * (WG_A_STAGE / 2) + j * 2048 + t * 1024,
                             At + (size_t)t * a.CA * 256 + (size_t)pl * a.CA * 128 + (size_t)(k0 / 8 + j) * 1024, 1024, full + s);
                } else if (lane >= 16 && lane < 18) {
                    const int pl = lane - 16;
                    bulk_g2s(Bs + pl * (WG_B_STAGE / 2), Wp + (size_t)pl * a.K * NTw * 2 + (size_t)(k0 / 8) * NTw * 16,
                             (uint32_t)(kc * NTw * 2), full + s);
                }
            }
        }
    } else if (warp == 1) {
        // ---------------- MMA issue ----------------
        const uint32_t lead = elect_one();
        uint32_t it = 0, tile_it = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x, ++tile_it) {
            const int nt = item % NTL;
            const int NTw = min(WG_NT, a.Nfull - nt * WG_NT);
            const uint32_t buf = tile_it & 1;
            mbar_wait(tempty + buf, ((tile_it >> 1) & 1) ^ 1);
            fence_after_sync();
            const uint32_t d_tmem = tmem + buf * 256;
            const uint32_t idesc = instr_desc_bf16(128, NTw, 0, MODE == 1 ? 1 : 0);
            const uint32_t bstep = 2u * NTw * 16u;
            for (int st = 0; st < nstages; ++st, ++it) {
                const int s = it % WG_STAGES;
                mbar_wait(full + s, (it / WG_STAGES) & 1);
                fence_after_sync();
                const int kc = min(WG_KC, a.K - st * WG_KC);
                const uint32_t As = smem_u32(smem + s * WG_STAGE), Bs = As + WG_A_STAGE;
                for (int ks = 0; ks < kc / 16; ++ks) {
                    const uint64_t ah = desc_kmajor(As + ks * 4096, 128, 2048);
                    const uint64_t al = desc_kmajor(As + WG_A_STAGE / 2 + ks * 4096, 128, 2048);
                    uint64_t bh, bl;
                    if (MODE == 0) {
                        bh = desc_kmajor(Bs + ks * bstep, 128, NTw * 16);
                        bl = desc_kmajor(Bs + WG_B_STAGE / 2 + ks * bstep, 128, NTw * 16);
                    } else {
                        bh = desc_mnmajor(Bs + ks * bstep, NTw * 16, 128);
                        bl = desc_mnmajor(Bs + WG_B_STAGE / 2 + ks * bstep, NTw * 16, 128);
                    }
                    wmma3(d_tmem, ah, al, bh, bl, idesc, st > 0 || ks > 0, lead);
                }
                mma_commit(empty + s, lead);
            }
            mma_commit(tfull + buf, lead);
        }
    } else {
        // ---------------- epilogue: TMEM -> bias/ReLU or mask -> bf16 hi/lo operand tiles in global memory ----------------
        const int q = warp & 3;
        uint32_t tile_it = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x, ++tile_it) {
            const int nt = item % NTL, rb = (item / NTL) % RB, p = item / (NTL * RB);
            const int NTw = min(WG_NT, a.Nfull - nt * WG_NT);
            const uint32_t buf = tile_it & 1;
            const int row = q * 32 + lane, r = row & 63;
            const size_t tile = (size_t)p * a.T64 + 2 * rb + (row >> 6);
            unsigned char *Ot = a.Out + tile * a.Nfull * 256;
            const unsigned char *Mt = (MODE == 1 && a.Mask) ? a.Mask + tile * a.Nfull * 256 : nullptr;
            const float *Wb = a.weights + (size_t)p * a.n_params;
            const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + buf * 256;
            // per-tile vector (bias of the forward layer / skip-connection weights of the first layer's input gradient),
            // zero beyond the real width: staged once per tile instead of one global load per element
            float *sv = svec + buf * WG_NT;
            const bool use_vec = MODE == 0 || a.gbar != nullptr;
            if (use_vec) {
                const int et = threadIdx.x - 64, voff = MODE == 0 ? a.bias_off : a.skip_off, vn = MODE == 0 ? a.n_valid : a.skip_n;
                for (int c = et; c < NTw; c += 128) {
                    const int col = nt * WG_NT + c;
                    sv[c] = col < vn ? __ldg(Wb + voff + col) : 0.f;
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
            }
            float gb = 0.f;
            if (MODE == 1 && a.gbar) gb = a.gbar[((size_t)p * a.T64 + 2 * rb) * 64 + row];
            mbar_wait(tfull + buf, (tile_it >> 1) & 1);
            fence_after_sync();
            for (int c0 = 0; c0 < NTw; c0 += 16) {
                float v[16];
                tmem_ld16(taddr + c0, v);
                const int col0 = nt * WG_NT + c0;
                float sb[16];
                if (use_vec) {
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4) *reinterpret_cast<float4 *>(sb + 4 * i4) = *reinterpret_cast<const float4 *>(sv + c0 + 4 * i4);
                }
                if (MODE == 0) {
                    // padded columns: the weights are zero there, so D = 0 and relu(0 + 0) = 0
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i] + sb[i], 0.f);
                } else {
                    if (a.gbar) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] = fmaf(gb, sb[i], v[i]);
                    }
                    if (Mt) {
                        const uint4 m0 = *reinterpret_cast<const uint4 *>(Mt + (col0 / 8) * 1024 + r * 16);
                        const uint4 m1 = *reinterpret_cast<const uint4 *>(Mt + (col0 / 8 + 1) * 1024 + r * 16);
                        const uint32_t mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const uint32_t bits = (mw[i >> 1] >> ((i & 1) * 16)) & 0x7FFFu;
                            v[i] = bits != 0 ? v[i] : 0.f;
                        }
                    }
                }
                uint4 h0, l0, h1, l1;
                split8(v, h0, l0);
                split8(v + 8, h1, l1);
                unsigned char *o = Ot + (col0 / 8) * 1024 + r * 16;
                *reinterpret_cast<uint4 *>(o) = h0;
                *reinterpret_cast<uint4 *>(o + 1024) = h1;
                *reinterpret_cast<uint4 *>(o + a.Nfull * 128) = l0;
                *reinterpret_cast<uint4 *>(o + a.Nfull * 128 + 1024) = l1;
            }
            fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty + buf);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---- weight-gradient GEMM: dst[m, j] (+)= sum_pt A[pt, m] B[pt, j] ------------------------------------------------
struct WideWgradArgs {
    const unsigned char *A;
    int CA;                      // columns of A (M, walked in blocks of 128)
    const unsigned char *B;
    int CB;                      // columns of B (N, walked in tiles of 256)
    float *dst;                  // [P] x dst_pstride floats
    size_t dst_pstride;
    int dst_off, dst_ld, row_lo, rows_valid, cols_valid;   // rows row_lo <= m < rows_valid, columns j < cols_valid are stored
    int bias_off;                // >= 0: one more N tile with B = ones -> dst[bias_off + m] = sum_pt A[pt, m]
    int P, T64, splits;          // T64 = operand tiles per problem, TP points each
    int TP;
    // splits > 1: split s of the point range stores its partial sums at partial + s * partial_sstride (same offsets as
    // dst); wide_reduce_kernel adds the splits in fixed order afterwards - deterministic, no atomics
    float *partial;
    size_t partial_sstride;
};

__global__ void __launch_bounds__(WG_THREADS, 1) wide_wgrad_kernel(const WideWgradArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char *ones = smem + WW_STAGES * WW_STAGE;
    uint64_t *bars = reinterpret_cast<uint64_t *>(ones + WW_ONES);
    uint64_t *full = bars, *empty = bars + WW_STAGES, *tfull = bars + 2 * WW_STAGES, *tempty = tfull + 1;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int MB = (a.CA + 127) / 128, NTL = (a.CB + WG_NT - 1) / WG_NT, NTT = NTL + (a.bias_off >= 0 ? 1 : 0);
    const int TP = a.TP;                                      // points per operand tile (64 per-layer path, 32 chain path)
    const int tps = (a.T64 + a.splits - 1) / a.splits;        // tiles per split
    const int items = a.P * MB * NTT * a.splits;
    for (int i = threadIdx.x; i < WW_ONES / 4; i += WG_THREADS) reinterpret_cast<uint32_t *>(ones)[i] = 0x3F803F80u;
    if (threadIdx.x == 0) {
        for (int s = 0; s < WW_STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
        mbar_init(tfull, 1);
        mbar_init(tempty, 4);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 256);
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    // item -> (p, mb, nt, split); split fastest so the splits of one output tile run concurrently on different SMs
    auto decode = [&](int item, int &p, int &mb, int &nt, int &t0, int &t1, int *split = nullptr) {
        const int sp = item % a.splits;
        if (split) *split = sp;
        nt = (item / a.splits) % NTT;
        mb = (item / (a.splits * NTT)) % MB;
        p = item / (a.splits * NTT * MB);
        t0 = sp * tps;
        t1 = min(a.T64, t0 + tps);
    };

    if (warp == 0) {
        uint32_t it = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x) {
            int p, mb, nt, t0, t1;
            decode(item, p, mb, nt, t0, t1);
            const bool bias_tile = nt == NTL;
            const int MA = min(128, a.CA - 128 * mb), NB = bias_tile ? 0 : min(WG_NT, a.CB - nt * WG_NT);
            for (int t = t0; t < t1; ++t, ++it) {
                const int s = it % WW_STAGES;
                mbar_wait(empty + s, ((it / WW_STAGES) & 1) ^ 1);
                unsigned char *As = smem + s * WW_STAGE, *Bs = As + WW_A_STAGE;
                const size_t tile = (size_t)p * a.T64 + t;
                if (lane == 0) mbar_expect_tx(full + s, (uint32_t)((MA + NB) * TP * 4));
                __syncwarp();
                if (lane < 2) {
                    bulk_g2s(As + lane * (WW_A_STAGE / 2), a.A + tile * a.CA * TP * 4 + (size_t)lane * a.CA * TP * 2 + (size_t)mb * 16 * TP * 16,
                             (uint32_t)(MA * TP * 2), full + s);
                } else if (lane < 4 && NB > 0) {
                    const int pl = lane - 2;
                    bulk_g2s(Bs + pl * (WW_B_STAGE / 2), a.B + tile * a.CB * TP * 4 + (size_t)pl * a.CB * TP * 2 + (size_t)nt * 32 * TP * 16,
                             (uint32_t)(NB * TP * 2), full + s);
                }
            }
        }
    } else if (warp == 1) {
        const uint32_t lead = elect_one();
        uint32_t it = 0, tile_it = 0;
        const uint32_t ones_u = smem_u32(ones);
        for (int item = blockIdx.x; item < items; item += gridDim.x, ++tile_it) {
            int p, mb, nt, t0, t1;
            decode(item, p, mb, nt, t0, t1);
            const bool bias_tile = nt == NTL;
            const int NB = bias_tile ? 16 : min(WG_NT, a.CB - nt * WG_NT);
            mbar_wait(tempty, (tile_it & 1) ^ 1);
            fence_after_sync();
            const uint32_t idesc = instr_desc_bf16(128, NB, 1, 1);
            bool started = false;
            for (int t = t0; t < t1; ++t, ++it) {
                const int s = it % WW_STAGES;
                mbar_wait(full + s, (it / WW_STAGES) & 1);
                fence_after_sync();
                const uint32_t As = smem_u32(smem + s * WW_STAGE), Bs = As + WW_A_STAGE;
                const uint32_t scol = (uint32_t)a.TP * 16u;
                for (int ks = 0; ks < a.TP / 16; ++ks) {
                    const uint64_t ah = desc_mnmajor(As + ks * 256, 128, scol);
                    const uint64_t al = desc_mnmajor(As + WW_A_STAGE / 2 + ks * 256, 128, scol);
                    if (bias_tile) {
                        const uint64_t bo = desc_mnmajor(ones_u + ks * 256, 128, scol);
                        mma_bf16(tmem, ah, bo, idesc, started ? 1u : 0u, lead);
                        mma_bf16(tmem, al, bo, idesc, 1u, lead);
                    } else {
                        const uint64_t bh = desc_mnmajor(Bs + ks * 256, 128, scol);
                        const uint64_t bl = desc_mnmajor(Bs + WW_B_STAGE / 2 + ks * 256, 128, scol);
                        wmma3(tmem, ah, al, bh, bl, idesc, started, lead);
                    }
                    started = true;
                }
                mma_commit(empty + s, lead);
            }
            mma_commit(tfull, lead);
        }
    } else {
        const int q = warp & 3;
        uint32_t tile_it = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x, ++tile_it) {
            int p, mb, nt, t0, t1, sp;
            decode(item, p, mb, nt, t0, t1, &sp);
            const bool bias_tile = nt == NTL;
            const int NB = bias_tile ? 16 : min(WG_NT, a.CB - nt * WG_NT);
            mbar_wait(tfull, tile_it & 1);
            fence_after_sync();
            const int m = mb * 128 + q * 32 + lane;
            float *D = (a.splits > 1 ? a.partial + (size_t)sp * a.partial_sstride : a.dst) + (size_t)p * a.dst_pstride;
            const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16);
            const bool has_work = t1 > t0;          // an empty split leaves its (zeroed) partial untouched
            for (int c0 = 0; c0 < NB; c0 += 16) {
                float v[16];
                tmem_ld16(taddr + c0, v);
                if (m < a.row_lo || m >= a.rows_valid || !has_work) continue;
                if (bias_tile) {
                    if (c0 == 0) D[a.bias_off + m] = v[0];
                    continue;
                }
                const int j0 = nt * WG_NT + c0;
                float *drow = D + a.dst_off + (size_t)m * a.dst_ld + j0;
                if ((a.dst_ld & 3) == 0 && (a.dst_off & 3) == 0) {
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4)
                        if (j0 + 4 * i4 + 3 < a.cols_valid)
                            *reinterpret_cast<float4 *>(drow + 4 * i4) = make_float4(v[4 * i4], v[4 * i4 + 1], v[4 * i4 + 2], v[4 * i4 + 3]);
                        else
                            for (int i = 4 * i4; i < 4 * i4 + 4; ++i)
                                if (j0 + i < a.cols_valid) drow[i] = v[i];
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (j0 + i < a.cols_valid) drow[i] = v[i];
                }
            }
            fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 256);
}

// ---- logit, dL/dz and the gradient of the last hidden layer ------------------------------------------------------
struct WideOutArgs {
    OnfDims d;
    const float *weights;
    const int32_t *count;
    const unsigned char *HL, *E;
    int Hp, K1, n, T64;
    int want_bwd, gbar_mode;
    const float *gz, *aux;
    float wc, beta;
    float *z_out;
    float *gbar;                 // [P][T64*64]
    unsigned char *G16;          // dL/dz as a 16-column operand (column 0), or null
    unsigned char *GL;           // gradient w.r.t. the last hidden layer's output (post-ReLU mask applied)
};

__global__ void __launch_bounds__(256) wide_out_kernel(const WideOutArgs a) {
    const int p = blockIdx.y, t = blockIdx.x, tid = threadIdx.x, pt = tid & 63, part = tid >> 6;
    const nfopp_onf_layout &L = a.d.lay;
    const int H = a.d.H;
    const float *W = a.weights + (size_t)p * L.n_params;
    const float *w3 = W + L.off_w_out;
    const int cnt = a.count ? a.count[p] : a.n;
    const int idx = t * 64 + pt;
    const bool valid = idx < cnt;
    const size_t tile = (size_t)p * a.T64 + t;
    const unsigned char *Ht = a.HL + tile * a.Hp * 256, *Et = a.E + tile * a.K1 * 256;
    __shared__ float zs[4][64];
    __shared__ float gs[64];
    float acc = 0.f;
    for (int co = part; co < a.Hp / 8; co += 4) {
        const uint4 hi = *reinterpret_cast<const uint4 *>(Ht + co * 1024 + pt * 16);
        const uint4 lo = *reinterpret_cast<const uint4 *>(Ht + a.Hp * 128 + co * 1024 + pt * 16);
        float v[8];
        unpack8(hi, lo, v);
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (co * 8 + i < H) acc = fmaf(v[i], __ldg(w3 + co * 8 + i), acc);
    }
    for (int co = part; co < a.K1 / 8; co += 4) {
        const uint4 hi = *reinterpret_cast<const uint4 *>(Et + co * 1024 + pt * 16);
        const uint4 lo = *reinterpret_cast<const uint4 *>(Et + a.K1 * 128 + co * 1024 + pt * 16);
        float v[8];
        unpack8(hi, lo, v);
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (co * 8 + i < L.f_in) acc = fmaf(v[i], __ldg(w3 + H + co * 8 + i), acc);
    }
    zs[part][pt] = acc;
    __syncthreads();
    if (part == 0) {
        const float z = (zs[0][pt] + zs[1][pt]) + (zs[2][pt] + zs[3][pt]) + W[L.off_b_out];
        const size_t gi = (size_t)p * a.n + idx;
        if (valid && a.z_out) a.z_out[gi] = z;
        float gb = 0.f;
        if (valid && a.want_bwd) {                      // dL/dz, as in onf_tc.cu
            if (a.gbar_mode == 0) gb = a.gz[gi];
            else if (a.gbar_mode == 1) gb = (sigmoid_f(z) - a.aux[gi]) / (float)cnt;
            else {
                const float zb = z * a.beta;
                const float ez = expf(zb);
                gb = a.wc * (zb > 20.f ? 1.f : ez / (ez + 1.f));
                if (a.gbar_mode == 3) {
                    const float th = tanhf(z);
                    gb += a.aux[gi] * (1.f - th * th);
                }
            }
        }
        gs[pt] = gb;
        if (a.want_bwd) {
            a.gbar[tile * 64 + pt] = gb;
            if (a.G16) {
                float e[8] = {gb, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                uint4 hi, lo;
                split8(e, hi, lo);
                const uint4 zero = make_uint4(0, 0, 0, 0);
                unsigned char *Gt = a.G16 + tile * 16 * 256;
                *reinterpret_cast<uint4 *>(Gt + pt * 16) = hi;
                *reinterpret_cast<uint4 *>(Gt + 1024 + pt * 16) = zero;
                *reinterpret_cast<uint4 *>(Gt + 2048 + pt * 16) = lo;
                *reinterpret_cast<uint4 *>(Gt + 3072 + pt * 16) = zero;
            }
        }
    }
    if (!a.want_bwd) return;
    __syncthreads();
    unsigned char *Gt = a.GL + tile * a.Hp * 256;
    const float gb = gs[pt];
    for (int co = part; co < a.Hp / 8; co += 4) {
        const uint4 hi = *reinterpret_cast<const uint4 *>(Ht + co * 1024 + pt * 16);
        const uint32_t mw[4] = {hi.x, hi.y, hi.z, hi.w};
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int col = co * 8 + i;
            const uint32_t bits = (mw[i >> 1] >> ((i & 1) * 16)) & 0x7FFFu;
            v[i] = (col < H && bits != 0) ? gb * __ldg(w3 + col) : 0.f;
        }
        uint4 oh, ol;
        split8(v, oh, ol);
        *reinterpret_cast<uint4 *>(Gt + co * 1024 + pt * 16) = oh;
        *reinterpret_cast<uint4 *>(Gt + a.Hp * 128 + co * 1024 + pt * 16) = ol;
    }
}

// ---- encoding backward: dL/dE -> dL/d(arg) (the A operand of the encoder-gradient job) and dL/dx ------------------
struct WideEbwdArgs {
    OnfDims d;
    const float *weights, *x;
    const int32_t *count;
    unsigned char *GE;           // in: dL/dE tiles; out (write_g0): dL/d(arg) tiles, in place
    int K1, n, x_stride, T64, write_g0;
    float *gx_out;               // [P][n][x_stride] or null
};

__global__ void __launch_bounds__(256) wide_ebwd_kernel(const WideEbwdArgs a) {
    const int p = blockIdx.y, t = blockIdx.x, tid = threadIdx.x, pt = tid & 63, part = tid >> 6;
    const nfopp_onf_layout &L = a.d.lay;
    const float *W = a.weights + (size_t)p * L.n_params;
    const int cnt = a.count ? a.count[p] : a.n;
    const int idx = t * 64 + pt;
    const bool valid = idx < cnt;
    float x0 = 0.f, x1 = 0.f, x2 = 0.f;
    if (valid) {
        const float *xp = a.x + ((size_t)p * a.n + idx) * a.x_stride;
        x0 = xp[0]; x1 = xp[1];
        if (L.in_dim == 3) x2 = xp[2];
    }
    const float u0 = (x0 - a.d.mean) / a.d.sigma, u1 = (x1 - a.d.mean) / a.d.sigma, th = x2;
    unsigned char *Gt = a.GE + ((size_t)p * a.T64 + t) * a.K1 * 256;
    float g0 = 0.f, g1 = 0.f, g2 = 0.f;
    for (int co = part; co < a.K1 / 8; co += 4) {
        unsigned char *ph = Gt + co * 1024 + pt * 16, *pl = ph + a.K1 * 128;
        float v[8];
        unpack8(*reinterpret_cast<const uint4 *>(ph), *reinterpret_cast<const uint4 *>(pl), v);
        FeatParams fp;
        load_feat_params(W, a.d, co, fp);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int f = co * 8 + i;
            const bool fourier = f < L.f_pos;
            const float arg = fmaf(fourier ? u1 : th, fp.pb[i], fmaf(u0, fp.pa[i], fp.pc[i]));
            const float gp = (valid && f < L.f_in) ? v[i] * __cosf(w_reduce_2pi(arg)) : 0.f;
            v[i] = gp;
            if (fourier) { g0 = fmaf(gp, fp.pa[i], g0); g1 = fmaf(gp, fp.pb[i], g1); }
            else g2 = fmaf(gp, fp.pb[i], g2);
        }
        if (a.write_g0) {
            uint4 hi, lo;
            split8(v, hi, lo);
            *reinterpret_cast<uint4 *>(ph) = hi;
            *reinterpret_cast<uint4 *>(pl) = lo;
        }
    }
    if (!a.gx_out) return;
    __shared__ float red[3][4][64];
    red[0][part][pt] = g0; red[1][part][pt] = g1; red[2][part][pt] = g2;
    __syncthreads();
    if (part == 0 && valid) {
        float s[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) s[c] = (red[c][0][pt] + red[c][1][pt]) + (red[c][2][pt] + red[c][3][pt]);
        float *out = a.gx_out + ((size_t)p * a.n + idx) * a.x_stride;
        out[0] = s[0] / a.d.sigma;
        out[1] = s[1] / a.d.sigma;
        if (L.in_dim == 3) out[2] = s[2];
        for (int c = L.in_dim; c < a.x_stride; ++c) out[c] = 0.f;
    }
}

// ---- encoder gradients: raw[f][0..3] = sum_pt gp[pt,f] * [u0, u1, theta, 1] -> blob order ---------------------------
struct WideAsmArgs {
    OnfDims d;
    const float *weights, *raw;   // raw [P][K1][16]
    float *grad;
    int K1;
};

__global__ void __launch_bounds__(256) wide_asm_kernel(const WideAsmArgs a) {
    const int p = blockIdx.y, f = blockIdx.x * blockDim.x + threadIdx.x;
    const nfopp_onf_layout &L = a.d.lay;
    if (f >= L.f_in) return;
    const float *r = a.raw + ((size_t)p * a.K1 + f) * 16;
    const float *W = a.weights + (size_t)p * L.n_params;
    float *G = a.grad + (size_t)p * L.n_params;
    if (f < L.f_pos) {
        G[L.off_enc_w + 2 * f] = r[0];
        G[L.off_enc_w + 2 * f + 1] = r[1];
        G[L.off_enc_b + f] = a.d.bias ? r[3] : 0.f;
    } else {
        const int k = f - L.f_pos;
        G[L.off_ang_b + k] = W[L.off_ang_f + k] * r[3];                  // d/d b_k = f_k * sum gp
        G[L.off_ang_f + k] = fmaf(W[L.off_ang_b + k], r[3], r[2]);       // d/d f_k = sum gp * (theta + b_k)   (trained: SURVEY 8 a3)
    }
}

// ======================================================================================================================
// Fused chain (hidden <= 256): encode -> L forward layers -> logit, dL/dz -> L input-gradient layers -> encoding backward
// in ONE persistent kernel.  A 128-point row block stays in shared memory as the bf16 hi/lo A operand of every GEMM
// (written in place by the epilogue of the previous one); only the weights stream (bulk copies, 5-stage ring of K=16
// chunks).  Activations and gradients are still written to global memory once - they are the operands of the
// weight-gradient GEMMs - but never read back by the chain (except the ReLU masks), which halves the HBM traffic of the
// per-layer path.  Operand tiles here hold 32 points: [P][n/32][hi|lo][C/8][32 points][8] (one wgrad stage = one tile).
// ======================================================================================================================
constexpr int WC_PARTS = 4;              // epilogue warps per TMEM lane quarter (each takes 1/WC_PARTS of the columns)
constexpr int WC_EPI = 128 * WC_PARTS;
constexpr int WC_THREADS = 64 + WC_EPI;  // warp 0 producer, warp 1 MMA issue + TMEM, then the epilogue warps
// device-side bounds asserts of the ring / operand-tile extents (make -C csrc debug)
#ifdef NFOPP_DEBUG
#define WIDE_ASSERT(cond)                                                                                                 \
    do {                                                                                                                \
        if (!(cond)) {                                                                                                  \
            printf("onf_wide.cu:%d: assertion failed: %s (block %d thread %d)\n", __LINE__, #cond, (int)blockIdx.x, (int)threadIdx.x); \
            __trap();                                                                                                   \
        }                                                                                                               \
    } while (0)
#else
#define WIDE_ASSERT(cond) do { } while (0)
#endif
constexpr int WC_STAGES = 5;
constexpr int WC_STAGE = 16 * WG_NT * 4;             // K = 16 x N <= 256 x (hi + lo)
constexpr int WC_ACT = 128 * 256 * 4;                // 128 points x K <= 256 x (hi + lo)
constexpr int WC_ACT_PLANE = WC_ACT / 2;
constexpr int WC_MISC_FLOATS = 5 * 256 + WC_PARTS * 128 + 3 * WC_PARTS * 128;   // pa pb pc svecA svecB | zpart | red
constexpr int WC_SMEM = WC_ACT + WC_STAGES * WC_STAGE + 256 + WC_MISC_FLOATS * 4;

struct WideChainArgs {
    OnfDims d;
    WideDims w;
    const float *weights;
    const unsigned char *packed;
    size_t w_stride;
    const float *x;
    const int32_t *count;
    int n, x_stride, P, T32, RB, CL;
    int want_bwd, want_wgrad, want_gx, gbar_mode;
    const float *gz, *aux;
    float wc, beta;
    float *z_out, *gx_out;
    unsigned char *E, *U16, *H[NFOPP_MAX_LAYERS], *G[NFOPP_MAX_LAYERS], *GE, *G16;
    // per (problem, row block, lane quarter) partial sums over the quarter's 32 points of everything that is a plain
    // reduction over points: bias gradients, output-layer gradient, encoder gradients (summed in fixed order afterwards)
    float *auxp;
};

// layout of one aux vector (floats): db[L][Hp] | dw_out_h[Hp] | dw_out_e[K1] | enc0[K1] | enc1[K1] | enc2[K1] | db_out[16]
__host__ __device__ inline int aux_db(const WideDims &w, int l) { return l * w.Hp; }
__host__ __device__ inline int aux_dw3h(const WideDims &w) { return w.L * w.Hp; }
__host__ __device__ inline int aux_dw3e(const WideDims &w) { return (w.L + 1) * w.Hp; }
__host__ __device__ inline int aux_enc(const WideDims &w, int k) { return (w.L + 1) * w.Hp + (1 + k) * w.K1; }
__host__ __device__ inline int aux_db3(const WideDims &w) { return (w.L + 1) * w.Hp + 4 * w.K1; }
__host__ __device__ inline int aux_len(const WideDims &w) { return (w.L + 1) * w.Hp + 4 * w.K1 + 16; }

// column sums over the 32 lanes of a warp for 16 columns: every lane returns sum_lanes v[lane >> 1] (v is clobbered)
__device__ __forceinline__ float wcolsum16(float (&v)[16], int lane) {
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

// optional phase trace of the chain kernel (build with -DNFOPP_TC_TRACE: `make trace`; scripts/wide_trace.py)
#ifdef NFOPP_TC_TRACE
__device__ long long g_wc_trace[96];
#define WC_TRACE(cond, slot) do { if ((cond) && blockIdx.x == 0 && (slot) < 96) g_wc_trace[(slot)] = clock64(); } while (0)
#else
#define WC_TRACE(cond, slot) do { } while (0)
#endif

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(WC_EPI) : "memory"); }

// phase ph of an item: forward layer ph (ph < L) or input-gradient layer 2L-1-ph
__device__ __forceinline__ void chain_phase(const WideDims &w, int ph, bool &dg, int &l, int &K, int &N) {
    dg = ph >= w.L;
    l = dg ? 2 * w.L - 1 - ph : ph;
    K = dg ? w.Hp : w.kin[l];
    N = dg ? w.kin[l] : w.Hp;
}

__global__ void __launch_bounds__(WC_THREADS, 1) wide_chain_kernel(const WideChainArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char *act = smem, *ring = smem + WC_ACT;
    uint64_t *bars = reinterpret_cast<uint64_t *>(ring + WC_STAGES * WC_STAGE);
    uint64_t *full = bars, *empty = bars + WC_STAGES, *tfull = bars + 2 * WC_STAGES, *act_ready = tfull + 1;   // act_ready[4]: one per 64-column group
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(act_ready + 4);
    float *fm = reinterpret_cast<float *>(ring + WC_STAGES * WC_STAGE + 256);
    float *s_pa = fm, *s_pb = fm + 256, *s_pc = fm + 512, *svecA = fm + 768, *svecB = fm + 1024, *zpart = fm + 1280, *red = fm + 1280 + WC_PARTS * 128;
    const WideDims &w = a.w;
    const nfopp_onf_layout &L = a.d.lay;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // clusters of CL CTAs walk the same problem (CL consecutive row blocks): every weight chunk is read from L2 once per
    // cluster and multicast into the shared memory of all its CTAs (each CTA issues 1/CL of every stage)
    const int CL = a.CL, rank = CL > 1 ? (int)cluster_ctarank() : 0;
    const uint16_t cmask = (uint16_t)((1u << CL) - 1u);
    const int RBG = (a.RB + CL - 1) / CL;              // row-block groups per problem
    const int items = a.P * RBG, cid = blockIdx.x / CL, ncl = gridDim.x / CL;
    const bool bwd = a.want_bwd != 0;
    const int nph = bwd ? ((a.want_gx || a.want_wgrad) ? 2 * w.L : 2 * w.L - 1) : w.L;
    if (threadIdx.x == 0) {
        for (int s = 0; s < WC_STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, CL); }
        mbar_init(tfull, 1);
        for (int g = 0; g < 4; ++g) mbar_init(act_ready + g, WC_EPI / 32);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);         // two 256-column accumulators, alternating by phase
    fence_before_sync();
    __syncthreads();
    if (CL > 1) cluster_sync_all();                    // every CTA's barriers exist before anyone multicasts into them
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ---------------- producer: the weight stream ----------------
        uint32_t it = 0;
        for (int item = cid; item < items; item += ncl) {
            const int p = item / RBG;
            const unsigned char *Wp = a.packed + (size_t)p * a.w_stride;
            for (int ph = 0; ph < nph; ++ph) {
                bool dg; int l, K, N;
                chain_phase(w, ph, dg, l, K, N);
                const unsigned char *Wl = Wp + (dg ? w.wd_off[l] : w.wf_off[l]);
                for (int st = 0; st < K / 16; ++st, ++it) {
                    const int s = it % WC_STAGES;
                    WIDE_ASSERT(N * 32 <= WC_STAGE / 2 && N % 16 == 0 && K % 16 == 0 && (size_t)(dg ? w.wd_off[l] : w.wf_off[l]) + (size_t)2 * K * N * 2 <= a.w_stride);
                    mbar_wait(empty + s, ((it / WC_STAGES) & 1) ^ 1);
                    if (lane == 0) mbar_expect_tx(full + s, (uint32_t)(N * 64));
                    __syncwarp();
                    if (CL == 1) {
                        if (lane < 2)
                            bulk_g2s(ring + s * WC_STAGE + lane * (WC_STAGE / 2), Wl + (size_t)lane * K * N * 2 + (size_t)st * N * 32,
                                     (uint32_t)(N * 32), full + s);
                    } else if (lane == 0) {
                        // this CTA's slice of the stage: plane = slice / (CL/2), part of the plane's chunk = slice % (CL/2)
                        const int per = CL / 2, pl = rank / per, part = rank - pl * per, bytes = N * 32 / per;
                        bulk_g2s_multicast(ring + s * WC_STAGE + pl * (WC_STAGE / 2) + part * bytes,
                                           Wl + (size_t)pl * K * N * 2 + (size_t)st * N * 32 + (size_t)part * bytes, (uint32_t)bytes, full + s,
                                           cmask);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ---------------- MMA issue ----------------
        const uint32_t lead = elect_one();
        const uint32_t act_u = smem_u32(act), ring_u = smem_u32(ring);
        uint32_t it = 0, gph = 0, apar = 0;                 // apar: phase parity bit of each act_ready barrier
        for (int item = cid; item < items; item += ncl) {
            for (int ph = 0; ph < nph; ++ph, ++gph) {
                bool dg; int l, K, N;
                chain_phase(w, ph, dg, l, K, N);
                const uint32_t idesc = instr_desc_bf16(128, N, 0, dg ? 1 : 0);
                const uint32_t acc = tmem + (gph & 1) * 256;   // the epilogue may still be reading the other accumulator
                for (int st = 0; st < K / 16; ++st, ++it) {
                    if ((st & 3) == 0) {                    // the 64 A-operand columns of this K group have been written
                        const int g = st >> 2;
                        mbar_wait(act_ready + g, (apar >> g) & 1);
                        apar ^= 1u << g;
                        fence_after_sync();
                        WC_TRACE(lane == 0 && item == cid + 2 * ncl && g == 0, 64 + 2 * ph);
                    }
                    const int s = it % WC_STAGES;
                    mbar_wait(full + s, (it / WC_STAGES) & 1);
                    fence_after_sync();
                    const uint32_t Bs = ring_u + s * WC_STAGE;
                    const uint64_t ah = desc_kmajor(act_u + st * 4096, 128, 2048);
                    const uint64_t al = desc_kmajor(act_u + WC_ACT_PLANE + st * 4096, 128, 2048);
                    const uint64_t bh = dg ? desc_mnmajor(Bs, N * 16, 128) : desc_kmajor(Bs, 128, N * 16);
                    const uint64_t bl = dg ? desc_mnmajor(Bs + WC_STAGE / 2, N * 16, 128) : desc_kmajor(Bs + WC_STAGE / 2, 128, N * 16);
                    wmma3(acc, ah, al, bh, bl, idesc, st > 0, lead);
                    if (CL == 1) mma_commit(empty + s, lead);
                    else mma_commit_multicast(empty + s, cmask, lead);
                }
                mma_commit(tfull, lead);
                WC_TRACE(lane == 0 && item == cid + 2 * ncl, 65 + 2 * ph);
            }
        }
    } else {
        // ---------------- epilogue warps: encoding, bias/ReLU, logit, dL/dz, masks, encoding backward ----------------
        const int et = threadIdx.x - 64, q = warp & 3, part = (warp - 2) >> 2, row = q * 32 + lane, r32 = row & 31;
        const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16);
        const bool wgrad = a.want_wgrad != 0;
        // the epilogue warps sweep the columns in order (chunk = 16 columns, group = 4 chunks, one chunk per `part`): as soon
        // as a 64-column group of the next A operand is in shared memory the MMA warp starts the next phase's K steps on it
        auto signal_group = [&](int g) {
            fence_proxy_async();
            fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(act_ready + g);
        };
        // 16 columns (col0..col0+15) of this thread's point: bf16 hi/lo split, A operand in shared memory, 32-point tile in
        // global memory.  The global stores are issued AFTER the group has been signalled (the proxy fence of the signal
        // otherwise waits for them too)
        auto split16 = [&](const float *v, uint4 (&s)[4]) {
            split8(v, s[0], s[1]);
            split8(v + 8, s[2], s[3]);
        };
        auto put_act = [&](const uint4 (&s)[4], int col0) {
            unsigned char *o = act + (col0 / 8) * 2048 + row * 16;
            *reinterpret_cast<uint4 *>(o) = s[0];
            *reinterpret_cast<uint4 *>(o + 2048) = s[2];
            *reinterpret_cast<uint4 *>(o + WC_ACT_PLANE) = s[1];
            *reinterpret_cast<uint4 *>(o + WC_ACT_PLANE + 2048) = s[3];
        };
        auto put_tile = [&](const uint4 (&s)[4], int col0, unsigned char *gt, int C) {
            unsigned char *o = gt + (col0 / 8) * 512 + r32 * 16;
            *reinterpret_cast<uint4 *>(o) = s[0];
            *reinterpret_cast<uint4 *>(o + 512) = s[2];
            *reinterpret_cast<uint4 *>(o + C * 64) = s[1];
            *reinterpret_cast<uint4 *>(o + C * 64 + 512) = s[3];
        };
        uint32_t gph = 0;
        int cur_p = -1;
        for (int item = cid; item < items; item += ncl) {
            const int p = item / RBG, rb = (item - p * RBG) * CL + rank;
            const bool live = rb < a.RB;                   // a cluster's last group may hold fewer than CL row blocks: the spare
                                                           // CTAs keep the weight-stream protocol going on dummy (all-invalid) rows
            const float *W = a.weights + (size_t)p * L.n_params;
            const int cnt = a.count ? a.count[p] : a.n;
            const int idx = rb * 128 + row;
            const bool valid = live && idx < cnt;
            const size_t t32 = live ? (size_t)p * a.T32 + rb * 4 + (row >> 5) : 0;
            float *auxv = (wgrad && live) ? a.auxp + (((size_t)p * a.RB + rb) * 4 + q) * aux_len(w) : nullptr;
            // column sums of 16 values per lane over the warp's 32 points -> aux[off + col]
            auto put_colsum = [&](float (&t)[16], int off) {
                const float sum = wcolsum16(t, lane);
                if (auxv && (lane & 1) == 0) auxv[off + (lane >> 1)] = sum;
            };
            // ---- per-problem feature parameters, skip-connection weights ----
            epi_bar();                                      // everyone is done with the previous item's vectors
            if (p != cur_p) {
                for (int f = et; f < w.K1; f += WC_EPI) {
                    float pa = 0.f, pb = 0.f, pc = 0.f;
                    if (f < L.f_pos) {
                        pa = W[L.off_enc_w + 2 * f];
                        pb = W[L.off_enc_w + 2 * f + 1];
                        pc = W[L.off_enc_b + f] + ((a.d.use_cos && f >= 100) ? 1.57079632679489662f : 0.f);
                    } else if (f < L.f_in) {
                        const int k = f - L.f_pos;
                        const float b = W[L.off_ang_b + k], fr = W[L.off_ang_f + k];
                        pb = fr;
                        pc = fmaf(b, fr, k >= L.n_angle / 2 ? 1.57079632679489662f : 0.f);
                    }
                    s_pa[f] = pa; s_pb[f] = pb; s_pc[f] = pc;
                }
                cur_p = p;
            }
            for (int f = et; f < w.K1; f += WC_EPI) svecA[f] = f < L.f_in ? W[L.off_w_out + w.H + f] : 0.f;
            epi_bar();
            float x0 = 0.f, x1 = 0.f, x2 = 0.f;
            if (valid) {
                const float *xp = a.x + ((size_t)p * a.n + idx) * a.x_stride;
                x0 = xp[0]; x1 = xp[1];
                if (L.in_dim == 3) x2 = xp[2];
            }
            const float u0 = (x0 - a.d.mean) / a.d.sigma, u1 = (x1 - a.d.mean) / a.d.sigma, th = x2;
            // ---- encoding (onf_model.py:37-46) -> A operand of layer 0, skip term of the logit ----
            float zp = 0.f;
            {
                const int nch = w.K1 / 16;
                unsigned char *Et = (bwd && wgrad && live) ? a.E + t32 * w.K1 * 128 : nullptr;
                for (int g = 0; 4 * g < nch; ++g) {
                    const int ch = 4 * g + part;
                    if (ch < nch) {
                        float e[16];
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const int f = ch * 16 + i;
                            const float arg = fmaf(f < L.f_pos ? u1 : th, s_pb[f], fmaf(u0, s_pa[f], s_pc[f]));
                            e[i] = (valid && f < L.f_in) ? __sinf(w_reduce_2pi(arg)) : 0.f;
                            zp = fmaf(e[i], svecA[f], zp);
                        }
                        uint4 sp[4];
                        split16(e, sp);
                        put_act(sp, ch * 16);
                        signal_group(g);
                        if (Et) put_tile(sp, ch * 16, Et, w.K1);
                    } else {
                        signal_group(g);
                    }
                }
            }
            float gb = 0.f;
            for (int ph = 0; ph < nph; ++ph, ++gph) {
                bool dg; int l, K, N;
                chain_phase(w, ph, dg, l, K, N);
                const bool tr = threadIdx.x == 64 && item == cid + 2 * ncl;
                WC_TRACE(tr, 8 * ph);
                // ---- per-phase vectors ----
                epi_bar();                                  // previous phase's vectors are dead
                if (!dg) {
                    for (int c = et; c < N; c += WC_EPI) svecA[c] = c < w.H ? W[L.off_b[l] + c] : 0.f;
                    if (l == w.L - 1)
                        for (int c = et; c < N; c += WC_EPI) svecB[c] = c < w.H ? W[L.off_w_out + c] : 0.f;
                } else if (l == 0) {
                    for (int c = et; c < N; c += WC_EPI) svecA[c] = c < L.f_in ? W[L.off_w_out + w.H + c] : 0.f;
                }
                epi_bar();
                const int nch = N / 16;
                const bool more = ph + 1 < nph;             // a next phase consumes what this one writes to shared memory
                const uint32_t tacc = taddr + (gph & 1) * 256;
                WC_TRACE(tr, 8 * ph + 1);
                mbar_wait(tfull, gph & 1);
                fence_after_sync();
                WC_TRACE(tr, 8 * ph + 2);
                if (!dg) {
                    const bool last = l == w.L - 1;
                    unsigned char *Ht = (bwd && live) ? a.H[l] + t32 * w.Hp * 128 : nullptr;     // masks (and wgrad operands)
                    for (int g = 0; 4 * g < nch; ++g) {
                        const int ch = 4 * g + part;
                        if (ch < nch) {
                            float v[16];
                            tmem_ld16(tacc + ch * 16, v);
#pragma unroll
                            for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i] + svecA[ch * 16 + i], 0.f);
                            if (last) {
#pragma unroll
                                for (int i = 0; i < 16; ++i) zp = fmaf(v[i], svecB[ch * 16 + i], zp);
                            }
                            if (!last) {                                                // H_L itself is not needed again
                                uint4 sp[4];
                                split16(v, sp);
                                put_act(sp, ch * 16);
                                if (more) signal_group(g);
                                if (Ht) put_tile(sp, ch * 16, Ht, w.Hp);
                            }
                        } else if (!last && more) {
                            signal_group(g);
                        }
                        WC_TRACE(tr, 8 * ph + 3 + g);
                    }
                    if (last) {
                        // ---- logit (onf_model.py:48-49), dL/dz ----
                        zpart[part * 128 + row] = zp;
                        epi_bar();
                        float z = W[L.off_b_out];
#pragma unroll
                        for (int k = 0; k < WC_PARTS; ++k) z += zpart[k * 128 + row];
                        const size_t gi = (size_t)p * a.n + idx;
                        if (part == 0 && valid && a.z_out) a.z_out[gi] = z;
                        if (bwd) {
                            if (valid) {
                                if (a.gbar_mode == 0) gb = a.gz[gi];
                                else if (a.gbar_mode == 1) gb = (sigmoid_f(z) - a.aux[gi]) / (float)cnt;
                                else {
                                    const float zb = z * a.beta;
                                    const float ez = expf(zb);
                                    gb = a.wc * (zb > 20.f ? 1.f : ez / (ez + 1.f));
                                    if (a.gbar_mode == 3) {
                                        const float tz = tanhf(z);
                                        gb += a.aux[gi] * (1.f - tz * tz);
                                    }
                                }
                            }
                            if (wgrad && part == 0) {
                                const float sgb = warp_sum(gb);
                                if (lane == 0 && auxv) auxv[aux_db3(w)] = sgb;
                            }
                            // second pass over the accumulator: gradient w.r.t. the last hidden layer's pre-activation
                            unsigned char *Gt = (wgrad && live) ? a.G[l] + t32 * w.Hp * 128 : nullptr;
                            for (int g = 0; 4 * g < nch; ++g) {
                                const int ch = 4 * g + part;
                                if (ch < nch) {
                                    float v[16], t[16];
                                    tmem_ld16(tacc + ch * 16, v);
#pragma unroll
                                    for (int i = 0; i < 16; ++i) {
                                        const float h = fmaxf(v[i] + svecA[ch * 16 + i], 0.f);
                                        t[i] = gb * h;                                        // d w_out[:H] = sum_pt gbar * H_L
                                        v[i] = h > 0.f ? gb * svecB[ch * 16 + i] : 0.f;
                                    }
                                    uint4 sp[4];
                                    split16(v, sp);
                                    put_act(sp, ch * 16);
                                    if (more) signal_group(g);                            // (before the column sums: MMA starts earlier)
                                    if (Gt) put_tile(sp, ch * 16, Gt, w.Hp);
                                    if (wgrad) {
                                        put_colsum(t, aux_dw3h(w) + ch * 16);
                                        put_colsum(v, aux_db(w, l) + ch * 16);                // d b_L = sum_pt G_L
                                    }
                                } else if (more) {
                                    signal_group(g);
                                }
                            }
                        }
                    }
                } else if (l > 0) {
                    // ---- input gradient of layer l -> ReLU mask of layer l-1 -> A operand of the next phase ----
                    const unsigned char *Mt = a.H[l - 1] + t32 * w.Hp * 128;        // dummy rows read tile 0: harmless
                    unsigned char *Gt = (wgrad && live) ? a.G[l - 1] + t32 * w.Hp * 128 : nullptr;
                    for (int g = 0; 4 * g < nch; ++g) {
                        const int ch = 4 * g + part;
                        if (ch < nch) {
                            float v[16];
                            const uint4 m0 = *reinterpret_cast<const uint4 *>(Mt + (ch * 2) * 512 + r32 * 16);
                            const uint4 m1 = *reinterpret_cast<const uint4 *>(Mt + (ch * 2 + 1) * 512 + r32 * 16);
                            tmem_ld16(tacc + ch * 16, v);
                            const uint32_t mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                const uint32_t bits = (mw[i >> 1] >> ((i & 1) * 16)) & 0x7FFFu;
                                v[i] = bits != 0 ? v[i] : 0.f;
                            }
                            uint4 sp[4];
                            split16(v, sp);
                            put_act(sp, ch * 16);
                            if (more) signal_group(g);
                            if (Gt) put_tile(sp, ch * 16, Gt, w.Hp);
                            if (wgrad) put_colsum(v, aux_db(w, l - 1) + ch * 16);             // d b_{l-1} = sum_pt G_{l-1}
                        } else if (more) {
                            signal_group(g);
                        }
                        WC_TRACE(tr, 8 * ph + 3 + g);
                    }
                } else {
                    // ---- dL/dE (+ skip connection) -> encoding backward: dL/d(arg) for the encoder gradients, dL/dx ----
                    float g0 = 0.f, g1 = 0.f, g2 = 0.f;
                    for (int ch = part; ch < nch; ch += WC_PARTS) {
                        float v[16], te[16], t0[16], t1[16];
                        tmem_ld16(tacc + ch * 16, v);
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const int f = ch * 16 + i;
                            const bool fourier = f < L.f_pos;
                            const float pa = s_pa[f], pb = s_pb[f];
                            const float xx = fourier ? u1 : th;
                            const float rr = w_reduce_2pi(fmaf(xx, pb, fmaf(u0, pa, s_pc[f])));
                            const bool on = valid && f < L.f_in;
                            const float ge = fmaf(gb, svecA[f], v[i]);
                            const float gp = on ? ge * __cosf(rr) : 0.f;
                            te[i] = on ? gb * __sinf(rr) : 0.f;              // d w_out[H:] = sum_pt gbar * E
                            t0[i] = gp * u0;                                  // encoder: sum gp*u0, sum gp*x1, sum gp
                            t1[i] = gp * xx;
                            v[i] = gp;
                            if (fourier) { g0 = fmaf(gp, pa, g0); g1 = fmaf(gp, pb, g1); }
                            else g2 = fmaf(gp, pb, g2);
                        }
                        if (wgrad) {
                            put_colsum(te, aux_dw3e(w) + ch * 16);
                            put_colsum(t0, aux_enc(w, 0) + ch * 16);
                            put_colsum(t1, aux_enc(w, 1) + ch * 16);
                            put_colsum(v, aux_enc(w, 2) + ch * 16);
                        }
                    }
                    if (a.gx_out) {
                        red[(0 * WC_PARTS + part) * 128 + row] = g0;
                        red[(1 * WC_PARTS + part) * 128 + row] = g1;
                        red[(2 * WC_PARTS + part) * 128 + row] = g2;
                        epi_bar();
                        if (part == 0 && valid) {
                            float *out = a.gx_out + ((size_t)p * a.n + idx) * a.x_stride;
                            float sg[3] = {0.f, 0.f, 0.f};
#pragma unroll
                            for (int c = 0; c < 3; ++c)
#pragma unroll
                                for (int k = 0; k < WC_PARTS; ++k) sg[c] += red[(c * WC_PARTS + k) * 128 + row];
                            out[0] = sg[0] / a.d.sigma;
                            out[1] = sg[1] / a.d.sigma;
                            if (L.in_dim == 3) out[2] = sg[2];
                            for (int c = L.in_dim; c < a.x_stride; ++c) out[c] = 0.f;
                        }
                    }
                }
                WC_TRACE(tr, 8 * ph + 7);
                if (!more) fence_before_sync();
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (CL > 1) cluster_sync_all();                    // no peer may still multicast into / arrive on this CTA's shared memory
    if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---- fixed-order sum of the aux partials over (row block, lane quarter) -> gradient blob --------------------------------
struct WideAuxArgs {
    OnfDims d;
    WideDims w;
    const float *weights, *aux;
    float *grad;
    int RB;
};

__global__ void __launch_bounds__(256) wide_aux_kernel(const WideAuxArgs a) {
    const int p = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    const WideDims &w = a.w;
    const nfopp_onf_layout &L = a.d.lay;
    const int n = aux_len(w);
    if (i >= n) return;
    const float *src = a.aux + (size_t)p * a.RB * 4 * n + i;
    float s = 0.f;
    for (int r = 0; r < a.RB * 4; ++r) s += src[(size_t)r * n];
    float *G = a.grad + (size_t)p * L.n_params;
    const float *W = a.weights + (size_t)p * L.n_params;
    if (i < aux_dw3h(w)) {
        const int l = i / w.Hp, c = i - l * w.Hp;
        if (c < w.H) G[L.off_b[l] + c] = s;
    } else if (i < aux_dw3e(w)) {
        const int c = i - aux_dw3h(w);
        if (c < w.H) G[L.off_w_out + c] = s;
    } else if (i < aux_enc(w, 0)) {
        const int f = i - aux_dw3e(w);
        if (f < L.f_in) G[L.off_w_out + w.H + f] = s;
    } else if (i < aux_db3(w)) {
        const int k = (i - aux_enc(w, 0)) / w.K1, f = i - aux_enc(w, k);
        if (f < L.f_pos) {
            if (k == 0) G[L.off_enc_w + 2 * f] = s;
            else if (k == 1) G[L.off_enc_w + 2 * f + 1] = s;
            else G[L.off_enc_b + f] = a.d.bias ? s : 0.f;
        } else if (f < L.f_in && k >= 1) {
            // heading features: d b_k = f_k * sum gp ; d f_k = sum gp * (theta + b_k)   (two aux entries -> two-term sums)
            const int kk = f - L.f_pos;
            const float *s2 = a.aux + (size_t)p * a.RB * 4 * n + aux_enc(w, 2) + f;
            float sgp = 0.f;
            for (int r = 0; r < a.RB * 4; ++r) sgp += s2[(size_t)r * n];
            if (k == 1) G[L.off_ang_f + kk] = fmaf(W[L.off_ang_b + kk], sgp, s);
            else G[L.off_ang_b + kk] = W[L.off_ang_f + kk] * sgp;
        }
    } else if (i == aux_db3(w)) {
        G[L.off_b_out] = s;
    }
}

// ---- weight gradients of ALL hidden layers in one launch, one 256 x N tile per CTA (TMEM: 2 x 256 columns) -------------
// item = (problem, layer, split of the point range); A = G_l (M = out), B = H_{l-1} or E (N = in), both MN-major views of
// 32-point operand tiles; every operand byte is read once per point (the 128-row-per-CTA kernel reads B twice).
constexpr int W2_STAGES = 3;
constexpr int W2_HALF = 256 * 32 * 2;                // one plane of one operand: 256 columns x 32 points x bf16
constexpr int W2_STAGE = 4 * W2_HALF;                // A hi | A lo | B hi | B lo
constexpr int W2_SMEM = W2_STAGES * W2_STAGE + 256;

struct WideWgrad2Args {
    WideDims w;
    const unsigned char *G[NFOPP_MAX_LAYERS], *Hin[NFOPP_MAX_LAYERS];    // A = G[l] (C = Hp), B = Hin[l] (C = kin[l])
    int off_w[NFOPP_MAX_LAYERS];
    float *dst;
    size_t dst_pstride;
    int P, T32, splits;
    float *partial;
    size_t partial_sstride;
};

__global__ void __launch_bounds__(WG_THREADS, 1) wide_wgrad2_kernel(const WideWgrad2Args a) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + W2_STAGES * W2_STAGE);
    uint64_t *full = bars, *empty = bars + W2_STAGES, *tfull = bars + 2 * W2_STAGES, *tempty = tfull + 1;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 1);
    const WideDims &w = a.w;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tps = (a.T32 + a.splits - 1) / a.splits;
    const int items = a.P * w.L * a.splits;
    const int MB = (w.Hp + 127) / 128;
    if (threadIdx.x == 0) {
        for (int s = 0; s < W2_STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
        mbar_init(tfull, 1);
        mbar_init(tempty, 4);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    auto decode = [&](int item, int &p, int &l, int &sp, int &t0, int &t1) {
        sp = item % a.splits;
        l = (item / a.splits) % w.L;
        p = item / (a.splits * w.L);
        t0 = sp * tps;
        t1 = min(a.T32, t0 + tps);
    };
    if (warp == 0) {
        uint32_t it = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x) {
            int p, l, sp, t0, t1;
            decode(item, p, l, sp, t0, t1);
            const int CB = w.kin[l];
            for (int t = t0; t < t1; ++t, ++it) {
                const int s = it % W2_STAGES;
                WIDE_ASSERT(w.Hp * 64 <= W2_HALF && CB * 64 <= W2_HALF);
                mbar_wait(empty + s, ((it / W2_STAGES) & 1) ^ 1);
                unsigned char *st = smem + s * W2_STAGE;
                const size_t tile = (size_t)p * a.T32 + t;
                if (lane == 0) mbar_expect_tx(full + s, (uint32_t)((w.Hp + CB) * 128));
                __syncwarp();
                if (lane < 2) bulk_g2s(st + lane * W2_HALF, a.G[l] + tile * w.Hp * 128 + (size_t)lane * w.Hp * 64, (uint32_t)(w.Hp * 64), full + s);
                else if (lane < 4)
                    bulk_g2s(st + lane * W2_HALF, a.Hin[l] + tile * CB * 128 + (size_t)(lane - 2) * CB * 64, (uint32_t)(CB * 64), full + s);
            }
        }
    } else if (warp == 1) {
        const uint32_t lead = elect_one();
        uint32_t it = 0, tile_it = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x, ++tile_it) {
            int p, l, sp, t0, t1;
            decode(item, p, l, sp, t0, t1);
            const int NB = w.kin[l];
            mbar_wait(tempty, (tile_it & 1) ^ 1);
            fence_after_sync();
            const uint32_t idesc = instr_desc_bf16(128, NB, 1, 1);
            bool started = false;
            for (int t = t0; t < t1; ++t, ++it) {
                const int s = it % W2_STAGES;
                mbar_wait(full + s, (it / W2_STAGES) & 1);
                fence_after_sync();
                const uint32_t st = smem_u32(smem + s * W2_STAGE);
#pragma unroll
                for (int ks = 0; ks < 2; ++ks) {
                    const uint64_t bh = desc_mnmajor(st + 2 * W2_HALF + ks * 256, 128, 512);
                    const uint64_t bl = desc_mnmajor(st + 3 * W2_HALF + ks * 256, 128, 512);
                    for (int mb = 0; mb < MB; ++mb) {
                        const uint64_t ah = desc_mnmajor(st + mb * 16 * 512 + ks * 256, 128, 512);
                        const uint64_t al = desc_mnmajor(st + W2_HALF + mb * 16 * 512 + ks * 256, 128, 512);
                        wmma3(tmem + mb * 256, ah, al, bh, bl, idesc, started, lead);
                    }
                    started = true;
                }
                mma_commit(empty + s, lead);
            }
            mma_commit(tfull, lead);
        }
    } else {
        const int q = warp & 3;
        uint32_t tile_it = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x, ++tile_it) {
            int p, l, sp, t0, t1;
            decode(item, p, l, sp, t0, t1);
            const int NB = w.kin[l], ld = w.kin_real[l];
            mbar_wait(tfull, tile_it & 1);
            fence_after_sync();
            float *D = (a.splits > 1 ? a.partial + (size_t)sp * a.partial_sstride : a.dst) + (size_t)p * a.dst_pstride + a.off_w[l];
            const bool has_work = t1 > t0;
            for (int mb = 0; mb < MB; ++mb) {
                const int m = mb * 128 + q * 32 + lane;
                const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + mb * 256;
                for (int c0 = 0; c0 < NB; c0 += 16) {
                    float v[16];
                    tmem_ld16(taddr + c0, v);
                    if (m >= w.H || !has_work) continue;
                    float *drow = D + (size_t)m * ld + c0;
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4)
                        if (c0 + 4 * i4 + 3 < ld)
                            *reinterpret_cast<float4 *>(drow + 4 * i4) = make_float4(v[4 * i4], v[4 * i4 + 1], v[4 * i4 + 2], v[4 * i4 + 3]);
                        else
                            for (int i = 4 * i4; i < 4 * i4 + 4; ++i)
                                if (c0 + i < ld) drow[i] = v[i];
                }
            }
            fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

// dst[i] = sum over splits (fixed order) of partial[s][i]
__global__ void __launch_bounds__(256) wide_reduce_kernel(const float *partial, size_t sstride, int splits, float *dst, size_t n4) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 acc = reinterpret_cast<const float4 *>(partial)[i];
        for (int s = 1; s < splits; ++s) {
            const float4 v = reinterpret_cast<const float4 *>(partial + (size_t)s * sstride)[i];
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        reinterpret_cast<float4 *>(dst)[i] = acc;
    }
}

size_t align256(size_t x) { return (x + 255) / 256 * 256; }

int g_sm_count = 0;
int sm_count() {
    if (!g_sm_count) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count <= 0) g_sm_count = 148;
    }
    return g_sm_count;
}

struct WideWorkspace {
    size_t E, U16, H[NFOPP_MAX_LAYERS], G[NFOPP_MAX_LAYERS], GE, G16, gbar, raw, partial, aux, total;
    int splits;                  // point-range splits of every weight-gradient job of the call
    size_t partial_sstride;      // floats between splits: [P][n_params] gradient partial, then [P][K1*16] encoder partial
};

// the fused chain kernel holds a 128-point row block of <= 256 columns in shared memory
bool wide_use_chain(const WideDims &w) {
    static int enabled = -1;
    if (enabled < 0) {
        const char *e = getenv("NFOPP_WIDE_CHAIN");
        enabled = (e && e[0] == '0') ? 0 : 1;
    }
    return enabled && w.Hp <= 256 && w.K1 <= 256;
}

void wide_workspace(const WideDims &w, int n_params, int P, int n, bool bwd, WideWorkspace *ws) {
    const bool chain = wide_use_chain(w);
    const size_t T64 = (size_t)(n + 127) / 128 * 2, tiles = (size_t)P * T64;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align256(bytes); return o; };
    ws->E = take(tiles * w.K1 * 256);
    const int nh = bwd ? w.L : (w.L < 2 ? w.L : 2);
    for (int l = 0; l < NFOPP_MAX_LAYERS; ++l) ws->H[l] = 0;
    for (int l = 0; l < nh; ++l) ws->H[l] = take(tiles * w.Hp * 256);
    if (!bwd)
        for (int l = 2; l < w.L; ++l) ws->H[l] = ws->H[l & 1];
    if (bwd) {
        ws->U16 = take(tiles * 16 * 256);
        for (int l = 0; l < NFOPP_MAX_LAYERS; ++l) ws->G[l] = 0;
        for (int l = 0; l < (chain ? w.L : 2); ++l) ws->G[l] = take(tiles * w.Hp * 256);    // chain: every layer's gradient is kept for wgrad
        ws->GE = take(tiles * w.K1 * 256);
        ws->G16 = take(tiles * 16 * 256);
        ws->gbar = take(tiles * 64 * 4);
        ws->raw = take((size_t)P * w.K1 * 16 * 4);
        // enough work items to fill the GPU twice over when there are few problems: split the point range
        const int wmax = w.K1 > w.Hp ? w.K1 : w.Hp;
        const int base = P * ((w.Hp + 127) / 128) * ((wmax + WG_NT - 1) / WG_NT + 1);
        int splits = (2 * sm_count() + base - 1) / base;
        if (splits > (int)T64) splits = (int)T64;
        ws->splits = splits < 1 ? 1 : splits;
        ws->partial_sstride = (size_t)P * ((size_t)n_params + (size_t)w.K1 * 16);
        ws->partial = ws->splits > 1 ? take((size_t)ws->splits * ws->partial_sstride * 4) : 0;
        ws->aux = chain ? take((size_t)P * (T64 / 2) * 4 * aux_len(w) * 4) : 0;
    } else {
        ws->U16 = ws->GE = ws->G16 = ws->gbar = ws->raw = ws->partial = ws->aux = 0;
        for (int l = 0; l < NFOPP_MAX_LAYERS; ++l) ws->G[l] = 0;
        ws->splits = 1;
        ws->partial_sstride = 0;
    }
    ws->total = off;
}

}  // namespace

// ---- host side -----------------------------------------------------------------------------------------------------
bool onf_wide_supported(const OnfDims &d) {
    WideDims w;
    return make_wide_dims(d, &w);
}

size_t onf_wide_packed_bytes(const OnfDims &d) {
    WideDims w;
    return make_wide_dims(d, &w) ? align256(w.packed_bytes) : 0;
}

size_t onf_wide_workspace_bytes(const OnfDims &d, int P, int n, bool bwd) {
    WideDims w;
    if (!make_wide_dims(d, &w)) return 0;
    WideWorkspace ws;
    wide_workspace(w, d.lay.n_params, P, n, bwd, &ws);
    return ws.total;
}

int onf_wide_prepare(const OnfDims &d) {
    WideDims w;
    NFOPP_REQUIRE(make_wide_dims(d, &w), "onf(wide): hidden=%d layers=%d unsupported (hidden %% 4 == 0, <= 1024)", d.H, d.L);
    NFOPP_CHECK_CUDA(cudaFuncSetAttribute(wide_gemm_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM));
    NFOPP_CHECK_CUDA(cudaFuncSetAttribute(wide_gemm_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM));
    NFOPP_CHECK_CUDA(cudaFuncSetAttribute(wide_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WW_SMEM));
    NFOPP_CHECK_CUDA(cudaFuncSetAttribute(wide_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WC_SMEM));
    NFOPP_CHECK_CUDA(cudaFuncSetAttribute(wide_wgrad2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, W2_SMEM));
    sm_count();
    return NFOPP_OK;
}

int onf_wide_pack(const OnfDims &d, const float *weights, unsigned char *packed, int P, cudaStream_t stream) {
    WidePackArgs a{};
    a.d = d;
    NFOPP_REQUIRE(make_wide_dims(d, &a.w), "onf(wide): unsupported network shape");
    a.weights = weights; a.packed = packed;
    if (P == 0) return NFOPP_OK;
    int max_tiles = 0;
    for (int l = 0; l < a.w.L; ++l) max_tiles = max(max_tiles, ((a.w.Hp + 31) / 32) * ((a.w.kin[l] / 8 + 7) / 8));
    dim3 grid(max_tiles, P);
    wide_pack_kernel<<<grid, 256, 0, stream>>>(a);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

// NFOPP_CHAIN_CLUSTER = 1 | 2 | 4 (default 1): CTAs per cluster of the fused chain kernel
static int chain_cluster_size() {
    static int cl = 0;
    if (!cl) {
        const char *e = getenv("NFOPP_CHAIN_CLUSTER");
        cl = e ? atoi(e) : 1;      // measured on B200: the weight stream is not the bottleneck, multicast does not pay
        if (cl != 1 && cl != 2 && cl != 4) cl = 1;
    }
    return cl;
}
// co-resident clusters of the chain kernel (1 CTA per SM): asked from the occupancy API once per cluster size
static int chain_max_clusters(int CL) {
    static int cached[5] = {0, 0, 0, 0, 0};
    if (!cached[CL]) {
        int n = 0;
        if (CL == 1) n = sm_count();
        else {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(sm_count() / CL * CL); cfg.blockDim = dim3(WC_THREADS); cfg.dynamicSmemBytes = WC_SMEM;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            if (cudaOccupancyMaxActiveClusters(&n, wide_chain_kernel, &cfg) != cudaSuccess || n <= 0) {
                cudaGetLastError();
                n = sm_count() / CL / 2;
            }
        }
        cached[CL] = n > 0 ? n : 1;
    }
    return cached[CL];
}

// engine: Adam on the whole blob fused with the refresh of the operand copies (one pass over w/m/v/grad)
int onf_wide_adam_pack(const OnfDims &d, float *weights, float *m, float *v, const float *grad, unsigned char *packed,
                       const StepParams *sp, int warm, float w1, float beta2, float w2, float eps, int P, cudaStream_t stream,
                       const int32_t *count) {
    WidePackArgs a{};
    a.d = d;
    NFOPP_REQUIRE(make_wide_dims(d, &a.w), "onf(wide): unsupported network shape");
    NFOPP_REQUIRE(weights && m && v && grad && packed && sp, "onf_wide_adam_pack: bad arguments");
    a.weights = weights; a.packed = packed; a.adam_w = weights; a.adam_m = m; a.adam_v = v; a.grad = grad; a.sp = sp; a.warm = warm;
    a.count = count;
    a.w1 = w1; a.beta2 = beta2; a.w2 = w2; a.eps = eps;
    int max_tiles = 0;
    for (int l = 0; l < a.w.L; ++l) max_tiles = max(max_tiles, ((a.w.Hp + 31) / 32) * ((a.w.kin[l] / 8 + 7) / 8));
    dim3 grid(max_tiles, P);
    wide_pack_kernel<<<grid, 256, 0, stream>>>(a);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

static int launch_wgrad(const WideWgradArgs &g, cudaStream_t s) {
    const int MB = (g.CA + 127) / 128, NTT = (g.CB + WG_NT - 1) / WG_NT + (g.bias_off >= 0 ? 1 : 0);
    const int items = g.P * MB * NTT * g.splits;
    wide_wgrad_kernel<<<items < sm_count() ? items : sm_count(), WG_THREADS, WW_SMEM, s>>>(g);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

// kernels one onf_wide_launch issues (memsets not counted)
int onf_wide_launch_count(const OnfDims &d, int P, int n, bool want_bwd, bool want_wgrad) {
    WideDims w;
    if (!make_wide_dims(d, &w)) return 0;
    WideWorkspace ws;
    wide_workspace(w, d.lay.n_params, P, n, want_bwd, &ws);
    const int red = ws.splits > 1 ? 1 : 0;
    if (wide_use_chain(w)) return 1 + (want_bwd && want_wgrad ? 2 + red : 0);
    if (!want_bwd) return 2 + d.L;
    return 2 + d.L + d.L + 1 + (want_wgrad ? d.L + 5 + 2 * red : 0);
}

// One ONF call (forward, or forward + backward): a fixed sequence of launches on `stream`.
int onf_wide_launch(const OnfArgs &o, const OnfWideExtra &x, int P, cudaStream_t stream) {
    WideDims w;
    NFOPP_REQUIRE(make_wide_dims(o.d, &w), "onf(wide): unsupported network shape");
    NFOPP_REQUIRE(x.packed && x.workspace, "onf(wide): packed operand copy / workspace missing");
    if (o.n == 0 || P == 0) return NFOPP_OK;
    const bool bwd = o.want_bwd != 0, wgrad = bwd && x.grad_out != nullptr, want_gx = bwd && o.gx_out != nullptr;
    const nfopp_onf_layout &L = o.d.lay;
    WideWorkspace ws;
    wide_workspace(w, L.n_params, P, o.n, bwd, &ws);
    unsigned char *base = x.workspace;
    const int T64 = (o.n + 127) / 128 * 2;
    const int x_stride = o.x_stride ? o.x_stride : L.in_dim;
    const int grid_g = sm_count();

    if (wide_use_chain(w)) {
        // ---- fused chain: ONE kernel for encode -> forward -> dL/dz -> input gradients -> encoding backward ----
        float *raw = reinterpret_cast<float *>(base + ws.raw);
        float *partial = reinterpret_cast<float *>(base + ws.partial);
        const int T32 = 2 * T64;
        if (wgrad) {
            NFOPP_CHECK_CUDA(cudaMemsetAsync(x.grad_out, 0, (size_t)P * L.n_params * sizeof(float), stream));
            NFOPP_CHECK_CUDA(cudaMemsetAsync(raw, 0, (size_t)P * w.K1 * 16 * sizeof(float), stream));
            if (ws.splits > 1)
                NFOPP_CHECK_CUDA(cudaMemsetAsync(partial, 0, (size_t)ws.splits * ws.partial_sstride * sizeof(float), stream));
        }
        WideChainArgs c{};
        c.d = o.d; c.w = w; c.weights = o.weights; c.packed = x.packed; c.w_stride = align256(w.packed_bytes);
        c.x = o.x; c.count = o.count; c.n = o.n; c.x_stride = x_stride; c.P = P; c.T32 = T32; c.RB = T64 / 2;
        c.want_bwd = bwd; c.want_wgrad = wgrad; c.want_gx = want_gx; c.gbar_mode = o.gbar_mode; c.gz = o.gz; c.aux = o.aux;
        c.wc = o.wc; c.beta = o.beta; c.z_out = o.z_out; c.gx_out = o.gx_out;
        c.E = base + ws.E; c.U16 = base + ws.U16; c.GE = base + ws.GE; c.G16 = base + ws.G16;
        c.auxp = reinterpret_cast<float *>(base + ws.aux);
        for (int l = 0; l < w.L; ++l) { c.H[l] = base + ws.H[l]; c.G[l] = base + ws.G[l]; }
        {
            // cluster size: as many CTAs sharing one weight stream as the problem has row blocks (<= 4)
            int CL = chain_cluster_size();
            while (CL > 1 && c.RB < CL) CL /= 2;
            c.CL = CL;
            const int groups = P * ((c.RB + CL - 1) / CL);
            cudaLaunchConfig_t cfg{};
            cfg.blockDim = dim3(WC_THREADS); cfg.dynamicSmemBytes = WC_SMEM; cfg.stream = stream;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int max_clusters = chain_max_clusters(CL);
            const int ncl = groups < max_clusters ? groups : max_clusters;
            cfg.gridDim = dim3(ncl * CL);
            NFOPP_CHECK_CUDA(cudaLaunchKernelEx(&cfg, wide_chain_kernel, c));
        }
        if (!wgrad) return NFOPP_OK;
        {   // hidden-layer weight gradients: one launch, 256 x N tile per CTA
            WideWgrad2Args g{};
            g.w = w; g.dst = x.grad_out; g.dst_pstride = L.n_params; g.P = P; g.T32 = T32; g.splits = ws.splits;
            g.partial = partial; g.partial_sstride = ws.partial_sstride;
            for (int l = 0; l < w.L; ++l) {
                g.G[l] = base + ws.G[l]; g.Hin[l] = l == 0 ? base + ws.E : base + ws.H[l - 1]; g.off_w[l] = L.off_w[l];
            }
            const int witems = P * w.L * ws.splits;
            wide_wgrad2_kernel<<<witems < grid_g ? witems : grid_g, WG_THREADS, W2_SMEM, stream>>>(g);
            NFOPP_CHECK_CUDA(cudaGetLastError());
        }
        if (ws.splits > 1) {
            const size_t n4g = (size_t)P * L.n_params / 4;
            wide_reduce_kernel<<<(unsigned)((n4g + 255) / 256 < 4096 ? (n4g + 255) / 256 : 4096), 256, 0, stream>>>(
                partial, ws.partial_sstride, ws.splits, x.grad_out, n4g);
            NFOPP_CHECK_CUDA(cudaGetLastError());
        }
        // biases, output layer, encoder: fixed-order sum of the chain kernel's per-row-block partials
        WideAuxArgs ax{};
        ax.d = o.d; ax.w = w; ax.weights = o.weights; ax.aux = reinterpret_cast<const float *>(base + ws.aux); ax.grad = x.grad_out;
        ax.RB = T64 / 2;
        wide_aux_kernel<<<dim3((aux_len(w) + 255) / 256, P), 256, 0, stream>>>(ax);
        NFOPP_CHECK_CUDA(cudaGetLastError());
        return NFOPP_OK;
    }

    {
        WideEncArgs e{};
        e.d = o.d; e.weights = o.weights; e.x = o.x; e.count = o.count; e.n = o.n; e.x_stride = x_stride; e.T64 = T64; e.K1 = w.K1;
        e.E = base + ws.E; e.U16 = wgrad ? base + ws.U16 : nullptr;
        wide_enc_kernel<<<dim3(T64, P), 256, 0, stream>>>(e);
        NFOPP_CHECK_CUDA(cudaGetLastError());
    }
    auto gemm_items = [&](int Nfull) { return P * (T64 / 2) * ((Nfull + WG_NT - 1) / WG_NT); };
    for (int l = 0; l < w.L; ++l) {
        WideGemmArgs g{};
        g.A = l == 0 ? base + ws.E : base + ws.H[l - 1]; g.CA = w.kin[l];
        g.Wpk = x.packed + w.wf_off[l]; g.w_stride = align256(w.packed_bytes);
        g.K = w.kin[l]; g.Nfull = w.Hp; g.Out = base + ws.H[l];
        g.weights = o.weights; g.n_params = L.n_params; g.bias_off = L.off_b[l]; g.n_valid = w.H;
        g.P = P; g.T64 = T64;
        const int items = gemm_items(g.Nfull);
        wide_gemm_kernel<0><<<items < grid_g ? items : grid_g, WG_THREADS, WG_SMEM, stream>>>(g);
        NFOPP_CHECK_CUDA(cudaGetLastError());
    }
    float *raw = reinterpret_cast<float *>(base + ws.raw);
    float *partial = reinterpret_cast<float *>(base + ws.partial);
    if (wgrad) {
        // every element of the gradient is stored exactly once (by a wgrad job or by the encoder assembly) except the
        // alignment padding of the blob, and with splits > 1 a split with no points stores nothing: start from zero
        NFOPP_CHECK_CUDA(cudaMemsetAsync(x.grad_out, 0, (size_t)P * L.n_params * sizeof(float), stream));
        NFOPP_CHECK_CUDA(cudaMemsetAsync(raw, 0, (size_t)P * w.K1 * 16 * sizeof(float), stream));
        if (ws.splits > 1) NFOPP_CHECK_CUDA(cudaMemsetAsync(partial, 0, (size_t)ws.splits * ws.partial_sstride * sizeof(float), stream));
    }
    int cur = 0;
    {
        WideOutArgs u{};
        u.d = o.d; u.weights = o.weights; u.count = o.count; u.HL = base + ws.H[w.L - 1]; u.E = base + ws.E;
        u.Hp = w.Hp; u.K1 = w.K1; u.n = o.n; u.T64 = T64; u.want_bwd = bwd; u.gbar_mode = o.gbar_mode; u.gz = o.gz; u.aux = o.aux;
        u.wc = o.wc; u.beta = o.beta; u.z_out = o.z_out;
        if (bwd) {
            u.gbar = reinterpret_cast<float *>(base + ws.gbar);
            u.G16 = wgrad ? base + ws.G16 : nullptr;
            u.GL = base + ws.G[cur];
        }
        wide_out_kernel<<<dim3(T64, P), 256, 0, stream>>>(u);
        NFOPP_CHECK_CUDA(cudaGetLastError());
    }
    if (!bwd) return NFOPP_OK;
    int rc;
    WideWgradArgs wg{};
    wg.P = P; wg.T64 = T64; wg.dst = x.grad_out; wg.dst_pstride = L.n_params;
    wg.splits = ws.splits; wg.partial = partial; wg.partial_sstride = ws.partial_sstride; wg.row_lo = 0; wg.TP = 64;
    if (wgrad) {
        // output layer: d w_out[:H] = sum_pt H_L * gbar ; d w_out[H:] = sum_pt E * gbar     (onf_model.py:48-49)
        WideWgradArgs j = wg;
        j.A = base + ws.H[w.L - 1]; j.CA = w.Hp; j.B = base + ws.G16; j.CB = 16;
        j.dst_off = L.off_w_out; j.dst_ld = 1; j.rows_valid = w.H; j.cols_valid = 1; j.bias_off = -1;
        if ((rc = launch_wgrad(j, stream))) return rc;
        j.A = base + ws.E; j.CA = w.K1; j.dst_off = L.off_w_out + w.H; j.rows_valid = w.FIN;
        if ((rc = launch_wgrad(j, stream))) return rc;
        // d b_out = sum_pt 1 * gbar: row 3 of U16^T G16 (U16 column 3 is the ones column)
        j.A = base + ws.U16; j.CA = 16; j.dst_off = L.off_b_out - 3; j.row_lo = 3; j.rows_valid = 4;
        if ((rc = launch_wgrad(j, stream))) return rc;
    }
    for (int l = w.L - 1; l >= 0; --l) {
        if (wgrad) {
            WideWgradArgs j = wg;
            j.A = base + ws.G[cur]; j.CA = w.Hp; j.B = l == 0 ? base + ws.E : base + ws.H[l - 1]; j.CB = w.kin[l];
            j.dst_off = L.off_w[l]; j.dst_ld = w.kin_real[l]; j.rows_valid = w.H; j.cols_valid = w.kin_real[l];
            j.bias_off = L.off_b[l];
            if ((rc = launch_wgrad(j, stream))) return rc;
        }
        if (l == 0 && !want_gx && !wgrad) break;
        WideGemmArgs g{};
        g.A = base + ws.G[cur]; g.CA = w.Hp;
        g.Wpk = x.packed + w.wd_off[l]; g.w_stride = align256(w.packed_bytes);
        g.K = w.Hp; g.Nfull = w.kin[l];
        g.weights = o.weights; g.n_params = L.n_params;
        g.P = P; g.T64 = T64;
        if (l > 0) {
            g.Out = base + ws.G[cur ^ 1]; g.Mask = base + ws.H[l - 1];
        } else {
            g.Out = base + ws.GE; g.gbar = reinterpret_cast<const float *>(base + ws.gbar);
            g.skip_off = L.off_w_out + w.H; g.skip_n = w.FIN;
        }
        const int items = gemm_items(g.Nfull);
        wide_gemm_kernel<1><<<items < grid_g ? items : grid_g, WG_THREADS, WG_SMEM, stream>>>(g);
        NFOPP_CHECK_CUDA(cudaGetLastError());
        cur ^= 1;
    }
    {
        WideEbwdArgs e{};
        e.d = o.d; e.weights = o.weights; e.x = o.x; e.count = o.count; e.GE = base + ws.GE; e.K1 = w.K1; e.n = o.n;
        e.x_stride = x_stride; e.T64 = T64; e.write_g0 = wgrad; e.gx_out = o.gx_out;
        wide_ebwd_kernel<<<dim3(T64, P), 256, 0, stream>>>(e);
        NFOPP_CHECK_CUDA(cudaGetLastError());
    }
    if (wgrad) {
        WideWgradArgs j = wg;
        j.A = base + ws.GE; j.CA = w.K1; j.B = base + ws.U16; j.CB = 16;
        j.dst = raw; j.dst_pstride = (size_t)w.K1 * 16;
        j.partial = partial + (size_t)P * L.n_params;             // the encoder partials follow the gradient partials
        j.dst_off = 0; j.dst_ld = 16; j.rows_valid = w.K1; j.cols_valid = 16; j.bias_off = -1;
        if ((rc = launch_wgrad(j, stream))) return rc;
        if (ws.splits > 1) {
            const size_t n4g = (size_t)P * L.n_params / 4, n4r = (size_t)P * w.K1 * 16 / 4;
            wide_reduce_kernel<<<(unsigned)((n4g + 255) / 256 < 4096 ? (n4g + 255) / 256 : 4096), 256, 0, stream>>>(
                partial, ws.partial_sstride, ws.splits, x.grad_out, n4g);
            wide_reduce_kernel<<<(unsigned)((n4r + 255) / 256), 256, 0, stream>>>(partial + (size_t)P * L.n_params, ws.partial_sstride,
                                                                                   ws.splits, raw, n4r);
            NFOPP_CHECK_CUDA(cudaGetLastError());
        }
        WideAsmArgs s{};
        s.d = o.d; s.weights = o.weights; s.raw = reinterpret_cast<const float *>(base + ws.raw); s.grad = x.grad_out; s.K1 = w.K1;
        wide_asm_kernel<<<dim3((w.FIN + 255) / 256, P), 256, 0, stream>>>(s);
        NFOPP_CHECK_CUDA(cudaGetLastError());
    }
    return NFOPP_OK;
}

}  // namespace nfopp

#ifdef NFOPP_TC_TRACE
extern "C" int nfopp_debug_wc_trace(long long *host_out) {
    return (int)cudaMemcpyFromSymbol(host_out, nfopp::g_wc_trace, sizeof(nfopp::g_wc_trace));
}
#endif
