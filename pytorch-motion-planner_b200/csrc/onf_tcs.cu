// onf_tcs.cu - K2, fused tensor path, second generation: 112-point tiles, weights STREAMED, every product exact.
//
// onf_tc.cu keeps a problem's hidden-layer weights resident in shared memory (150 KB), which leaves room for only two
// 32-point tiles in flight: ~1000 tcgen05.mma of N = 32 per problem and 14 latency-bound handshakes per 32 points.  Here
// the operand roles are the same (HIDDEN index on the UMMA M axis = TMEM lanes, POINT index on N) but
//   * a tile is 112 points (a problem's 209 training samples = 2 tiles, its 99 path points = 1), so every GEMM of
//     onf_model.py:33-50 and of its backward is a handful of 128 x 112 x 16 MMAs and every epilogue stage runs once per
//     tile on all 16 worker warps;
//   * the weights are not resident: a producer warp streams the packed operand copy (L2-resident, 150 KB per problem)
//     through an 8-stage ring of K = 16 chunks for the forward GEMMs and as hi / lo blocks of 25 KB for the backward GEMMs;
//   * operands are fp16 hi/lo pairs and EVERY product is three MMAs (hi*hi + lo*hi + hi*lo, fp32 accumulate, ~1e-7):
//     a one-pass backward was built and measured first (profiles/r02a_tcs_bwd1_*: gradients 2e-4 .. 1.5e-3 of the max
//     norm, four parity tests red) and the emulation in profiles/r02_precision_emulation.json shows that every backward
//     operand (gradient, weight, activation) needs its lo part for a 3x margin under the 1e-3 bar;
//   * the upstream gradient is scaled by a power of two chosen BEFORE the forward from a bound of |dL/dz| (BCE: 1 / N;
//     path terms: collision weight + max multiplier; external: max |gz|), so fp16 keeps its range and the weight-gradient
//     accumulators of all tiles share one scale; epilogues unscale in fp32;
//   * weight-gradient accumulators (dW2: 112, dW1: 224 TMEM columns) persist across the tiles of a problem next to ONE
//     112-column work accumulator, so any point count is supported.
//
//   fwd1   D1[j,n]  = sum_k W1p[j,k] E[n,k]     A = W1p chunk (K-major, ring)    B = E chunk  (K-major, ring)
//   fwd2   D2[i,n]  = sum_j W2p[i,j] H1[j,n]    A = W2p chunk (K-major, ring)    B = H1       (MN-major)
//   wgrad2 dW2[i,j]+= sum_n gH2[i,n] H1[j,n]    A = gH2       (K-major)          B = H1       (K-major)
//   dgrad2 gD1[j,n] = sum_i W2p[i,j] gH2[i,n]   A = W2p       (MN-major, slots)  B = gH2      (MN-major)
//   dgrad1 gE[k,n]  = sum_j W1p[j,k] gH1[j,n]   A = W1p       (MN-major, slots)  B = gH1      (MN-major)   2 feature blocks
//   wgrad1 dW1[j,k]+= sum_n gH1[j,n] E[n,k]     A = gH1       (K-major)          B = E^T      (K-major)    2 feature blocks
//
// Row H of W1p holds the skip weights w_out[H:] (onf_model.py:48-49) exactly as in onf_tc.cu, so the skip term, its input
// gradient and its weight gradient fall out of fwd1 / dgrad1 / wgrad1.  Biases, ReLU, logit, dL/dz (BCE
// nerf_opt_planner.py:85-90, path terms constrained...:87-97), the Fourier / heading encoding and its backward run on
// the CUDA cores, one TMEM lane (hidden unit or feature) per thread.  The epilogue stages the gradient blob in shared
// memory and streams it through Adam (planner_factory.py:68), rewriting the fp16 operand copy - as onf_tc.cu does.
//
// Supported: 2 hidden layers, hidden <= 111, f_in <= 256 (the reference's network shapes), any point count.
#include <stdlib.h>

#include "onf.cuh"
#include "tc05.cuh"

namespace nfopp {
using namespace tc05;

// optional phase trace (build with -DNFOPP_TC_TRACE; scripts/tcs_trace.py): clock64 stamps of worker thread 0, the issue
// warp and the weight-stream warp of the first and last CTA
#ifdef NFOPP_TC_TRACE
__device__ long long g_ts_trace[3][2][4][64];   // [launch kind: fwd | train | path][first / last CTA][worker | issuer | producer | optimiser][slot]
#define TS_TRACE(role, slot)                                                                                              \
    do {                                                                                                                \
        if ((threadIdx.x & 31) == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1) && (slot) < 64 &&               \
            ((role) != 0 || threadIdx.x == 0))                                                                          \
            g_ts_trace[!a.want_bwd ? 0 : ((a.grad_out || a.adam_w) ? 1 : 2)][blockIdx.x == 0 ? 0 : 1][(role)][(slot)] = clock64(); \
    } while (0)
#else
#define TS_TRACE(role, slot) do { } while (0)
#endif
#ifdef NFOPP_TC_TRACE
#define TS_TRACE_VAL(role, slot, val)                                                                                    \
    do {                                                                                                                \
        if ((threadIdx.x & 31) == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1))                                \
            g_ts_trace[!a.want_bwd ? 0 : ((a.grad_out || a.adam_w) ? 1 : 2)][blockIdx.x == 0 ? 0 : 1][(role)][(slot)] = (val); \
    } while (0)
#define TS_CLOCK() clock64()
#define TS_TWAIT(acc, stmt) do { const long long _t0 = clock64(); stmt; (acc) += clock64() - _t0; } while (0)
#else
#define TS_TRACE_VAL(role, slot, val) do { } while (0)
#define TS_CLOCK() 0LL
#define TS_TWAIT(acc, stmt) do { stmt; } while (0)
#endif

// device-side bounds asserts (make -C csrc debug; compute-sanitizer is closed on the GPU pool): a violated extent traps with
// the line number instead of corrupting a neighbouring buffer
#ifdef NFOPP_DEBUG
#define TS_ASSERT(cond)                                                                                                   \
    do {                                                                                                                \
        if (!(cond)) {                                                                                                  \
            printf("onf_tcs.cu:%d: assertion failed: %s (block %d thread %d)\n", __LINE__, #cond, (int)blockIdx.x, (int)threadIdx.x); \
            __trap();                                                                                                   \
        }                                                                                                               \
    } while (0)
#else
#define TS_ASSERT(cond) do { } while (0)
#endif

constexpr int TS_WORKERS = 512;                 // 16 epilogue / encoding warps
constexpr int TS_OPT_WARPS = 6;                 // optimiser warps (Adam + operand-copy refresh of the PREVIOUS problem)
constexpr int TS_OPT = 32 * TS_OPT_WARPS;
constexpr int TS_THREADS = TS_WORKERS + 64 + TS_OPT;   // + MMA-issue warp + weight-stream warp + optimiser warps
constexpr int TS_GBUF = 1;                      // gradient hand-over buffers per CTA in use (of the 2 allocated): the optimiser has read
                                                // a problem's gradient long before the next one is ready, and 148 x 133 KB less live data in L2 pays
constexpr int TS_REGS_WORKER = 96, TS_REGS_SERVICE = 48;   // setmaxnreg split of the 768 x 80 registers the CTA starts with: 512 x 88 + 256 x 64
constexpr int TS_NT = 112;                      // points per tile
constexpr int TS_HALF = 64;                     // ... handled as two halves (0..63 | 64..111) from the first epilogue on
constexpr uint32_t TS_SPT = TS_NT * 16;         // S_row of the point-indexed operand buffers [hidden / 8][point / 8][8][8]
constexpr int TS_WST = 8;                       // weight ring stages (K = 16 chunk of W, hi + lo)
constexpr int TS_EF = 64;                       // features per encoding slot (2 slots)
constexpr uint32_t TS_TM_A = 0, TS_TM_DW2 = TS_NT, TS_TM_DW1 = 2 * TS_NT;   // TMEM columns: work | dW2 [KH] | dW1 [K1]
constexpr int TS_GX_FLOATS = 4 * TS_NT * 3;     // per-lane-quarter partials of the input gradient
constexpr int TS_AST = 3;                       // cp.async stages of an updating warp
constexpr int TS_ASTAGE = 4 * 32 * 16 + 32 * 8; // one stage: 32 lanes x ([g | w | m | v] x 16 B + the lane's two offsets)
constexpr int TS_ARING = TS_AST * TS_ASTAGE;    // bytes of one warp's ring
constexpr int TS_SROW = 68;                     // row stride (floats) of the 64-column gradient staging buffer

struct TsDims {
    int F, A, FIN, H, K1, KH, MB, FB0;           // MB feature blocks of FB0 (last one shorter) for dgrad1 / wgrad1
    int S_colW, w1_half, w2_half;                // packed operand copy: [W1 hi | W1 lo | W2 hi | W2 lo], each [k / 8][row][8]
    int h1_half, e_half, w_stage, b_slot, S_ET, et_half;
    int off_h1h, off_h1l, off_r, off_w, off_misc, misc_floats, off_stage, off_red, off_aring, off_praw, praw_floats, off_bar, smem_bytes, packed_bytes;
};

__host__ __device__ inline int ts_ru16(int x) { return (x + 15) / 16 * 16; }
__host__ __device__ inline int ts_max(int a, int b) { return a > b ? a : b; }
__host__ __device__ inline int ts_min(int a, int b) { return a < b ? a : b; }

__host__ __device__ inline bool ts_make_dims(const OnfDims &d, TsDims *out) {
    TsDims t{};
    t.F = d.lay.f_pos; t.A = d.lay.n_angle; t.FIN = d.lay.f_in; t.H = d.H;
    t.K1 = ts_ru16(t.FIN); t.KH = ts_ru16(t.H + 1);
    t.MB = t.K1 > 128 ? 2 : 1;
    t.FB0 = t.MB == 2 ? ts_ru16(t.K1 / 2) : t.K1;
    t.S_colW = t.KH * 16;
    t.w1_half = (t.K1 / 8) * t.S_colW;
    t.w2_half = (t.KH / 8) * t.S_colW;
    t.h1_half = (t.KH / 8) * (int)TS_SPT;
    t.S_ET = t.FB0 * 16;
    t.et_half = (TS_NT / 8) * t.S_ET;            // E^T of one feature block: [point / 8][feature][8 points]
    t.e_half = (TS_EF / 8) * (int)TS_SPT;        // one encoding slot, hi (or lo) part
    t.w_stage = 4 * t.S_colW;                    // 2 column octets x (hi, lo)
    t.b_slot = 16 * t.S_colW;                    // a 128-row MN-major A operand never reads past its slot
    t.off_h1h = 0;
    t.off_h1l = t.h1_half;
    t.off_r = 2 * t.h1_half;                     // encoding ring (2 slots, hi + lo) | gH2 hi, lo | E^T hi, lo | per-unit partial sums
    t.off_w = t.off_r + ts_max(ts_max(4 * t.e_half, 2 * t.h1_half), ts_max(2 * t.et_half, 10 * 512 * 4));
    t.off_misc = t.off_w + ts_max(TS_WST * t.w_stage, 2 * t.b_slot);
    t.misc_floats = 3 * t.K1 + 32 + 3 * t.KH + 8 + 10 * TS_NT + TS_GX_FLOATS + 32;
    t.off_stage = t.off_h1h;                     // [128 rows][TS_SROW] floats: transposes the weight gradient when the problem's
                                                 // MMAs are done (H1 / gH1 are dead until the next problem's first epilogue)
    t.off_red = ts_max(t.off_r, t.off_stage + 128 * TS_SROW * 4);   // per-unit partial sums [10][4][128] floats, behind it
    t.off_aring = (t.off_misc + t.misc_floats * 4 + 127) / 128 * 128;   // cp.async rings of the optimiser warps
    // raw fp32 parameters outside W1 / W2 / the skip row (encoder | b1 | b2, w_out[:H] | b_out) of the problem being set up and
    // of the next one: bulk-copied by the weight-stream warp while the previous problem computes
    t.praw_floats = d.lay.off_w[0] + t.H + (d.lay.off_w_out + t.H - d.lay.off_b[1]) + 4;
    t.off_praw = t.off_aring + TS_OPT_WARPS * TS_ARING;
    t.off_bar = t.off_praw + 2 * t.praw_floats * 4;
    t.smem_bytes = t.off_bar + 64 * 8 + 16;
    t.packed_bytes = 2 * t.w1_half + 2 * t.w2_half;
    *out = t;
    return d.L == 2 && t.KH <= TS_NT && t.K1 <= 256 && t.H >= 8 && t.H % 4 == 0 && t.FIN % 4 == 0 &&
           (int)TS_TM_DW1 + t.K1 <= 512 && t.off_red + 10 * 512 * 4 <= t.off_w && t.smem_bytes <= 227 * 1024;
}

struct OnfTsArgs {
    OnfDims d;
    TsDims t;
    const float *weights;          // [P, n_params] fp32 blob
    const unsigned char *packed;   // [P, packed_bytes] fp16 hi/lo operand copy of W1p / W2p
    const float *x;
    const int32_t *count;
    const StepParams *sp;
    int P, n, x_stride;
    int want_bwd;
    int gbar_mode;
    const float *gz, *aux;
    float wc, beta;
    float *z_out, *gx_out;
    float *grad_out;
    float *adam_w, *adam_m, *adam_v;
    float *gscratch;               // fused Adam: [gridDim.x, 2, n_params] gradient hand-over buffers (workers -> optimiser warps)
    unsigned char *packed_out;
    float w1, beta2, w2, eps;
    int warm;
    int dbg;                       // trace builds: NFOPP_TCS_DBG bit 0 skips the operand-copy stores, bit 1 the w / m / v stores (timing only)
};

struct TsPackArgs {
    OnfDims d;
    TsDims t;
    const float *weights;
    unsigned char *packed;
};

// fp32 blob -> packed fp16 hi/lo operands (API entry points and engine start; the training kernel keeps the copy current)
__global__ void __launch_bounds__(256) onf_tcs_pack_kernel(const TsPackArgs a) {
    const int p = blockIdx.y;
    const TsDims &t = a.t;
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
        split8h(v, hi, lo);
        unsigned char *base = out + (second ? 2 * t.w1_half : 0);
        const int half = second ? t.w2_half : t.w1_half;
        const int off = oct * t.S_colW + (r / 8) * 128 + (r % 8) * 16;
        *reinterpret_cast<uint4 *>(base + off) = hi;
        *reinterpret_cast<uint4 *>(base + half + off) = lo;
    }
}

// ---- device helpers ------------------------------------------------------------------------------------
// column sums over the 32 lanes of a warp for 16 columns: lane 2c (and 2c+1) returns sum_lanes v[c]
__device__ __forceinline__ float ts_colsum16(float (&v)[16], int lane) {
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
// 2^-floor(log2 m) for m > 0 (1 for zero / non-finite m): a bound m of |dL/dz| maps to [1, 2)
__device__ __forceinline__ float ts_pow2_scale(float m) {
    int e = (int)((__float_as_uint(m) >> 23) & 0xFFu) - 127;
    if (!(m > 0.f) || e > 100) e = 0;
    if (e < -100) e = -100;
    return __uint_as_float((uint32_t)(127 - e) << 23);
}

struct TsShared {
    unsigned char *h1h, *h1l, *r, *w, *aring;
    // per input feature f < K1: arg = x1 * pb + u0 * pa + pc  (x1 = u1 for Fourier features, theta for heading features;
    // pc carries the bias and +pi/2 for cosine features, so every feature is sin(arg) with derivative cos(arg))
    float *pa, *pb, *pc, *ang_b, *b1, *b2, *w3h, *scal, *in0, *aux, *skip, *gbar, *zp, *gxp, *red, *praw;
    uint64_t *wfull, *wempty, *efull, *eempty, *bfull, *bempty, *ready, *done, *gfull, *gempty, *pfull;
    uint32_t *tmem_slot;
    __device__ TsShared(unsigned char *base, const TsDims &t) {
        h1h = base + t.off_h1h; h1l = base + t.off_h1l; r = base + t.off_r; w = base + t.off_w;
        float *f = reinterpret_cast<float *>(base + t.off_misc);
        pa = f; f += t.K1;
        pb = f; f += t.K1;
        pc = f; f += t.K1;
        ang_b = f; f += 32;
        b1 = f; f += t.KH;
        b2 = f; f += t.KH;
        w3h = f; f += t.KH;
        scal = f; f += 8;                           // [0] b3
        in0 = f; f += 3 * TS_NT;                    // u0 | u1 | theta
        aux = f; f += TS_NT;
        skip = f; f += TS_NT;
        gbar = f; f += TS_NT;
        zp = f; f += 4 * TS_NT;
        gxp = f; f += TS_GX_FLOATS;
        red = f; f += 32;
        praw = reinterpret_cast<float *>(base + t.off_praw);
        aring = base + t.off_aring;
        uint64_t *b = reinterpret_cast<uint64_t *>(base + t.off_bar);
        wfull = b; b += TS_WST;
        wempty = b; b += TS_WST;
        efull = b; b += 2;
        eempty = b; b += 2;
        bfull = b; b += 2;
        bempty = b++;
        ready = b; b += 2;                          // per half tile (points 0..63 | 64..111)
        done = b; b += 2;
        gfull = b; b += 2;
        gempty = b; b += 2;
        pfull = b; b += 2;
        tmem_slot = reinterpret_cast<uint32_t *>(b);
    }
};

// K loop of a three-pass split product, descriptors advanced by constant strides:
// D (+)= sum_k Ahi*Bhi + Alo*Bhi + Ahi*Blo.  (Unrolling it with descriptors formed as base + ks * stride was measured: no
// gain in the issue warp's ~65 cycles per tcgen05.mma - uniform-datapath latency, not instruction count - and the larger
// code cost 3 %.)
__device__ __forceinline__ void ts_kloop3(uint32_t tmem_d, uint64_t ah, uint64_t al, uint64_t bh, uint64_t bl, uint32_t a_step,
                                          uint32_t b_step, int ksteps, uint32_t idesc, bool accumulate, uint32_t issue) {
    const uint64_t da = a_step >> 4, db = b_step >> 4;
#pragma unroll 1
    for (int ks = 0; ks < ksteps; ++ks) {
        mma_bf16(tmem_d, ah, bh, idesc, (accumulate || ks > 0) ? 1u : 0u, issue);
        mma_bf16(tmem_d, al, bh, idesc, 1u, issue);
        mma_bf16(tmem_d, ah, bl, idesc, 1u, issue);
        ah += da; al += da; bh += db; bl += db;
    }
}

__device__ __forceinline__ void ts_worker_sync() { asm volatile("bar.sync 1, %0;" ::"r"(TS_WORKERS) : "memory"); }
// a worker warp has finished writing MMA operands (generic proxy) / reading TMEM: publish to the async side
__device__ __forceinline__ void ts_signal(uint64_t *bar) {
    fence_proxy_async();
    fence_before_sync();
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(bar);
}
__device__ __forceinline__ void ts_wait(uint64_t *bar, uint32_t use) {     // wait for completion number `use` (0-based)
#ifdef NFOPP_TS_SPIN
    mbar_wait_spin(bar, use & 1u);
#else
    mbar_wait(bar, use & 1u);
#endif
    fence_after_sync();
}
__device__ __forceinline__ void ts_wait_relaxed(uint64_t *bar, uint32_t use, uint32_t sleep_ns) {   // long, non-critical waits
    if (sleep_ns == 0) mbar_wait(bar, use & 1u);
    else mbar_wait_relaxed(bar, use & 1u, sleep_ns);
    fence_after_sync();
}

// store 16 consecutive columns (col0 .. col0+15) of row `row` into [row / 8][col / 8][row % 8][col % 8] operand buffers
__device__ __forceinline__ void ts_store_row16(unsigned char *hi_buf, unsigned char *lo_buf, int row, int col0, const float *v) {
    uint4 h0, l0, h1, l1;
    split8h(v, h0, l0);
    split8h(v + 8, h1, l1);
    const int off = (row >> 3) * TS_SPT + (col0 >> 3) * 128 + (row & 7) * 16;
    *reinterpret_cast<uint4 *>(hi_buf + off) = h0;
    *reinterpret_cast<uint4 *>(hi_buf + off + 128) = h1;
    *reinterpret_cast<uint4 *>(lo_buf + off) = l0;
    *reinterpret_cast<uint4 *>(lo_buf + off + 128) = l1;
}

// ---- the weight-stream warp (one elected lane): per problem and tile the forward ring, then the backward blocks (hi ->
// slot 0, lo -> slot 1; the slots overlay the ring).  The CTA is persistent: the ring runs on into the next problem's weights as soon as the last
// backward block of this one has been consumed, and the operand copy + the small fp32 parameters of the problem after
// that are pulled into L2 while this one computes.
__device__ __forceinline__ void ts_producer(const OnfTsArgs &a, const TsShared &s, bool want_bwd) {
    const TsDims &t = a.t;
    const nfopp_onf_layout &L = a.d.lay;
    const int c1 = t.K1 / 16, c2 = t.KH / 16;
    uint32_t wc = 0, nbe = 0;
    auto block = [&](const unsigned char *hi, int half, int bytes) {      // one weight block, hi and lo parts
        TS_ASSERT(bytes > 0 && bytes <= t.b_slot && t.off_w + 2 * t.b_slot <= t.off_misc && (bytes & 15) == 0);
        TS_ASSERT(hi + half + bytes <= a.packed + (size_t)a.P * t.packed_bytes);
        mbar_expect_tx(s.bfull + 0, (uint32_t)bytes);
        bulk_g2s(s.w, hi, bytes, s.bfull + 0);
        mbar_expect_tx(s.bfull + 1, (uint32_t)bytes);
        bulk_g2s(s.w + t.b_slot, hi + half, bytes, s.bfull + 1);
    };
    // raw fp32 parameters of problem q -> buffer (use % 2) of s.praw (set-up of the worker warps)
    auto stage_params = [&](int q, uint32_t use) {
        const float *Wq = a.weights + (size_t)q * L.n_params;
        float *dst = s.praw + (size_t)(use & 1u) * t.praw_floats;
        uint64_t *bar = s.pfull + (use & 1u);
        const int l0 = L.off_w[0], l1 = t.H, l2 = L.off_w_out + t.H - L.off_b[1];
        mbar_expect_tx(bar, (uint32_t)(l0 + l1 + l2 + 4) * 4u);
        bulk_g2s(dst, Wq, (uint32_t)l0 * 4u, bar);
        bulk_g2s(dst + l0, Wq + L.off_b[0], (uint32_t)l1 * 4u, bar);
        bulk_g2s(dst + l0 + l1, Wq + L.off_b[1], (uint32_t)l2 * 4u, bar);
        bulk_g2s(dst + l0 + l1 + l2, Wq + L.off_b_out, 16u, bar);
    };
    auto next_problem = [&](int q) {                    // the next problem of this CTA with points (or >= a.P)
        for (q += (int)gridDim.x; q < a.P; q += (int)gridDim.x)
            if ((a.count ? a.count[q] : a.n) > 0) break;
        return q;
    };
    uint32_t np = 0;
    {
        const int p0 = next_problem((int)blockIdx.x - (int)gridDim.x);
        if (p0 < a.P) stage_params(p0, 0);
    }
    for (int p = blockIdx.x; p < a.P; p += gridDim.x) {
        const int cnt = a.count ? a.count[p] : a.n;
        if (cnt <= 0) continue;
        const int n_tiles = (cnt + TS_NT - 1) / TS_NT;
        const unsigned char *src = a.packed + (size_t)p * t.packed_bytes;
        ++np;
        for (int tile = 0; tile < n_tiles; ++tile) {
            for (int c = 0; c < c1 + c2; ++c, ++wc) {                 // fwd1, fwd2: K = 16 chunks of W1p, then of W2p (hi + lo)
                const uint32_t st = wc % TS_WST, k = wc / TS_WST;
                if (k > 0) ts_wait_relaxed(s.wempty + st, k - 1, 64);
                unsigned char *dst = s.w + st * t.w_stage;
                TS_ASSERT(st < (uint32_t)TS_WST && t.off_w + (int)(st + 1) * t.w_stage <= t.off_misc);
                const bool second = c >= c1;
                const unsigned char *g = src + (second ? 2 * t.w1_half : 0) + (size_t)(second ? c - c1 : c) * 2 * t.S_colW;
                const int half = second ? t.w2_half : t.w1_half;
                mbar_expect_tx(s.wfull + st, (uint32_t)t.w_stage);
                bulk_g2s(dst, g, 2 * t.S_colW, s.wfull + st);
                bulk_g2s(dst + 2 * t.S_colW, g + half, 2 * t.S_colW, s.wfull + st);
            }
            TS_TRACE(2, 1);
            if (tile == n_tiles - 1) {
                // (prefetching the CTA's FIRST problem here for the step's next ONF launch was measured: no effect)
                // L2 prefetches, issued late so that the lines are still resident when they are used (ncu: issued at the
                // start of the problem, a wave's 148 x 0.8 MB outlived the L2 and was read from HBM twice):
                const int pn = next_problem(p);
                if (pn < a.P) {
                    // this CTA's next problem: operand copy -> L2, the fp32 parameters its set-up reads -> shared memory (the
                    // buffer was last read by the set-up of the previous problem)
                    bulk_prefetch_l2(a.packed + (size_t)pn * t.packed_bytes, (uint32_t)t.packed_bytes);
                    stage_params(pn, np);
                }
            }
            if (!want_bwd) continue;
            // every forward MMA that reads the ring has completed before the ring is reused as two big slots
            for (uint32_t c = wc > TS_WST ? wc - TS_WST : 0; c < wc; ++c) ts_wait_relaxed(s.wempty + c % TS_WST, c / TS_WST, 64);
            TS_TRACE(2, 2);
            block(src + 2 * t.w1_half, t.w2_half, t.w2_half);                           // W2p -> dgrad2
            ts_wait_relaxed(s.bempty, nbe++, 64);
            const int len0 = t.FB0 / 8 * t.S_colW;
            block(src, t.w1_half, len0);                                                // W1, feature block 0 -> dgrad1
            ts_wait_relaxed(s.bempty, nbe++, 64);
            if (t.MB == 2) {
                block(src + len0, t.w1_half, (t.K1 - t.FB0) / 8 * t.S_colW);            // W1, feature block 1
                ts_wait_relaxed(s.bempty, nbe++, 64);
            }
            TS_TRACE(2, 3);
        }
    }
}

// ---- the MMA-issue warp ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void ts_issuer(const OnfTsArgs &a, const TsShared &s, uint32_t tmem, bool want_bwd, bool want_wgrad) {
    const TsDims &t = a.t;
    const uint32_t lead = elect_one();
    TS_TRACE(1, 0);
    const uint32_t h1h = smem_u32(s.h1h), h1l = smem_u32(s.h1l), rr = smem_u32(s.r), ww = smem_u32(s.w);
    const uint32_t s0 = ww, s1 = ww + t.b_slot;                      // backward weight slots: hi, lo
    const uint32_t tmA = tmem + TS_TM_A, tmDW2 = tmem + TS_TM_DW2, tmDW1 = tmem + TS_TM_DW1;
    const int c2 = t.KH / 16;
    const int n_slots = (t.K1 + TS_EF - 1) / TS_EF;
    uint32_t wc = 0, ec = 0, nbf = 0;
    uint32_t nready0 = 0, nready1 = 0;
    long long t_ef = 0, t_wf = 0, t_bf = 0, t_rd = 0, t_tile0 = 0;
    for (int p = blockIdx.x; p < a.P; p += gridDim.x) {
    const int cnt = a.count ? a.count[p] : a.n;
    if (cnt <= 0) continue;
    const int n_tiles = (cnt + TS_NT - 1) / TS_NT;
    for (int tile = 0; tile < n_tiles; ++tile) {
        const int Nt = ts_ru16(min(TS_NT, cnt - tile * TS_NT));
        TS_ASSERT(Nt >= 16 && Nt <= TS_NT && (int)TS_TM_DW1 + t.K1 <= 512 && (int)TS_TM_DW2 + t.KH <= (int)TS_TM_DW1);
        TS_ASSERT(t.off_r + 2 * t.et_half <= t.off_w && t.off_r + 4 * t.e_half <= t.off_w && 2 * t.h1_half <= t.off_r);
        const bool acc = tile > 0;                                   // weight-gradient accumulators persist across tiles
        t_ef = t_wf = t_bf = t_rd = 0;
        t_tile0 = TS_CLOCK();
        // From the first epilogue on the tile is handled as two HALVES (points 0..63 | 64..Nt-1) with their own ready / done
        // barriers: the epilogue of one half runs under the MMAs of the other (same operand buffers and the same work
        // accumulator - the halves are disjoint point columns of both).
        const int nh = (want_bwd && Nt > TS_HALF) ? 2 : 1;           // (forward-only launches have nothing to overlap: one half)
        // fwd1: D1[j,n] (+)= W1p[j,k] E[n,k]: one encoding slot (64 features) = up to 4 ring stages of K = 16
        for (int sl = 0; sl < n_slots; ++sl, ++ec) {
            const uint32_t es = ec & 1u;
            TS_TWAIT(t_ef, ts_wait(s.efull + es, ec >> 1));
            const uint32_t eb = rr + es * 2 * t.e_half;
            const int ksl = min(TS_EF, t.K1 - sl * TS_EF) / 16;
            for (int ks = 0; ks < ksl; ++ks, ++wc) {
                const uint32_t st = wc % TS_WST;
                TS_TWAIT(t_wf, ts_wait(s.wfull + st, wc / TS_WST));
                const uint32_t wb = ww + st * t.w_stage;
                const uint32_t idesc = instr_desc_f16(128, Nt, 0, 0);
                const bool first = sl == 0 && ks == 0;
                mma_bf16(tmA, desc_kmajor(wb, 128, t.S_colW), desc_kmajor(eb + ks * 2 * TS_SPT, 128, TS_SPT), idesc, first ? 0u : 1u, lead);
                mma_bf16(tmA, desc_kmajor(wb + 2 * t.S_colW, 128, t.S_colW), desc_kmajor(eb + ks * 2 * TS_SPT, 128, TS_SPT), idesc, 1u, lead);
                mma_bf16(tmA, desc_kmajor(wb, 128, t.S_colW), desc_kmajor(eb + t.e_half + ks * 2 * TS_SPT, 128, TS_SPT), idesc, 1u, lead);
                mma_commit(s.wempty + st, lead);
            }
            mma_commit(s.eempty + es, lead);
        }
        for (int h = 0; h < nh; ++h) mma_commit(s.done + h, lead);
        TS_TRACE(1, 1);
        // fwd2: D2[i,n] = W2p[i,j] H1[j,n]      (both halves at once: its weights stream through the ring behind fwd1's)
        for (int h = 0; h < nh; ++h) TS_TWAIT(t_rd, ts_wait(s.ready + h, (h ? nready1++ : nready0++)));
        TS_TRACE(1, 2);
        for (int ks = 0; ks < c2; ++ks, ++wc) {
            const uint32_t st = wc % TS_WST;
            TS_TWAIT(t_wf, ts_wait(s.wfull + st, wc / TS_WST));
            const uint32_t wb = ww + st * t.w_stage;
            ts_kloop3(tmA, desc_kmajor(wb, 128, t.S_colW), desc_kmajor(wb + 2 * t.S_colW, 128, t.S_colW),
                      desc_mnmajor(h1h + ks * 2 * TS_SPT, TS_SPT, 128), desc_mnmajor(h1l + ks * 2 * TS_SPT, TS_SPT, 128), 0, 0, 1,
                      instr_desc_f16(128, Nt, 0, 1), ks > 0, lead);
            mma_commit(s.wempty + st, lead);
        }
        for (int h = 0; h < nh; ++h) mma_commit(s.done + h, lead);
        TS_TRACE(1, 3);
        if (!want_bwd) continue;
        // wgrad2: dW2[i,j] += gH2[i,n] H1[j,n];   dgrad2: gD1[j,n] = W2p[i,j] gH2[i,n]     (gH2 hi, lo in the old encoding ring)
        for (int h = 0; h < nh; ++h) {
            const int Nh = nh == 1 ? Nt : (h == 0 ? TS_HALF : Nt - TS_HALF);
            const uint32_t po = (uint32_t)h * (TS_HALF / 8) * 128u;
            TS_TWAIT(t_rd, ts_wait(s.ready + h, (h ? nready1++ : nready0++)));
            TS_TRACE(1, 4);
            if (want_wgrad)
                ts_kloop3(tmDW2, desc_kmajor(rr + po, TS_SPT, 128), desc_kmajor(rr + t.h1_half + po, TS_SPT, 128),
                          desc_kmajor(h1h + po, TS_SPT, 128), desc_kmajor(h1l + po, TS_SPT, 128), 256, 256, Nh / 16,
                          instr_desc_f16(128, t.KH, 0, 0), acc || h > 0, lead);
            if (h == 0) {
                TS_TWAIT(t_bf, ts_wait(s.bfull + 0, nbf));
                TS_TWAIT(t_bf, ts_wait(s.bfull + 1, nbf));
                ++nbf;
            }
            ts_kloop3(tmA + h * TS_HALF, desc_mnmajor(s0, 128, t.S_colW), desc_mnmajor(s1, 128, t.S_colW), desc_mnmajor(rr + po, TS_SPT, 128),
                      desc_mnmajor(rr + t.h1_half + po, TS_SPT, 128), 256, 2 * TS_SPT, c2, instr_desc_f16(128, Nh, 1, 1), false, lead);
            if (h == nh - 1) mma_commit(s.bempty, lead);
            mma_commit(s.done + h, lead);
        }
        TS_TRACE(1, 6);
        // dgrad1, feature block 0: gE[k,n] = W1p[j,k] gH1[j,n]      (gH1 hi, lo in place over H1 hi, lo)
        for (int h = 0; h < nh; ++h) {
            const int Nh = nh == 1 ? Nt : (h == 0 ? TS_HALF : Nt - TS_HALF);
            const uint32_t po = (uint32_t)h * (TS_HALF / 8) * 128u;
            TS_TWAIT(t_rd, ts_wait(s.ready + h, (h ? nready1++ : nready0++)));
            TS_TRACE(1, 7);
            if (h == 0) {
                TS_TWAIT(t_bf, ts_wait(s.bfull + 0, nbf));
                TS_TWAIT(t_bf, ts_wait(s.bfull + 1, nbf));
                ++nbf;
            }
            ts_kloop3(tmA + h * TS_HALF, desc_mnmajor(s0, 128, t.S_colW), desc_mnmajor(s1, 128, t.S_colW), desc_mnmajor(h1h + po, TS_SPT, 128),
                      desc_mnmajor(h1l + po, TS_SPT, 128), 256, 2 * TS_SPT, c2, instr_desc_f16(128, Nh, 1, 1), false, lead);
            if (h == nh - 1) mma_commit(s.bempty, lead);
            mma_commit(s.done + h, lead);
        }
        for (int mb = 0; mb < t.MB; ++mb) {
            for (int h = 0; h < nh; ++h) {
                const int Nh = nh == 1 ? Nt : (h == 0 ? TS_HALF : Nt - TS_HALF);
                const uint32_t po = (uint32_t)h * (TS_HALF / 8) * 128u;
                TS_TWAIT(t_rd, ts_wait(s.ready + h, (h ? nready1++ : nready0++)));          // block mb of gE consumed, E^T of block mb written (hi, lo)
                TS_TRACE(1, 9 + 2 * mb);
                if (want_wgrad)                             // wgrad1: dW1[j, block mb] += gH1[j,n] E[n, block mb]
                    ts_kloop3(tmDW1 + mb * t.FB0, desc_kmajor(h1h + po, TS_SPT, 128), desc_kmajor(h1l + po, TS_SPT, 128),
                              desc_kmajor(rr + h * (TS_HALF / 8) * t.S_ET, 128, t.S_ET),
                              desc_kmajor(rr + t.et_half + h * (TS_HALF / 8) * t.S_ET, 128, t.S_ET), 256, 2 * t.S_ET, Nh / 16,
                              instr_desc_f16(128, min(t.FB0, t.K1 - mb * t.FB0), 0, 0), acc || h > 0, lead);
                if (mb + 1 < t.MB) {
                    if (h == 0) {
                        TS_TWAIT(t_bf, ts_wait(s.bfull + 0, nbf));
                        TS_TWAIT(t_bf, ts_wait(s.bfull + 1, nbf));
                        ++nbf;
                    }
                    ts_kloop3(tmA + h * TS_HALF, desc_mnmajor(s0, 128, t.S_colW), desc_mnmajor(s1, 128, t.S_colW),
                              desc_mnmajor(h1h + po, TS_SPT, 128), desc_mnmajor(h1l + po, TS_SPT, 128), 256, 2 * TS_SPT, c2,
                              instr_desc_f16(128, Nh, 1, 1), false, lead);
                    if (h == nh - 1) mma_commit(s.bempty, lead);
                }
                if (want_wgrad || mb + 1 < t.MB) mma_commit(s.done + h, lead);
                TS_TRACE(1, 10 + 2 * mb);
            }
        }
        TS_TRACE_VAL(1, 40, t_ef); TS_TRACE_VAL(1, 41, t_wf); TS_TRACE_VAL(1, 42, t_bf); TS_TRACE_VAL(1, 43, t_rd);
        TS_TRACE_VAL(1, 44, TS_CLOCK() - t_tile0);
    }
    }
}

// one octet of encoded features for one point: e[i] = sin(x1 * pb + u0 * pa + pc), x1 = u1 (Fourier) or theta (heading)
__device__ __forceinline__ void ts_encode8(const TsShared &s, int f0, int F, float u0, float u1, float th, uint4 &hi, uint4 &lo) {
    float e[8], pa[8], pb[8], pc[8];
    *reinterpret_cast<float4 *>(pa) = *reinterpret_cast<const float4 *>(s.pa + f0);
    *reinterpret_cast<float4 *>(pa + 4) = *reinterpret_cast<const float4 *>(s.pa + f0 + 4);
    *reinterpret_cast<float4 *>(pb) = *reinterpret_cast<const float4 *>(s.pb + f0);
    *reinterpret_cast<float4 *>(pb + 4) = *reinterpret_cast<const float4 *>(s.pb + f0 + 4);
    *reinterpret_cast<float4 *>(pc) = *reinterpret_cast<const float4 *>(s.pc + f0);
    *reinterpret_cast<float4 *>(pc + 4) = *reinterpret_cast<const float4 *>(s.pc + f0 + 4);
    if (f0 + 8 <= F) {
#pragma unroll
        for (int i = 0; i < 8; ++i) e[i] = __sinf(fmaf(u1, pb[i], fmaf(u0, pa[i], pc[i])));
    } else if (f0 >= F) {
#pragma unroll
        for (int i = 0; i < 8; ++i) e[i] = __sinf(fmaf(th, pb[i], pc[i]));
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) e[i] = __sinf(fmaf(f0 + i < F ? u1 : th, pb[i], fmaf(u0, pa[i], pc[i])));
    }
    split8h(e, hi, lo);
}

// ---- the optimiser warps: torch.optim.Adam (_single_tensor_adam, planner_factory.py:68) of the problem the workers have
// just finished, while they compute the next one.  The gradient arrives in a per-CTA global buffer (L2), w / m / v come
// straight from HBM (prefetched into L2 under the problem's tiles); the fp16 hi/lo operand copy of the new weights is written
// in the same pass.
__device__ __forceinline__ void ts_adam4(float4 &w, float4 &m, float4 &v, const float4 g, float w1, float beta2, float w2, float eps,
                                         float step_size, float inv_bc2) {
    float gg[4] = {g.x, g.y, g.z, g.w}, ww[4] = {w.x, w.y, w.z, w.w}, mm[4] = {m.x, m.y, m.z, m.w}, v2[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        mm[c] = fmaf(w1, gg[c] - mm[c], mm[c]);
        v2[c] = fmaf(w2 * gg[c], gg[c], v2[c] * beta2);
        // (ftz forms: a denormal second moment is below eps^2 anyway and the denominator is at least eps, so the results
        // are those of sqrt.approx / __fdividef without their denormal-range fix-up code)
        float sq, rc;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(v2[c]));
        const float denom = fmaf(sq, inv_bc2, eps);
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(denom));
        ww[c] = fmaf(-step_size, mm[c] * rc, ww[c]);
    }
    w = make_float4(ww[0], ww[1], ww[2], ww[3]);
    m = make_float4(mm[0], mm[1], mm[2], mm[3]);
    v = make_float4(v2[0], v2[1], v2[2], v2[3]);
}

// the four ranges of the blob outside W1 / W2 / the skip row
__device__ __forceinline__ void ts_misc_ranges(const TsDims &t, const nfopp_onf_layout &L, int (&off)[4], int (&len)[4]) {
    off[0] = 0; len[0] = L.off_w[0];
    off[1] = L.off_b[0]; len[1] = L.off_w[1] - L.off_b[0];
    off[2] = L.off_b[1]; len[2] = L.off_w_out + t.H - L.off_b[1];
    off[3] = L.off_b_out; len[3] = L.n_params - L.off_b_out;
}

__device__ __forceinline__ void cp_async16(void *dst_smem, const void *src_gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src_gmem) : "memory");
}
// L2 eviction-priority hints: the Adam state streams through L2 exactly once per step - marked evict_first it stops pushing
// the operand copies (read four times per problem by the weight stream) out of L2
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void cp_async16_hint(void *dst_smem, const void *src_gmem, uint64_t pol) {
    asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_global_hint4(float *dst, const float4 v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(dst), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }


// The warps that apply a problem's update share its parameters by interleaving 32-lane groups of "half-octets" (4 consecutive
// weights of one row): lane pair (2i, 2i + 1) owns one column octet of one row of W1p / W2, so the fp32 loads and stores run
// through the rows linearly (coalesced) and one shuffle per lane completes the 16 B of hi parts (even lane) / lo parts (odd
// lane) at [octet][row] of the operand copy; a row's last octet may be half (or all) padding - zeros in the operand copy.
// Gradient (hand-over buffer, L2), w / m / v (HBM / L2) travel through a per-warp cp.async ring in shared memory, TS_AST - 1
// groups ahead of the arithmetic (a lane reads back only what it copied itself: no barrier; measured with loads into
// registers one group ahead the sweep was latency-bound at ~1.3 k cycles per group).  Warp `rank` of `nranks` takes groups
// rank, rank + nranks, ...
__device__ __forceinline__ void ts_adam_problem(const OnfTsArgs &a, const float *G, int p, int rank, int nranks, int lane,
                                                unsigned char *ring) {
    const TsDims &t = a.t;
    const nfopp_onf_layout &L = a.d.lay;
    const int hf = lane & 1;
    const float step_size = a.sp->onf_step_size, inv_bc2 = 1.f / a.sp->onf_bc2_sqrt;
    int moff[4], mlen[4];
    ts_misc_ranges(t, L, moff, mlen);
    const int nhs1 = 2 * ((t.FIN + 7) / 8), nhs2 = 2 * ((t.H + 7) / 8);       // half-octet slots per row (incl. padding)
    const int n1 = (t.H + 1) * nhs1, n2 = t.H * nhs2;
    const int n3 = (mlen[0] + mlen[1] + mlen[2] + mlen[3]) / 4;
    const int ntot = n1 + n2 + n3;
    const float inv1 = 1.f / (float)nhs1, inv2 = 1.f / (float)nhs2;
    const size_t po = (size_t)p * L.n_params;
    float *gw = a.adam_w + po, *gm = a.adam_m + po, *gv = a.adam_v + po;
    unsigned char *pk = a.packed_out + (size_t)p * t.packed_bytes;
    // flat index i of a half-octet: [0, n1) rows of W1p, [n1, n1 + n2) rows of W2, then the float4s of the other ranges
    // -> go: float offset in the blob (-1: nothing to update), poff: byte offset of the lane pair's 16 B in the operand copy (-1: none)
    auto locate = [&](int i, int &go, int &poff) {
        go = -1; poff = -1;
        if (i < n1 + n2) {
            const bool second = i >= n1;
            const int j = second ? i - n1 : i;
            const int r = (int)(((float)j + 0.5f) * (second ? inv2 : inv1));
            const int hs = j - r * (second ? nhs2 : nhs1);
            const int rowlen = second ? t.H : t.FIN;
            const int row_off = second ? L.off_w[1] + r * t.H : (r < t.H ? L.off_w[0] + r * t.FIN : L.off_w_out + t.H);
            if (4 * hs < rowlen) go = row_off + 4 * hs;
            poff = (second ? 2 * t.w1_half : 0) + (hf ? (second ? t.w2_half : t.w1_half) : 0) + (hs >> 1) * t.S_colW +
                   (r >> 3) * 128 + (r & 7) * 16;
        } else if (i < ntot) {
            int j = 4 * (i - n1 - n2);
            go = j < mlen[0] ? moff[0] + j
                 : (j -= mlen[0]) < mlen[1] ? moff[1] + j
                 : (j -= mlen[1]) < mlen[2] ? moff[2] + j : moff[3] + j - mlen[2];
        }
    };
    const int stride = 32 * nranks;
    if (rank * 32 >= ntot) return;
    const uint64_t pol = l2_policy_evict_first();
    unsigned char *mine = ring + lane * 16;                                  // stage s, array k at mine + s * TS_ASTAGE + k * 512
    int2 *mine_off = reinterpret_cast<int2 *>(ring + 2048) + lane;          // (go, poff) of stage s at byte s * TS_ASTAGE
    auto issue = [&](int base, int sl) {                                     // group `base` -> ring stage sl (one commit group)
        int go = -1, poff = -1;
        if (base < ntot) locate(base + lane, go, poff);
        TS_ASSERT(sl >= 0 && sl < TS_AST && go + 4 <= L.n_params && (go < 0 || (go & 3) == 0));
        TS_ASSERT(poff < 0 || (poff + 16 <= t.packed_bytes && (poff & 15) == 0));
        *reinterpret_cast<int2 *>(reinterpret_cast<unsigned char *>(mine_off) + sl * TS_ASTAGE) = make_int2(go, poff);
        if (go >= 0) {
            unsigned char *dst = mine + sl * TS_ASTAGE;
            cp_async16(dst, G + go);
            cp_async16_hint(dst + 512, gw + go, pol);
            cp_async16_hint(dst + 1024, gm + go, pol);
            cp_async16_hint(dst + 1536, gv + go, pol);
        }
        cp_async_commit();
    };
#pragma unroll
    for (int j = 0; j < TS_AST - 1; ++j) issue(rank * 32 + j * stride, j);
    int sl = 0;
    for (int base = rank * 32; base < ntot; base += stride) {                // warp-uniform trip count (shuffles inside)
        issue(base + (TS_AST - 1) * stride, sl == 0 ? TS_AST - 1 : sl - 1);
        cp_async_wait<TS_AST - 1>();
        const int2 gp = *reinterpret_cast<const int2 *>(reinterpret_cast<const unsigned char *>(mine_off) + sl * TS_ASTAGE);
        const int go = gp.x, poff = gp.y;
        const bool upd = go >= 0;
        float4 w4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (upd) {
            const unsigned char *srcp = mine + sl * TS_ASTAGE;
            const float4 g4 = *reinterpret_cast<const float4 *>(srcp);
            float4 m4 = *reinterpret_cast<const float4 *>(srcp + 1024), v4 = *reinterpret_cast<const float4 *>(srcp + 1536);
            w4 = *reinterpret_cast<const float4 *>(srcp + 512);
            ts_adam4(w4, m4, v4, g4, a.w1, a.beta2, a.w2, a.eps, step_size, inv_bc2);
#ifdef NFOPP_TC_TRACE
            if (!(a.dbg & 2))
#endif
            {
                st_global_hint4(gw + go, w4, pol);
                st_global_hint4(gm + go, m4, pol);
                st_global_hint4(gv + go, v4, pol);
            }
        }
        if (base < n1 + n2) {                                              // (warp-uniform; lanes past the matrices have poff < 0)
            uint32_t h0, l0, h1, l1;
            split2h(w4.x, w4.y, h0, l0);
            split2h(w4.z, w4.w, h1, l1);
            // even lane keeps its hi parts and receives the partner's; odd lane keeps its lo parts and receives the partner's
            const uint32_t s0 = hf ? h0 : l0, s1 = hf ? h1 : l1;
            const uint32_t r0 = __shfl_xor_sync(0xffffffffu, s0, 1), r1 = __shfl_xor_sync(0xffffffffu, s1, 1);
#ifdef NFOPP_TC_TRACE
            if (poff >= 0 && !(a.dbg & 1))
#else
            if (poff >= 0)
#endif
                *reinterpret_cast<uint4 *>(pk + poff) = hf ? make_uint4(r0, r1, l0, l1) : make_uint4(h0, h1, r0, r1);
        }
        sl = sl + 1 == TS_AST ? 0 : sl + 1;
    }
    cp_async_wait<0>();
}

// The CTA's last problem has no further tiles to hide its update under: the sixteen worker warps join the optimiser warps for
// it (ranks TS_OPT_WARPS ..).  With every SM in the same launch that burst is HBM-bound and the extra warps gain nothing
// (measured), but the engine's lanes run 2-4 launches side by side on a share of the SMs each, whose tails fall at different
// times.  NFOPP_TCS_DBG bit 7 (128) switches it off for launches with several problems per CTA.
__device__ __forceinline__ bool ts_join_last(const OnfTsArgs &a, int p) {
    if (a.P <= (int)gridDim.x) return true;
    return !(a.dbg & 128) && p + (int)gridDim.x >= a.P;
}

// the optimiser warps: problem after problem as the workers hand the gradients over.  (Letting the sixteen worker warps join
// in for the CTA's last problem, which has nothing to hide under, does not shorten it: all CTAs reach that point together and
// the 148 x 1.1 MB burst is HBM-bound - measured.)
__device__ __forceinline__ void ts_optimizer(const OnfTsArgs &a, const TsShared &s, int otid) {
    const nfopp_onf_layout &L = a.d.lay;
    const int wo = otid >> 5, lane = otid & 31;
    int it = 0;
    for (int p = blockIdx.x; p < a.P; p += gridDim.x) {
        const int cnt = a.count ? a.count[p] : a.n;
        if (cnt <= 0) continue;
        const int gb = it % TS_GBUF;
        const float *G = a.gscratch + ((size_t)blockIdx.x * 2 + gb) * L.n_params;
        if (otid == 0) {
            // this problem's Adam state: pull it into L2 now - the previous problem's update has just finished, the tiles
            // of this one are under way (measured: issued at the problem's first tile the lines do not survive in L2 until
            // they are used, issued at its last tile they arrive late)
            const size_t po = (size_t)p * L.n_params;
            const uint32_t bytes = (uint32_t)(L.n_params * sizeof(float)) & ~15u;
            bulk_prefetch_l2(a.adam_w + po, bytes);
            bulk_prefetch_l2(a.adam_m + po, bytes);
            bulk_prefetch_l2(a.adam_v + po, bytes);
        }
        ts_wait_relaxed(s.gfull + gb, (uint32_t)(it / TS_GBUF), 256);
        if (otid == 0) TS_TRACE(3, 1);
        // (a CTA with a single problem has nothing to overlap the update with: the worker warps join in, see the kernel)
#ifdef NFOPP_TC_TRACE
        if (!(a.dbg & 4))
#endif
        ts_adam_problem(a, G, p, wo, ts_join_last(a, p) ? TS_OPT_WARPS + TS_WORKERS / 32 : TS_OPT_WARPS, lane, s.aring + wo * TS_ARING);
        // every gradient value this warp needs is in its registers: once all warps say so the workers may overwrite the
        // hand-over buffer
        __syncwarp();
        if (lane == 0) mbar_arrive(s.gempty + gb);
        if (otid == 0) TS_TRACE(3, 2);
        ++it;
    }
}

// Thread layout: 24 warps.  Warps 0..15 are workers: warp = 4*cq + q, q = TMEM lane quarter (rows 32q..32q+31), cq = which
// 16-column chunks of the tile (cq, cq+4); a thread owns one hidden unit / feature (`row`).  Warp 16 issues every
// tcgen05.mma, warp 17 streams the weights, warps 18..23 are the optimiser.  The CTA is persistent: problems blockIdx.x,
// blockIdx.x + gridDim.x, ...; all mbarrier phase counters run on across tiles and problems.
__global__ void __launch_bounds__(TS_THREADS, 1) onf_tcs_kernel(const OnfTsArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const OnfDims &d = a.d;
    const TsDims &t = a.t;
    const nfopp_onf_layout &L = d.lay;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (a.sp && !a.warm && !a.sp->do_train) return;    // engine: not a field-training step
    const bool want_bwd = a.want_bwd != 0;
    const bool want_wgrad = want_bwd && (a.grad_out != nullptr || a.adam_w != nullptr);
    const bool want_gx = want_bwd && a.gx_out != nullptr;
    const bool fused_adam = want_bwd && a.adam_w != nullptr;
    const int F = t.F, FIN = t.FIN, H = t.H, K1 = t.K1, KH = t.KH;
    const bool is_issuer = warp == TS_WORKERS / 32, is_producer = warp == TS_WORKERS / 32 + 1;
    const bool is_worker = warp < TS_WORKERS / 32;

    TsShared s(smem_raw, t);
    TS_TRACE(0, 0);

    // ---- one-time setup: barriers, TMEM ----
    if (tid == 0) {
        for (int i = 0; i < TS_WST; ++i) { mbar_init(s.wfull + i, 1); mbar_init(s.wempty + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(s.efull + i, TS_WORKERS / 32); mbar_init(s.eempty + i, 1); mbar_init(s.bfull + i, 1); }
        mbar_init(s.bempty, 1);
        for (int i = 0; i < 2; ++i) { mbar_init(s.ready + i, TS_WORKERS / 32); mbar_init(s.done + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(s.gfull + i, 1); mbar_init(s.gempty + i, TS_OPT_WARPS); mbar_init(s.pfull + i, 1); }
        fence_mbar_init();
    }
    if (is_issuer) tmem_alloc(s.tmem_slot, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *s.tmem_slot;
    TS_TRACE(0, 1);

    // worker identity (the service warps compute harmless values)
    const int q = warp & 3, cq = (warp >> 2) & 3;
    const int row = q * 32 + lane;
    const uint32_t tm_lane = tmem + ((uint32_t)(q * 32) << 16);

    // register split (the CTA starts with 65536 / 768 -> 80 registers per thread): the two service warpgroups (issuer, weight
    // stream, optimiser) give registers back, the four worker warpgroups take them
    if (!is_worker) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(TS_REGS_SERVICE));
        if (is_issuer) {
            ts_issuer(a, s, tmem, want_bwd, want_wgrad);
        } else if (is_producer) {
            if (lane == 0) ts_producer(a, s, want_bwd);
        } else if (fused_adam) {
            ts_optimizer(a, s, tid - (TS_WORKERS + 64));
        }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(TS_REGS_WORKER));
        const uint32_t tA = tm_lane + TS_TM_A;
        uint32_t ndone0 = 0, ndone1 = 0, ec = 0;
        int it = 0;
        float *u0s = s.in0, *u1s = s.in0 + TS_NT, *ths = s.in0 + 2 * TS_NT;
        const int n_slots = (K1 + TS_EF - 1) / TS_EF;
        auto next_problem = [&](int q, int &c) {             // the next problem of this CTA with points (or >= a.P)
            c = 0;
            for (q += (int)gridDim.x; q < a.P; q += (int)gridDim.x)
                if ((c = (a.count ? a.count[q] : a.n)) > 0) break;
            return q;
        };
        // inputs (and label / upstream gradient) of point i of problem q, for thread tid < TS_NT; fetched one tile ahead
        float cx0 = 0.f, cx1 = 0.f, cx2 = 0.f, cax = 0.f;
        auto fetch_inputs = [&](int q, int i, int c) {
            cx0 = 0.f; cx1 = 0.f; cx2 = 0.f; cax = 0.f;
            if (tid < TS_NT && i < c) {
                const size_t gi = (size_t)q * a.n + i;
                const float *xp = a.x + gi * a.x_stride;
                cx0 = xp[0]; cx1 = xp[1];
                if (L.in_dim == 3) cx2 = xp[2];
                if (want_bwd) {
                    if (a.gbar_mode == 0) cax = a.gz[gi];
                    else if (a.gbar_mode != 2) cax = a.aux[gi];
                }
            }
        };
        int cnt = 0;
        int p = next_problem((int)blockIdx.x - (int)gridDim.x, cnt);
        if (p < a.P) fetch_inputs(p, tid, cnt);
        uint32_t np = 0;
        while (p < a.P) {
            int cnt_next;
            const int p_next = next_problem(p, cnt_next);
            if (np == 1) TS_TRACE(0, 29);
            TS_TRACE(0, 30);
            const int n_tiles = (cnt + TS_NT - 1) / TS_NT;
            // ---- per-problem setup: per-feature / per-unit fp32 parameters (raw values staged by the weight-stream warp),
            // gradient operand scale ----
            {
                ts_wait(s.pfull + (np & 1u), np >> 1);
                const float *R = s.praw + (size_t)(np & 1u) * t.praw_floats;      // [encoder | b1 | b2, w_out[:H] | b_out]
                const float *Rb1 = R + L.off_w[0], *Rb2 = Rb1 + H, *Rw3 = Rb2 + (L.off_w_out - L.off_b[1]), *Rb3 = Rw3 + H;
                for (int f = tid; f < K1; f += TS_WORKERS) {
                    float pa = 0.f, pb = 0.f, pc = 0.f;
                    if (f < F) {
                        pa = R[L.off_enc_w + 2 * f];
                        pb = R[L.off_enc_w + 2 * f + 1];
                        pc = R[L.off_enc_b + f] + ((d.use_cos && f >= 100) ? 1.57079632679489662f : 0.f);
                    } else if (f < FIN) {
                        const int k = f - F;
                        const float b = R[L.off_ang_b + k], fr = R[L.off_ang_f + k];
                        s.ang_b[k] = b;
                        pb = fr;
                        pc = fmaf(b, fr, k >= t.A / 2 ? 1.57079632679489662f : 0.f);
                    }
                    s.pa[f] = pa; s.pb[f] = pb; s.pc[f] = pc;
                }
                for (int i = tid; i < KH; i += TS_WORKERS) {
                    s.b1[i] = i < H ? Rb1[i] : 0.f;
                    s.b2[i] = i < H ? Rb2[i] : 0.f;
                    s.w3h[i] = i < H ? Rw3[i] : 0.f;
                }
                if (tid == 0) s.scal[0] = Rb3[0];
                ++np;
            }
            {
                // bound of |dL/dz| over the problem -> power-of-two scale of the fp16 gradient operands (one scale for all
                // tiles: the weight-gradient accumulators are shared): external gz / multipliers are scanned (a single tile's
                // values are already in the registers), BCE is bounded by 1 / N
                float m = 0.f;
                if (want_bwd && (a.gbar_mode == 0 || a.gbar_mode == 3)) {
                    if (n_tiles == 1) m = fabsf(cax);
                    else {
                        const float *src = (a.gbar_mode == 0 ? a.gz : a.aux) + (size_t)p * a.n;
                        for (int i = tid; i < cnt; i += TS_WORKERS) m = fmaxf(m, fabsf(src[i]));
                    }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                if (lane == 0) s.red[warp] = m;
            }
            ts_worker_sync();
            float S = 1.f;
            if (want_bwd) {
                float m = 0.f;
#pragma unroll
                for (int i = 0; i < TS_WORKERS / 32; ++i) m = fmaxf(m, s.red[i]);
                if (a.gbar_mode == 1) m = 1.f / (float)cnt;
                else if (a.gbar_mode == 2) m = fabsf(a.wc);
                else if (a.gbar_mode == 3) m += fabsf(a.wc);
                S = ts_pow2_scale(m);
            }
            const float invS = 1.f / S;
            const float my_b1 = row < H ? s.b1[row] : 0.f, my_b2 = row < H ? s.b2[row] : 0.f, my_w3 = row < H ? s.w3h[row] : 0.f;
            // thread-local gradient accumulators (hidden unit `row`; features row of block 0 / block 1; point tid)
            float acc_db1 = 0.f, acc_db2 = 0.f, acc_dw3 = 0.f, acc_db3 = 0.f;
            float fa0 = 0.f, fb0 = 0.f, fc0 = 0.f, fa1 = 0.f, fb1 = 0.f, fc1 = 0.f;
            TS_TRACE(0, 1);
            if (np == 1) TS_TRACE(0, 31);
            for (int tile = 0; tile < n_tiles; ++tile) {
                const int n0 = tile * TS_NT;
                const int nv = min(TS_NT, cnt - n0);
                const int Nt = ts_ru16(nv);
                const int nch = Nt / 16;
                // ---- P0: inputs of the tile (onf_model.py:37-38), labels / upstream gradients: fetched during the previous
                // tile; the loads of the next tile (of this or of the next problem) are issued right away ----
                if (tid < TS_NT) {
                    u0s[tid] = (cx0 - d.mean) / d.sigma;
                    u1s[tid] = (cx1 - d.mean) / d.sigma;
                    ths[tid] = cx2;
                    s.aux[tid] = cax;
                }
                if (tile + 1 < n_tiles) fetch_inputs(p, n0 + TS_NT + tid, cnt);
                else if (p_next < a.P) fetch_inputs(p_next, tid, cnt_next);
                ts_worker_sync();
                TS_TRACE(0, 2);
                if (np == 1 && tile == 0) TS_TRACE(0, 32);
                // ---- P1: encoding (CUDA cores) feeding fwd1 in slots of 64 features: thread = (point, feature octets g, g+4) ----
                {
                    const int g = tid / Nt, n = tid - g * Nt;
                    const bool e_on = g < 4;
                    const float u0 = e_on ? u0s[n] : 0.f, u1 = e_on ? u1s[n] : 0.f, th = e_on ? ths[n] : 0.f;
    #pragma unroll 1
                    for (int sl = 0; sl < n_slots; ++sl, ++ec) {
                        const uint32_t es = ec & 1u;
                        const int fa = TS_EF * sl + 8 * g, fb = fa + 32;
                        const bool on_a = e_on && fa < K1, on_b = e_on && fb < K1;
                        uint4 ah = {}, al = {}, bh = {}, bl = {};
                        if (on_a) ts_encode8(s, fa, F, u0, u1, th, ah, al);
                        if (on_b) ts_encode8(s, fb, F, u0, u1, th, bh, bl);
                        if (ec >= 2) ts_wait(s.eempty + es, (ec >> 1) - 1);        // fwd1 of the slot's previous use is done
                        unsigned char *eb = s.r + es * 2 * t.e_half + g * TS_SPT + n * 16;
                        if (on_a) {
                            *reinterpret_cast<uint4 *>(eb) = ah;
                            *reinterpret_cast<uint4 *>(eb + t.e_half) = al;
                        }
                        if (on_b) {
                            *reinterpret_cast<uint4 *>(eb + 4 * TS_SPT) = bh;
                            *reinterpret_cast<uint4 *>(eb + 4 * TS_SPT + t.e_half) = bl;
                        }
                        ts_signal(s.efull + es);
                    }
                }
                TS_TRACE(0, 3);
                if (np == 1 && tile == 0) TS_TRACE(0, 33);
                // From here on the tile is two halves (points 0..63 | 64..Nt-1) with their own done / ready barriers: a warp owns
                // ONE 16-point chunk per half (chunk 4 h + cq), and the epilogue of one half runs under the MMAs of the other.
                const int nh = (want_bwd && Nt > TS_HALF) ? 2 : 1;
                // chunks of half h for this warp: 4 h + cq (two halves) or cq, cq + 4 (one)
    #define TS_HALF_CHUNKS(ch) for (int ch = 4 * h + cq, ch##_end = (nh == 2 && h == 0) ? 4 : nch; ch < ch##_end; ch += 4)
                // ---- P2: epilogue 1: bias + ReLU -> H1 operand (hi, lo); row H = skip term of the logit ----
    #pragma unroll 1
                for (int h = 0; h < nh; ++h) {
                    ts_wait(s.done + h, (h ? ndone1++ : ndone0++));
                    TS_TRACE(0, 4);
                    if (np == 1 && tile == 0 && h == 0) TS_TRACE(0, 34);
                    TS_HALF_CHUNKS(ch) {
                        const int c0 = 16 * ch;
                        float v[16];
                        tmem_ld16(tA + c0, v);
                        if (row < H) {
    #pragma unroll
                            for (int i = 0; i < 16; ++i) v[i] = fminf(fmaxf(v[i] + my_b1, 0.f), 60000.f);
                        } else {
                            if (row == H) {
    #pragma unroll
                                for (int i = 0; i < 16; ++i) s.skip[c0 + i] = v[i];
                            }
    #pragma unroll
                            for (int i = 0; i < 16; ++i) v[i] = 0.f;
                        }
                        if (row < KH) ts_store_row16(s.h1h, s.h1l, row, c0, v);
                    }
                    ts_signal(s.ready + h);                   // -> fwd2 of this half
                }
                TS_TRACE(0, 5);
                if (np == 1 && tile == 0) TS_TRACE(0, 35);
                // ---- P3 / P4: epilogue 2: logit partial sums over the hidden units of each lane quarter, z and dL/dz per
                // point, then gH2 = dL/dz * w3 * 1[h2 > 0] (scaled, fp16 hi + lo) -> old encoding ring ----
    #pragma unroll 1
                for (int h = 0; h < nh; ++h) {
                    ts_wait(s.done + h, (h ? ndone1++ : ndone0++));
                    TS_TRACE(0, 6);
                    TS_HALF_CHUNKS(ch) {
                        const int c0 = 16 * ch;
                        float pz[16];
                        tmem_ld16(tA + c0, pz);
    #pragma unroll
                        for (int i = 0; i < 16; ++i) pz[i] = row < H ? my_w3 * fmaxf(pz[i] + my_b2, 0.f) : 0.f;
                        const float zs = ts_colsum16(pz, lane);
                        if ((lane & 1) == 0) s.zp[q * TS_NT + c0 + (lane >> 1)] = zs;
                    }
                    ts_worker_sync();
                    TS_TRACE(0, 7);
                    if (tid >= h * TS_HALF && tid < ((nh == 2 && h == 0) ? TS_HALF : Nt)) {
                        const int n = tid;
                        float gb = 0.f;
                        if (n < nv) {
                            const float z = s.zp[n] + s.zp[TS_NT + n] + s.zp[2 * TS_NT + n] + s.zp[3 * TS_NT + n] + s.skip[n] + s.scal[0];
                            const size_t gi = (size_t)p * a.n + n0 + n;
                            if (a.z_out) a.z_out[gi] = z;
                            if (want_bwd) {                                 // dL/dz
                                const float ax = s.aux[n];
                                if (a.gbar_mode == 0) gb = ax;
                                else if (a.gbar_mode == 1) gb = (sigmoid_f(z) - ax) / (float)cnt;
                                else {
                                    const float zb = z * a.beta;
                                    const float ez = expf(zb);
                                    gb = a.wc * (zb > 20.f ? 1.f : ez / (ez + 1.f));
                                    if (a.gbar_mode == 3) {
                                        const float tz = tanhf(z);
                                        gb += ax * (1.f - tz * tz);
                                    }
                                }
                            }
                        }
                        s.gbar[n] = gb;
                        acc_db3 += gb;
                    }
                    if (!want_bwd) continue;
                    ts_worker_sync();
                    TS_TRACE(0, 8);
                    TS_HALF_CHUNKS(ch) {
                        const int c0 = 16 * ch;
                        float r2[16], v[16];
                        tmem_ld16(tA + c0, r2);
    #pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float h2 = row < H ? fmaxf(r2[i] + my_b2, 0.f) : 0.f;
                            const float gb = s.gbar[c0 + i];
                            acc_dw3 = fmaf(gb, h2, acc_dw3);
                            const float gg = h2 > 0.f ? gb * my_w3 : 0.f;
                            acc_db2 += gg;
                            v[i] = gg * S;
                        }
                        if (row < KH) ts_store_row16(s.r, s.r + t.h1_half, row, c0, v);
                    }
                    ts_signal(s.ready + h);                   // -> wgrad2 + dgrad2 of this half
                }
                if (!want_bwd) {
                    ts_worker_sync();                         // zp / skip / inputs are rewritten by the next tile
                    continue;
                }
                TS_TRACE(0, 9);
                // ---- P5: epilogue 3: ReLU mask of layer 1 -> gH1 operand (in place over H1 hi, lo); row H := dL/dz ----
    #pragma unroll 1
                for (int h = 0; h < nh; ++h) {
                    ts_wait(s.done + h, (h ? ndone1++ : ndone0++));
                    TS_TRACE(0, 10);
                    TS_HALF_CHUNKS(ch) {
                        const int c0 = 16 * ch;
                        float v[16];
                        tmem_ld16(tA + c0, v);
                        if (row < H) {
                            const int off = (row >> 3) * TS_SPT + (c0 >> 3) * 128 + (row & 7) * 16;
                            const uint4 m0 = *reinterpret_cast<const uint4 *>(s.h1h + off);
                            const uint4 m1 = *reinterpret_cast<const uint4 *>(s.h1h + off + 128);
                            const uint32_t mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
                            float sum = 0.f;
    #pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                const uint32_t bits = (mw[i >> 1] >> ((i & 1) * 16)) & 0x7FFFu;
                                v[i] = bits != 0 ? fminf(fmaxf(v[i], -60000.f), 60000.f) : 0.f;
                                sum += v[i];
                            }
                            acc_db1 = fmaf(sum, invS, acc_db1);
                        } else if (row == H) {
    #pragma unroll
                            for (int i = 0; i < 16; ++i) v[i] = s.gbar[c0 + i] * S;
                        } else {
    #pragma unroll
                            for (int i = 0; i < 16; ++i) v[i] = 0.f;
                        }
                        if (row < KH) ts_store_row16(s.h1h, s.h1l, row, c0, v);
                    }
                    ts_signal(s.ready + h);                   // -> dgrad1 (block 0) of this half
                }
                TS_TRACE(0, 11);
                // ---- P6: epilogue 4 per feature block: encoding backward (+ input gradient), E^T operand (hi, lo), wgrad1 ----
    #pragma unroll 1
                for (int mb = 0; mb < t.MB; ++mb) {
                    const int fb = mb * t.FB0;
                    const int pf = min(t.FB0, K1 - fb);       // features of this block that are MMA rows
                    const int f = fb + row;
                    const bool f_on = row < pf;
                    const bool fourier = f < F;
                    const float ppa = f_on ? s.pa[f] : 0.f, ppb = f_on ? s.pb[f] : 0.f, ppc = f_on ? s.pc[f] : 0.f;
                    const float ew0 = (f_on && fourier) ? ppa : 0.f, ew1 = (f_on && fourier) ? ppb : 0.f,
                                eaf = (f_on && !fourier && f < FIN) ? ppb : 0.f;
                    const float *x1src = fourier ? u1s : ths;
                    float ta = 0.f, tb = 0.f, tc = 0.f;      // sum gp*u0, sum gp*x1, sum gp  (all still scaled by S)
    #pragma unroll 1
                    for (int h = 0; h < nh; ++h) {
                        ts_wait(s.done + h, (h ? ndone1++ : ndone0++));      // gE of block mb (and wgrad1 of block mb - 1: E^T is free again)
                        TS_TRACE(0, 12 + 2 * mb);
                        TS_HALF_CHUNKS(ch) {
                            const int c0 = 16 * ch;
                            float v[16], ev[16], xu[16], x1[16];
                            tmem_ld16(tA + c0, v);
    #pragma unroll
                            for (int i4 = 0; i4 < 4; ++i4) {
                                *reinterpret_cast<float4 *>(xu + 4 * i4) = *reinterpret_cast<const float4 *>(u0s + c0 + 4 * i4);
                                *reinterpret_cast<float4 *>(x1 + 4 * i4) = *reinterpret_cast<const float4 *>(x1src + c0 + 4 * i4);
                            }
    #pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                const float arg = fmaf(x1[i], ppb, fmaf(xu[i], ppa, ppc));
                                const float gp = (f_on ? v[i] : 0.f) * __cosf(arg);      // rows beyond the block hold garbage
                                ev[i] = __sinf(arg);                                       // (= 0 there: pa = pb = pc = 0)
                                v[i] = gp;
                                ta = fmaf(gp, xu[i], ta);
                                tb = fmaf(gp, x1[i], tb);
                                tc += gp;
                            }
                            if (want_gx) {
                                // gx[n] = (sum_f gp[f,n] * We[f,c]) / sigma ; g_theta[n] = sum_k gp[k,n] * f_k   (reduce over lanes)
                                float cv[16];
                                float *dst = s.gxp + ((size_t)q * TS_NT + c0 + (lane >> 1)) * 3;
    #pragma unroll
                                for (int c = 0; c < 3; ++c) {
                                    const float wgt = c == 0 ? ew0 : c == 1 ? ew1 : eaf;
    #pragma unroll
                                    for (int i = 0; i < 16; ++i) cv[i] = v[i] * wgt;
                                    const float sum = ts_colsum16(cv, lane);
                                    if ((lane & 1) == 0) dst[c] = mb == 0 ? sum : dst[c] + sum;
                                }
                            }
                            if (want_wgrad && row < t.FB0) {     // E^T operand [point / 8][feature][8 points], hi and lo
    #pragma unroll
                                for (int o = 0; o < 2; ++o) {
                                    const int off = ((c0 >> 3) + o) * t.S_ET + (row >> 3) * 128 + (row & 7) * 16;
                                    uint4 eh, el;
                                    split8h(ev + 8 * o, eh, el);
                                    *reinterpret_cast<uint4 *>(s.r + off) = eh;
                                    *reinterpret_cast<uint4 *>(s.r + t.et_half + off) = el;
                                }
                            }
                        }
                        ts_signal(s.ready + h);               // -> wgrad1(mb), dgrad1(mb + 1) of this half
                    }
                    ta *= invS; tb *= invS; tc *= invS;
                    if (f_on && !fourier) {                   // heading features: d ang_b = f_k * sum gp,
                        const float b = f < FIN ? s.ang_b[f - F] : 0.f;   // d ang_f = sum gp * (theta + b_k)  (trained)
                        tb = fmaf(b, tc, tb);
                        ta = eaf * tc;
                    }
                    if (mb == 0) { fa0 += ta; fb0 += tb; fc0 += tc; }
                    else { fa1 += ta; fb1 += tb; fc1 += tc; }
                    TS_TRACE(0, 13 + 2 * mb);
                }
                if (want_gx) {
                    ts_worker_sync();
                    if (tid < nv) {
                        const int n = tid;
                        float g0 = 0.f, g1 = 0.f, g2 = 0.f;
    #pragma unroll
                        for (int qq = 0; qq < 4; ++qq) {
                            const float *src = s.gxp + ((size_t)qq * TS_NT + n) * 3;
                            g0 += src[0]; g1 += src[1]; g2 += src[2];
                        }
                        float *out = a.gx_out + ((size_t)p * a.n + n0 + n) * a.x_stride;
                        out[0] = g0 * invS / d.sigma;
                        out[1] = g1 * invS / d.sigma;
                        if (L.in_dim == 3) out[2] = g2 * invS;
                        for (int c = L.in_dim; c < a.x_stride; ++c) out[c] = 0.f;   // pose columns the ONF does not read
                    }
                }
                if (want_wgrad)                               // the last wgrad1: E^T (= the encoding ring) and gH1 are free again
                    for (int h = 0; h < nh; ++h) ts_wait(s.done + h, (h ? ndone1++ : ndone0++));
                TS_TRACE(0, 16);
                ts_worker_sync();                             // zp, gx, skip, gbar, inputs are rewritten by the next tile
            }
            TS_TRACE(0, 20);
            if (!want_wgrad) { p = p_next; cnt = cnt_next; continue; }
            // ---- weight gradients: TMEM accumulators -> global hand-over buffer (blob order) -> optimiser warps / caller ----
            // every MMA of the problem has completed (the wait after the last wgrad1): the encoding ring is dead until the next
            // problem's encoding and holds the per-unit partial sums [10][4][128]
            {
                float *red = reinterpret_cast<float *>(smem_raw + t.off_red);
                float *r0 = red + cq * 128 + row;
                r0[0 * 512] = acc_db1; r0[1 * 512] = acc_db2; r0[2 * 512] = acc_dw3;
                r0[3 * 512] = fa0; r0[4 * 512] = fb0; r0[5 * 512] = fc0;
                r0[6 * 512] = fa1; r0[7 * 512] = fb1; r0[8 * 512] = fc1;
                if (tid < 128) {                                           // (acc_db3 lives in threads 0 .. TS_NT-1, zero elsewhere)
                    const float s3 = warp_sum(tid < TS_NT ? acc_db3 : 0.f);
                    if (lane == 0) red[9 * 512 + warp] = s3;
                }
                const int gb = it % TS_GBUF;
                float *G = fused_adam ? a.gscratch + ((size_t)blockIdx.x * 2 + gb) * L.n_params : a.grad_out + (size_t)p * L.n_params;
                if (fused_adam && it >= TS_GBUF) ts_wait(s.gempty + gb, (uint32_t)(it / TS_GBUF - 1));   // the optimiser has read this buffer
                TS_TRACE(0, 22);
                {
                    // everything outside W1 / W2 / the skip row: zero first (alignment padding stays zero), totals below
                    int moff[4], mlen[4];
                    ts_misc_ranges(t, L, moff, mlen);
#pragma unroll
                    for (int r = 0; r < 4; ++r)
                        for (int i = tid; i < mlen[r]; i += TS_WORKERS) G[moff[r] + i] = 0.f;
                }
                ts_worker_sync();
                // dW1: rows j < H -> mlp.0.weight[j,:], row H -> mlp2.0.weight[0, H:];  dW2: rows i < H -> mlp.2.weight[i,:].
                // A thread holds 16 consecutive columns of ONE row (TMEM lane): 64 columns at a time go through a shared-memory
                // buffer so that the global stores run along the rows (a warp writes 2 x 256 contiguous bytes per instruction
                // instead of 32 x 16)
                float *S = reinterpret_cast<float *>(smem_raw + t.off_stage);
                auto drain = [&](uint32_t tm_col, int ncols, int width, bool first) {
                    for (int c0 = 0; c0 < ncols; c0 += 64) {
                        const int cl = c0 + 16 * cq;
                        TS_ASSERT(t.off_stage + (row * TS_SROW + 16 * cq + 16) * 4 <= t.off_red && t.off_red + 10 * 512 * 4 <= t.off_w);
                        if (cl < ncols) {
                            float g[16];
                            tmem_ld16(tm_lane + tm_col + cl, g);
#pragma unroll
                            for (int c4 = 0; c4 < 4; ++c4)
                                *reinterpret_cast<float4 *>(S + row * TS_SROW + 16 * cq + 4 * c4) =
                                    make_float4(g[4 * c4] * invS, g[4 * c4 + 1] * invS, g[4 * c4 + 2] * invS, g[4 * c4 + 3] * invS);
                        }
                        ts_worker_sync();
#pragma unroll
                        for (int k2 = 0; k2 < 4; ++k2) {
                            const int r = 8 * warp + 2 * k2 + (lane >> 4), c = c0 + 4 * (lane & 15);
                            const bool row_on = first ? r <= H : r < H;
                            if (row_on && c < width) {
                                float *dst = G + (first ? (r < H ? L.off_w[0] + r * FIN : L.off_w_out + H) : L.off_w[1] + r * H) + c;
                                *reinterpret_cast<float4 *>(dst) = *reinterpret_cast<const float4 *>(S + r * TS_SROW + 4 * (lane & 15));
                            }
                        }
                        ts_worker_sync();
                    }
                };
                TS_TRACE(0, 24);
                drain(TS_TM_DW1, K1, FIN, true);
                TS_TRACE(0, 25);
                drain(TS_TM_DW2, KH, H, false);
                TS_TRACE(0, 26);
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
                        const int f = mb * t.FB0 + tid;
                        const bool on = tid < min(t.FB0, K1 - mb * t.FB0);
                        const float ga = tot(3 + 3 * mb), gb = tot(4 + 3 * mb), gc = tot(5 + 3 * mb);
                        if (on && f < F) {
                            G[L.off_enc_w + 2 * f] = ga;
                            G[L.off_enc_w + 2 * f + 1] = gb;
                            G[L.off_enc_b + f] = d.bias ? gc : 0.f;
                        } else if (on && f < FIN) {
                            G[L.off_ang_b + f - F] = ga;
                            G[L.off_ang_f + f - F] = gb;
                        }
                    }
                    if (tid == 0) {
                        float g3 = 0.f;
                        for (int i = 0; i < 4; ++i) g3 += red[9 * 512 + i];
                        G[L.off_b_out] = g3;
                    }
                }
                fence_before_sync();
                ts_worker_sync();                         // (also: `red` and the per-problem parameters are free again)
                if (fused_adam && tid == 0) mbar_arrive(s.gfull + gb);
                // one problem per CTA (small batches, the single-problem planner API): nothing left to overlap, all 22 warps
                // share the update.  (With several problems per CTA the last update is HBM-bound - every CTA reaches it at the
                // same time - and the extra warps measured no gain.)
                // (their cp.async rings sit in the encoding-ring / weight-ring regions: no further problem, nothing streams)
                if (fused_adam && ts_join_last(a, p))
                    ts_adam_problem(a, G, p, TS_OPT_WARPS + warp, TS_OPT_WARPS + TS_WORKERS / 32, lane, s.r + warp * TS_ARING);
                ++it;
            }
            TS_TRACE(0, 27);
            p = p_next; cnt = cnt_next;
        }
    }
    fence_before_sync();
    __syncthreads();
    TS_TRACE(0, 28);
    if (is_issuer) tmem_dealloc(tmem, 512);
}

// ---- host side ---------------------------------------------------------------------------------------------
bool onf_tcs_supported(const OnfDims &d) {
    TsDims t;
    return ts_make_dims(d, &t);
}

static int g_tcs_sms = 0;

int onf_tcs_prepare(const OnfDims &d) {
    TsDims t;
    NFOPP_REQUIRE(ts_make_dims(d, &t), "onf(tensor, streamed): needs 2 hidden layers, hidden <= 111, f_in <= 256 (hidden=%d layers=%d)",
                  d.H, d.L);
    NFOPP_CHECK_CUDA(cudaFuncSetAttribute(onf_tcs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, t.smem_bytes));
    int dev = 0;
    NFOPP_CHECK_CUDA(cudaGetDevice(&dev));
    NFOPP_CHECK_CUDA(cudaDeviceGetAttribute(&g_tcs_sms, cudaDevAttrMultiProcessorCount, dev));
    return NFOPP_OK;
}

// persistent grid: one CTA per SM; launches that are meant to run side by side (the engine's lanes) take an equal share of
// the SMs each, so that they do run side by side instead of queueing CTA by CTA behind one another (measured, 1024 corridor
// problems: 0.860 ms/step with 2 lanes x 148 CTAs, 0.844 with 2 x 74, 0.838 with 4 x 37).  NFOPP_TCS_GRID overrides the cap
// (0 = one CTA per problem).
int onf_tcs_grid(int P, int slots, int sm_share) {
    const int sms = g_tcs_sms > 0 ? g_tcs_sms : 148;
    int cap = sm_share > 1 ? sms / sm_share : sms;
    if (const char *e = getenv("NFOPP_TCS_GRID")) {
        const int v = atoi(e);
        cap = v > 0 ? v : P;
    }
    if (cap < 1) cap = 1;
    int grid = P < cap ? P : cap;
    if (slots > 0 && grid > slots) grid = slots;
    return grid < 1 ? 1 : grid;
}

int onf_tcs_pack(const OnfDims &d, const float *weights, unsigned char *packed, int P, cudaStream_t stream) {
    TsPackArgs a{};
    a.d = d;
    NFOPP_REQUIRE(ts_make_dims(d, &a.t), "onf(tensor, streamed): unsupported network shape");
    a.weights = weights; a.packed = packed;
    if (P == 0) return NFOPP_OK;
    const int items = (a.t.K1 / 8 + a.t.KH / 8) * a.t.KH;
    dim3 grid((items + 255) / 256, P);
    onf_tcs_pack_kernel<<<grid, 256, 0, stream>>>(a);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

int onf_tcs_launch(const OnfArgs &o, const OnfTcExtra &x, int P, cudaStream_t stream) {
    OnfTsArgs a{};
    a.d = o.d;
    NFOPP_REQUIRE(ts_make_dims(o.d, &a.t), "onf(tensor, streamed): unsupported network shape");
    a.weights = o.weights; a.packed = x.packed; a.x = o.x; a.count = o.count; a.sp = o.sp;
    a.P = P; a.n = o.n; a.x_stride = o.x_stride ? o.x_stride : o.d.lay.in_dim;
    a.want_bwd = o.want_bwd; a.gbar_mode = o.gbar_mode; a.gz = o.gz; a.aux = o.aux; a.wc = o.wc; a.beta = o.beta;
    a.z_out = o.z_out; a.gx_out = o.gx_out; a.grad_out = x.grad_out;
    a.adam_w = x.adam_w; a.adam_m = x.adam_m; a.adam_v = x.adam_v; a.packed_out = x.packed_out; a.gscratch = x.gscratch;
    a.w1 = x.w1; a.beta2 = x.beta2; a.w2 = x.w2; a.eps = x.eps; a.warm = x.warm;
    if (const char *e = getenv("NFOPP_TCS_DBG")) a.dbg = atoi(e);
    NFOPP_REQUIRE(a.packed, "onf(tensor): packed operand copy missing");
    NFOPP_REQUIRE(!a.adam_w || (a.sp && a.packed_out), "onf(tensor): fused Adam needs step parameters and the packed copy");
    NFOPP_REQUIRE(!a.adam_w || (a.gscratch && x.gscratch_slots > 0), "onf(tensor, streamed): fused Adam needs the gradient hand-over buffers");
    if (a.n == 0 || P == 0) return NFOPP_OK;
    const int grid = onf_tcs_grid(P, a.adam_w ? x.gscratch_slots : 0, x.sm_share);
    onf_tcs_kernel<<<grid, TS_THREADS, a.t.smem_bytes, stream>>>(a);
    NFOPP_CHECK_CUDA(cudaGetLastError());
    return NFOPP_OK;
}

}  // namespace nfopp

#ifdef NFOPP_TC_TRACE
extern "C" int nfopp_debug_tcs_trace(long long *host_out) {
    return (int)cudaMemcpyFromSymbol(host_out, nfopp::g_ts_trace, sizeof(nfopp::g_ts_trace));
}
#endif
