// umma_probe.cu - stand-alone check (run on a B200) of the tcgen05 conventions onf_tc.cu relies on:
// SWIZZLE_NONE core-block operand storage viewed K-major and MN-major, descriptor advance per K = 16 step,
// bf16 hi/lo three-pass products accumulated in TMEM, tcgen05.ld read-back.  Each case mirrors one GEMM of
// the fused ONF kernel (fwd1, fwd2, wgrad2, dgrad2, dgrad1-second-M-block).  Prints max errors; exit 0 = pass.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 umma_probe.cu -o build/umma_probe
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "tc05.cuh"

using namespace tc05;

struct Operand {
    int mn_major;         // 0: K-major view (r = mn, c = k); 1: MN-major view (r = k, c = mn)
    int S_row, S_col;     // core-block strides in bytes
    int mn_valid;         // logical rows available in the source matrix
    int mn_start;         // first logical mn index the MMA starts at (multiple of 8)
};
struct Case {
    Operand a, b;
    int N, K;             // MMA N (multiple of 16), K (multiple of 16)
    const float *A, *B;   // A [a.mn_valid][K], B [b.mn_valid][K] row-major fp32
    float *D;             // [128][N]
};

constexpr int BUF = 48 * 1024;

__device__ void pack_operand(unsigned char *hi, unsigned char *lo, const Operand &o, const float *src, int K) {
    for (int idx = threadIdx.x; idx < o.mn_valid * K; idx += blockDim.x) {
        int mn = idx / K, k = idx % K;
        int r = o.mn_major ? k : mn, c = o.mn_major ? mn : k;
        int off = (r / 8) * o.S_row + (c / 8) * o.S_col + (r % 8) * 16 + (c % 8) * 2;
        __nv_bfloat16 h, l;
        split_bf16(src[idx], h, l);
        *reinterpret_cast<__nv_bfloat16 *>(hi + off) = h;
        *reinterpret_cast<__nv_bfloat16 *>(lo + off) = l;
    }
}

__global__ void __launch_bounds__(128, 1) probe_kernel(Case cs) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    unsigned char *a_hi = smem, *a_lo = smem + BUF, *b_hi = smem + 2 * BUF, *b_lo = smem + 3 * BUF;
    for (int i = threadIdx.x; i < 4 * BUF / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0xFFFFFFFFu;
    __syncthreads();
    pack_operand(a_hi, a_lo, cs.a, cs.A, cs.K);
    pack_operand(b_hi, b_lo, cs.b, cs.B, cs.K);
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    if (threadIdx.x < 32) tmem_alloc(&tmem_base_s, 256);
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_base_s;
    if (threadIdx.x == 0) {
        const uint32_t idesc = instr_desc_bf16(128, cs.N, cs.a.mn_major, cs.b.mn_major);
        auto start = [](unsigned char *buf, const Operand &o) {
            // offset of logical index mn_start along the mn dimension
            uint32_t oct = o.mn_start / 8;
            return smem_u32(buf) + oct * (o.mn_major ? o.S_col : o.S_row);
        };
        auto kstep = [](const Operand &o) { return 2u * (o.mn_major ? o.S_row : o.S_col); };
        auto mk = [](uint32_t addr, const Operand &o) {
            return o.mn_major ? desc_mnmajor(addr, o.S_row, o.S_col) : desc_kmajor(addr, o.S_row, o.S_col);
        };
        for (int ks = 0; ks < cs.K / 16; ++ks) {
            uint64_t ah = mk(start(a_hi, cs.a) + ks * kstep(cs.a), cs.a), al = mk(start(a_lo, cs.a) + ks * kstep(cs.a), cs.a);
            uint64_t bh = mk(start(b_hi, cs.b) + ks * kstep(cs.b), cs.b), bl = mk(start(b_lo, cs.b) + ks * kstep(cs.b), cs.b);
            mma_bf16(tmem, ah, bh, idesc, ks > 0);
            mma_bf16(tmem, al, bh, idesc, 1);
            mma_bf16(tmem, ah, bl, idesc, 1);
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    fence_after_sync();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int c0 = 0; c0 < cs.N; c0 += 16) {
        float v[16];
        tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
        for (int i = 0; i < 16; ++i) cs.D[(size_t)(warp * 32 + lane) * cs.N + c0 + i] = v[i];
    }
    fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(tmem, 256);
}

static float frand() { return (float)rand() / RAND_MAX * 2.f - 1.f; }

static int run_case(const char *name, Operand a, Operand b, int N, int K) {
    std::vector<float> A((size_t)a.mn_valid * K), B((size_t)b.mn_valid * K), D(128 * (size_t)N);
    for (auto &x : A) x = frand();
    for (auto &x : B) x = frand();
    Case cs{};
    cs.a = a; cs.b = b; cs.N = N; cs.K = K;
    float *dA, *dB, *dD;
    cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, D.size() * 4);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, D.size() * 4);
    cs.A = dA; cs.B = dB; cs.D = dD;
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * BUF);
    probe_kernel<<<1, 128, 4 * BUF>>>(cs);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        printf("%-28s CUDA error: %s\n", name, cudaGetErrorString(e));
        return 1;
    }
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    double worst = 0, scale = 0;
    int rows = a.mn_valid - a.mn_start; if (rows > 128) rows = 128;
    int cols = b.mn_valid - b.mn_start; if (cols > N) cols = N;
    for (int m = 0; m < rows; ++m)
        for (int n = 0; n < cols; ++n) {
            double ref = 0;
            for (int k = 0; k < K; ++k) ref += (double)A[(size_t)(a.mn_start + m) * K + k] * B[(size_t)(b.mn_start + n) * K + k];
            double d = fabs(ref - D[(size_t)m * N + n]);
            if (!(d <= worst)) worst = d;     // NaN-propagating max
            if (fabs(ref) > scale) scale = fabs(ref);
        }
    int bad = !(worst <= 2e-4 * scale);
    printf("%-28s M=128 N=%3d K=%3d rows=%3d cols=%3d  max|err|=%.3e  max|ref|=%.3e  %s\n", name, N, K, rows, cols, worst,
           scale, bad ? "FAIL" : "ok");
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
    return bad;
}

int main() {
    srand(1);
    int bad = 0;
    // weights W [rows j][K k] stored K-major with 104 allocated rows: S_row = 128, S_col = 104*16
    Operand W1{0, 128, 104 * 16, 101, 0};
    // activations as B (N = point, K = feature), k contiguous: rows = n, S_row = 128, S_col = 1024 (64 points)
    Operand E{0, 128, 1024, 64, 0};
    bad += run_case("fwd1  A=K-major B=K-major", W1, E, 64, 224);
    // H1 stored [j/8][n/8][j%8][n%8]: r = j (K), c = n (N): MN-major view, S_row = 1024, S_col = 128
    Operand W2{0, 128, 104 * 16, 100, 0};
    Operand H1{1, 1024, 128, 64, 0};
    bad += run_case("fwd2  A=K-major B=MN-major", W2, H1, 64, 112);
    bad += run_case("fwd2  (N=48)", W2, Operand{1, 1024, 128, 48, 0}, 48, 112);
    // wgrad2: A = gH2 (M = i, K = n) K-major view of [i/8][n/8][i%8][n%8]; B = H1 (N = j, K = n) K-major view
    Operand gA{0, 1024, 128, 112, 0};
    Operand hB{0, 1024, 128, 112, 0};
    bad += run_case("wgrad A=K-major B=K-major", gA, hB, 112, 64);
    // dgrad2: A (M = j, K = i) = MN-major view of W2 storage (r = i, c = j); B = gH2 (N = n, K = i) MN-major
    Operand W2T{1, 128, 104 * 16, 112, 0};
    Operand gB{1, 1024, 128, 64, 0};
    bad += run_case("dgrad A=MN-major B=MN-major", W2T, gB, 64, 96);
    // dgrad1, second M block: mn (feature) starts at 128, 224 valid
    Operand W1T{1, 128, 104 * 16, 224, 128};
    bad += run_case("dgrad (M block 1)", W1T, gB, 64, 96);
    printf(bad ? "PROBE FAILED (%d)\n" : "PROBE OK\n", bad);
    return bad ? 1 : 0;
}
