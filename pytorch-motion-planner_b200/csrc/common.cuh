// common.cuh - shared host/device helpers for libnfopp (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/nfopp.h"

namespace nfopp {

// ---- error plumbing -----------------------------------------------------------------------------
void set_error(const char *fmt, ...);
#define NFOPP_CHECK_CUDA(expr)                                                                  \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess) {                                                                \
            ::nfopp::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,             \
                               cudaGetErrorString(_e));                                         \
            return NFOPP_ERR_CUDA;                                                              \
        }                                                                                       \
    } while (0)
#define NFOPP_REQUIRE(cond, ...)                                                                \
    do {                                                                                        \
        if (!(cond)) {                                                                          \
            ::nfopp::set_error(__VA_ARGS__);                                                    \
            return NFOPP_ERR_INVALID;                                                           \
        }                                                                                       \
    } while (0)

// ---- ONF dimensions handed to kernels by value ---------------------------------------------------
struct OnfDims {
    nfopp_onf_layout lay;
    float mean, sigma;
    int use_cos, angle, H, L, bias;
};
int make_onf_dims(const nfopp_onf_config *cfg, OnfDims *out);

// ---- small device helpers ----------------------------------------------------------------------
#define NFOPP_PI_F 3.14159274101257324f      // float32(np.pi), what torch uses for `angles + np.pi`
#define NFOPP_2PI_F 6.28318548202514648f     // float32(2*np.pi)

// torch_math.py:5-7 wrap_angle in float32: (a + pi) % (2 pi) - pi with python-sign modulo
__device__ __forceinline__ float wrap_angle_f(float a) {
    float s = __fadd_rn(a, NFOPP_PI_F);
    float r = fmodf(s, NFOPP_2PI_F);
    if (r != 0.f && r < 0.f) r = __fadd_rn(r, NFOPP_2PI_F);
    return __fsub_rn(r, NFOPP_PI_F);
}

__device__ __forceinline__ float sigmoid_f(float z) { return 1.f / (1.f + expf(-z)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- Philox4x32-10 counter-based RNG (throughput mode; one stream per problem/step/purpose) -----
struct Philox {
    uint32_t key[2];
    __device__ __forceinline__ Philox(uint64_t seed) {
        key[0] = (uint32_t)seed;
        key[1] = (uint32_t)(seed >> 32);
    }
    __device__ __forceinline__ uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
        uint32_t k0 = key[0], k1 = key[1];
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
            c0 = n0; c1 = n1; c2 = n2; c3 = n3;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};
// 53-bit uniform in [0,1) exactly as numpy's legacy random_sample builds it from two 32-bit words
__device__ __forceinline__ double u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) / 9007199254740992.0;
}
// stream ids
enum { RNG_T_SAMPLE = 0, RNG_COARSE = 1, RNG_FINE = 2, RNG_CHOICE = 3, RNG_FIELD = 4, RNG_T_PATH = 5, RNG_WARM = 6 };
// two uniforms for (problem, step, stream, index)
__device__ __forceinline__ void rng_uniform2(uint64_t seed, uint32_t problem, uint32_t step, uint32_t stream,
                                             uint32_t index, double *u0, double *u1) {
    uint4 r = Philox(seed)(index, stream, step, problem);
    *u0 = u53(r.x, r.y);
    *u1 = u53(r.z, r.w);
}
// two standard normals (Box-Muller, float64)
__device__ __forceinline__ void rng_normal2(uint64_t seed, uint32_t problem, uint32_t step, uint32_t stream,
                                            uint32_t index, double *z0, double *z1) {
    double u0, u1;
    rng_uniform2(seed, problem, step, stream, index, &u0, &u1);
    double r = sqrt(-2.0 * log(1.0 - u0));
    double s, c;
    sincospi(2.0 * u1, &s, &c);
    *z0 = r * c;
    *z1 = r * s;
}

}  // namespace nfopp
