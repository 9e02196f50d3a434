// onf_static.cu - nfopp_onf_forward_static: ONF forward on the streamed tensor path WITHOUT caller-supplied scratch and
// without any allocation.  No counterpart in the reference; added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// nfopp_onf_forward (impl = tensor) keeps the fp16 hi/lo operand copy of the hidden-layer weights in `scratch_dev`.  Here that
// copy lives in ONE file-scope device array of NFOPP_STATIC_SCRATCH_BYTES (4 MiB; every GPU that loads the library holds it
// until the process exits).  Consequences, all documented in include/nfopp.h:
//   * P * packed_bytes(cfg) must fit the array (reference network: 150 528 B per weight set -> P <= 27), else NFOPP_ERR_INVALID;
//   * every call shares the array and it keeps the last call's contents: the pack kernel rewrites every byte the forward kernel
//     reads (all K octets x all KH rows of both matrices, padding included), so nothing has to start at zero and nothing is
//     cleared; calls that overlap on the GPU (different streams, or different host threads) race on it;
//   * the array's name is a host-side stand-in in host code: its device address is taken once per device with
//     cudaGetSymbolAddress (that device current) and only that pointer is handed to the kernels.
#include <mutex>

#include "onf.cuh"

namespace nfopp {

__device__ __align__(256) static unsigned char g_static_scratch[NFOPP_STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
static unsigned char *g_static_ptr[STATIC_MAX_DEVICES];      // device address of g_static_scratch on device i (null: not asked yet)
static std::mutex g_static_mutex;

// device address of this device's copy of the array (set-up: once per device, with that device current)
static int static_scratch_ptr(unsigned char **out) {
    int dev = -1;
    NFOPP_CHECK_CUDA(cudaGetDevice(&dev));
    NFOPP_REQUIRE(dev >= 0 && dev < STATIC_MAX_DEVICES, "onf_forward_static: device index %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (g_static_ptr[dev] == nullptr) {
        void *p = nullptr;
        NFOPP_CHECK_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        NFOPP_REQUIRE(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, "onf_forward_static: bad scratch address %p", p);
        g_static_ptr[dev] = static_cast<unsigned char *>(p);
    }
    *out = g_static_ptr[dev];
    return NFOPP_OK;
}

}  // namespace nfopp

using namespace nfopp;

extern "C" int nfopp_onf_forward_static(const nfopp_onf_config *cfg, const float *weights, const float *x, const int32_t *count,
                                        int P, int n, float *z, void *stream) {
    OnfDims d;
    int rc = make_onf_dims(cfg, &d);
    if (rc) return rc;
    NFOPP_REQUIRE(weights && x && z && P > 0 && n >= 0, "onf_forward_static: bad arguments");
    NFOPP_REQUIRE(onf_tc_supported(d) && onf_tcs_supported(d),
                  "onf_forward_static: needs the streamed tensor path (2 hidden layers, hidden <= 111, f_in <= 256; hidden=%d layers=%d)",
                  d.H, d.L);
    // operand copy of one weight set on the streamed path: [W1p hi | W1p lo | W2p hi | W2p lo], each [k / 8][KH rows][8] fp16
    // (onf_tcs.cu: ts_make_dims)
    const size_t K1 = ((size_t)d.lay.f_in + 15) / 16 * 16, KH = ((size_t)d.H + 1 + 15) / 16 * 16;
    const size_t per_problem = 2 * (K1 / 8) * KH * 16 + 2 * (KH / 8) * KH * 16;
    const size_t need = (size_t)P * per_problem;
    NFOPP_REQUIRE(need <= (size_t)NFOPP_STATIC_SCRATCH_BYTES,
                  "onf_forward_static: %d weight sets x %zu bytes of operand copy = %zu bytes exceed the static scratch of %d bytes "
                  "(at most %d weight sets of this network shape per call; use nfopp_onf_forward with caller-owned scratch beyond that)",
                  P, per_problem, need, (int)NFOPP_STATIC_SCRATCH_BYTES, (int)((size_t)NFOPP_STATIC_SCRATCH_BYTES / per_problem));
    unsigned char *scratch = nullptr;
    if ((rc = static_scratch_ptr(&scratch))) return rc;
    // kernel attribute for this network shape's shared-memory size (cudaFuncSetAttribute + a device attribute query, as
    // nfopp_onf_forward does on every call: no allocation)
    if ((rc = onf_tcs_prepare(d))) return rc;
    if (n == 0) return NFOPP_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if ((rc = onf_tcs_pack(d, weights, scratch, P, s))) return rc;
    OnfArgs a{};
    a.d = d; a.weights = weights; a.x = x; a.count = count; a.n = n; a.want_bwd = 0; a.z_out = z;
    OnfTcExtra e{};
    e.packed = scratch;
    e.variant = 1;
    return onf_tcs_launch(a, e, P, s);
}
