// onf.cuh - argument blocks of the ONF kernels, shared by onf_fma.cu (CUDA-core path), onf_tc.cu (tcgen05
// path) and engine.cu.
#pragma once
#include "engine.cuh"

namespace nfopp {

struct OnfArgs {
    OnfDims d;
    const float *weights;     // [P, n_params]
    const float *x;           // [P, n, x_stride] (first in_dim columns are used)
    const int32_t *count;     // [P] valid points per problem, or null (= n)
    const StepParams *sp;     // non-null: skip the launch on the device when !sp->do_train
    int n, x_stride;
    int want_bwd;
    int gbar_mode;            // 0 external, 1 bce, 2 path, 3 path_mu (include/nfopp.h)
    const float *gz;          // [P, n]
    const float *aux;         // [P, n]
    float wc, beta;
    float *z_out;             // [P, n] or null
    float *gx_out;            // [P, n, x_stride] or null
    float *gpart;             // [P, tiles, n_params] or null
    int tiles;                // filled by the launcher
};

struct AdamArgs {
    float *w, *m, *v;
    const float *gpart;       // [P, tiles, n_params]
    const int32_t *count;
    const StepParams *sp;     // non-null: step_size / bc2_sqrt / do_train come from device memory
    int warm;                 // with sp: warm-up iteration (ignore do_train)
    int n, tp, tiles, n_params;
    float w1, beta2, w2, step_size, bc2_sqrt, eps;
    float *grad_out;          // reduce-only mode when non-null
};

int onf_fma_tile_points(const OnfDims &d);          // 32, 16 or 0 (does not fit)
int onf_fma_prepare(const OnfDims &d);               // one-time kernel attributes (outside graph capture)
int onf_fma_launch(OnfArgs a, int P, cudaStream_t stream);
int onf_adam_launch(const AdamArgs &a, int P, cudaStream_t stream);

}  // namespace nfopp
