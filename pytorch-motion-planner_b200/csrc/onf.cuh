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

// extra arguments of the tensor-core path (onf_tc.cu)
struct OnfTcExtra {
    const unsigned char *packed;   // [P, onf_tc_packed_bytes] bf16 hi/lo operand copy of the hidden-layer weights
    float *grad_out;               // [P, n_params] reduced gradient, or null
    float *adam_w, *adam_m, *adam_v;   // fused Adam on [P, n_params] (engine), or null
    unsigned char *packed_out;     // operand copy rewritten by the fused Adam
    float w1, beta2, w2, eps;      // 1-beta1, beta2, 1-beta2, eps; step size / bias correction come from StepParams
    int warm;                      // warm-up iteration: ignore StepParams::do_train
    int variant;                   // 0: first generation (onf_tc.cu, bf16 operands, resident weights); 1: streamed (onf_tcs.cu, fp16)
    float *gscratch;               // streamed kernel with fused Adam: [gscratch_slots, 2, n_params] gradient hand-over buffers
    int gscratch_slots;            //   (one pair per persistent CTA: the grid is capped at this many CTAs)
    int sm_share;                  // streamed kernel: launches that run side by side (engine lanes) - each gets #SMs / sm_share CTAs
};

bool onf_tc_supported(const OnfDims &d);            // 2 hidden layers, hidden <= 111, f_in <= 224
size_t onf_tc_packed_bytes(const OnfDims &d);       // per problem
int onf_tc_prepare(const OnfDims &d);
int onf_tc_pack(const OnfDims &d, const float *weights, unsigned char *packed, int P, cudaStream_t stream, int variant);
int onf_tc_launch(const OnfArgs &a, const OnfTcExtra &x, int P, cudaStream_t stream);
// second generation (onf_tcs.cu): 112-point tiles, weights streamed, fp16 hi/lo operands, every product three passes, any
// point count.  Same packed layout and size as the first generation, different element type: the operand copy
// must be packed by the variant that reads it.
bool onf_tcs_supported(const OnfDims &d);
int onf_tcs_prepare(const OnfDims &d);
int onf_tcs_grid(int P, int slots, int sm_share);   // CTAs of the persistent kernel for P problems (slots > 0: cap)
constexpr int ONF_TCS_SCRATCH_SLOTS = 320;           // hand-over buffer pairs an engine owns (split evenly between its lanes)
int onf_tcs_pack(const OnfDims &d, const float *weights, unsigned char *packed, int P, cudaStream_t stream);
int onf_tcs_launch(const OnfArgs &a, const OnfTcExtra &x, int P, cudaStream_t stream);
// which generation serves a tensor-path call whose weight-gradient launches see at most n_wgrad points per problem
// (0: no weight gradients).  NFOPP_TC_LEGACY=1 in the environment forces the first generation.
int onf_tc_variant(const OnfDims &d, int n_wgrad);

// wide / deep networks and large point sets (onf_wide.cu): per-layer grouped GEMMs on tcgen05 with streamed weights
struct OnfWideExtra {
    const unsigned char *packed;   // [P, onf_wide_packed_bytes] bf16 hi/lo operand copies (forward and input-gradient layouts)
    unsigned char *workspace;      // onf_wide_workspace_bytes(d, P, n, want_bwd): activation / gradient operand tiles
    float *grad_out;               // [P, n_params] weight gradient, or null
};
bool onf_wide_supported(const OnfDims &d);
size_t onf_wide_packed_bytes(const OnfDims &d);     // per problem (256-byte multiple)
size_t onf_wide_workspace_bytes(const OnfDims &d, int P, int n, bool want_bwd);
int onf_wide_prepare(const OnfDims &d);
int onf_wide_pack(const OnfDims &d, const float *weights, unsigned char *packed, int P, cudaStream_t stream);
int onf_wide_launch(const OnfArgs &a, const OnfWideExtra &x, int P, cudaStream_t stream);
int onf_wide_launch_count(const OnfDims &d, int P, int n, bool want_bwd, bool want_wgrad);
int onf_wide_adam_pack(const OnfDims &d, float *weights, float *m, float *v, const float *grad, unsigned char *packed,
                       const StepParams *sp, int warm, float w1, float beta2, float w2, float eps, int P, cudaStream_t stream,
                       const int32_t *count = nullptr);

int onf_fma_tile_points(const OnfDims &d);          // 32, 16 or 0 (does not fit)
int onf_fma_prepare(const OnfDims &d);               // one-time kernel attributes (outside graph capture)
int onf_fma_launch(OnfArgs a, int P, cudaStream_t stream);
int onf_adam_launch(const AdamArgs &a, int P, cudaStream_t stream);

}  // namespace nfopp
