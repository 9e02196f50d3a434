// engine.cu - the C ABI of libnfopp.so (include/nfopp.h): stateless ONF / Adam entry points and the planner
// engine that sequences one NFOPP step as a handful of launches (optionally replayed as a CUDA graph).
#include <math.h>
#include <stdarg.h>

#include <string>
#include <vector>

#include "onf.cuh"

namespace nfopp {

static thread_local char g_error[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

static int align4(int n) { return (n + 3) / 4 * 4; }

int make_onf_dims(const nfopp_onf_config *c, OnfDims *out) {
    NFOPP_REQUIRE(c && out, "onf config missing");
    NFOPP_REQUIRE(c->hidden >= 4 && c->hidden % 4 == 0, "onf: hidden=%d must be a positive multiple of 4", c->hidden);
    NFOPP_REQUIRE(c->layers >= 1 && c->layers <= NFOPP_MAX_LAYERS, "onf: layers=%d out of range", c->layers);
    NFOPP_REQUIRE(c->sigma != 0.f, "onf: sigma must be non-zero");
    OnfDims d{};
    nfopp_onf_layout &L = d.lay;
    L.f_pos = c->use_cos ? 200 : 100;          // onf_model.py:14
    L.n_angle = c->angle ? 20 : 0;             // angle_encoder.py:7,20-22
    L.f_in = L.f_pos + L.n_angle;
    L.in_dim = c->angle ? 3 : 2;
    int off = 0;
    L.off_enc_w = off; off = align4(off + L.f_pos * 2);
    L.off_enc_b = off; off = align4(off + L.f_pos);
    L.off_ang_b = off; off = align4(off + L.n_angle);
    L.off_ang_f = off; off = align4(off + L.n_angle);
    for (int l = 0; l < c->layers; ++l) {
        int k = l == 0 ? L.f_in : c->hidden;
        L.off_w[l] = off; off = align4(off + c->hidden * k);
        L.off_b[l] = off; off = align4(off + c->hidden);
    }
    L.off_w_out = off; off = align4(off + c->hidden + L.f_in);
    L.off_b_out = off; off = align4(off + 1);
    L.n_params = off;
    d.mean = c->mean; d.sigma = c->sigma; d.use_cos = c->use_cos; d.angle = c->angle;
    d.H = c->hidden; d.L = c->layers; d.bias = c->bias;
    *out = d;
    return NFOPP_OK;
}

// from sample.cu / traj.cu
int launch_sample_pre(const EngineDev &e, cudaStream_t s);
int launch_warm_sample(const EngineDev &e, int iteration, cudaStream_t s);
int launch_resample_label(const EngineDev &e, int resample, cudaStream_t s);
int resample_prepare(const EngineDev &e);
int traj_prepare(const EngineDev &e);
int launch_traj_pre(const EngineDev &e, cudaStream_t s);
int launch_traj_post(const EngineDev &e, cudaStream_t s);
int launch_tick(const EngineDev &e, bool warm, cudaStream_t s);
int launch_paths(const EngineDev &e, float *out, cudaStream_t s);
int launch_reparam(const EngineDev &e, cudaStream_t s);
int launch_evaluate_paths(const EngineDev &e, int k, uint8_t *collides, float *length, int32_t *n_hit, cudaStream_t s);
int launch_path_statistics(const EngineDev &e, double cusp_threshold, int window, float *out, cudaStream_t s);
size_t astar_workspace_bytes(int P, int X, int Y);
int launch_astar_init(const EngineDev &e, int X, int Y, double resolution, const double *origin_xy, const int32_t *start_cell,
                      const int32_t *goal_cell, unsigned char *ws, size_t ws_bytes, int32_t *path_len, cudaStream_t s);

// impl 0 = auto: the tensor path whenever the network shape fits it
static int resolve_impl(const OnfDims &d, int impl, int *out) {
    NFOPP_REQUIRE(impl >= 0 && impl <= 3, "impl=%d (0 auto, 1 FMA, 2 tensor, 3 tensor-wide)", impl);
    if (impl == 0) impl = onf_tc_supported(d) ? 2 : (onf_wide_supported(d) ? 3 : 1);
    NFOPP_REQUIRE(impl != 3 || onf_wide_supported(d), "onf(wide): hidden=%d layers=%d unsupported", d.H, d.L);
    NFOPP_REQUIRE(impl != 2 || onf_tc_supported(d),
                  "onf(tensor): needs 2 hidden layers, hidden <= 111, f_in <= 224 (hidden=%d layers=%d)", d.H, d.L);
    NFOPP_REQUIRE(impl != 1 || onf_fma_tile_points(d) != 0, "onf(fma): hidden=%d layers=%d does not fit shared memory", d.H, d.L);
    *out = impl;
    return NFOPP_OK;
}

// gradient slots per problem
static int tiles_for(const OnfDims &d, int n, int impl) {
    if (impl >= 2) return 1;
    int tp = onf_fma_tile_points(d);
    return tp ? (n + tp - 1) / tp : 0;
}

}  // namespace nfopp

using namespace nfopp;

extern "C" int nfopp_abi_version(void) { return NFOPP_ABI_VERSION; }
extern "C" const char *nfopp_last_error(void) { return g_error; }

extern "C" int nfopp_onf_get_layout(const nfopp_onf_config *cfg, nfopp_onf_layout *out) {
    OnfDims d;
    int rc = make_onf_dims(cfg, &d);
    if (rc) return rc;
    NFOPP_REQUIRE(out, "layout out pointer missing");
    *out = d.lay;
    return NFOPP_OK;
}

extern "C" int nfopp_onf_tiles(const nfopp_onf_config *cfg, int n, int impl) {
    OnfDims d;
    if (make_onf_dims(cfg, &d)) return NFOPP_ERR_INVALID;
    if (resolve_impl(d, impl, &impl)) return NFOPP_ERR_INVALID;
    return tiles_for(d, n, impl);
}

extern "C" int nfopp_onf_scratch_bytes(const nfopp_onf_config *cfg, int P, int impl, size_t *bytes) {
    OnfDims d;
    int rc = make_onf_dims(cfg, &d);
    if (rc) return rc;
    NFOPP_REQUIRE(bytes && P >= 0, "onf_scratch_bytes: bad arguments");
    if ((rc = resolve_impl(d, impl, &impl))) return rc;
    NFOPP_REQUIRE(impl != 3, "onf(wide): the scratch size depends on the point count - use nfopp_onf_scratch_bytes_n");
    *bytes = impl == 2 ? (size_t)P * onf_tc_packed_bytes(d) : 0;
    return NFOPP_OK;
}

extern "C" int nfopp_onf_scratch_bytes_n(const nfopp_onf_config *cfg, int P, int n, int impl, int want_bwd, size_t *bytes) {
    OnfDims d;
    int rc = make_onf_dims(cfg, &d);
    if (rc) return rc;
    NFOPP_REQUIRE(bytes && P >= 0 && n >= 0, "onf_scratch_bytes_n: bad arguments");
    if ((rc = resolve_impl(d, impl, &impl))) return rc;
    if (impl == 3) *bytes = (size_t)P * onf_wide_packed_bytes(d) + onf_wide_workspace_bytes(d, P, n, want_bwd != 0);
    else *bytes = impl == 2 ? (size_t)P * onf_tc_packed_bytes(d) : 0;
    return NFOPP_OK;
}

// one ONF launch through the chosen implementation (stateless API: the tensor path packs its operand copy first)
static int onf_api_launch(const OnfDims &d, OnfArgs &a, int P, int impl, void *scratch, float *grad_out, cudaStream_t s) {
    int rc;
    if (impl == 1) {
        if ((rc = onf_fma_prepare(d))) return rc;
        a.gpart = grad_out;
        return onf_fma_launch(a, P, s);
    }
    NFOPP_REQUIRE(scratch, "onf(tensor): scratch_dev missing (nfopp_onf_scratch_bytes)");
    if (impl == 3) {
        if ((rc = onf_wide_prepare(d))) return rc;
        unsigned char *packed = static_cast<unsigned char *>(scratch);
        if ((rc = onf_wide_pack(d, a.weights, packed, P, s))) return rc;
        OnfWideExtra x{};
        x.packed = packed;
        x.workspace = packed + (size_t)P * onf_wide_packed_bytes(d);
        x.grad_out = grad_out;
        return onf_wide_launch(a, x, P, s);
    }
    if ((rc = onf_tc_prepare(d))) return rc;
    const int variant = onf_tc_variant(d, grad_out ? a.n : 0);
    if ((rc = onf_tc_pack(d, a.weights, static_cast<unsigned char *>(scratch), P, s, variant))) return rc;
    OnfTcExtra x{};
    x.packed = static_cast<const unsigned char *>(scratch);
    x.grad_out = grad_out;
    x.variant = variant;
    if (grad_out) NFOPP_CHECK_CUDA(cudaMemsetAsync(grad_out, 0, (size_t)P * d.lay.n_params * sizeof(float), s));
    return onf_tc_launch(a, x, P, s);
}

extern "C" int nfopp_onf_forward(const nfopp_onf_config *cfg, const float *weights, const float *x,
                                 const int32_t *count, int P, int n, float *z, int impl, void *scratch, void *stream) {
    OnfDims d;
    int rc = make_onf_dims(cfg, &d);
    if (rc) return rc;
    NFOPP_REQUIRE(weights && x && z && P > 0 && n >= 0, "onf_forward: bad arguments");
    if ((rc = resolve_impl(d, impl, &impl))) return rc;
    OnfArgs a{};
    a.d = d; a.weights = weights; a.x = x; a.count = count; a.n = n; a.want_bwd = 0; a.z_out = z;
    return onf_api_launch(d, a, P, impl, scratch, nullptr, (cudaStream_t)stream);
}

extern "C" int nfopp_onf_backward(const nfopp_onf_config *cfg, const float *weights, const float *x,
                                  const int32_t *count, int P, int n, int gbar_mode, const float *gz,
                                  const float *aux, float collision_weight, float beta, float *z, float *gx,
                                  float *gpart, int impl, void *scratch, void *stream) {
    OnfDims d;
    int rc = make_onf_dims(cfg, &d);
    if (rc) return rc;
    NFOPP_REQUIRE(weights && x && P > 0 && n >= 0, "onf_backward: bad arguments");
    NFOPP_REQUIRE(gbar_mode >= 0 && gbar_mode <= 3, "onf_backward: gbar_mode=%d", gbar_mode);
    NFOPP_REQUIRE(gbar_mode != 0 || gz, "onf_backward: external mode needs gz");
    NFOPP_REQUIRE((gbar_mode != 1 && gbar_mode != 3) || aux, "onf_backward: mode %d needs aux", gbar_mode);
    NFOPP_REQUIRE(gx || gpart, "onf_backward: neither gx nor gpart requested");
    if ((rc = resolve_impl(d, impl, &impl))) return rc;
    OnfArgs a{};
    a.d = d; a.weights = weights; a.x = x; a.count = count; a.n = n; a.want_bwd = 1; a.gbar_mode = gbar_mode;
    a.gz = gz; a.aux = aux; a.wc = collision_weight; a.beta = beta; a.z_out = z; a.gx_out = gx;
    return onf_api_launch(d, a, P, impl, scratch, gpart, (cudaStream_t)stream);
}

extern "C" int nfopp_onf_reduce_grad(const nfopp_onf_config *cfg, const float *gpart, const int32_t *count, int P,
                                     int n, int impl, float *grad, void *stream) {
    OnfDims d;
    int rc = make_onf_dims(cfg, &d);
    if (rc) return rc;
    NFOPP_REQUIRE(gpart && grad && P > 0 && n > 0, "onf_reduce_grad: bad arguments");
    if ((rc = resolve_impl(d, impl, &impl))) return rc;
    AdamArgs a{};
    a.gpart = gpart; a.count = count; a.n = n; a.tp = impl >= 2 ? (1 << 30) : onf_fma_tile_points(d);
    a.tiles = tiles_for(d, n, impl);
    a.n_params = d.lay.n_params; a.grad_out = grad;
    return onf_adam_launch(a, P, (cudaStream_t)stream);
}

static void adam_scalars(double lr, double b1, double b2, double eps, int64_t step, AdamArgs *a) {
    a->w1 = (float)(1.0 - b1);
    a->beta2 = (float)b2;
    a->w2 = (float)(1.0 - b2);
    a->step_size = (float)(lr / (1.0 - pow(b1, (double)step)));
    a->bc2_sqrt = (float)sqrt(1.0 - pow(b2, (double)step));
    a->eps = (float)eps;
}

extern "C" int nfopp_onf_adam(const nfopp_onf_config *cfg, float *w, float *m, float *v, const float *gpart,
                              const int32_t *count, int P, int n, int impl, int step, double lr, double beta1,
                              double beta2, double eps, void *stream) {
    OnfDims d;
    int rc = make_onf_dims(cfg, &d);
    if (rc) return rc;
    NFOPP_REQUIRE(w && m && v && gpart && P > 0 && n > 0 && step >= 1, "onf_adam: bad arguments");
    if ((rc = resolve_impl(d, impl, &impl))) return rc;
    AdamArgs a{};
    a.w = w; a.m = m; a.v = v; a.gpart = gpart; a.count = count; a.n = n;
    a.tp = impl >= 2 ? (1 << 30) : onf_fma_tile_points(d);
    a.tiles = tiles_for(d, n, impl); a.n_params = d.lay.n_params;
    adam_scalars(lr, beta1, beta2, eps, step, &a);
    return onf_adam_launch(a, P, (cudaStream_t)stream);
}

// =====================================================================================================
// planner engine
// =====================================================================================================
struct BufferInfo {
    std::string name;
    void **slot;
    size_t bytes;
    char dtype;
    int32_t dims[3];
    size_t offset;
};

struct nfopp_engine {
    nfopp_planner_config cfg;
    OnfDims onf;
    EngineDev dev;
    std::vector<BufferInfo> buffers;
    unsigned char *workspace;
    size_t workspace_bytes;
    int impl;                 // 1 FMA, 2 tensor (resolved)
    int tc_variant;           // tensor path: 0 first generation (onf_tc.cu), 1 streamed single-tile kernel (onf_tcs.cu)
    bool pack_dirty;          // tensor path: onf_wp must be rebuilt from onf_w before the next launch
    int tile_points, tiles;
    StepParams host_sp;
    bool sp_dirty;
    int64_t launches;
    int launches_per_step, launches_per_step_1;     // with the current lane count / with one lane
    cudaGraph_t graph;
    cudaGraphExec_t graph_exec;
    // wide path with optimize_collision_model_freq > 1: its ONF launches cannot be switched off by a device-side flag (dozens of
    // kernels), so steps that do not train replay a second graph captured WITHOUT the candidates-forward and train launches;
    // the host mirrors the step counter (host_sp) and picks the graph
    cudaGraph_t graph_skip;
    cudaGraphExec_t graph_exec_skip;
    int launches_per_step_skip;
    int warm_iterations_done;
    // lanes: the problems of a graphed step split into contiguous ranges, each enqueued on its own stream inside the one
    // CUDA graph (fork / join): independent problems have no cross-lane dependency, so the latency-bound small kernels and
    // the HBM-bound Adam phase of one lane overlap the tensor phases of another
    int lanes;
    std::vector<cudaStream_t> lane_streams;
    std::vector<cudaEvent_t> lane_events;     // [0] fork, [l] join of lane l
    size_t lane_ws_stride;                    // wide path: bytes between the lanes' activation workspaces
};

static void drop_graphs(nfopp_engine *e) {
    if (e->graph_exec) { cudaGraphExecDestroy(e->graph_exec); e->graph_exec = nullptr; }
    if (e->graph) { cudaGraphDestroy(e->graph); e->graph = nullptr; }
    if (e->graph_exec_skip) { cudaGraphExecDestroy(e->graph_exec_skip); e->graph_exec_skip = nullptr; }
    if (e->graph_skip) { cudaGraphDestroy(e->graph_skip); e->graph_skip = nullptr; }
}

static void host_step_params(nfopp_engine *e) {
    StepParams &sp = e->host_sp;
    const nfopp_planner_config &c = e->cfg;
    sp.do_train = (sp.step_count % c.optimize_collision_model_freq) == 0;
    sp.do_reparam = (sp.step_count % c.reparametrize_trajectory_freq) == 0;
    const double ko = sp.onf_step + 1, kt = sp.traj_step + 1;
    sp.onf_step_size = (float)((double)c.onf_lr / (1.0 - pow((double)c.onf_beta1, ko)));
    sp.onf_bc2_sqrt = (float)sqrt(1.0 - pow((double)c.onf_beta2, ko));
    sp.traj_step_size = (float)((double)c.traj_lr / (1.0 - pow((double)c.traj_beta1, kt)));
    sp.traj_bc2_sqrt = (float)sqrt(1.0 - pow((double)c.traj_beta2, kt));
}

template <typename T>
static void add_buffer(nfopp_engine *e, const char *name, T **slot, char dtype, size_t d0, size_t d1 = 0, size_t d2 = 0) {
    BufferInfo b;
    b.name = name;
    b.slot = reinterpret_cast<void **>(slot);
    size_t count = d0 * (d1 ? d1 : 1) * (d2 ? d2 : 1);
    b.bytes = count * sizeof(T);
    b.dtype = dtype;
    b.dims[0] = (int32_t)d0; b.dims[1] = (int32_t)d1; b.dims[2] = (int32_t)d2;
    b.offset = 0;
    e->buffers.push_back(b);
}

// kernel launches of one step over P problems (incl. the counter tick)
static int step_launches(const nfopp_engine *e, int P, bool train = true) {
    if (e->impl != 3) return e->impl == 2 ? 8 : 9;
    if (!train)                                                 // wide path, step that does not train: path launches only
        return 5 + onf_wide_launch_count(e->onf, P, e->dev.T - 1, true, false);
    // 5 planner kernels + candidates forward + train (backward, Adam + re-pack) + path forward/input-gradient
    const EngineDev &d = e->dev;
    return 5 + onf_wide_launch_count(e->onf, P, d.NC, false, false) + onf_wide_launch_count(e->onf, P, d.NTR, true, true) + 1 +
           onf_wide_launch_count(e->onf, P, d.T - 1, true, false);
}

// device view of problems [p0, p0 + count): every per-problem buffer advanced by p0 rows, P = count, RNG streams keep their
// global problem index (p_base)
static EngineDev slice_problems(const nfopp_engine *e, int p0, int count, int lane) {
    EngineDev d = e->dev;
    const int P = e->dev.P;
    for (const BufferInfo &b : e->buffers) {
        const bool shared = b.name == "step_params" || b.name == "inv_hessian_t" || b.name == "lin_s" || b.name == "onf_ws" ||
                            (b.name == "onf_gpart" && e->impl == 2) ||
                            ((b.name == "obstacles" || b.name == "obstacle_n" || b.name == "grid") && e->dev.env_shared);
        if (shared || b.dims[0] != P) continue;
        const size_t field = (size_t)(reinterpret_cast<const char *>(b.slot) - reinterpret_cast<const char *>(&e->dev));
        void **dst = reinterpret_cast<void **>(reinterpret_cast<char *>(&d) + field);
        *dst = static_cast<char *>(*b.slot) + (size_t)p0 * (b.bytes / P);
    }
    d.P = count;
    d.p_base = e->dev.p_base + p0;
    if (e->impl == 3) d.onf_ws = e->dev.onf_ws + (size_t)lane * e->lane_ws_stride;   // wide path: one activation workspace per lane
    if (e->impl == 2 && e->dev.gslots > 0) {
        // streamed tensor kernel: every lane owns an equal share of the gradient hand-over buffers
        const int lanes = e->lanes > 0 ? e->lanes : 1;
        d.gslots = e->dev.gslots / lanes > 0 ? e->dev.gslots / lanes : 1;
        d.onf_gpart = e->dev.onf_gpart + (size_t)lane * d.gslots * 2 * e->onf.lay.n_params;
    }
    return d;
}

static int validate(const nfopp_planner_config *c, OnfDims *onf, int *impl) {
    NFOPP_REQUIRE(c, "planner config missing");
    int rc = make_onf_dims(&c->onf, onf);
    if (rc) return rc;
    NFOPP_REQUIRE(c->problems >= 1, "problems=%d", c->problems);
    NFOPP_REQUIRE(c->trajectory_length >= 2 && c->trajectory_length <= 4096, "trajectory_length=%d", c->trajectory_length);
    NFOPP_REQUIRE(c->constrained || !c->onf.angle, "holonomic planner cannot feed an angle-encoding ONF");
    NFOPP_REQUIRE(c->collision_point_count >= 1 && c->random_field_points >= 0 && c->init_collision_points >= 0,
                  "sample counts out of range");
    NFOPP_REQUIRE(c->reparametrize_trajectory_freq >= 1 && c->optimize_collision_model_freq >= 1, "freq must be >= 1");
    NFOPP_REQUIRE(c->checker_kind >= 0 && c->checker_kind <= 3, "checker_kind=%d", c->checker_kind);
    NFOPP_REQUIRE(c->checker_kind != 2 || c->constrained, "rectangle checker needs poses with heading");
    NFOPP_REQUIRE(c->checker_kind != 3 || (c->grid_rows > 1 && c->grid_cols > 1 && c->grid_cell > 0), "grid shape missing");
    NFOPP_REQUIRE(c->rng_mode == 0 || c->rng_mode == 1, "rng_mode=%d", c->rng_mode);
    if ((rc = resolve_impl(*onf, c->onf_impl, impl))) return rc;
    return NFOPP_OK;
}

static int build_layout(nfopp_engine *e) {
    const nfopp_planner_config &c = e->cfg;
    EngineDev &d = e->dev;
    const size_t P = c.problems, T = c.trajectory_length, D = c.constrained ? 3 : 2;
    const size_t C = c.collision_point_count, NF = c.random_field_points, NW = c.init_collision_points;
    const size_t NC = C + T - 1;
    size_t NTR = (T - 1) + C + NF;
    if (NW > NTR) NTR = NW;
    const size_t NP = e->onf.lay.n_params;
    e->tile_points = e->impl >= 2 ? 64 : onf_fma_tile_points(e->onf);
    e->tiles = e->impl >= 2 ? 1 : (int)((NTR + e->tile_points - 1) / e->tile_points);
    const bool inj = c.rng_mode == 1;
    const size_t MU = inj ? (c.max_choice_uniforms > 0 ? c.max_choice_uniforms : 16 * C) : 0;
    d.P = (int)P; d.T = (int)T; d.D = (int)D; d.C = (int)C; d.NF = (int)NF; d.NC = (int)NC; d.NTR = (int)NTR; d.NW = (int)NW;
    d.max_choice_u = (int)MU;
    const size_t PE = c.env_shared ? 1 : P;

    add_buffer(e, "step_params", &d.sp, 'i', 1);
    e->buffers.back().dims[0] = (int32_t)(sizeof(StepParams) / 4);
    add_buffer(e, "onf_w", &d.onf_w, 'f', P, NP);
    add_buffer(e, "onf_m", &d.onf_m, 'f', P, NP);
    add_buffer(e, "onf_v", &d.onf_v, 'f', P, NP);
    if (e->impl == 2) {
        // the tensor path accumulates weight gradients in TMEM and applies Adam in the same kernel: no partials
        // (streamed kernel: its persistent CTAs hand each problem's gradient to their optimiser warps through a pair of
        // L2-resident buffers per CTA)
        d.gslots = onf_tc_variant(e->onf, (int)NTR) == 1 ? (int)(P < (size_t)ONF_TCS_SCRATCH_SLOTS ? P : (size_t)ONF_TCS_SCRATCH_SLOTS) : 0;
        if (d.gslots > 0) add_buffer(e, "onf_gpart", &d.onf_gpart, 'f', d.gslots, 2, NP);
        else add_buffer(e, "onf_gpart", &d.onf_gpart, 'f', 1, 1, 4);
        add_buffer(e, "onf_wp", &d.onf_wp, 'b', P, onf_tc_packed_bytes(e->onf));
    } else if (e->impl == 3) {
        // wide path: per-layer grouped GEMMs; one reduced gradient per problem, operand copies of the weights and
        // one activation workspace shared by the three ONF calls of a step
        size_t nmax = NTR > NC ? NTR : NC;
        if (T - 1 > nmax) nmax = T - 1;
        add_buffer(e, "onf_gpart", &d.onf_gpart, 'f', P, 1, NP);
        add_buffer(e, "onf_wp", &d.onf_wp, 'b', P, onf_wide_packed_bytes(e->onf));
        // [MiB, 2^20] so that the int32 dims survive workspaces beyond 2 GiB
        const size_t ws_bytes = onf_wide_workspace_bytes(e->onf, (int)P, (int)nmax, true);
        add_buffer(e, "onf_ws", &d.onf_ws, 'b', (ws_bytes + (1u << 20) - 1) >> 20, (size_t)1 << 20);
    } else {
        add_buffer(e, "onf_gpart", &d.onf_gpart, 'f', P, e->tiles, NP);
    }
    add_buffer(e, "traj", &d.traj, 'f', P, T, D);
    add_buffer(e, "traj_m", &d.traj_m, 'f', P, T, D);
    add_buffer(e, "traj_v", &d.traj_v, 'f', P, T, D);
    add_buffer(e, "prev_traj", &d.prev_traj, 'f', P, T, D);
    add_buffer(e, "lam", &d.lam, 'f', P, T + 1);
    add_buffer(e, "mu", &d.mu, 'f', P, T);
    add_buffer(e, "start", &d.start, 'f', P, D);
    add_buffer(e, "goal", &d.goal, 'f', P, D);
    add_buffer(e, "buf_pos", &d.buf_pos, 'd', P, C, D);
    add_buffer(e, "buf_age", &d.buf_age, 'd', P, C);
    add_buffer(e, "buf_n", &d.buf_n, 'i', P);
    add_buffer(e, "cand_pos", &d.cand_pos, 'd', P, NC, D);
    add_buffer(e, "cand_age", &d.cand_age, 'd', P, NC);
    add_buffer(e, "cand_x", &d.cand_x, 'f', P, NC, D);
    add_buffer(e, "cand_z", &d.cand_z, 'f', P, NC);
    add_buffer(e, "cand_n", &d.cand_n, 'i', P);
    add_buffer(e, "train_pos", &d.train_pos, 'd', P, NTR, D);
    add_buffer(e, "train_x", &d.train_x, 'f', P, NTR, D);
    add_buffer(e, "train_y", &d.train_y, 'f', P, NTR);
    add_buffer(e, "train_z", &d.train_z, 'f', P, NTR);
    add_buffer(e, "train_n", &d.train_n, 'i', P);
    add_buffer(e, "path_q", &d.path_q, 'f', P, T - 1, D);
    add_buffer(e, "path_mu", &d.path_mu, 'f', P, T - 1);
    add_buffer(e, "path_z", &d.path_z, 'f', P, T - 1);
    add_buffer(e, "path_gq", &d.path_gq, 'f', P, T - 1, D);
    add_buffer(e, "path_t", &d.path_t, 'f', P, T - 1);
    add_buffer(e, "path_n", &d.path_n, 'i', P);
    add_buffer(e, "choice_idx", &d.choice_idx, 'i', P, C);
    add_buffer(e, "choice_used", &d.choice_used, 'i', P);
    add_buffer(e, "status", &d.status, 'i', P);
    add_buffer(e, "frozen", &d.frozen, 'i', P);
    add_buffer(e, "inv_hessian_t", &d.inv_hessian_t, 'f', T, T);
    add_buffer(e, "lin_s", &d.lin_s, 'f', T);
    if (c.checker_kind == 1 || c.checker_kind == 2) {
        add_buffer(e, "obstacles", &d.obstacles, 'd', PE, c.max_obstacles > 0 ? c.max_obstacles : 1, 2);
        add_buffer(e, "obstacle_n", &d.obstacle_n, 'i', PE);
    }
    if (c.checker_kind == 3) add_buffer(e, "grid", &d.grid, 'b', PE, c.grid_rows, c.grid_cols);
    if (inj) {
        add_buffer(e, "var_t_sample", &d.var_t_sample, 'd', P, T - 1);
        add_buffer(e, "var_coarse_xy", &d.var_coarse_xy, 'd', P, T - 1, 2);
        add_buffer(e, "var_coarse_th", &d.var_coarse_th, 'd', P, T - 1);
        add_buffer(e, "var_fine_xy", &d.var_fine_xy, 'd', P, T - 1, 2);
        add_buffer(e, "var_fine_th", &d.var_fine_th, 'd', P, T - 1);
        add_buffer(e, "var_choice_u", &d.var_choice_u, 'd', P, MU);
        add_buffer(e, "var_field_xy", &d.var_field_xy, 'd', P, NF > 0 ? NF : 1, 2);
        add_buffer(e, "var_field_th", &d.var_field_th, 'd', P, NF > 0 ? NF : 1);
        add_buffer(e, "var_warm_xy", &d.var_warm_xy, 'd', P, NW > 0 ? NW : 1, 2);
        add_buffer(e, "var_warm_th", &d.var_warm_th, 'd', P, NW > 0 ? NW : 1);
        add_buffer(e, "var_t_path", &d.var_t_path, 'f', P, T - 1);
    }
    size_t off = 0;
    for (auto &b : e->buffers) {
        b.offset = off;
        off += (b.bytes + 255) / 256 * 256;
    }
    e->workspace_bytes = off;
    return NFOPP_OK;
}

static void fill_dev_config(nfopp_engine *e) {
    const nfopp_planner_config &c = e->cfg;
    EngineDev &d = e->dev;
    d.constrained = c.constrained; d.rng_mode = c.rng_mode; d.seed = c.seed;
    d.fine_offset = c.trajectory_random_offset; d.coarse_offset = c.course_random_offset; d.angle_offset = c.angle_offset;
    for (int i = 0; i < 4; ++i) { d.border[i] = c.border[i]; d.border_f[i] = (float)c.border[i]; }
    d.collision_weight = c.collision_weight; d.angle_weight = c.angle_weight;
    d.constraint_weight = c.constraint_deltas_weight; d.multipliers_lr = c.multipliers_lr;
    d.boundary_weight = c.boundary_weight; d.mu_lr = c.collision_multipliers_lr; d.beta = c.collision_beta;
    d.direction_weight = c.direction_delta_weight;
    d.traj_w1 = (float)(1.0 - (double)c.traj_beta1); d.traj_beta2 = (float)c.traj_beta2;
    d.traj_w2 = (float)(1.0 - (double)c.traj_beta2); d.traj_eps = (float)c.adam_eps;
    d.onf_lr_d = c.onf_lr; d.onf_b1_d = c.onf_beta1; d.onf_b2_d = c.onf_beta2;
    d.traj_lr_d = c.traj_lr; d.traj_b1_d = c.traj_beta1; d.traj_b2_d = c.traj_beta2;
    d.reparam_freq = c.reparametrize_trajectory_freq; d.train_freq = c.optimize_collision_model_freq;
    CheckerParams &k = d.checker;
    k.kind = c.checker_kind; k.has_bounds = c.has_bounds;
    for (int i = 0; i < 4; ++i) { k.bounds[i] = c.bounds[i]; k.box[i] = c.box[i]; }
    k.radius = c.robot_radius; k.radius_sq = circle_threshold(c.robot_radius); k.rows = c.grid_rows; k.cols = c.grid_cols;
    k.ox = c.grid_origin_x; k.oy = c.grid_origin_y; k.cell = c.grid_cell;
    d.env_shared = c.env_shared; d.m_max = c.max_obstacles > 0 ? c.max_obstacles : 1;
}

extern "C" int nfopp_engine_workspace_bytes(const nfopp_planner_config *cfg, size_t *bytes) {
    nfopp_engine tmp{};
    int rc = validate(cfg, &tmp.onf, &tmp.impl);
    if (rc) return rc;
    NFOPP_REQUIRE(bytes, "bytes pointer missing");
    tmp.cfg = *cfg;
    rc = build_layout(&tmp);
    if (rc) return rc;
    *bytes = tmp.workspace_bytes;
    return NFOPP_OK;
}

static int upload_step_params(nfopp_engine *e, cudaStream_t s) {
    host_step_params(e);
    NFOPP_CHECK_CUDA(cudaMemcpyAsync(e->dev.sp, &e->host_sp, sizeof(StepParams), cudaMemcpyHostToDevice, s));
    NFOPP_CHECK_CUDA(cudaStreamSynchronize(s));
    e->sp_dirty = false;
    return NFOPP_OK;
}

extern "C" int nfopp_engine_create(const nfopp_planner_config *cfg, void *workspace, size_t workspace_bytes,
                                   nfopp_engine **out) {
    NFOPP_REQUIRE(out, "engine out pointer missing");
    nfopp_engine *e = new nfopp_engine{};
    int rc = validate(cfg, &e->onf, &e->impl);
    if (rc) { delete e; return rc; }
    e->cfg = *cfg;
    build_layout(e);
    if (!workspace || workspace_bytes < e->workspace_bytes || (reinterpret_cast<uintptr_t>(workspace) & 255)) {
        set_error("workspace: need %zu bytes, 256-byte aligned (got %zu at %p)", e->workspace_bytes, workspace_bytes, workspace);
        delete e;
        return NFOPP_ERR_WORKSPACE;
    }
    e->workspace = static_cast<unsigned char *>(workspace);
    for (auto &b : e->buffers) *b.slot = e->workspace + b.offset;
    fill_dev_config(e);
    cudaError_t ce = cudaMemset(e->workspace, 0, e->workspace_bytes);
    if (ce != cudaSuccess) {
        set_error("cudaMemset(workspace) failed: %s", cudaGetErrorString(ce));
        delete e;
        return NFOPP_ERR_CUDA;
    }
    if ((rc = e->impl == 3 ? onf_wide_prepare(e->onf) : e->impl == 2 ? onf_tc_prepare(e->onf) : onf_fma_prepare(e->onf)) || (rc = resample_prepare(e->dev)) ||
        (rc = traj_prepare(e->dev))) {
        delete e;
        return rc;
    }
    e->tc_variant = e->impl == 2 ? onf_tc_variant(e->onf, e->dev.NTR) : 0;
    e->host_sp = StepParams{};
    e->sp_dirty = true;
    e->pack_dirty = true;
    e->launches_per_step = step_launches(e, e->dev.P);
    e->launches_per_step_1 = e->launches_per_step;
    e->launches_per_step_skip = step_launches(e, e->dev.P, false);
    e->lanes = 1;
    *out = e;
    return NFOPP_OK;
}

extern "C" int nfopp_engine_destroy(nfopp_engine *e) {
    if (!e) return NFOPP_OK;
    drop_graphs(e);
    for (cudaStream_t st : e->lane_streams) cudaStreamDestroy(st);
    for (cudaEvent_t ev : e->lane_events) cudaEventDestroy(ev);
    delete e;
    return NFOPP_OK;
}

extern "C" int nfopp_engine_buffer_count(nfopp_engine *e) { return e ? (int)e->buffers.size() : 0; }
extern "C" const char *nfopp_engine_buffer_name(nfopp_engine *e, int index) {
    if (!e || index < 0 || index >= (int)e->buffers.size()) return nullptr;
    return e->buffers[index].name.c_str();
}

extern "C" int nfopp_engine_buffer(nfopp_engine *e, const char *name, void **ptr, size_t *bytes, char *dtype,
                                   int32_t dims[3]) {
    NFOPP_REQUIRE(e && name, "engine_buffer: bad arguments");
    for (auto &b : e->buffers) {
        if (b.name == name) {
            if (b.name == "onf_w") e->pack_dirty = true;      // the caller may write the weights
            if (ptr) *ptr = *b.slot;
            if (bytes) *bytes = b.bytes;
            if (dtype) *dtype = b.dtype;
            if (dims) for (int i = 0; i < 3; ++i) dims[i] = b.dims[i];
            return NFOPP_OK;
        }
    }
    set_error("unknown engine buffer '%s'", name);
    return NFOPP_ERR_NAME;
}

extern "C" int nfopp_engine_invalidate_weights(nfopp_engine *e) {
    NFOPP_REQUIRE(e, "invalidate_weights: bad arguments");
    e->pack_dirty = true;
    return NFOPP_OK;
}

extern "C" int nfopp_engine_get_counter(nfopp_engine *e, int which, int64_t *value) {
    NFOPP_REQUIRE(e && value && which >= 0 && which <= 3, "get_counter: bad arguments");
    const StepParams &sp = e->host_sp;
    *value = which == 0 ? sp.step_count : which == 1 ? sp.onf_step : which == 2 ? sp.traj_step : sp.has_prev;
    return NFOPP_OK;
}

extern "C" int nfopp_engine_set_counter(nfopp_engine *e, int which, int64_t value) {
    NFOPP_REQUIRE(e && which >= 0 && which <= 3 && value >= 0, "set_counter: bad arguments");
    StepParams &sp = e->host_sp;
    if (which == 0) sp.step_count = (int32_t)value;
    else if (which == 1) sp.onf_step = (int32_t)value;
    else if (which == 2) sp.traj_step = (int32_t)value;
    else sp.has_prev = value != 0;
    e->sp_dirty = true;
    return NFOPP_OK;
}

// tensor path: (re)build the bf16 operand copy of the weights if the host may have written them
static int ensure_packed(nfopp_engine *e, cudaStream_t s) {
    if (e->impl < 2 || !e->pack_dirty) return NFOPP_OK;
    int rc = e->impl == 3 ? onf_wide_pack(e->onf, e->dev.onf_w, e->dev.onf_wp, e->dev.P, s)
                          : onf_tc_pack(e->onf, e->dev.onf_w, e->dev.onf_wp, e->dev.P, s, e->tc_variant);
    if (rc) return rc;
    e->pack_dirty = false;
    e->launches += 1;
    return NFOPP_OK;
}

// forward / input-gradient ONF launch of the engine through the chosen implementation
static int engine_onf(nfopp_engine *e, const EngineDev &d, OnfArgs &a, cudaStream_t s) {
    if (e->impl == 1) return onf_fma_launch(a, d.P, s);
    if (e->impl == 3) {
        OnfWideExtra x{};
        x.packed = d.onf_wp; x.workspace = d.onf_ws;
        a.sp = nullptr;
        return onf_wide_launch(a, x, d.P, s);
    }
    OnfTcExtra x{};
    x.packed = d.onf_wp;
    x.variant = e->tc_variant;
    x.sm_share = d.P < e->dev.P ? e->lanes : 1;        // a lane's launch shares the GPU with the other lanes'
    return onf_tc_launch(a, x, d.P, s);
}

// one ONF training update on train_x/train_y/train_n (nerf_opt_planner.py:83-90)
static int enqueue_train(nfopp_engine *e, const EngineDev &d, bool warm, cudaStream_t s, cudaEvent_t *ev = nullptr) {
    OnfArgs a{};
    a.d = e->onf; a.weights = d.onf_w; a.x = d.train_x; a.count = d.train_n; a.n = d.NTR; a.x_stride = d.D;
    a.want_bwd = 1; a.gbar_mode = 1; a.aux = d.train_y; a.z_out = d.train_z;
    const float w1 = (float)(1.0 - (double)e->cfg.onf_beta1), beta2 = (float)e->cfg.onf_beta2,
                w2 = (float)(1.0 - (double)e->cfg.onf_beta2), eps = (float)e->cfg.adam_eps;
    int rc;
    if (e->impl == 2) {
        // fwd + bwd + weight gradients in TMEM + Adam + operand-copy refresh: ONE kernel, no gradient in HBM
        a.sp = d.sp;
        OnfTcExtra x{};
        x.packed = d.onf_wp; x.adam_w = d.onf_w; x.adam_m = d.onf_m; x.adam_v = d.onf_v; x.packed_out = d.onf_wp;
        x.gscratch = d.onf_gpart; x.gscratch_slots = d.gslots;
        x.sm_share = d.P < e->dev.P ? e->lanes : 1;
        x.w1 = w1; x.beta2 = beta2; x.w2 = w2; x.eps = eps; x.warm = warm;
        x.variant = e->tc_variant;
        if ((rc = onf_tc_launch(a, x, d.P, s))) return rc;
        if (ev) {
            NFOPP_CHECK_CUDA(cudaEventRecord(ev[0], s));
            NFOPP_CHECK_CUDA(cudaEventRecord(ev[1], s));
        }
        return NFOPP_OK;
    }
    a.gpart = d.onf_gpart;
    a.sp = warm ? nullptr : d.sp;
    if (e->impl == 3) {
        OnfWideExtra x{};
        x.packed = d.onf_wp; x.workspace = d.onf_ws; x.grad_out = d.onf_gpart;
        a.sp = nullptr;
        rc = onf_wide_launch(a, x, d.P, s);
    } else {
        rc = onf_fma_launch(a, d.P, s);
    }
    if (rc) return rc;
    if (ev) NFOPP_CHECK_CUDA(cudaEventRecord(ev[0], s));
    if (e->impl == 3) {
        // Adam + operand-copy refresh in one pass over w/m/v/grad
        if ((rc = onf_wide_adam_pack(e->onf, d.onf_w, d.onf_m, d.onf_v, d.onf_gpart, d.onf_wp, d.sp, warm, w1, beta2, w2, eps, d.P, s, d.train_n)))
            return rc;
        if (ev) NFOPP_CHECK_CUDA(cudaEventRecord(ev[1], s));
        return NFOPP_OK;
    }
    AdamArgs ad{};
    ad.w = d.onf_w; ad.m = d.onf_m; ad.v = d.onf_v; ad.gpart = d.onf_gpart; ad.count = d.train_n; ad.n = d.NTR;
    ad.tp = e->tile_points; ad.tiles = e->tiles; ad.n_params = e->onf.lay.n_params;
    ad.w1 = w1; ad.beta2 = beta2; ad.w2 = w2; ad.eps = eps;
    ad.sp = d.sp;
    ad.warm = warm;
    rc = onf_adam_launch(ad, d.P, s);
    if (rc) return rc;
    if (ev) NFOPP_CHECK_CUDA(cudaEventRecord(ev[1], s));
    return NFOPP_OK;
}

static const char *const kStepKernelNames[9] = {"sample_pre", "onf_forward_candidates", "resample_label",
                                                "onf_train_fwd_bwd", "onf_reduce_adam", "traj_pre",
                                                "onf_path_fwd_dgrad", "traj_post", "tick"};

// train_host: the host's copy of StepParams::do_train for this step.  Only the wide path looks at it (its ONF launches for
// the candidates and the training update are left out of a step that does not train); the other paths skip on the device.
static int enqueue_step(nfopp_engine *e, const EngineDev &d, cudaStream_t s, cudaEvent_t *ev = nullptr, bool tick = true,
                        bool train_host = true) {
    int rc;
    const bool skip_onf_train = e->impl == 3 && !train_host;
    int evi = 0;
#define NFOPP_MARK() do { if (ev) NFOPP_CHECK_CUDA(cudaEventRecord(ev[evi++], s)); } while (0)
    NFOPP_MARK();
    // (1) near-path + uniform samples, candidates            nerf_opt_planner.py:101-120
    if ((rc = launch_sample_pre(d, s))) return rc;
    NFOPP_MARK();
    // (2) ONF forward on the candidates (resampling weights)  :123-125   [skipped on device when !do_train]
    if (!skip_onf_train) {
        OnfArgs a{};
        a.d = e->onf; a.weights = d.onf_w; a.x = d.cand_x; a.count = d.cand_n; a.n = d.NC; a.x_stride = d.D;
        a.z_out = d.cand_z; a.sp = d.sp;
        if ((rc = engine_onf(e, d, a, s))) return rc;
    }
    NFOPP_MARK();
    // (3) weighted resampling of the replay buffer + ground-truth labels  :126-133, :86
    if ((rc = launch_resample_label(d, 1, s))) return rc;
    NFOPP_MARK();
    // (4,5) BCE training step of the ONF: fwd+bwd per tile, then reduce + Adam   :83-90
    if (!skip_onf_train) {
        if ((rc = enqueue_train(e, d, false, s, ev ? &ev[evi] : nullptr))) return rc;
    } else if (ev) {
        NFOPP_CHECK_CUDA(cudaEventRecord(ev[evi], s));
        NFOPP_CHECK_CUDA(cudaEventRecord(ev[evi + 1], s));
    }
    if (ev) evi += 2;
    // (6) collision sample points on the path                 :113-117 / constrained:78-83
    if ((rc = launch_traj_pre(d, s))) return rc;
    NFOPP_MARK();
    // (7) ONF forward + input gradient at the path points     :166-169 / constrained:85-89
    {
        OnfArgs a{};
        a.d = e->onf; a.weights = d.onf_w; a.x = d.path_q; a.count = d.path_n; a.n = d.T - 1; a.x_stride = d.D;
        a.want_bwd = 1;
        a.gbar_mode = d.constrained ? 3 : 2; a.aux = d.path_mu; a.wc = d.collision_weight;
        a.beta = d.constrained ? d.beta : 1.f; a.z_out = d.path_z; a.gx_out = d.path_gq;
        if ((rc = engine_onf(e, d, a, s))) return rc;
    }
    NFOPP_MARK();
    // (8) trajectory gradient, preconditioner, Adam, multipliers, re-parametrisation  :143-155,224-244
    if ((rc = launch_traj_post(d, s))) return rc;
    NFOPP_MARK();
    // (9) counters
    if (tick && (rc = launch_tick(d, false, s))) return rc;
    NFOPP_MARK();
#undef NFOPP_MARK
    return NFOPP_OK;
}

// one step of all problems inside a stream capture: lanes > 1 fork the problem ranges onto the lane streams and join them
// before the (single) counter tick
static int enqueue_step_lanes(nfopp_engine *e, cudaStream_t s, bool train_host = true) {
    const int L = e->lanes, P = e->dev.P;
    if (L <= 1) return enqueue_step(e, e->dev, s, nullptr, true, train_host);
    int rc = NFOPP_OK;
    NFOPP_CHECK_CUDA(cudaEventRecord(e->lane_events[0], s));
    for (int l = 0; l < L && !rc; ++l) {
        const int p0 = (int)((long long)P * l / L), p1 = (int)((long long)P * (l + 1) / L);
        const EngineDev d = slice_problems(e, p0, p1 - p0, l);
        cudaStream_t ls = l == 0 ? s : e->lane_streams[l];
        if (l > 0) NFOPP_CHECK_CUDA(cudaStreamWaitEvent(ls, e->lane_events[0], 0));
        rc = enqueue_step(e, d, ls, nullptr, false, train_host);
        if (l > 0 && !rc) NFOPP_CHECK_CUDA(cudaEventRecord(e->lane_events[l], ls));
    }
    for (int l = 1; l < L; ++l) NFOPP_CHECK_CUDA(cudaStreamWaitEvent(s, e->lane_events[l], 0));   // (joins even after an error)
    if (rc) return rc;
    return launch_tick(e->dev, false, s);
}

extern "C" int nfopp_engine_set_lanes(nfopp_engine *e, int lanes) {
    NFOPP_REQUIRE(e && lanes >= 1 && lanes <= 16, "set_lanes: lanes=%d", lanes);
    if (lanes > e->dev.P) lanes = e->dev.P;
    if (lanes == e->lanes) return NFOPP_OK;
    if (e->impl == 3 && lanes > 1) {
        // every lane needs its own activation workspace inside the one that was sized for all problems
        const EngineDev &d = e->dev;
        int nmax = d.NTR > d.NC ? d.NTR : d.NC;
        if (d.T - 1 > nmax) nmax = d.T - 1;
        const int pmax = (d.P + lanes - 1) / lanes;
        const size_t stride = (onf_wide_workspace_bytes(e->onf, pmax, nmax, true) + 255) / 256 * 256;
        size_t have = 0;
        for (const BufferInfo &b : e->buffers) if (b.name == "onf_ws") have = b.bytes;
        NFOPP_REQUIRE(stride * lanes <= have, "set_lanes: %d lanes need %zu bytes of wide-ONF workspace, the engine has %zu", lanes,
                      stride * lanes, have);
        e->lane_ws_stride = stride;
    }
    drop_graphs(e);
    while ((int)e->lane_streams.size() < lanes) {
        cudaStream_t st = nullptr;
        cudaEvent_t ev = nullptr;
        NFOPP_CHECK_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        NFOPP_CHECK_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        e->lane_streams.push_back(st);
        e->lane_events.push_back(ev);
    }
    e->lanes = lanes;
    e->launches_per_step = 1;                                  // the tick; everything else runs once per lane
    e->launches_per_step_skip = 1;
    for (int l = 0; l < lanes; ++l) {
        const int P = e->dev.P, p0 = (int)((long long)P * l / lanes), p1 = (int)((long long)P * (l + 1) / lanes);
        e->launches_per_step += step_launches(e, p1 - p0) - 1;
        e->launches_per_step_skip += step_launches(e, p1 - p0, false) - 1;
    }
    return NFOPP_OK;
}

extern "C" int nfopp_engine_set_bounds(nfopp_engine *e, const double *sample_border, int checker_bounds_mode,
                                       const double *checker_bounds) {
    NFOPP_REQUIRE(e && checker_bounds_mode >= -1 && checker_bounds_mode <= 1, "set_bounds: bad arguments");
    NFOPP_REQUIRE(checker_bounds_mode != 1 || checker_bounds, "set_bounds: checker_bounds missing");
    if (sample_border)
        for (int i = 0; i < 4; ++i) e->cfg.border[i] = sample_border[i];
    if (checker_bounds_mode >= 0) {
        e->cfg.has_bounds = checker_bounds_mode;
        for (int i = 0; i < 4; ++i) e->cfg.bounds[i] = checker_bounds_mode ? checker_bounds[i] : 0.0;
    }
    fill_dev_config(e);
    // kernel arguments are captured by value: the next graphed step re-captures
    drop_graphs(e);
    return NFOPP_OK;
}

extern "C" int nfopp_engine_warmup(nfopp_engine *e, int iterations, void *stream) {
    NFOPP_REQUIRE(e && iterations >= 0, "warmup: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    int rc;
    if (e->sp_dirty && (rc = upload_step_params(e, s))) return rc;
    if ((rc = ensure_packed(e, s))) return rc;
    for (int it = 0; it < iterations; ++it) {
        if ((rc = launch_warm_sample(e->dev, 1000000 + e->warm_iterations_done, s))) return rc;
        if ((rc = launch_resample_label(e->dev, 0, s))) return rc;
        if ((rc = enqueue_train(e, e->dev, true, s))) return rc;
        if ((rc = launch_tick(e->dev, true, s))) return rc;
        e->host_sp.onf_step += 1;
        e->warm_iterations_done += 1;
        e->launches += e->impl == 2 ? 4 : e->impl == 3 ? 4 + onf_wide_launch_count(e->onf, e->dev.P, e->dev.NTR, true, true) : 5;
    }
    host_step_params(e);
    return NFOPP_OK;
}

extern "C" int nfopp_engine_step(nfopp_engine *e, int n_steps, int use_graph, void *stream) {
    NFOPP_REQUIRE(e && n_steps >= 0, "step: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    int rc;
    if (e->sp_dirty && (rc = upload_step_params(e, s))) return rc;
    if ((rc = ensure_packed(e, s))) return rc;
    const bool two_graphs = e->impl == 3 && e->cfg.optimize_collision_model_freq > 1;
    auto capture = [&](bool train_host, cudaGraph_t *g, cudaGraphExec_t *ge) -> int {
        NFOPP_REQUIRE(s != nullptr, "step: CUDA-graph capture needs a non-default stream");
        NFOPP_CHECK_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        const int rc2 = enqueue_step_lanes(e, s, train_host);
        cudaError_t ce = cudaStreamEndCapture(s, g);
        if (rc2) return rc2;
        NFOPP_CHECK_CUDA(ce);
        NFOPP_CHECK_CUDA(cudaGraphInstantiate(ge, *g, 0));
        return NFOPP_OK;
    };
    if (use_graph && !e->graph_exec && (rc = capture(true, &e->graph, &e->graph_exec))) return rc;
    if (use_graph && two_graphs && !e->graph_exec_skip && (rc = capture(false, &e->graph_skip, &e->graph_exec_skip))) return rc;
    for (int it = 0; it < n_steps; ++it) {
        const bool train_host = e->host_sp.do_train != 0;
        const bool skip = two_graphs && !train_host;
        if (use_graph) {
            NFOPP_CHECK_CUDA(cudaGraphLaunch(skip ? e->graph_exec_skip : e->graph_exec, s));
        } else if ((rc = enqueue_step(e, e->dev, s, nullptr, true, train_host))) {
            return rc;
        }
        StepParams &sp = e->host_sp;
        if (sp.do_train) { sp.onf_step += 1; sp.has_prev = 1; }
        sp.traj_step += 1;
        sp.step_count += 1;
        host_step_params(e);
        e->launches += skip ? e->launches_per_step_skip : e->launches_per_step;
    }
    return NFOPP_OK;
}

extern "C" int nfopp_engine_profile_step(nfopp_engine *e, float *ms_out, int n_out, void *stream) {
    NFOPP_REQUIRE(e && ms_out && n_out >= 9, "profile_step: need room for 9 kernel times");
    cudaStream_t s = (cudaStream_t)stream;
    int rc;
    if (e->sp_dirty && (rc = upload_step_params(e, s))) return rc;
    if ((rc = ensure_packed(e, s))) return rc;
    cudaEvent_t ev[10];
    for (auto &x : ev) NFOPP_CHECK_CUDA(cudaEventCreate(&x));
    const bool prof_train = e->host_sp.do_train != 0;
    rc = enqueue_step(e, e->dev, s, ev, true, prof_train);
    if (!rc) {
        cudaError_t ce = cudaStreamSynchronize(s);
        if (ce != cudaSuccess) { set_error("profile_step: %s", cudaGetErrorString(ce)); rc = NFOPP_ERR_CUDA; }
    }
    if (!rc)
        for (int i = 0; i < 9; ++i) cudaEventElapsedTime(&ms_out[i], ev[i], ev[i + 1]);
    for (auto &x : ev) cudaEventDestroy(x);
    if (rc) return rc;
    StepParams &sp = e->host_sp;
    if (sp.do_train) { sp.onf_step += 1; sp.has_prev = 1; }
    sp.traj_step += 1;
    sp.step_count += 1;
    host_step_params(e);
    e->launches += e->launches_per_step;
    return NFOPP_OK;
}

extern "C" const char *nfopp_engine_kernel_name(int index) {
    return index >= 0 && index < 9 ? kStepKernelNames[index] : nullptr;
}

extern "C" int nfopp_engine_get_paths(nfopp_engine *e, float *paths, void *stream) {
    NFOPP_REQUIRE(e && paths, "get_paths: bad arguments");
    e->launches += 1;
    return launch_paths(e->dev, paths, (cudaStream_t)stream);
}

extern "C" int nfopp_engine_reparametrize(nfopp_engine *e, void *stream) {
    NFOPP_REQUIRE(e, "reparametrize: bad arguments");
    e->launches += 1;
    return launch_reparam(e->dev, (cudaStream_t)stream);
}

extern "C" int nfopp_engine_astar_workspace_bytes(nfopp_engine *e, int x_cells, int y_cells, size_t *bytes) {
    NFOPP_REQUIRE(e && bytes && x_cells >= 1 && y_cells >= 1, "astar_workspace_bytes: bad arguments");
    *bytes = astar_workspace_bytes(e->dev.P, x_cells, y_cells);
    return NFOPP_OK;
}

extern "C" int nfopp_engine_init_astar(nfopp_engine *e, int x_cells, int y_cells, double resolution, const double *origin_xy,
                                       const int32_t *start_cells, const int32_t *goal_cells, void *workspace, size_t workspace_bytes,
                                       int32_t *path_cells, void *stream) {
    NFOPP_REQUIRE(e, "init_astar: engine missing");
    e->launches += 3;
    return launch_astar_init(e->dev, x_cells, y_cells, resolution, origin_xy, start_cells, goal_cells,
                             static_cast<unsigned char *>(workspace), workspace_bytes, path_cells, (cudaStream_t)stream);
}

extern "C" int nfopp_engine_path_statistics(nfopp_engine *e, double cusp_angle_threshold, int clearing_window_cells, float *stats,
                                            void *stream) {
    NFOPP_REQUIRE(e && stats, "path_statistics: bad arguments");
    e->launches += 1;
    return launch_path_statistics(e->dev, cusp_angle_threshold, clearing_window_cells, stats, (cudaStream_t)stream);
}

extern "C" int nfopp_engine_evaluate_paths(nfopp_engine *e, int samples_per_segment, uint8_t *collides, float *length,
                                           int32_t *n_colliding, void *stream) {
    NFOPP_REQUIRE(e && collides && length, "evaluate_paths: bad arguments");
    e->launches += 1;
    return launch_evaluate_paths(e->dev, samples_per_segment, collides, length, n_colliding, (cudaStream_t)stream);
}

extern "C" int nfopp_engine_launch_count(nfopp_engine *e, int64_t *launches) {
    NFOPP_REQUIRE(e && launches, "launch_count: bad arguments");
    *launches = e->launches;
    return NFOPP_OK;
}
