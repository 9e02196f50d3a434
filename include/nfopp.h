/*
 * nfopp.h - C ABI of libnfopp.so: the B200-native (sm_100a) NFOPP optimisation inner loop.
 *
 * The reference (MisterMap/pytorch-motion-planner) is pure Python/PyTorch and has no FFI for this path;
 * each entry point below therefore cites the reference *Python* interface it replaces
 * (paths relative to the reference root, neural_field_optimal_planner/...).  The host side that
 * mirrors the reference's classes lives in pytorch-motion-planner_b200/nfopp_b200 and binds this
 * header through ctypes (INTEGRATION.md shows the stub a reference maintainer would add).
 *
 * Conventions
 *  - every function returns 0 on success or a negative nfopp_status; nfopp_last_error() gives the text
 *    (thread-local).  No exceptions, no torch types: plain pointers and sizes only.
 *  - every pointer named *_dev is a DEVICE pointer on the current CUDA device; the library never
 *    allocates device memory - the caller (PyTorch, for device memory and streams) owns all buffers.
 *  - `stream` is a cudaStream_t passed as void*; all work is asynchronous with respect to it.
 *  - "problems" (P) are independent planning problems, each with its own ONF weights, trajectory,
 *    environment and RNG stream; arrays are problem-major and contiguous.
 */
#ifndef NFOPP_H
#define NFOPP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NFOPP_ABI_VERSION 5
#define NFOPP_MAX_LAYERS 8

typedef enum {
    NFOPP_OK = 0,
    NFOPP_ERR_INVALID = -1,     /* bad argument / unsupported configuration */
    NFOPP_ERR_CUDA = -2,        /* a CUDA runtime call failed */
    NFOPP_ERR_WORKSPACE = -3,   /* workspace too small or misaligned */
    NFOPP_ERR_NAME = -4         /* unknown buffer name */
} nfopp_status;

/* ---- occupancy neural field (onf_model.py:8-31, angle_encoder.py:7-13) ---------------------------- */
typedef struct {
    float mean, sigma;          /* onf_model.py:27-28,38 */
    int32_t use_cos;            /* onf_model.py:14,40-43: 200 Fourier features (sin|cos) instead of 100 (sin) */
    int32_t angle;              /* onf_model.py:11-13: 20 heading features, input is (x,y,theta) */
    int32_t hidden;             /* width of the hidden layers (100 in the reference); multiple of 4 */
    int32_t layers;             /* number of hidden Linear+ReLU layers (2 in the reference), <= NFOPP_MAX_LAYERS */
    int32_t bias;               /* onf_model.py:29: encoding_layer bias present */
} nfopp_onf_config;

/* Packed float32 parameter blob of one ONF (every segment starts on a multiple of 4 floats):
 *   enc_w [F_pos,2] | enc_b [F_pos] | ang_b [A] | ang_f [A] | W_0 [H,F_in] | b_0 [H] | W_l [H,H] | b_l [H] ...
 *   | w_out [H+F_in] | b_out [1]              (row-major, same element order as torch's nn.Linear.weight) */
typedef struct {
    int32_t f_pos, n_angle, f_in, in_dim;
    int32_t off_enc_w, off_enc_b, off_ang_b, off_ang_f;
    int32_t off_w[NFOPP_MAX_LAYERS], off_b[NFOPP_MAX_LAYERS];
    int32_t off_w_out, off_b_out;
    int32_t n_params;           /* blob length in floats (multiple of 4) */
} nfopp_onf_layout;

int nfopp_abi_version(void);
const char *nfopp_last_error(void);
int nfopp_onf_get_layout(const nfopp_onf_config *cfg, nfopp_onf_layout *out);

/* ---- K1: ground-truth collision labelling (float64, bit-exact w.r.t. numpy) ----------------------- */
/* poses_dev: [P, n, pose_dim] float64 (pose_dim 2 or 3); labels_dev: [P, n] uint8 (0/1).
 * obstacles_dev: [P_env, m_max, 2] float64 with obstacle_count_dev [P_env] int32 (NULL => all m_max);
 * P_env is 1 (shared by all problems) or P.  bounds: (x0,x1,y0,y1) or NULL for "no boundaries". */

/* collision_checker/collision_checker.py:9-19 (base class: boundary only) */
int nfopp_label_boundary(const double *poses_dev, int P, int n, int pose_dim, const double *bounds,
                         uint8_t *labels_dev, void *stream);
/* collision_checker/circle_collision_checker.py:11-14; circle_directed_collision_checker.py:5-6 */
int nfopp_label_circle(const double *poses_dev, int P, int n, int pose_dim, const double *obstacles_dev,
                       const int32_t *obstacle_count_dev, int P_env, int m_max, double radius,
                       const double *bounds, uint8_t *labels_dev, void *stream);
/* collision_checker/rectangle_collision_checker.py:11-26 + utils/position2.py:91-94; box=(x0,x1,y0,y1) */
int nfopp_label_rectangle(const double *poses_dev, int P, int n, const double *obstacles_dev,
                          const int32_t *obstacle_count_dev, int P_env, int m_max, const double *box,
                          const double *bounds, uint8_t *labels_dev, void *stream);
/* notebooks/onf_planner_image_map.ipynb:65-80 (MapCollisionChecker); grid_dev: [P_env, rows, cols] uint8 */
int nfopp_label_grid(const double *poses_dev, int P, int n, int pose_dim, const uint8_t *grid_dev, int P_env,
                     int rows, int cols, double origin_x, double origin_y, double cell, const double *bounds,
                     uint8_t *labels_dev, void *stream);

/* ---- K2: ONF forward / backward with per-problem weights (onf_model.py:33-50) ---------------------- */
/* weights_dev [P, n_params]; x_dev [P, n, in_dim] float32; count_dev [P] int32 = valid points per problem
 * (NULL => n); z_dev [P, n] logits.
 * `impl`: 0 = auto, 1 = FMA (CUDA-core fp32), 2 = tensor (tcgen05, ONE fused kernel per call: bf16 hi/lo split
 * operands, fp32 accumulation in TMEM, weights resident in shared memory; 2 hidden layers, hidden <= 111,
 * f_in <= 224 = the reference's network), 3 = tensor-wide (tcgen05 grouped GEMM per layer with streamed weights,
 * the same split-bf16 arithmetic: any hidden <= 1024 (multiple of 4), up to NFOPP_MAX_LAYERS layers, any point
 * count - BASELINE configs 4 and 5).  auto = 2 when the shape fits it, else 3.
 * `scratch_dev`: device scratch of nfopp_onf_scratch_bytes_n(cfg, P, n, impl, want_bwd) bytes: the tensor paths
 * keep a bf16 operand copy of the hidden-layer weights there, the wide path also its activation / gradient
 * operand tiles (~(2*f_in + 2*layers*hidden) * 4 bytes per point); 0 bytes / NULL for the FMA path.
 * nfopp_onf_scratch_bytes is the n-independent form (impl 1 and 2 only). */
int nfopp_onf_scratch_bytes(const nfopp_onf_config *cfg, int P, int impl, size_t *bytes);
int nfopp_onf_scratch_bytes_n(const nfopp_onf_config *cfg, int P, int n, int impl, int want_bwd, size_t *bytes);
int nfopp_onf_forward(const nfopp_onf_config *cfg, const float *weights_dev, const float *x_dev,
                      const int32_t *count_dev, int P, int n, float *z_dev, int impl, void *scratch_dev,
                      void *stream);

/* nfopp_onf_forward on the streamed tensor path WITHOUT scratch_dev and without any allocation.  NO COUNTERPART IN THE
 * REFERENCE: added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  The fp16 hi/lo operand copy of the hidden-layer weights lives
 * in one static device array of NFOPP_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it
 * until exit); the entry point calls no cudaMalloc / cudaFree / allocating routine.  Limits and rules:
 *   - network shapes of the streamed tensor path only (2 hidden layers, hidden <= 111, f_in <= 256), forward only;
 *   - P * (operand copy per weight set) <= NFOPP_STATIC_SCRATCH_BYTES, else NFOPP_ERR_INVALID with a message naming the limit
 *     (reference network, 150 528 bytes per weight set: P <= 27); n is unlimited;
 *   - every call shares the array, which keeps the last call's contents; each call rewrites every byte it reads, so nothing
 *     is cleared; calls that overlap on the GPU (other streams / threads) RACE on it - serialise them on one stream;
 *   - works on the current device; results are bit-identical to nfopp_onf_forward(impl = 2). */
#define NFOPP_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int nfopp_onf_forward_static(const nfopp_onf_config *cfg, const float *weights_dev, const float *x_dev,
                             const int32_t *count_dev, int P, int n, float *z_dev, void *stream);

/* gbar_mode selects dL/dz:
 *   0 external : gz_dev [P,n] given (what autograd hands ONF.backward)
 *   1 bce      : (sigmoid(z) - y)/count, y = aux_dev [P,n]           nerf_opt_planner.py:25,85-89
 *   2 path     : w_c*sigmoid(z)                                      nerf_opt_planner.py:159,166-169
 *   3 path_mu  : w_c*softplus_beta'(z) + aux*(1-tanh(z)^2), aux=mu~  constrained_nerf_opt_planner.py:82-89
 * Outputs (each may be NULL): z_dev [P,n]; gx_dev [P,n,in_dim]; gpart_dev [P, slots, n_params] weight
 * gradients, slots = nfopp_onf_tiles(cfg, n, impl): per-point-tile partials for the FMA path, ONE reduced
 * gradient for the tensor path (it accumulates over the problem's points in TMEM); reduced in fixed order by
 * nfopp_onf_adam / nfopp_onf_reduce_grad called with the same cfg, n and impl.  x_dev rows have in_dim columns. */
int nfopp_onf_tiles(const nfopp_onf_config *cfg, int n, int impl);   /* gradient slots per problem for n points */
int nfopp_onf_backward(const nfopp_onf_config *cfg, const float *weights_dev, const float *x_dev,
                       const int32_t *count_dev, int P, int n, int gbar_mode, const float *gz_dev,
                       const float *aux_dev, float collision_weight, float beta, float *z_dev, float *gx_dev,
                       float *gpart_dev, int impl, void *scratch_dev, void *stream);
/* sums the gradient slots [P, slots, n_params] into grad_dev [P, n_params] (fixed order) */
int nfopp_onf_reduce_grad(const nfopp_onf_config *cfg, const float *gpart_dev, const int32_t *count_dev, int P,
                          int n, int impl, float *grad_dev, void *stream);
/* torch.optim.Adam single-tensor update (planner_factory.py:54,56,68,70) over [P, n_params]; `step` is the
 * 1-based index of this update; gpart_dev is what nfopp_onf_backward wrote for the same cfg, n and impl. */
int nfopp_onf_adam(const nfopp_onf_config *cfg, float *weights_dev, float *m_dev, float *v_dev,
                   const float *gpart_dev, const int32_t *count_dev, int P, int n, int impl, int step, double lr,
                   double beta1, double beta2, double eps, void *stream);

/* numpy legacy RandomState.choice(n, size, replace, p) on given uniforms (nerf_opt_planner.py:131).
 * p_dev [P, n] float64 (normalised), u_dev [P, n_u] float64; idx_dev [P, size] int32; used_dev [P] int32 =
 * uniforms consumed (negative => stream exhausted). */
int nfopp_choice(const double *p_dev, int P, int n, int size, int replace, const double *u_dev, int n_u,
                 int32_t *idx_dev, int32_t *used_dev, void *stream);

/* ---- the planner engine: P independent NFOPP problems stepped together -------------------------- */
typedef struct {
    nfopp_onf_config onf;
    int32_t problems;                   /* P */
    int32_t constrained;                /* 0: NERFOptPlanner (x,y)   1: ConstrainedNERFOptPlanner (x,y,theta) */
    int32_t trajectory_length;          /* T (planner_factory.py:15,55,69) */
    /* optimisers (planner_factory.py:54-56,68-70) */
    double onf_lr, onf_beta1, onf_beta2, traj_lr, traj_beta1, traj_beta2, adam_eps;
    /* NERFOptPlanner.__init__ (nerf_opt_planner.py:11-14) */
    double trajectory_random_offset, course_random_offset;
    float collision_weight;
    int32_t init_collision_points, reparametrize_trajectory_freq, optimize_collision_model_freq,
        random_field_points, collision_point_count;
    /* ConstrainedNERFOptPlanner.__init__ (constrained_nerf_opt_planner.py:13-18) */
    double angle_offset;
    float angle_weight, constraint_deltas_weight, multipliers_lr, boundary_weight, collision_multipliers_lr,
        collision_beta, direction_delta_weight;
    /* collision checker (collision_checker/*.py): 0 boundary only, 1 circle, 2 rectangle, 3 grid */
    int32_t checker_kind;
    int32_t env_shared;                 /* 1: one obstacle set / grid for all problems, 0: one per problem */
    int32_t max_obstacles;              /* m_max (circle / rectangle) */
    int32_t grid_rows, grid_cols;       /* grid checker */
    double grid_origin_x, grid_origin_y, grid_cell;
    double robot_radius;                /* circle */
    double box[4];                      /* rectangle */
    int32_t has_bounds;
    double bounds[4];                   /* checker boundaries (x0,x1,y0,y1) */
    double border[4];                   /* sampling / boundary-loss border = init(..., boundaries) */
    /* randomness: 0 = on-device Philox4x32-10 (throughput), 1 = variates injected by the host (parity) */
    int32_t rng_mode;
    uint64_t seed;
    int32_t max_choice_uniforms;        /* per problem per step, injected mode */
    int32_t onf_impl;                   /* 0 auto, 1 FMA, 2 tensor (fused, reference shape), 3 tensor-wide */
} nfopp_planner_config;

typedef struct nfopp_engine nfopp_engine;

/* Bytes of device workspace the engine needs (256-byte aligned).  The caller allocates it (torch.empty). */
int nfopp_engine_workspace_bytes(const nfopp_planner_config *cfg, size_t *bytes);
/* Replaces the constructors NERFOptPlanner.__init__ / ConstrainedNERFOptPlanner.__init__
 * (nerf_opt_planner.py:11-43, constrained_nerf_opt_planner.py:13-40) for P problems at once. */
int nfopp_engine_create(const nfopp_planner_config *cfg, void *workspace_dev, size_t workspace_bytes,
                        nfopp_engine **out);
int nfopp_engine_destroy(nfopp_engine *e);
/* Named state / IO buffers inside the workspace (device pointer, byte size, dtype tag 'f','d','i','b' and up
 * to 3 dims).  The host wraps them as tensors: this is export_state()/import_state() and the place where
 * environments, start/goal, initial weights and (parity mode) variates are written.  Names: see DESIGN.md. */
int nfopp_engine_buffer(nfopp_engine *e, const char *name, void **ptr_dev, size_t *bytes, char *dtype,
                        int32_t dims[3]);
int nfopp_engine_buffer_count(nfopp_engine *e);
const char *nfopp_engine_buffer_name(nfopp_engine *e, int index);
/* The tensor path works on a bf16 operand copy of the hidden-layer weights that its training kernel keeps
 * current.  nfopp_engine_buffer(e, "onf_w", ...) marks the copy stale (the caller may write the weights);
 * call this after writing through a pointer obtained earlier (e.g. an nn.Module view of the buffer).  The copy
 * is rebuilt at the start of the next warmup / step / profile_step. */
int nfopp_engine_invalidate_weights(nfopp_engine *e);
/* Host-side counters (the reference keeps them as Python ints / optimizer state):
 * which: 0 = step_count (nerf_opt_planner.py:35), 1 = ONF Adam step, 2 = trajectory Adam step,
 *        3 = has_previous_trajectory (nerf_opt_planner.py:37,79-82) */
int nfopp_engine_get_counter(nfopp_engine *e, int which, int64_t *value);
int nfopp_engine_set_counter(nfopp_engine *e, int which, int64_t value);
/* NERFOptPlanner._init_collision_model (nerf_opt_planner.py:197-200): `iterations` ONF training steps on
 * init_collision_points uniform field points each. */
int nfopp_engine_warmup(nfopp_engine *e, int iterations, void *stream);
/* NERFOptPlanner.step (nerf_opt_planner.py:60-71) / the constrained overrides, `n_steps` times for all P
 * problems.  use_graph != 0 replays a captured CUDA graph of one step (Philox mode only). */
int nfopp_engine_step(nfopp_engine *e, int n_steps, int use_graph, void *stream);
/* One step() launched kernel by kernel with CUDA events in between (no graph): ms_out[0..9) = device time of the
 * nine kernels of a step, in launch order; nfopp_engine_kernel_name(i) names them.  bench.py's live roofline. */
int nfopp_engine_profile_step(nfopp_engine *e, float *ms_out, int n_out, void *stream);
const char *nfopp_engine_kernel_name(int index);
/* NERFOptPlanner.get_path / full_trajectory (nerf_opt_planner.py:73-74,178-179): [P, T+2, D] float32 */
int nfopp_engine_get_paths(nfopp_engine *e, float *paths_dev, void *stream);
/* NERFOptPlanner.reparametrize_trajectory (nerf_opt_planner.py:224-244; constrained:132-171) on the stored
 * trajectories and multipliers of all problems, outside a step (update_goal_point / update_start_point). */
int nfopp_engine_reparametrize(nfopp_engine *e, void *stream);
/* Path evaluation for the stop criterion of scripts/run_bench_mr.py:119-126 (benchmark.evaluate_path ->
 * BenchmarkAdapter.cpp:201-208 evaluatePath: interpolate the path, validity-check every interpolated pose, path
 * length), for all P problems: `samples_per_segment` equally spaced poses on each of the T+1 segments of the
 * full path (x, y linear; heading along the wrapped difference; float64) plus the goal, each labelled by the
 * engine's own collision checker.  collides_dev [P] uint8 (1 = some pose collides), length_dev [P] float32
 * (sum of x,y segment lengths), n_colliding_dev [P] int32 or NULL. */
int nfopp_engine_evaluate_paths(nfopp_engine *e, int samples_per_segment, uint8_t *collides_dev, float *length_dev,
                                int32_t *n_colliding_dev, void *stream);
/* AstarTrajectoryInitializer.initialize_trajectory (astar/astar_trajectory_initializer.py:15-46) for all P problems:
 * the engine's collision checker labels the centres of an x_cells x y_cells grid over the checker boundaries (cell size
 * `resolution`, heading 3 pi / 4), A* (astar/jps.py search with jps=False: 8-connected, sqrt(2) diagonals, Euclidean
 * heuristic, heap order (key, (row, col))) runs from every problem's start cell to its goal cell, and the cell path
 * (+ start, goal) is re-parametrised to the T trajectory points by utils/math.py:57-65 (quadratic interpolating
 * spline).  Writes x, y of the stored trajectories; headings are left as they are.  start_cells_dev / goal_cells_dev
 * [P, 2] int32 (row = y cell, column = x cell; computed by the caller with the reference's float32 floor division),
 * workspace_dev from nfopp_engine_astar_workspace_bytes, path_cells_dev [P] int32 (cells on each A* path) or NULL.
 * origin_xy (host, 2 doubles: x_min, y_min of collision_checker.get_boundaries(), astar_trajectory_initializer.py:29-34) places
 * the cell grid; NULL = the labelling boundaries of the engine's checker.  The two differ for a checker that reports
 * boundaries but does not OR them into its labels (the notebook's MapCollisionChecker). */
int nfopp_engine_astar_workspace_bytes(nfopp_engine *e, int x_cells, int y_cells, size_t *bytes);
int nfopp_engine_init_astar(nfopp_engine *e, int x_cells, int y_cells, double resolution, const double *origin_xy,
                            const int32_t *start_cells_dev, const int32_t *goal_cells_dev, void *workspace_dev,
                            size_t workspace_bytes, int32_t *path_cells_dev, void *stream);
/* Path-quality statistics of every problem's current path, the numbers BenchmarkAdapter::evaluate
 * (benchmark/bench_mr_binding/BenchmarkAdapter.cpp:68-103) files per plan: stats_dev [P, 8] float32 =
 * {path length (x,y), max curvature, normalised curvature (sum of curvature x arc length), angle over length,
 * OMPL smoothness, cusps (computeCusps, BenchmarkAdapter.cpp:105-135, yaw change > cusp_angle_threshold),
 * min clearing, mean clearing} on the T+2 path nodes in float64.  Clearing = distance from the pose to the nearest
 * obstacle of the engine's checker (disc / box / occupied cell within +-clearing_window_cells / boundary), 0 in collision. */
int nfopp_engine_path_statistics(nfopp_engine *e, double cusp_angle_threshold, int clearing_window_cells,
                                 float *stats_dev, void *stream);
/* Lanes of the graphed step (default 1): the problems are split into `lanes` contiguous ranges, each enqueued on its own
 * stream inside the step's CUDA graph (fork / join around the single counter tick).  The reference's problems are
 * independent (one planner object each), so there is no cross-lane dependency; results are bit-identical for any lane
 * count (per-problem RNG streams keep their global problem index; on the wide ONF path every lane takes its own slice of
 * the activation workspace and the call fails if the slices do not fit). */
int nfopp_engine_set_lanes(nfopp_engine *e, int lanes);
/* NERFOptPlanner.set_boundaries (nerf_opt_planner.py:220-222: the rectangle the uniform field points are drawn from) and
 * CollisionChecker.update_boundaries (collision_checker/collision_checker.py:24-25: the rectangle outside of which every pose
 * collides) on a live engine, without re-creating it: both are part of the kernels' arguments, so the captured step graph is
 * dropped and re-captured by the next nfopp_engine_step.  sample_border (x0,x1,y0,y1) or NULL = unchanged;
 * checker_bounds_mode: -1 unchanged, 0 the checker has no boundaries, 1 use checker_bounds (x0,x1,y0,y1).  State, counters
 * and the ONF operand copy are untouched. */
int nfopp_engine_set_bounds(nfopp_engine *e, const double *sample_border, int checker_bounds_mode, const double *checker_bounds);
/* PathPostprocessor.process (ros/path_postprocessor.py:13-71) for P paths of n nodes each (what the ROS adapter does to a
 * path before publishing it): nodes closer than minimal_distance to their successor are dropped, the path is
 * re-sampled every distance_step metres along its chord-length parametrisation with scipy's quadratic interpolating
 * spline (headings unfolded first, utils/math.py:38-43), and the leading poses up to a direction flip within the first six
 * segments are removed.  paths_dev [P, n, 3] float32 (nfopp_engine_get_paths), out_dev [P, max_points, 3] float64,
 * count_dev [P] int32: poses written; -1 = fewer than 3 distinct nodes (scipy raises there); -k (k > max_points) = the
 * path needs k poses, nothing written.  workspace_dev from nfopp_postprocess_workspace_bytes.  Not engine-bound. */
int nfopp_postprocess_workspace_bytes(int P, int n, size_t *bytes);
int nfopp_postprocess_paths(const float *paths_dev, int P, int n, double minimal_distance, double distance_step,
                            int max_points, void *workspace_dev, size_t workspace_bytes, double *out_dev,
                            int32_t *count_dev, void *stream);
/* number of kernel launches issued by this engine so far (bench.py's gpu_launches) */
int nfopp_engine_launch_count(nfopp_engine *e, int64_t *launches);

#ifdef __cplusplus
}
#endif
#endif /* NFOPP_H */
