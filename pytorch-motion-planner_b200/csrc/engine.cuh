// engine.cuh - device-side view of one planner engine (P problems) shared by sample.cu / traj.cu / engine.cu
#pragma once
#include "common.cuh"
#include "label.cuh"

namespace nfopp {

// per-step scalars kept in device memory so a captured CUDA graph of one step can be replayed: the last
// kernel of every step ("tick") advances the counters and precomputes the next step's Adam bias terms.
struct StepParams {
    int32_t step_count;      // nerf_opt_planner.py:35
    int32_t onf_step;        // Adam step counter of the ONF optimiser (updates done so far)
    int32_t traj_step;       // Adam step counter of the trajectory optimiser
    int32_t has_prev;        // nerf_opt_planner.py:37,79-82
    int32_t do_train;        // step_count % optimize_collision_model_freq == 0   (:62)
    int32_t do_reparam;      // step_count % reparametrize_trajectory_freq == 0   (:66)
    float onf_step_size, onf_bc2_sqrt;     // lr/(1-b1^k), sqrt(1-b2^k) for k = onf_step+1
    float traj_step_size, traj_bc2_sqrt;   // same for k = traj_step+1
};

// obstacle sets up to this many points per problem are staged in shared memory by the kernels that label poses; larger
// ones (ros/grid_map.py:14-20 turns an occupancy grid into up to ~1e5 occupied-cell centres) are read through L1 / L2
constexpr int OBS_SMEM_POINTS = 2048;
__host__ __device__ inline int obs_staged_points(int m_max) { return m_max <= OBS_SMEM_POINTS ? (m_max > 0 ? m_max : 1) : 1; }

struct EngineDev {
    // configuration (by value)
    int p_base;                            // global index of this view's problem 0 (lanes: RNG streams stay per global problem)
    int P, T, D, C, NF, NC, NTR, NW;       // C = collision_point_count, NF = random_field_points,
                                           // NC = C + T-1 (candidates), NTR = train capacity, NW = warm-up points
    int constrained, rng_mode, max_choice_u;
    uint64_t seed;
    double fine_offset, coarse_offset, angle_offset;
    double border[4];
    float border_f[4];
    float collision_weight, angle_weight, constraint_weight, multipliers_lr, boundary_weight, mu_lr, beta,
        direction_weight;
    float traj_w1, traj_beta2, traj_w2, traj_eps;
    double onf_lr_d, onf_b1_d, onf_b2_d, traj_lr_d, traj_b1_d, traj_b2_d;
    int reparam_freq, train_freq;
    CheckerParams checker;
    int env_shared, m_max;
    int gslots;                            // streamed tensor kernel: gradient hand-over buffer pairs behind onf_gpart
    // state
    StepParams *sp;
    float *onf_w, *onf_m, *onf_v, *onf_gpart;
    unsigned char *onf_wp;                  // tensor paths: bf16 hi/lo operand copy of the hidden-layer weights
    unsigned char *onf_ws;                  // wide path: activation / gradient operand tiles
    float *traj, *traj_m, *traj_v, *prev_traj, *lam, *mu, *start, *goal;
    double *buf_pos, *buf_age;
    int32_t *buf_n;
    // per-step scratch
    double *cand_pos, *cand_age;
    float *cand_x, *cand_z;
    int32_t *cand_n;
    double *train_pos;
    float *train_x, *train_y, *train_z;
    int32_t *train_n;
    float *path_q, *path_mu, *path_z, *path_gq, *path_t;
    int32_t *path_n;
    int32_t *choice_idx, *choice_used, *status;
    int32_t *frozen;                        // [P] != 0: the problem is finished - no kernel of a step touches its state any more
    // constants
    float *inv_hessian_t, *lin_s;
    double *obstacles;
    int32_t *obstacle_n;
    uint8_t *grid;
    // injected variates (rng_mode == 1)
    double *var_t_sample, *var_coarse_xy, *var_coarse_th, *var_fine_xy, *var_fine_th, *var_choice_u, *var_field_xy,
        *var_field_th, *var_warm_xy, *var_warm_th;
    float *var_t_path;
};

enum { STATUS_CHOICE_EXHAUSTED = 1, STATUS_NONFINITE = 2 };

}  // namespace nfopp
