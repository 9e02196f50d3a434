"""ConstrainedNERFOptPlanner - the reference's SE(2) non-holonomic planner API
(constrained_nerf_opt_planner.py:12-194) over the batched CUDA engine (P=1)."""
import torch

from .nerf_opt_planner import NERFOptPlanner
from .utils.position2 import Position2


class ConstrainedNERFOptPlanner(NERFOptPlanner):
    _constrained = True

    def __init__(self, trajectory, collision_model, collision_checker, collision_optimizer, trajectory_optimizer,
                 trajectory_initializer, trajectory_random_offset, collision_weight, velocity_hessian_weight,
                 init_collision_iteration=100, init_collision_points=100, reparametrize_trajectory_freq=10,
                 optimize_collision_model_freq=1, random_field_points=10, angle_weight=0.5, constraint_deltas_weight=20,
                 multipliers_lr=1e-1, boundary_weight=1, collision_multipliers_lr=1e-3, angle_offset=0,
                 collision_beta=1, direction_delta_weight=0, rng_mode="philox", seed=None, use_graph=True):
        # like the reference (:19-22), course_random_offset / collision_point_count / collision_loss_koef are
        # not configurable here and stay 1.5 / 100 / 1
        super().__init__(trajectory, collision_model, collision_checker, collision_optimizer, trajectory_optimizer,
                         trajectory_random_offset, collision_weight, velocity_hessian_weight, init_collision_iteration,
                         init_collision_points, reparametrize_trajectory_freq, optimize_collision_model_freq,
                         random_field_points, rng_mode=rng_mode, seed=seed, use_graph=use_graph)
        self._angle_weight = angle_weight
        self._constraint_delta_weight = constraint_deltas_weight
        self._multipliers_lr = multipliers_lr
        self._collision_multipliers_lr = collision_multipliers_lr
        self._boundary_weight = boundary_weight
        self._angle_offset = angle_offset
        self._collision_beta = collision_beta
        self._trajectory_initializer = trajectory_initializer
        self._direction_delta_weight = direction_delta_weight
        t = trajectory.shape[0]
        self._lam0 = torch.zeros(t + 1, device=self._device)
        self._mu0 = torch.zeros(t, device=self._device)

    def _extra_config(self):
        return dict(angle_weight=self._angle_weight, constraint_deltas_weight=self._constraint_delta_weight,
                    multipliers_lr=self._multipliers_lr, boundary_weight=self._boundary_weight,
                    collision_multipliers_lr=self._collision_multipliers_lr, angle_offset=self._angle_offset,
                    collision_beta=self._collision_beta, direction_delta_weight=self._direction_delta_weight)

    @property
    def _constraint_multipliers(self):
        return self._lam0 if self._engine is None else self._engine["lam"][0]

    @property
    def _collision_multipliers(self):
        return self._mu0 if self._engine is None else self._engine["mu"][0]

    @property
    def checked_positions(self):
        if self._engine is None:
            return super().checked_positions
        return Position2.from_vec(self._checked()[0])
