"""Oracle: trajectory loss gradient (analytic), preconditioner, multiplier ascent, re-parametrisation.

torch float32 on CPU; restates nerf_opt_planner.py:143-176,224-248 and
constrained_nerf_opt_planner.py:63-171 without autograd.
"""
import numpy as np
import torch


def wrap_angle(a: torch.Tensor) -> torch.Tensor:
    """torch_math.py:5-7: (a + pi) % (2 pi) - pi, python-sign modulo, float32."""
    return (a + np.pi) % (2 * np.pi) - np.pi


def inv_hessian(n: int, weight: float) -> torch.Tensor:
    """nerf_opt_planner.py:45-58: inverse of weight*tridiag(-2,4,-2) + I, float64 inverse cast to float32."""
    k = np.zeros((n, n), dtype=np.float32)
    idx = np.arange(n)
    k[idx, idx] = 4
    k[idx[1:], idx[:-1]] = -2
    k[idx[:-1], idx[1:]] = -2
    return torch.tensor(np.linalg.inv(weight * k + np.eye(n)).astype(np.float32))


def path_points(traj: torch.Tensor, t: torch.Tensor, constrained: bool):
    """Collision sample points on the current path.
    holonomic  nerf_opt_planner.py:113-117,158 : X[i+1](1-t) + X[i] t
    constrained constrained_nerf_opt_planner.py:78-81: X[i+1] + t * (X[i]-X[i+1]) with wrapped theta delta."""
    t = t.to(torch.float32).reshape(-1, 1)
    if not constrained:
        return traj[1:] * (1 - t) + traj[:-1] * t
    delta = traj[:-1] - traj[1:]
    delta[:, 2] = wrap_angle(delta[:, 2])
    return traj[1:] + t * delta


def interp_multipliers(mu: torch.Tensor, t: torch.Tensor):
    """constrained_nerf_opt_planner.py:82-83."""
    t = t.to(torch.float32).reshape(-1)
    return mu[1:] * (1 - t) + mu[:-1] * t


def path_gbar(z: torch.Tensor, mu_i, collision_weight: float, beta: float):
    """dL/dz for the path term.  holonomic: w_c*sigmoid(z) (nerf_opt_planner.py:159,166-169).
    constrained: w_c*softplus_beta'(z) + mu_i*(1-tanh(z)^2) (constrained_nerf_opt_planner.py:87-89,94-97);
    softplus uses torch's threshold 20 on beta*z."""
    zb = z * beta
    ez = torch.exp(zb)
    sp = torch.where(zb > 20, torch.ones_like(z), ez / (ez + 1))
    g = collision_weight * sp
    if mu_i is not None:
        th = torch.tanh(z)
        g = g + mu_i * (1 - th * th)
    return g


def softplus_beta(z, beta):
    zb = z * beta
    return torch.where(zb > 20, z, torch.log1p(torch.exp(zb)) / beta)


def holonomic_gradient(traj, start, goal, t, gq):
    """Gradient of nerf_opt_planner.py:157-169 w.r.t. the trajectory given gq = dL/dq (already
    including collision_weight through path_gbar)."""
    full = torch.cat([start[None], traj, goal[None]], 0)
    d = full[1:] - full[:-1]
    g = torch.zeros_like(full)
    g[1:] += 2 * d
    g[:-1] -= 2 * d
    g = g[1:-1].clone()
    t = t.to(torch.float32).reshape(-1, 1)
    g[:-1] += t * gq
    g[1:] += (1 - t) * gq
    return g


def constrained_gradient(traj, start, goal, lam, mu, t, z, gq, *, angle_weight, constraint_weight,
                         boundary_weight, direction_weight, border):
    """Gradient of constrained_nerf_opt_planner.py:76-100 w.r.t. trajectory, lambda (constraint
    multipliers) and mu (collision multipliers).  z = ONF logits at the path points, gq = dL/dq."""
    full = torch.cat([start[None], traj, goal[None]], 0)
    d = full[1:] - full[:-1]
    g = torch.zeros_like(full)
    # distance term (:120-130)
    dth = d[:, 2].clone()
    s = wrap_angle(dth).sum() - full[-1, 2] + full[0, 2]
    dth[-1] = dth[-1] + s
    dth = dth * angle_weight
    g[1:, :2] += 2 * d[:, :2]
    g[:-1, :2] -= 2 * d[:, :2]
    g[1:, 2] += 2 * angle_weight * dth
    g[:-1, 2] -= 2 * angle_weight * dth
    # non-holonomic residual (:102-109) with Lagrangian + quadratic weight (:94-96)
    dx, dy, a = d[:, 0], d[:, 1], full[:, 2]
    abar = a[:-1] + wrap_angle(a[1:] - a[:-1]) / 2
    sn, cs = torch.sin(abar), torch.cos(abar)
    c = dx * sn - dy * cs
    kap = lam + 2 * constraint_weight * c
    g[1:, 0] += kap * sn
    g[:-1, 0] -= kap * sn
    g[1:, 1] -= kap * cs
    g[:-1, 1] += kap * cs
    dcda = kap * (dx * cs + dy * sn) * 0.5
    g[:-1, 2] += dcda
    g[1:, 2] += dcda
    # forward-only direction term (:111-118, :93, :98); note wrap(a_j - a_{j+1})
    if direction_weight != 0:
        abar2 = a[:-1] + wrap_angle(a[:-1] - a[1:]) / 2
        sn2, cs2 = torch.sin(abar2), torch.cos(abar2)
        dd = -(cs2 * dx + sn2 * dy)
        rho = 2 * direction_weight * torch.where(dd > 0, dd, torch.zeros_like(dd))
        g[1:, 0] -= rho * cs2
        g[:-1, 0] += rho * cs2
        g[1:, 1] -= rho * sn2
        g[:-1, 1] += rho * sn2
        ddda = rho * (sn2 * dx - cs2 * dy)
        g[:-1, 2] += 1.5 * ddda
        g[1:, 2] -= 0.5 * ddda
    g = g[1:-1].clone()
    # boundary term (nerf_opt_planner.py:171-176) on interior points
    x, y = traj[:, 0], traj[:, 1]
    g[:, 0] += boundary_weight * 2 * (torch.relu(x - border[1]) - torch.relu(-x + border[0]))
    g[:, 1] += boundary_weight * 2 * (torch.relu(y - border[3]) - torch.relu(-y + border[2]))
    # collision term through the path points (:79-89)
    tt = t.to(torch.float32).reshape(-1, 1)
    g[:-1] += tt * gq
    g[1:] += (1 - tt) * gq
    th = torch.tanh(z)
    g_mu = torch.zeros_like(mu)
    g_mu[:-1] += tt[:, 0] * th
    g_mu[1:] += (1 - tt[:, 0]) * th
    return g, c, g_mu


def constrained_loss_value(traj, start, goal, lam, mu, t, z, *, collision_weight, beta, angle_weight,
                           constraint_weight, boundary_weight, direction_weight, border):
    """Scalar value of constrained_nerf_opt_planner.py:94-98 (used only to cross-check the reference)."""
    full = torch.cat([start[None], traj, goal[None]], 0)
    d = full[1:] - full[:-1]
    dth = d[:, 2].clone()
    dth[-1] = dth[-1] + wrap_angle(dth).sum() - full[-1, 2] + full[0, 2]
    dist = (d[:, :2] ** 2).sum() + ((dth * angle_weight) ** 2).sum()
    a = full[:, 2]
    abar = a[:-1] + wrap_angle(a[1:] - a[:-1]) / 2
    c = d[:, 0] * torch.sin(abar) - d[:, 1] * torch.cos(abar)
    abar2 = a[:-1] + wrap_angle(a[:-1] - a[1:]) / 2
    dd = torch.relu(-(torch.cos(abar2) * d[:, 0] + torch.sin(abar2) * d[:, 1]))
    x, y = traj[:, 0], traj[:, 1]
    bnd = (torch.relu(-x + border[0]) ** 2 + torch.relu(x - border[1]) ** 2 + torch.relu(-y + border[2]) ** 2
           + torch.relu(y - border[3]) ** 2).sum()
    mu_i = interp_multipliers(mu, t)
    return (dist + collision_weight * softplus_beta(z, beta).sum() + (lam * c).sum()
            + constraint_weight * (c ** 2).sum() + boundary_weight * bnd + (mu_i * torch.tanh(z)).sum()
            + direction_weight * (dd ** 2).sum())


def reparam_indices(full_xy_or_all: torch.Tensor):
    """Shared part of nerf_opt_planner.py:224-243 / constrained:132-150: (lo, hi, tau)."""
    seg = torch.norm(full_xy_or_all[1:] - full_xy_or_all[:-1], dim=1)
    cdf = torch.cat([torch.zeros(1), torch.cumsum(seg / torch.sum(seg), 0)])
    n = len(full_xy_or_all)
    s = torch.linspace(0, 1, n)[1:-1]
    idx = torch.searchsorted(cdf, s)
    hi = torch.clamp(idx, max=n - 1)
    lo = torch.clamp(idx - 1, min=0)
    den = cdf[hi] - cdf[lo]
    den = torch.where(den < 1e-5, torch.full_like(den, 1e-5), den)
    return lo, hi, (s - cdf[lo]) / den


def reparam_holonomic(traj, start, goal):
    """nerf_opt_planner.py:224-248 (distances over all columns)."""
    full = torch.cat([start[None], traj, goal[None]], 0)
    lo, hi, tau = reparam_indices(full)
    return (1 - tau[:, None]) * full[lo] + tau[:, None] * full[hi]


def reparam_constrained(traj, start, goal, lam, mu):
    """constrained_nerf_opt_planner.py:132-171 (distances over x,y only, :45-47)."""
    full = torch.cat([start[None], traj, goal[None]], 0)
    lo, hi, tau = reparam_indices(full[:, :2])
    new = torch.empty_like(traj)
    new[:, :2] = (1 - tau[:, None]) * full[lo, :2] + tau[:, None] * full[hi, :2]
    new[:, 2] = full[lo, 2] + tau * wrap_angle(full[hi, 2] - full[lo, 2])
    mu_pad = torch.cat([torch.zeros(1), mu, torch.zeros(1)])
    new_mu = (1 - tau) * mu_pad[lo] + tau * mu_pad[hi]
    node = torch.cat([lam[:1], (lam[:-1] + lam[1:]) / 2, lam[-1:]])
    c = (1 - tau) * node[lo] + tau * node[hi]
    new_lam = torch.cat([c[:1], (c[:-1] + c[1:]) / 2, c[-1:]])
    return new, new_lam, new_mu
