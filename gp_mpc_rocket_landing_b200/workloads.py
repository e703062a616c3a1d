"""Synthetic workloads of BASELINE.json's configs: the GP residual data set and the Monte Carlo
initial-condition distribution.  Host-side input generation only (seeded numpy); no numerics of the hot path."""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


def residual_dataset(n: int = 1000, seed: int = 0):
    """Config 1 GP data: n synthetic residual points (X (n,7), U (n,3), D (n,3)); drag-like residual
    -0.5 rho(alt) C_D A |v| v / m (cf. experiments/dispersion.py:349-360) + N(0, 0.01^2)."""
    rng = np.random.default_rng(seed)
    m = rng.uniform(1.2, 2, n); alt = rng.uniform(0, 40, n)
    y = rng.normal(0, 3, n); z = rng.normal(0, 3, n)
    vx = rng.normal(-3, 1, n); vy = rng.normal(0, .5, n); vz = rng.normal(0, .5, n)
    X = np.stack([m, alt, y, z, vx, vy, vz], 1)
    U = np.stack([rng.uniform(.3, 5, n), rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)], 1)
    rho = 1.225 * np.exp(-alt / 8500)
    v = X[:, 4:7]; sp = np.linalg.norm(v, axis=1)
    D = -0.5 * rho[:, None] * 0.5 * 0.5 * sp[:, None] * v / m[:, None] + rng.normal(0, 0.01, (n, 3))
    return X, U, D


@dataclass
class SimulationConfig:
    """experiments/monte_carlo.py:107-130 fields used by the IC sampler, with the feasible values of
    scripts/run_experiments.py:359-371 as defaults (the class defaults are infeasible for the QP box)."""

    dt: float = 0.1
    max_time: float = 30.0
    altitude_mean: float = 30.0
    altitude_std: float = 5.0
    horizontal_std: float = 3.0
    velocity_mean: tuple = (-3.0, 0.0, 0.0)
    velocity_std: tuple = (1.0, 0.5, 0.5)
    mass_mean: float = 2.0
    mass_std: float = 0.1


def sample_initial_condition(seed: int, cfg: SimulationConfig) -> np.ndarray:
    """experiments/monte_carlo.py:368-399: np.random.seed(seed); 7 randn draws in fixed order; clips."""
    np.random.seed(seed)
    m = np.clip(cfg.mass_mean + np.random.randn() * cfg.mass_std, 1.5, 2.5)
    alt = np.clip(cfg.altitude_mean + np.random.randn() * cfg.altitude_std, 10, 100)
    hy = np.random.randn() * cfg.horizontal_std
    hz = np.random.randn() * cfg.horizontal_std
    vv = min(cfg.velocity_mean[0] + np.random.randn() * cfg.velocity_std[0], -1)
    vy = cfg.velocity_mean[1] + np.random.randn() * cfg.velocity_std[1]
    vz = cfg.velocity_mean[2] + np.random.randn() * cfg.velocity_std[2]
    return np.array([m, alt, hy, hz, vv, vy, vz])


def sample_initial_conditions(first: int, count: int, seed: int = 42, cfg: SimulationConfig | None = None):
    """Scenarios [first, first+count): x0 (count,7) with seed+i per scenario (monte_carlo.py:606-616), and the
    targets x_target = [m0, 0...] (monte_carlo.py:428-429).  Reproducible for any sharding."""
    cfg = cfg or SimulationConfig()
    state = np.random.get_state()
    X0 = np.stack([sample_initial_condition(seed + i, cfg) for i in range(first, first + count)]) if count else np.zeros((0, 7))
    np.random.set_state(state)
    XT = np.zeros_like(X0)
    XT[:, 0] = X0[:, 0]
    return X0, XT


def residual_dataset_6dof(n: int = 1000, seed: int = 0):
    """Config 4 GP data: n synthetic residual points on the 14-state model (X (n,14), U (n,3), D_v (n,3), D_omega (n,3)):
    the translational residual is the drag-like term of `residual_dataset`, the rotational one a damping-like term
    -0.05 |w| w + 0.02 (v_B x e_x) (cf. experiments/dispersion.py:349-360), both + N(0, 0.01^2)."""
    rng = np.random.default_rng(seed)
    m = rng.uniform(1.2, 2, n); alt = rng.uniform(0, 40, n)
    y = rng.normal(0, 3, n); z = rng.normal(0, 3, n)
    v = np.stack([rng.normal(-3, 1, n), rng.normal(0, .5, n), rng.normal(0, .5, n)], 1)
    q = np.concatenate([np.ones((n, 1)), rng.normal(0, 0.05, (n, 3))], 1)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w = rng.normal(0, 0.1, (n, 3))
    X = np.concatenate([m[:, None], alt[:, None], y[:, None], z[:, None], v, q, w], 1)
    U = np.stack([rng.uniform(1.5, 5, n), rng.uniform(-.5, .5, n), rng.uniform(-.5, .5, n)], 1)
    rho = 1.225 * np.exp(-alt / 8500)
    sp = np.linalg.norm(v, axis=1)
    Dv = -0.5 * rho[:, None] * 0.5 * 0.5 * sp[:, None] * v / m[:, None] + rng.normal(0, 0.01, (n, 3))
    wn = np.linalg.norm(w, axis=1)
    Dw = -0.05 * wn[:, None] * w + 0.02 * np.cross(v, np.array([1.0, 0.0, 0.0])) + rng.normal(0, 0.01, (n, 3))
    return X, U, Dv, Dw


def sample_initial_conditions_6dof(first: int, count: int, seed: int = 42, cfg: SimulationConfig | None = None):
    """The 3-DoF scenario draw embedded in the 14-state model: identity attitude perturbed by a small seeded rotation,
    zero body rates.  Targets: [m0, 0, 0, 0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0]."""
    X7, _ = sample_initial_conditions(first, count, seed, cfg)
    X0 = np.zeros((count, 14)); XT = np.zeros((count, 14))
    X0[:, :7] = X7
    for i in range(count):
        r = np.random.default_rng(seed + first + i)
        q = np.concatenate([[1.0], r.normal(0, 0.02, 3)])
        X0[i, 7:11] = q / np.linalg.norm(q)
    XT[:, 0] = X0[:, 0]
    XT[:, 7] = 1.0
    return X0, XT
