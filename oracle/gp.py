"""Oracle (test infrastructure): numpy restatement of the reference GP numerics.

Follows /root/reference/src/gp/{kernels,exact_gp,sparse_gp,fast_gp,features}.py.
Operation order matters for parity at 1e-9, so the Gram uses the same
expansion form ``s1 + s2^T - 2 X1s X2s^T`` as the reference.
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import cho_solve, solve_triangular

SE, M32, M52 = 0, 1, 2
KIND_NAMES = {SE: "se_ard", M32: "matern32", M52: "matern52"}


# --------------------------------------------------------------------------- kernels
def scaled_dist_sq(X1, X2, ls):
    """kernels.py:205-236 (_compute_scaled_distance_sq)."""
    X1s = X1 / ls
    X2s = X1s if X2 is None else X2 / ls
    s1 = np.sum(X1s**2, axis=1, keepdims=True)
    s2 = np.sum(X2s**2, axis=1, keepdims=True)
    d2 = s1 + s2.T - 2 * X1s @ X2s.T
    return np.maximum(d2, 0.0)


def gram(kind, X1, X2, ls, sv):
    """kernels.py:238-262 (SE-ARD), :532-545 (Matern32), :625-638 (Matern52)."""
    X1 = np.atleast_2d(X1)
    if X2 is not None:
        X2 = np.atleast_2d(X2)
    d2 = scaled_dist_sq(X1, X2, ls)
    if kind == SE:
        return sv * np.exp(-0.5 * d2)
    r = np.sqrt(d2)
    if kind == M32:
        s = np.sqrt(3) * r
        return sv * (1 + s) * np.exp(-s)
    if kind == M52:
        s = np.sqrt(5) * r
        return sv * (1 + s + 5 * r**2 / 3) * np.exp(-s)
    raise ValueError(kind)


def se_ard_gradients(X1, X2, ls, sv):
    """kernels.py:279-318: dK/dlog(sv)=K ; dK/dlog(l_i)=K*(x1_i-x2_i)^2/l_i^2."""
    X1 = np.atleast_2d(X1)
    X2 = X1 if X2 is None else np.atleast_2d(X2)
    K = gram(SE, X1, X2, ls, sv)
    out = {"log_signal_variance": K.copy()}
    for i in range(X1.shape[1]):
        diff = (X1[:, i : i + 1] - X2[:, i : i + 1].T) ** 2 / (ls[i] ** 2)
        out[f"log_lengthscale_{i}"] = K * diff
    return out


# --------------------------------------------------------------------------- exact GP
class ExactFit:
    """Result of exact_fit: mirrors ExactGP cached attributes (exact_gp.py:87-99)."""

    __slots__ = ("kind", "ls", "sv", "noise", "X", "y_n", "y_mean", "y_std", "L", "alpha", "lml", "jitter")


def exact_fit(kind, X, y, ls, sv, noise, normalize_y=True):
    """exact_gp.py:118-204 (fit + _compute_log_marginal_likelihood)."""
    X = np.atleast_2d(X)
    y = np.atleast_1d(y).flatten()
    n = X.shape[0]
    f = ExactFit()
    f.kind, f.ls, f.sv, f.noise, f.X = kind, np.asarray(ls, float), float(sv), float(noise), X
    if normalize_y:
        f.y_mean = np.mean(y)
        f.y_std = np.std(y)
        if f.y_std < 1e-10:
            f.y_std = 1.0
        f.y_n = (y - f.y_mean) / f.y_std
    else:
        f.y_mean, f.y_std, f.y_n = 0.0, 1.0, y
    K = gram(kind, X, None, f.ls, f.sv)
    Kn = K + noise * np.eye(n)
    f.jitter = 0.0
    try:
        f.L = np.linalg.cholesky(Kn)
    except np.linalg.LinAlgError:
        jitter = 1e-6
        while jitter < 1.0:
            try:
                f.L = np.linalg.cholesky(Kn + jitter * np.eye(n))
                f.jitter = jitter
                break
            except np.linalg.LinAlgError:
                jitter *= 10
        else:
            raise ValueError("Kernel matrix is not positive definite even with jitter")
    f.alpha = cho_solve((f.L, True), f.y_n)
    f.lml = -0.5 * np.dot(f.y_n, f.alpha) - np.sum(np.log(np.diag(f.L))) - 0.5 * n * np.log(2 * np.pi)
    return f


def exact_predict(f, Xq, return_cov=False):
    """exact_gp.py:213-268: returns (mean, var, std) or (mean, cov)."""
    Xq = np.atleast_2d(Xq)
    Ks = gram(f.kind, Xq, f.X, f.ls, f.sv)
    mean = Ks @ f.alpha
    mean = mean * f.y_std + f.y_mean
    v = solve_triangular(f.L, Ks.T, lower=True)
    if return_cov:
        Kss = gram(f.kind, Xq, None, f.ls, f.sv)
        return mean, (Kss - v.T @ v) * f.y_std**2
    var = np.full(Xq.shape[0], f.sv) - np.sum(v**2, axis=0)
    var = np.maximum(var, 1e-10) * f.y_std**2
    return mean, var, np.sqrt(var)


def multi_exact_fit(X, Y, noise, ls=None, sv=1.0, kind=SE):
    """exact_gp.py:476-499: P independent GPs on the same inputs (default SE-ARD, l=1)."""
    X = np.atleast_2d(X)
    Y = np.atleast_2d(Y)
    ls = np.ones(X.shape[1]) if ls is None else ls
    return [exact_fit(kind, X, Y[:, i], ls, sv, noise) for i in range(Y.shape[1])]


def multi_exact_predict(fits, Xq):
    """exact_gp.py:501-523 -> (means (Q,P), vars (Q,P))."""
    Xq = np.atleast_2d(Xq)
    m = np.zeros((Xq.shape[0], len(fits)))
    v = np.zeros_like(m)
    for i, f in enumerate(fits):
        m[:, i], v[:, i], _ = exact_predict(f, Xq)
    return m, v


def exact_lml_gradient(f):
    """Gradient of exact_fit's LML (exact_gp.py:186-204) w.r.t. [log sv, log l_0.., log noise], the parameter vector of
    optimize_hyperparameters (exact_gp.py:380-384), from the kernel derivatives of kernels.py:279-318:
    dLML/dtheta = 1/2 tr((alpha alpha' - K^{-1}) dK/dtheta).  The reference itself differences the LML numerically
    (scipy L-BFGS-B without jac); tests/ check this closed form against central differences of exact_fit."""
    n = f.X.shape[0]
    Kinv = np.linalg.solve(f.L.T, np.linalg.solve(f.L, np.eye(n)))
    W = np.outer(f.alpha, f.alpha) - Kinv
    G = se_ard_gradients(f.X, None, f.ls, f.sv)
    g = [0.5 * np.sum(W * G["log_signal_variance"])]
    for i in range(f.X.shape[1]):
        g.append(0.5 * np.sum(W * G[f"log_lengthscale_{i}"]))
    g.append(0.5 * f.noise * np.trace(W))
    return np.array(g)


# --------------------------------------------------------------------------- sparse GP
class SparseFit:
    __slots__ = ("kind", "ls", "sv", "noise", "method", "jitter", "Z", "X", "y_n", "y_mean", "y_std",
                 "L_uu", "L_B", "alpha", "lml", "Lambda")


def sparse_fit(kind, X, y, Z, ls, sv, noise, method="fitc", jitter=1e-6):
    """sparse_gp.py:150-253 with the inducing points Z injected."""
    X = np.atleast_2d(X)
    y = np.atleast_1d(y).flatten()
    f = SparseFit()
    f.kind, f.ls, f.sv, f.noise, f.method, f.jitter = kind, np.asarray(ls, float), float(sv), float(noise), method, jitter
    f.X, f.Z = X, np.atleast_2d(Z)
    N, M = X.shape[0], f.Z.shape[0]
    f.y_mean = np.mean(y)
    f.y_std = np.std(y)
    if f.y_std < 1e-10:
        f.y_std = 1.0
    f.y_n = (y - f.y_mean) / f.y_std
    K_uu = gram(kind, f.Z, None, f.ls, f.sv)
    K_uf = gram(kind, f.Z, X, f.ls, f.sv)
    f.L_uu = np.linalg.cholesky(K_uu + jitter * np.eye(M))
    A = solve_triangular(f.L_uu, K_uf, lower=True)
    if method == "fitc":
        Lam = np.full(N, f.sv) - np.sum(A**2, axis=0) + noise
        Lam = np.maximum(Lam, 1e-10)
        A_s = A * (1.0 / np.sqrt(Lam))
        B = np.eye(M) + A_s @ A_s.T
        f.L_B = np.linalg.cholesky(B)
        c = A @ (f.y_n / Lam)
        f.alpha = cho_solve((f.L_B, True), c)
        data_fit = -0.5 * (np.sum(f.y_n**2 / Lam) - np.dot(c, cho_solve((f.L_B, True), c)))
        complexity = -np.sum(np.log(np.diag(f.L_B))) - 0.5 * np.sum(np.log(Lam))
        f.lml = data_fit + complexity - 0.5 * N * np.log(2 * np.pi)
        f.Lambda = Lam
    else:
        s2 = noise
        B = K_uu + (1.0 / s2) * K_uf @ K_uf.T
        B = B + jitter * np.eye(M)
        f.L_B = np.linalg.cholesky(B)
        c = K_uf @ f.y_n / s2
        f.alpha = cho_solve((f.L_B, True), c)
        trace_term = (N * f.sv - np.sum(A**2)) / s2
        data_fit = (-0.5 / s2 * np.dot(f.y_n, f.y_n) + (1.0 / s2) * np.dot(f.y_n, K_uf.T @ f.alpha)
                    - 0.5 * np.dot(f.alpha, K_uu @ f.alpha))
        complexity = -np.sum(np.log(np.diag(f.L_B))) + np.sum(np.log(np.diag(f.L_uu))) - 0.5 * N * np.log(s2)
        f.lml = data_fit - 0.5 * trace_term + complexity - 0.5 * N * np.log(2 * np.pi)
        f.Lambda = None
    return f


def sparse_predict(f, Xq):
    """sparse_gp.py:255-305 (mean = K*u @ alpha, no L_uu^-T: reference quirk, replicated)."""
    Xq = np.atleast_2d(Xq)
    Ksu = gram(f.kind, Xq, f.Z, f.ls, f.sv)
    mean = Ksu @ f.alpha * f.y_std + f.y_mean
    v = solve_triangular(f.L_uu, Ksu.T, lower=True)
    w = solve_triangular(f.L_B, v, lower=True)
    var = np.full(Xq.shape[0], f.sv) - np.sum(v**2, axis=0) + np.sum(w**2, axis=0)
    var = np.maximum(var, 1e-10) * f.y_std**2
    return mean, var, np.sqrt(var)


# --------------------------------------------------------------------------- fast GP
def fast_fit(X, Y, ls, variance, noise, jitter=1e-6):
    """fast_gp.py:113-169: no y-normalisation, K += (noise+jitter) I, alpha by two trsv."""
    X = np.ascontiguousarray(X)
    Y = np.ascontiguousarray(Y.reshape(len(X), -1))
    n = len(X)
    K = gram(SE, X, X, ls, variance)
    K = K + (noise + jitter) * np.eye(n)
    try:
        L = np.linalg.cholesky(K)
    except np.linalg.LinAlgError:
        K = K + 1e-4 * np.eye(n)
        L = np.linalg.cholesky(K)
    alpha = np.zeros_like(Y)
    for i in range(Y.shape[1]):
        v = solve_triangular(L, Y[:, i], lower=True)
        alpha[:, i] = solve_triangular(L.T, v, lower=False)
    return L, alpha


def fast_predict(X, L, alpha, ls, variance, Xq):
    """fast_gp.py:171-213: var = max(variance - sum v^2, 0) broadcast to all outputs."""
    Xq = np.atleast_2d(Xq)
    Ks = gram(SE, Xq, X, ls, variance)
    mean = Ks @ alpha
    v = solve_triangular(L, Ks.T, lower=True)
    vd = np.maximum(variance - np.sum(v**2, axis=0), 0)
    return mean, np.repeat(vd[:, None], alpha.shape[1], axis=1)


# --------------------------------------------------------------------------- features
RHO0, HSCALE, VREF = 1.225, 8500.0, 10.0


def quat_to_dcm(q):
    """features.py:260-270 ([w,x,y,z] -> C_BI)."""
    w, x, y, z = q
    return np.array([
        [1 - 2 * (y**2 + z**2), 2 * (x * y + w * z), 2 * (x * z - w * y)],
        [2 * (x * y - w * z), 1 - 2 * (x**2 + z**2), 2 * (y * z + w * x)],
        [2 * (x * z + w * y), 2 * (y * z - w * x), 1 - 2 * (x**2 + y**2)],
    ])


def features_3dof(x, u):
    """features.py:403-444 (Simple3DoFFeatureExtractor.extract), 11 features."""
    v = x[4:7]
    alt = x[1]
    speed = np.linalg.norm(v)
    rho = RHO0 * np.exp(-alt / HSCALE)
    q_dyn = 0.5 * rho * speed**2
    T = np.linalg.norm(u)
    return np.array([v[0] / VREF, v[1] / VREF, v[2] / VREF, speed / VREF, q_dyn / (0.5 * RHO0 * VREF**2),
                     u[0] / 10.0, u[1] / 10.0, u[2] / 10.0, T / 10.0, alt / 100.0, rho / RHO0])


def features_6dof_trans(x, u):
    """features.py:194-258 (TranslationalFeatureExtractor.extract), 13 features."""
    r, v, q = x[1:4], x[4:7], x[7:11]
    alt = r[0]
    speed = np.linalg.norm(v)
    rho = RHO0 * np.exp(-alt / HSCALE)
    q_dyn = 0.5 * rho * speed**2
    vB = quat_to_dcm(q) @ v
    if speed > 1e-3:
        alpha = np.arctan2(-vB[2], vB[0])
        beta = np.arcsin(np.clip(vB[1] / speed, -1, 1))
    else:
        alpha = beta = 0.0
    T = np.linalg.norm(u)
    return np.array([v[0] / VREF, v[1] / VREF, v[2] / VREF, speed / VREF, q_dyn / (0.5 * RHO0 * VREF**2),
                     alpha, beta, u[0] / 10.0, u[1] / 10.0, u[2] / 10.0, T / 10.0, alt / 100.0, rho / RHO0])


def features_6dof_rot(x, u):
    """features.py:305-354 (RotationalFeatureExtractor.extract), 12 features."""
    r, v, q, om = x[1:4], x[4:7], x[7:11], x[11:14]
    alt = r[0]
    speed = np.linalg.norm(v)
    om_mag = np.linalg.norm(om)
    rho = RHO0 * np.exp(-alt / HSCALE)
    q_dyn = 0.5 * rho * speed**2
    vB = quat_to_dcm(q) @ v
    return np.array([om[0] / 1.0, om[1] / 1.0, om[2] / 1.0, om_mag / 1.0, u[0] / 10.0, u[1] / 10.0, u[2] / 10.0,
                     vB[0] / VREF, vB[1] / VREF, vB[2] / VREF, speed / VREF, q_dyn / (0.5 * RHO0 * VREF**2)])


def features_batch(fn, X, U):
    """features.py:129-146 (extract_batch: python loop of extract)."""
    return np.stack([fn(X[i], U[i]) for i in range(X.shape[0])])


# --------------------------------------------------------------------------- wrappers
class Simple3DoFGPOracle:
    """structured_gp.py:414-496 with use_sparse=False (exact) or injected Z (sparse)."""

    def __init__(self, noise_variance=1e-4, use_sparse=False, Z=None):
        self.noise, self.use_sparse, self.Z = noise_variance, use_sparse, Z
        self.X, self.U, self.D = [], [], []
        self.fits = None

    def add_data(self, X, U, D):
        X, U, D = np.atleast_2d(X), np.atleast_2d(U), np.atleast_2d(D)
        for i in range(X.shape[0]):
            self.X.append(X[i]); self.U.append(U[i]); self.D.append(D[i])
        self.fits = None

    def fit(self):
        if not self.X:
            raise RuntimeError("No data")
        Zf = features_batch(features_3dof, np.array(self.X), np.array(self.U))
        D = np.array(self.D)
        ls = np.ones(11)
        if self.use_sparse:
            self.fits = [sparse_fit(SE, Zf, D[:, i], self.Z, ls, 1.0, self.noise) for i in range(3)]
        else:
            self.fits = multi_exact_fit(Zf, D, self.noise)

    def predict(self, x, u):
        if self.fits is None:
            if self.X:
                self.fit()
            else:
                return np.zeros(3), np.ones(3) * 0.1
        z = features_3dof(x, u).reshape(1, -1)
        return self.predict_features(z)[0][0], self.predict_features(z)[1][0]

    def predict_features(self, Zq):
        if self.use_sparse:
            m = np.zeros((Zq.shape[0], 3)); v = np.zeros_like(m)
            for i, f in enumerate(self.fits):
                m[:, i], v[:, i], _ = sparse_predict(f, Zq)
            return m, v
        return multi_exact_predict(self.fits, Zq)

    def predict_batch(self, X, U):
        return self.predict_features(features_batch(features_3dof, X, U))


# --------------------------------------------------------------------------- arbiter
def longdouble_exact_predict(f, Xq):
    """Extended-precision arbiter: same formulas, np.longdouble throughout,
    including the Gram, the Cholesky and the solves.  O(n^3) python loop -- keep n <= ~300."""
    ld = np.longdouble
    X, ls = f.X.astype(ld), f.ls.astype(ld)

    def g(A, B):
        As, Bs = A / ls, B / ls
        d2 = np.sum(As**2, 1, keepdims=True) + np.sum(Bs**2, 1, keepdims=True).T - 2 * As @ Bs.T
        d2 = np.maximum(d2, ld(0))
        if f.kind == SE:
            return ld(f.sv) * np.exp(-ld(0.5) * d2)
        r = np.sqrt(d2)
        if f.kind == M32:
            s = np.sqrt(ld(3)) * r
            return ld(f.sv) * (1 + s) * np.exp(-s)
        s = np.sqrt(ld(5)) * r
        return ld(f.sv) * (1 + s + 5 * r**2 / 3) * np.exp(-s)

    n = X.shape[0]
    K = g(X, X) + ld(f.noise + f.jitter) * np.eye(n, dtype=ld)
    L = np.zeros((n, n), dtype=ld)
    for j in range(n):
        L[j, j] = np.sqrt(K[j, j] - L[j, :j] @ L[j, :j])
        L[j + 1:, j] = (K[j + 1:, j] - L[j + 1:, :j] @ L[j, :j]) / L[j, j]
    y = f.y_n.astype(ld)
    t = np.zeros(n, dtype=ld)
    for i in range(n):
        t[i] = (y[i] - L[i, :i] @ t[:i]) / L[i, i]
    a = np.zeros(n, dtype=ld)
    for i in range(n - 1, -1, -1):
        a[i] = (t[i] - L[i + 1:, i] @ a[i + 1:]) / L[i, i]
    Ks = g(np.atleast_2d(Xq).astype(ld), X)
    mean = (Ks @ a) * ld(f.y_std) + ld(f.y_mean)
    V = np.zeros((n, Ks.shape[0]), dtype=ld)
    for i in range(n):
        V[i] = (Ks[:, i] - L[i, :i] @ V[:i]) / L[i, i]
    var = np.maximum(ld(f.sv) - np.sum(V**2, 0), ld(1e-10)) * ld(f.y_std) ** 2
    return mean, var
