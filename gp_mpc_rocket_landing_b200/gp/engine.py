"""Device-side GP engine: thin Python over the C-ABI (include/gpmpc_b200.h).

Everything numeric happens in libgpmpc_b200.so; torch only owns the HBM buffers.  Layout conventions
(DESIGN.md "Data layout"): FP64 row-major; training/query counts padded to multiples of 128 rows, feature
dimension padded to a multiple of 4 with zeros; the Cholesky factor carries a unit diagonal in the padding.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _lib
from .._lib import call, ptr, stream

SE, M32, M52, DIST2 = 0, 1, 2, 3


def torch_():
    return _lib.require_cuda()


def pad_to(n: int, m: int) -> int:
    return ((max(n, 1) + m - 1) // m) * m


def to_dev(a, dtype=None):
    """numpy / sequence -> contiguous CUDA tensor (float64 by default); CUDA tensors pass through."""
    torch = torch_()
    if isinstance(a, torch.Tensor):
        t = a if a.is_cuda else a.cuda()
        if dtype is not None and t.dtype != dtype:
            t = t.to(dtype)
        return t.contiguous()
    arr = np.ascontiguousarray(np.asarray(a, dtype=np.float64 if dtype is None else None))
    t = torch.from_numpy(arr).cuda()
    if dtype is not None:
        t = t.to(dtype)
    return t


def scale_inputs(X, ls, n_pad=None, DP=None):
    """kernels.py:223-229: Xs = X / l (zero padded) and numpy-order row sums of squares."""
    torch = torch_()
    n, D = X.shape
    n_pad = pad_to(n, 128) if n_pad is None else n_pad
    DP = pad_to(D, 4) if DP is None else DP
    Xs = torch.empty((n_pad, DP), dtype=torch.float64, device=X.device)
    sq = torch.empty((n_pad,), dtype=torch.float64, device=X.device)
    call("gpmpc_scale_inputs", ptr(X), n, D, X.stride(0), ptr(ls), ptr(Xs), n_pad, DP, ptr(sq), stream())
    return Xs, sq


def gram_scaled(kind, Xs1, sq1, n1, Xs2, sq2, n2, sv, diag_add=0.0, pad_identity=0, lower_only=0, out=None):
    torch = torch_()
    n1_pad, DP = Xs1.shape
    n2_pad = Xs2.shape[0]
    K = out if out is not None else torch.empty((n1_pad, n2_pad), dtype=torch.float64, device=Xs1.device)
    call("gpmpc_gram", int(kind), ptr(Xs1), ptr(sq1), n1, n1_pad, ptr(Xs2), ptr(sq2), n2, n2_pad, DP, float(sv),
         float(diag_add), int(pad_identity), int(lower_only), ptr(K), K.stride(0), stream())
    return K


def gram(kind, X1, X2, ls, sv):
    """Kernel.__call__ (kernels.py:238-262, :532-545, :625-638) on the device -> (N1, N2) CUDA tensor."""
    X1 = to_dev(np.atleast_2d(X1) if not hasattr(X1, "is_cuda") else X1)
    lsd = to_dev(ls)
    Xs1, sq1 = scale_inputs(X1, lsd)
    if X2 is None:
        Xs2, sq2, n2 = Xs1, sq1, X1.shape[0]
    else:
        X2 = to_dev(np.atleast_2d(X2) if not hasattr(X2, "is_cuda") else X2)
        Xs2, sq2 = scale_inputs(X2, lsd)
        n2 = X2.shape[0]
    K = gram_scaled(kind, Xs1, sq1, X1.shape[0], Xs2, sq2, n2, sv)
    return K[: X1.shape[0], :n2]


def gram_grad_lengthscale(X1, X2, ls, sv, dim):
    """SE-ARD dK/dlog(l_dim), kernels.py:310-316."""
    torch = torch_()
    X1 = to_dev(np.atleast_2d(X1))
    lsd = to_dev(ls)
    Xs1, sq1 = scale_inputs(X1, lsd)
    if X2 is None:
        Xs2, sq2, n2 = Xs1, sq1, X1.shape[0]
    else:
        X2 = to_dev(np.atleast_2d(X2))
        Xs2, sq2 = scale_inputs(X2, lsd)
        n2 = X2.shape[0]
    G = torch.empty((Xs1.shape[0], Xs2.shape[0]), dtype=torch.float64, device=X1.device)
    call("gpmpc_gram_grad_lengthscale", ptr(Xs1), ptr(sq1), X1.shape[0], Xs1.shape[0], ptr(Xs2), ptr(sq2), n2,
         Xs2.shape[0], Xs1.shape[1], float(sv), int(dim), ptr(G), G.stride(0), stream())
    return G[: X1.shape[0], :n2]


def features(which, X, U):
    """features.py extract_batch on the device: (B,nx),(B,3) -> (B, 11|13|12)."""
    torch = torch_()
    X, U = to_dev(X), to_dev(U)
    nf = {0: 11, 1: 13, 2: 12}[which]
    Z = torch.empty((X.shape[0], nf), dtype=torch.float64, device=X.device)
    call("gpmpc_features", which, ptr(X), X.stride(0), ptr(U), U.stride(0), X.shape[0], ptr(Z), Z.stride(0), stream())
    return Z


def potrf(A, batch=1):
    """In-place batched Cholesky of (batch, n_pad, n_pad) (or (n_pad, n_pad)); returns (Winv, info)."""
    torch = torch_()
    n_pad = A.shape[-1]
    Winv = torch.empty((batch, n_pad // 64, 64, 64), dtype=torch.float64, device=A.device)
    info = torch.empty((batch,), dtype=torch.int32, device=A.device)
    call("gpmpc_potrf", ptr(A), n_pad, A.stride(-2), n_pad * A.stride(-2), batch, ptr(Winv), ptr(info), stream())
    return Winv, info


class ExactFactor:
    """K = k(X,X) + (noise+jitter) I = L L^T for one hyper-parameter set (exact_gp.py:157-175).

    kind in (SE, M32, M52): the Gram is built by the fused device kernel from (ls, sv).  kind None: `kernel` is any
    object with the reference Kernel API (Sum / Product / isotropic ...) evaluated on device tensors."""

    JITTERS = (0.0, 1e-6, 1e-5, 1e-4, 1e-3, 1e-2, 1e-1)  # exact_gp.py:163-175 (1e-6 * 10^j while < 1.0)

    def __init__(self, kind, ls, sv, noise, kernel=None):
        self.kind = None if kind is None else int(kind)
        self.sv, self.noise = float(sv), float(noise)
        self.ls_host = np.asarray(ls, dtype=np.float64).copy()
        self.kernel = kernel
        self.jitter = 0.0

    def _gram_noise(self, jit, out):
        if self.kind is not None:
            gram_scaled(self.kind, self.Xs, self.sq, self.n, self.Xs, self.sq, self.n, self.sv,
                        diag_add=self.noise + jit, pad_identity=1, lower_only=1, out=out)
            return out
        torch = torch_()
        out.zero_()
        out.diagonal().fill_(1.0)
        out[: self.n, : self.n] = self.kernel._gram_dev(self.X)
        out.diagonal()[: self.n].add_(self.noise + jit)
        return out

    def fit(self, X, jitters=None):
        torch = torch_()
        self.X = to_dev(X)
        self.n, self.D = self.X.shape
        self.n_pad = pad_to(self.n, 128)
        if self.kind is not None:
            self.ls = to_dev(self.ls_host)
            self.Xs, self.sq = scale_inputs(self.X, self.ls)
            self.DP = self.Xs.shape[1]
        L = torch.empty((self.n_pad, self.n_pad), dtype=torch.float64, device=self.X.device)
        for jit in (self.JITTERS if jitters is None else jitters):
            self._gram_noise(jit, L)
            self.Winv, info = potrf(L)
            if int(info.item()) == 0:
                self.L, self.jitter = L, jit
                self.Linv = None
                return self
        raise ValueError("Kernel matrix is not positive definite even with jitter")

    def append(self, Xnew):
        """Extend the factor by the rows of the new points instead of re-factorising (the reference refits from scratch
        for every batch of new observations: online_update.py:361-408, sparse_gp.py:328-353).  With K' = [[K, k], [k', c]]:
        L' = [[L, 0], [v', l]], v = L^{-1} k (the blocked DMMA TRSM of `predict`), l l' = c - v'v (a <= 128-point
        Cholesky).  O(n^2 m) instead of O(n^3 / 3); the existing rows of L are untouched, so the result equals the
        factor of a full refit up to the rounding of the last rows.  Raises ValueError when the Schur complement is not
        positive definite (the caller then refits with the jitter escalation of exact_gp.py:163-175)."""
        torch = torch_()
        if self.kind is None:
            raise ValueError("ExactFactor.append: needs one of the fused ARD kernels")
        Xnew = to_dev(Xnew)
        if Xnew.dim() == 1:
            Xnew = Xnew.reshape(1, -1)
        for s0 in range(0, Xnew.shape[0], 128):
            self._append_chunk(Xnew[s0:s0 + 128].contiguous())
        return self

    def _append_chunk(self, Xnew):
        torch = torch_()
        m, n = Xnew.shape[0], self.n
        n2 = n + m
        dev = self.L.device
        if n2 > self.n_pad:   # grow the padded buffers by whole 128-blocks (unit diagonal in the padding)
            np2 = pad_to(n2, 128)
            L2 = torch.eye(np2, dtype=torch.float64, device=dev)
            L2[: self.n_pad, : self.n_pad] = self.L
            W2 = torch.eye(64, dtype=torch.float64, device=dev).repeat(1, np2 // 64, 1, 1)
            W2[:, : self.n_pad // 64] = self.Winv
            self.L, self.Winv, self.n_pad = L2, W2.contiguous(), np2
        Xall = torch.cat([self.X, Xnew], dim=0)
        Xs_all, sq_all = scale_inputs(Xall, self.ls, n_pad=self.n_pad, DP=self.DP)
        Xs_new, sq_new = scale_inputs(Xnew, self.ls, DP=self.DP)
        Kx = gram_scaled(self.kind, Xs_new, sq_new, m, Xs_all, sq_all, n2, self.sv)        # (128, n_pad)
        V = Kx.clone()
        V[:, n:] = 0.0
        call("gpmpc_trsm_right", ptr(self.L), self.n_pad, self.L.stride(0), ptr(self.Winv), ptr(V), 128, V.stride(0),
             stream())                                                                    # rows: v' = (L^{-1} k)'
        VVt = torch.empty((128, 128), dtype=torch.float64, device=dev)
        call("gpmpc_dgemm_nt", ptr(V), V.stride(0), ptr(V), V.stride(0), ptr(VVt), 128, 128, 128, self.n_pad, 1.0, 0.0, 0,
             stream())
        S = torch.eye(128, dtype=torch.float64, device=dev)
        S[:m, :m] = Kx[:m, n:n2] - VVt[:m, :m]
        S.diagonal()[:m].add_(self.noise + self.jitter)
        _, info = potrf(S)
        if int(info.item()) != 0:
            raise ValueError("ExactFactor.append: Schur complement of the new points is not positive definite")
        self.L[n:n2, :n] = V[:m, :n]
        self.L[n:n2, n:n2] = S[:m, :m]
        call("gpmpc_trinv64_blocks", ptr(self.L), self.n_pad, self.L.stride(0), ptr(self.Winv), n // 64, (n2 - 1) // 64,
             stream())
        self.X, self.Xs, self.sq, self.n = Xall, Xs_all, sq_all, n2
        self.Linv = None

    def solve(self, Bm):
        """cho_solve for the rows of Bm (nrhs, n_pad), in place."""
        call("gpmpc_potrs", ptr(self.L), self.n_pad, self.L.stride(0), 0, ptr(self.Winv), ptr(Bm), Bm.shape[0], 1,
             stream())
        return Bm

    def cross(self, Xq, out=None):
        """K* rows for queries Xq (Q,D): returns (Ks (q_pad, n_pad), Q)."""
        Q = Xq.shape[0]
        if self.kind is not None:
            Xqs, sqq = scale_inputs(Xq, self.ls, DP=self.DP)
            return gram_scaled(self.kind, Xqs, sqq, Q, self.Xs, self.sq, self.n, self.sv, out=out), Q
        torch = torch_()
        Ks = torch.zeros((pad_to(Q, 128), self.n_pad), dtype=torch.float64, device=Xq.device) if out is None else out.zero_()
        Ks[:Q, : self.n] = self.kernel._gram_dev(Xq, self.X)
        return Ks, Q

    def prior_diag(self, Xq):
        """k(x*, x*) per query (kernels.py:264-277): a constant for the ARD kernels."""
        if self.kind is not None:
            return None
        return to_dev(self.kernel.diagonal(Xq))

    def sumsq_inplace(self, Ks):
        """v = L^{-1} K*^T in place (row-wise) and sum_i v_i^2 per query (exact_gp.py:260-261)."""
        torch = torch_()
        sumsq = torch.empty((Ks.shape[0],), dtype=torch.float64, device=Ks.device)
        call("gpmpc_trsm_sumsq", ptr(self.L), self.n_pad, self.L.stride(0), ptr(self.Winv), ptr(Ks), Ks.shape[0],
             Ks.stride(0), ptr(sumsq), stream())
        return sumsq


    def ensure_linv(self):
        """W = L^{-1} (lower triangular), formed once per fit for `sumsq_linv`."""
        if getattr(self, "Linv", None) is None:
            torch = torch_()
            self.Linv = torch.empty_like(self.L)
            call("gpmpc_tri_inverse", ptr(self.L), self.n_pad, self.L.stride(0), ptr(self.Winv), ptr(self.Linv),
                 self.Linv.stride(0), stream())
        return self.Linv

    def sumsq_linv(self, Ks, work=None, one_cta_per_sm=False):
        """sum_i v_i^2 with v = L^{-1} K*^T evaluated as ONE triangular GEMM against the explicit inverse factor (Ks is
        left untouched).  Same quantity as `sumsq_inplace` (exact_gp.py:260-261); the rounding differs by
        O(cond(L) eps) -- measured 1e-11 / 7e-10 relative on the two 1000-point fixtures -- so the reference-faithful
        blocked solve stays the default of the `ExactGP.predict` API and this path serves the horizon pipeline."""
        torch = torch_()
        W = self.ensure_linv()
        q_pad = Ks.shape[0]
        if work is None or work.numel() < (self.n_pad // 64) * q_pad:
            work = torch.empty(((self.n_pad // 64) * q_pad,), dtype=torch.float64, device=Ks.device)
        sumsq = torch.empty((q_pad,), dtype=torch.float64, device=Ks.device)
        call("gpmpc_linv_sumsq", ptr(W), self.n_pad, W.stride(0), ptr(Ks), q_pad, Ks.stride(0), ptr(work), ptr(sumsq),
             1 if one_cta_per_sm else 0, stream())
        return sumsq


class ExactModelDev:
    """P outputs on one ExactFactor: alpha, y normalisation, LML (exact_gp.py:118-204) and predict (:213-268)."""

    def __init__(self, factor: ExactFactor):
        self.f = factor

    def fit_targets(self, Y, normalize=True, prior_mean=None):
        """prior_mean: optional (n, P) values m(X) subtracted as m(X)/y_std (exact_gp.py:152-154)."""
        torch = torch_()
        f = self.f
        Y = to_dev(Y)
        if Y.dim() == 1:
            Y = Y[:, None].contiguous()
        self.P = Y.shape[1]
        dev = Y.device
        self.Yn = torch.empty((self.P, f.n_pad), dtype=torch.float64, device=dev)
        self.ymean = torch.empty((self.P,), dtype=torch.float64, device=dev)
        self.ystd = torch.empty((self.P,), dtype=torch.float64, device=dev)
        call("gpmpc_normalize_targets", ptr(Y), f.n, self.P, Y.stride(0), f.n_pad, int(normalize), ptr(self.Yn),
             ptr(self.ymean), ptr(self.ystd), stream())
        if prior_mean is not None:
            pm = to_dev(prior_mean)
            if pm.dim() == 1:
                pm = pm[:, None]
            self.Yn[:, : f.n] -= (pm / self.ystd[None, :]).T
        self.alpha = f.solve(self.Yn.clone())
        self.lml = torch.empty((self.P,), dtype=torch.float64, device=dev)
        call("gpmpc_lml", ptr(f.L), f.n, f.n_pad, f.L.stride(0), 0, ptr(self.Yn), ptr(self.alpha), self.P, 1,
             ptr(self.lml), stream())
        return self

    def lml_gradient(self):
        """d LML_p / d[log s2, log l_1..l_D, log noise] -> (P, D + 2) CUDA tensor, for the SE-ARD kernel.
        K^{-1} = L^{-T} L^{-1} comes from the explicit inverse factor (one triangular GEMM); the D + 2 contractions with
        W = alpha alpha' - K^{-1} run as one fused pass (gpmpc_lml_grad_se).  The LML is the one of the normalised
        targets (exact_gp.py:186-204), so this is the gradient of `self.lml`."""
        torch = torch_()
        f = self.f
        if f.kind != SE:
            raise ValueError("lml_gradient: analytic gradients are implemented for the SE-ARD kernel")
        if f.DP > 32:
            raise ValueError("lml_gradient: input dimension > 32")
        dev = f.L.device
        # K^{-1} = W'W with the explicit inverse factor W = L^{-1}: a right-looking DMMA TRSM of the identity (gives W'
        # row-major, the operand layout of the NT GEMM) and one lower-triangular GEMM, instead of n cho_solves
        Wt = torch.eye(f.n_pad, dtype=torch.float64, device=dev)   # rows solved in place: row q <- column q of L^{-1}
        call("gpmpc_trsm_right", ptr(f.L), f.n_pad, f.L.stride(0), ptr(f.Winv), ptr(Wt), f.n_pad, Wt.stride(0), stream())
        Kinv = torch.empty((f.n_pad, f.n_pad), dtype=torch.float64, device=dev)
        call("gpmpc_dgemm_nt", ptr(Wt), Wt.stride(0), ptr(Wt), Wt.stride(0), ptr(Kinv), Kinv.stride(0), f.n_pad, f.n_pad,
             f.n_pad, 1.0, 0.0, 1, stream())
        nwork = int(_lib.lib.gpmpc_lml_grad_se_work_doubles(f.n_pad, f.DP, self.P))
        work = torch.empty((nwork,), dtype=torch.float64, device=dev)
        grad = torch.empty((self.P, f.D + 2), dtype=torch.float64, device=dev)
        call("gpmpc_lml_grad_se", ptr(f.Xs), f.n, f.n_pad, f.DP, f.D, f.sv, f.noise, ptr(Kinv), Kinv.stride(0),
             ptr(self.alpha), self.alpha.stride(0), self.P, ptr(work), ptr(grad), stream())
        return grad

    def predict(self, Xq, want_var=True, chunk=65536, col=None, keep_v=False, var_floor=1e-10):
        """-> mean (Q,P), var (Q,P) CUDA tensors (col: a single output -> (Q,1)).  keep_v: also return the solved
        V = (L^{-1} K*^T)^T rows (Q_pad, n_pad) of the last chunk (full-covariance branch, exact_gp.py:247-254)."""
        torch = torch_()
        f = self.f
        Xq = to_dev(Xq)
        Q = Xq.shape[0]
        c0, P = (0, self.P) if col is None else (int(col), 1)
        alpha, ymean, ystd = self.alpha[c0:c0 + P], self.ymean[c0:c0 + P], self.ystd[c0:c0 + P]
        mean = torch.empty((Q, P), dtype=torch.float64, device=Xq.device)
        var = torch.empty((Q, P), dtype=torch.float64, device=Xq.device) if want_var else None
        V = None
        for s in range(0, Q, chunk):
            e = min(Q, s + chunk)
            Ks, q = f.cross(Xq[s:e])
            call("gpmpc_predict_mean", ptr(Ks), Ks.stride(0), q, f.n, ptr(alpha), f.n_pad, P,
                 ptr(ymean), ptr(ystd), ptr(mean[s:e]), mean.stride(0), 0, stream())
            if want_var or keep_v:
                sumsq = f.sumsq_inplace(Ks)
                V = Ks
            if want_var:
                prior = f.prior_diag(Xq[s:e])
                if prior is None:
                    call("gpmpc_predict_var", ptr(sumsq), None, q, P, f.sv, var_floor, ptr(ystd), ptr(var[s:e]),
                         var.stride(0), 0, stream())
                else:  # generic kernels: per-query prior variance
                    var[s:e] = torch.clamp(prior - sumsq[:q], min=var_floor)[:, None] * (ystd * ystd)[None, :]
        if keep_v:
            return mean, var, V
        return mean, var

    def struct(self):
        """GpmpcExactModel for the fused horizon kernels (keeps the tensors alive through self)."""
        f = self.f
        if f.kind is None:
            raise _lib.GpmpcError("the fused horizon kernels need an SE-ARD / Matern kernel (one fused Gram launch)")
        return _lib.ExactModel(kind=f.kind, n=f.n, n_pad=f.n_pad, D=f.D, DP=f.DP, P=self.P, signal_var=f.sv,
                               lengthscales=ptr(f.ls), Xs=ptr(f.Xs), sq=ptr(f.sq), alpha=ptr(self.alpha),
                               ymean=ptr(self.ymean), ystd=ptr(self.ystd), L=ptr(f.L), Winv=ptr(f.Winv))


# ====================================================================================================== sparse GP
class SparseFactor:
    """Inducing-point factor shared by every output with the same kernel, noise and Z (sparse_gp.py:150-232).

    FITC and VFE both need (L_uu, L_B); neither depends on the targets, so P outputs cost one factorisation (the
    reference refits them per output, sparse_gp.py:448-454).  All matrices are M_pad x M_pad with a unit diagonal in the
    padding (M_pad = 128 for M <= 128), so the blocked DMMA TRSM of the exact GP serves the predict path unchanged."""

    def __init__(self, kind, ls, sv, noise, method="fitc", jitter=1e-6):
        self.kind = int(kind)
        self.ls_host = np.asarray(ls, dtype=np.float64).copy()
        self.sv, self.noise, self.jitter = float(sv), float(noise), float(jitter)
        assert method in ("fitc", "vfe")
        self.method = method

    def fit(self, X, Z):
        torch = torch_()
        self.X, self.Z = to_dev(X), to_dev(Z)
        self.n, self.D = self.X.shape
        self.M = self.Z.shape[0]
        if self.M > 144:
            raise _lib.GpmpcError("SparseFactor: at most 144 inducing points (gpmpc_trsm_small keeps L_uu in shared memory)")
        self.n_pad, self.M_pad = pad_to(self.n, 128), pad_to(self.M, 128)
        dev = self.X.device
        self.ls = to_dev(self.ls_host)
        self.Xs, self.sq = scale_inputs(self.X, self.ls)
        self.DP = self.Xs.shape[1]
        self.Zs, self.sqz = scale_inputs(self.Z, self.ls, DP=self.DP)
        M, n, Mp, npad = self.M, self.n, self.M_pad, self.n_pad
        # K_uu (kept without jitter for VFE's B and the bound), L_uu = chol(K_uu + jitter I)
        self.Kuu = gram_scaled(self.kind, self.Zs, self.sqz, M, self.Zs, self.sqz, M, self.sv, pad_identity=1)
        self.L_uu = self.Kuu.clone()
        call("gpmpc_add_diag", ptr(self.L_uu), M, Mp, self.jitter, stream())
        self.W_uu, info = potrf(self.L_uu)
        if int(info.item()) != 0:
            raise np.linalg.LinAlgError("Matrix is not positive definite")   # np.linalg.cholesky at sparse_gp.py:187
        # K_uf (M_pad x n_pad) and A = L_uu^{-1} K_uf with q_ff = column sums of A^2 (sparse_gp.py:190,196)
        self.Kuf = gram_scaled(self.kind, self.Zs, self.sqz, M, self.Xs, self.sq, n, self.sv)
        self.A = self.Kuf.clone()
        self.qff = torch.zeros((npad,), dtype=torch.float64, device=dev)
        call("gpmpc_trsm_small", ptr(self.L_uu), M, Mp, ptr(self.A), npad, n, 0, ptr(self.qff), 1.0, stream())
        return self

    def factor_B(self, Yn):
        """(L_B, c) for the normalised targets Yn (P, n_pad): sparse_gp.py:198-210 (FITC) / :225-232 (VFE)."""
        torch = torch_()
        M, n, Mp, npad = self.M, self.n, self.M_pad, self.n_pad
        P = Yn.shape[0]
        dev = Yn.device
        self.Lam = None
        if self.method == "fitc":
            self.Lam = torch.empty((npad,), dtype=torch.float64, device=dev)
            As = torch.empty_like(self.A)
            Ylam = torch.empty_like(Yn)
            call("gpmpc_fitc_scale", ptr(self.A), npad, Mp, n, npad, ptr(self.qff), self.sv, self.noise, 1e-10, ptr(Yn), P,
                 ptr(self.Lam), ptr(As), npad, ptr(Ylam), stream())
            Bm = torch.eye(Mp, dtype=torch.float64, device=dev)
            call("gpmpc_dgemm_nt", ptr(As), npad, ptr(As), npad, ptr(Bm), Mp, Mp, Mp, npad, 1.0, 1.0, 0, stream())
            rows, rhs, scale = self.A, Ylam, None
        else:
            Bm = self.Kuu.clone()
            call("gpmpc_dgemm_nt", ptr(self.Kuf), npad, ptr(self.Kuf), npad, ptr(Bm), Mp, Mp, Mp, npad, 1.0 / self.noise,
                 1.0, 0, stream())
            call("gpmpc_add_diag", ptr(Bm), M, Mp, self.jitter, stream())
            rows, rhs, scale = self.Kuf, Yn, self.noise
        self.W_B, info = potrf(Bm)
        if int(info.item()) != 0:
            raise np.linalg.LinAlgError("Matrix is not positive definite")
        self.L_B = Bm
        # c[p, m] = sum_i rows[m, i] rhs[p, i]  (A @ (y / Lambda) or K_uf @ y / s2)
        cT = torch.zeros((Mp, P), dtype=torch.float64, device=dev)
        zeros, ones = torch.zeros((P,), dtype=torch.float64, device=dev), torch.ones((P,), dtype=torch.float64, device=dev)
        call("gpmpc_predict_mean", ptr(rows), npad, M, n, ptr(rhs), npad, P, ptr(zeros), ptr(ones), ptr(cT), P, 0, stream())
        c = cT.T.contiguous()
        if scale is not None:
            c = c / scale
        return c

    def solve_B(self, c):
        out = c.clone()
        call("gpmpc_potrs", ptr(self.L_B), self.M_pad, self.M_pad, 0, ptr(self.W_B), ptr(out), out.shape[0], 1, stream())
        return out


class SparseModelDev:
    """P outputs on one SparseFactor: alpha, normalisation, LML / ELBO, predict (sparse_gp.py:150-305)."""

    def __init__(self, factor: SparseFactor):
        self.f = factor

    def fit_targets(self, Y):
        torch = torch_()
        f = self.f
        Y = to_dev(Y)
        if Y.dim() == 1:
            Y = Y[:, None].contiguous()
        self.P = Y.shape[1]
        dev = Y.device
        self.Yn = torch.empty((self.P, f.n_pad), dtype=torch.float64, device=dev)
        self.ymean = torch.empty((self.P,), dtype=torch.float64, device=dev)
        self.ystd = torch.empty((self.P,), dtype=torch.float64, device=dev)
        call("gpmpc_normalize_targets", ptr(Y), f.n, self.P, Y.stride(0), f.n_pad, 1, ptr(self.Yn), ptr(self.ymean),
             ptr(self.ystd), stream())
        self.c = f.factor_B(self.Yn)
        self.alpha = f.solve_B(self.c)
        self.lml = torch.empty((self.P,), dtype=torch.float64, device=dev)
        call("gpmpc_sparse_lml", 0 if f.method == "fitc" else 1, f.M, f.n, f.n_pad, f.M_pad, ptr(f.Kuf), ptr(f.Kuu),
             ptr(f.L_B), ptr(f.L_uu), ptr(self.Yn), ptr(self.alpha), ptr(self.c), ptr(f.Lam) if f.Lam is not None else None,
             ptr(f.qff), f.sv, f.noise, self.P, ptr(self.lml), stream())
        return self

    def mean_struct(self):
        """GpmpcExactModel view of the posterior MEAN of this sparse model for the fused horizon kernels: the mean is
        K*u alpha (sparse_gp.py:281), i.e. an "exact" model whose training inputs are the M scaled inducing points."""
        f = self.f
        return _lib.ExactModel(kind=f.kind, n=f.M, n_pad=f.M_pad, D=f.D, DP=f.DP, P=self.P, signal_var=f.sv,
                               lengthscales=ptr(f.ls), Xs=ptr(f.Zs), sq=ptr(f.sqz), alpha=ptr(self.alpha),
                               ymean=ptr(self.ymean), ystd=ptr(self.ystd), L=None, Winv=None)

    def predict(self, Xq, want_var=True, chunk=262144, col=None):
        """mean = K*u alpha (no L_uu^-T: reference quirk, sparse_gp.py:281), var = max(sv - |v|^2 + |w|^2, 1e-10) ystd^2
        with v = L_uu^{-1} K_u*, w = L_B^{-1} v (:293-300).  -> (Q,P), (Q,P) CUDA tensors."""
        torch = torch_()
        f = self.f
        Xq = to_dev(Xq)
        Q = Xq.shape[0]
        c0, P = (0, self.P) if col is None else (int(col), 1)
        alpha, ymean, ystd = self.alpha[c0:c0 + P], self.ymean[c0:c0 + P], self.ystd[c0:c0 + P]
        mean = torch.empty((Q, P), dtype=torch.float64, device=Xq.device)
        var = torch.empty((Q, P), dtype=torch.float64, device=Xq.device) if want_var else None
        for s in range(0, Q, chunk):
            e = min(Q, s + chunk)
            q = e - s
            Xqs, sqq = scale_inputs(Xq[s:e], f.ls, DP=f.DP)
            Ks = gram_scaled(f.kind, Xqs, sqq, q, f.Zs, f.sqz, f.M, f.sv)        # (q_pad, M_pad)
            call("gpmpc_predict_mean", ptr(Ks), Ks.stride(0), q, f.M, ptr(alpha), f.M_pad, P, ptr(ymean), ptr(ystd),
                 ptr(mean[s:e]), mean.stride(0), 0, stream())
            if want_var:
                ssv = torch.empty((Ks.shape[0],), dtype=torch.float64, device=Xq.device)
                ssw = torch.empty_like(ssv)
                call("gpmpc_trsm_sumsq", ptr(f.L_uu), f.M_pad, f.M_pad, ptr(f.W_uu), ptr(Ks), Ks.shape[0], Ks.stride(0),
                     ptr(ssv), stream())
                call("gpmpc_trsm_sumsq", ptr(f.L_B), f.M_pad, f.M_pad, ptr(f.W_B), ptr(Ks), Ks.shape[0], Ks.stride(0),
                     ptr(ssw), stream())
                call("gpmpc_predict_var", ptr(ssv), ptr(ssw), q, P, f.sv, 1e-10, ptr(ystd), ptr(var[s:e]), var.stride(0),
                     0, stream())
        return mean, var


# ====================================================================================================== refit sweep
class ExactSweep:
    """Batched exact-GP refit over a hyper-parameter grid (BASELINE config 3; the loop `optimize_hyperparameters`
    (exact_gp.py:357-421) / `HyperparameterTuner._compute_nll` (learning/hyperparameter_tuner.py:361-395) runs one
    `ExactGP.fit` at a time).  G candidates theta_g = (lengthscales (D,), signal_var, noise) share the inputs X and the
    P target columns; the G Gram matrices are built into one (G, n_pad, n_pad) buffer and factorised by ONE batched
    blocked Cholesky (every panel / trailing-update launch covers all candidates), then cho_solve and LML for all
    (g, p).  Jitter escalation (exact_gp.py:163-175) is per candidate: failed ones are rebuilt with the next jitter."""

    def __init__(self, kind, X, Y, normalize=True):
        torch = torch_()
        self.kind = int(kind)
        self.X = to_dev(X)
        self.n, self.D = self.X.shape
        self.n_pad = pad_to(self.n, 128)
        Y = to_dev(Y)
        if Y.dim() == 1:
            Y = Y[:, None].contiguous()
        self.P = Y.shape[1]
        dev = self.X.device
        self.Yn = torch.empty((self.P, self.n_pad), dtype=torch.float64, device=dev)
        self.ymean = torch.empty((self.P,), dtype=torch.float64, device=dev)
        self.ystd = torch.empty((self.P,), dtype=torch.float64, device=dev)
        call("gpmpc_normalize_targets", ptr(Y), self.n, self.P, Y.stride(0), self.n_pad, int(normalize), ptr(self.Yn),
             ptr(self.ymean), ptr(self.ystd), stream())

    def _build(self, L, g, ls, sv, diag):
        Xs, sq = scale_inputs(self.X, to_dev(np.asarray(ls, dtype=np.float64)))
        gram_scaled(self.kind, Xs, sq, self.n, Xs, sq, self.n, sv, diag_add=diag, pad_identity=1, lower_only=1, out=L[g])

    def fit(self, lengthscales, signal_vars, noises, keep_factors=False, want_alpha=False):
        """lengthscales (G,D), signal_vars (G,), noises (G,) -> dict(lml (G,P), jitter (G,), info (G,)[, alpha (G,P,n)])
        (+ L (G,n_pad,n_pad) when keep_factors)."""
        torch = torch_()
        ls = np.atleast_2d(np.asarray(lengthscales, dtype=np.float64))
        G = ls.shape[0]
        sv = np.broadcast_to(np.asarray(signal_vars, dtype=np.float64), (G,))
        nz = np.broadcast_to(np.asarray(noises, dtype=np.float64), (G,))
        dev = self.X.device
        npad = self.n_pad
        L = torch.empty((G, npad, npad), dtype=torch.float64, device=dev)
        for g in range(G):
            self._build(L, g, ls[g], sv[g], nz[g])
        Winv = torch.empty((G, npad // 64, 64, 64), dtype=torch.float64, device=dev)
        info = torch.empty((G,), dtype=torch.int32, device=dev)
        call("gpmpc_potrf", ptr(L), npad, npad, npad * npad, G, ptr(Winv), ptr(info), stream())
        jitter = np.zeros(G)
        bad = np.nonzero(info.cpu().numpy())[0]
        for g in bad:                                   # exact_gp.py:163-175, candidate by candidate
            ok = False
            for jit in ExactFactor.JITTERS[1:]:
                self._build(L, g, ls[g], sv[g], nz[g] + jit)
                call("gpmpc_potrf", ptr(L[g]), npad, npad, 0, 1, ptr(Winv[g]), ptr(info[g:g + 1]), stream())
                if int(info[g].item()) == 0:
                    jitter[g], ok = jit, True
                    break
            if not ok:
                jitter[g] = np.nan
        # y' K^-1 y = |L^-1 y|^2: forward solve only, all candidates in one batched DMMA TRSM with the P targets as rows
        # (right-looking: few rows, so the parallelism has to come from the columns still to be updated)
        V = torch.zeros((G, 128, npad), dtype=torch.float64, device=dev)
        V[:, : self.P] = self.Yn[None]
        call("gpmpc_trsm_right_batched", ptr(L), npad, npad, npad * npad, ptr(Winv), ptr(V), 128, npad, 128 * npad, G,
             stream())
        Vp = V[:, : self.P].contiguous()
        lml = torch.empty((G, self.P), dtype=torch.float64, device=dev)
        call("gpmpc_lml", ptr(L), self.n, npad, npad, npad * npad, ptr(Vp), ptr(Vp), self.P, G, ptr(lml), stream())
        out = {"lml": lml, "jitter": jitter, "info": info}
        if want_alpha:                                  # alpha = K^-1 y (exact_gp.py:179) for every candidate
            alpha = self.Yn[None].expand(G, self.P, npad).contiguous()
            call("gpmpc_potrs", ptr(L), npad, npad, npad * npad, ptr(Winv), ptr(alpha), self.P, G, stream())
            out["alpha"] = alpha[:, :, : self.n]
        if keep_factors:
            out["L"], out["Winv"] = L, Winv
        return out
