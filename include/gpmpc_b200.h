/* gpmpc_b200.h -- C ABI of the B200-native GP-MPC hot path (libgpmpc_b200.so).
 *
 * The reference (shiivashaakeri/gp-mpc-rocket-landing) is pure Python and has no FFI; the boundary its hot path
 * sits behind is a set of duck-typed Python objects.  Each entry point below replaces the
 * numeric body of the reference method it cites; `gp_mpc_rocket_landing_b200/` (and the `src.gp` / `src.mpc`
 * alias package) is the Python side that binds them with ctypes -- INTEGRATION.md shows the binding.
 *
 * Conventions: every function returns 0 on success, <0 on error (text via gpmpc_last_error(), no exceptions cross
 * the boundary); all pointers are DEVICE pointers to row-major FP64 unless a name ends in `_host`; the caller owns
 * every buffer; the last argument is a cudaStream_t (passed as void*); kernels are stream-ordered and re-entrant
 * on different streams.  "pad" sizes are multiples of 128 (rows) / 4 (feature columns) and padding is zero.
 * Process model: one process per GPU (the Python side uses torch.distributed that way).  Kernel attributes are configured
 * lazily per device ordinal, so a host that switches devices with cudaSetDevice works too, but first calls from several
 * host threads at once need external locking.
 */
#ifndef GPMPC_B200_H
#define GPMPC_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define GPMPC_KERNEL_SE 0        /* src/gp/kernels.py:130-379  SquaredExponentialARD */
#define GPMPC_KERNEL_MATERN32 1  /* src/gp/kernels.py:482-576  Matern32 */
#define GPMPC_KERNEL_MATERN52 2  /* src/gp/kernels.py:579-668  Matern52 */
#define GPMPC_KERNEL_DIST2 3     /* max(scaled squared distance, 0): kernels.py:205-236; isotropic SE :417-421 */

#define GPMPC_FEATURES_3DOF 0        /* src/gp/features.py:403-444 (11 features) */
#define GPMPC_FEATURES_6DOF_TRANS 1  /* src/gp/features.py:194-258 (13 features) */
#define GPMPC_FEATURES_6DOF_ROT 2    /* src/gp/features.py:305-354 (12 features) */

int gpmpc_version(void);
const char* gpmpc_last_error(void);

/* ---- GP: Gram / fit / predict primitives -------------------------------------------------------------------- */

/* Xs = X / lengthscales (n_pad x DP, zero padded), sq[i] = sum_d Xs[i,d]^2 in numpy's pairwise order.
 * Replaces kernels.py:223-229 (SE-ARD), :522-526 (Matern32), :615-619 (Matern52). */
int gpmpc_scale_inputs(const double* X, int n, int D, long ldx, const double* lengthscales, double* Xs, int n_pad,
                       int DP, double* sq, void* stream);

/* K[i,j] = k(x1_i, x2_j) (+ diag_add where i == j); rows >= n1_valid / cols >= n2_valid are written as zero (or the
 * identity when pad_identity).  lower_only skips 128x128 tiles strictly above the diagonal.
 * Replaces Kernel.__call__: kernels.py:238-262, :532-545, :625-638 and the `+ noise*I` at exact_gp.py:160. */
int gpmpc_gram(int kind, const double* Xs1, const double* sq1, int n1_valid, int n1_pad, const double* Xs2,
               const double* sq2, int n2_valid, int n2_pad, int DP, double signal_var, double diag_add,
               int pad_identity, int lower_only, double* K, long ldk, void* stream);

/* G = K * (x1_dim - x2_dim)^2 / l_dim^2  (SE-ARD gradient w.r.t. log lengthscale `dim`), kernels.py:310-316. */
int gpmpc_gram_grad_lengthscale(const double* Xs1, const double* sq1, int n1_valid, int n1_pad, const double* Xs2,
                                const double* sq2, int n2_valid, int n2_pad, int DP, double signal_var, int dim,
                                double* G, long ldg, void* stream);

/* In-place batched lower Cholesky (A = L L^T, upper triangle zeroed) + inverses of the 64x64 diagonal blocks
 * (Winv: batch x n_pad/64 x 64 x 64).  info[b] = 0 or the order of the first non-positive pivot.
 * Replaces np.linalg.cholesky at exact_gp.py:164,170, sparse_gp.py:187,205,228, fast_gp.py:153,158. */
int gpmpc_potrf(double* A, int n_pad, long lda, long strideA, int batch, double* Winv, int* info, void* stream);

/* B[(f*nrhs + r)*n_pad + i] <- (L_f L_f^T)^{-1} b : cho_solve at exact_gp.py:179, sparse_gp.py:210,232 and the two
 * solve_triangular calls at fast_gp.py:163-165. */
int gpmpc_potrs(const double* L, int n_pad, long lda, long strideL, const double* Winv, double* B, int nrhs, int batch,
                void* stream);

/* V (q_pad x n_pad rows, ldv) <- (L^{-1} V')' in place, right-looking: the form for FEW rows (appending points to a
 * factor); gpmpc_trsm_sumsq is the form for many query rows.  Same mathematics as solve_triangular at exact_gp.py:260. */
int gpmpc_trsm_right(const double* L, int n_pad, long lda, const double* Winv, double* V, int q_pad, long ldv,
                     void* stream);
int gpmpc_trsm_right_batched(const double* L, int n_pad, long lda, long strideL, const double* Winv, double* V, int q_pad,
                             long ldv, long strideV, int batch, void* stream);

/* Winv[b] = (64x64 diagonal block b of the factored L)^{-1} for b in [block_first, block_last]: refreshes the cached
 * block inverses after rows were appended to a factor in place (incremental alternative to the full refit of
 * online_update.py:361-408 / sparse_gp.py:328-353, which re-factorise from scratch for every batch of new points). */
int gpmpc_trinv64_blocks(const double* L, int n_pad, long lda, double* Winv, int block_first, int block_last,
                         void* stream);

/* Target normalisation, exact_gp.py:141-150: Yn[p*n_pad + i] = (Y[i*ldy + p] - mean_p) / std_p. */
int gpmpc_normalize_targets(const double* Y, int n, int P, long ldy, int n_pad, int normalize, double* Yn,
                            double* ymean, double* ystd, void* stream);

/* Log marginal likelihood, exact_gp.py:186-204; one value per (factor, rhs). */
int gpmpc_lml(const double* L, int n, int n_pad, long lda, long strideL, const double* Yn, const double* alpha,
              int nrhs, int batch, double* lml, void* stream);

/* Gradient of the log marginal likelihood w.r.t. the SE-ARD log hyper-parameters, all D + 2 components in one pass:
 * grad[p*(D+2) + c], c = 0: log signal variance, 1..D: log lengthscales, D+1: log noise variance, for P outputs that
 * share one factor.  dLML/dtheta = 1/2 sum_ij (alpha_i alpha_j - K^{-1}_ij) (dK/dtheta)_ij with the kernel derivatives of
 * kernels.py:279-318 and d(noise I)/dlog noise = noise I.  Xs: inputs scaled by 1/l (gpmpc_scale_inputs), Kinv: the
 * n_pad x n_pad inverse, of which only the lower triangle (by 128-tiles) is read -- e.g. W'W with W = L^{-1} from
 * gpmpc_tri_inverse, or gpmpc_potrs on the identity; alpha: P x lda.  work: gpmpc_lml_grad_se_work_doubles.
 * Replaces the finite-difference gradient of scipy's L-BFGS-B in ExactGP.optimize_hyperparameters (exact_gp.py:357-421:
 * D + 3 refits per gradient) with one refit + one pass. */
size_t gpmpc_lml_grad_se_work_doubles(int n_pad, int DP, int P);
int gpmpc_lml_grad_se(const double* Xs, int n, int n_pad, int DP, int D, double signal_var, double noise_var,
                      const double* Kinv, long ldk, const double* alpha, long lda, int P, double* work, double* grad,
                      void* stream);

/* Ks (q_pad x n_pad) <- (L^{-1} Ks^T)^T in place and sumsq[q] = sum_i v_iq^2: solve_triangular + np.sum(v**2, axis=0)
 * at exact_gp.py:260-261 / fast_gp.py:194-199, as a blocked DMMA TRSM. */
int gpmpc_trsm_sumsq(const double* L, int n_pad, long lda, const double* Winv, double* Ks, int q_pad, long ldk,
                     double* sumsq, void* stream);

/* The same solve for `batch` factors at once (L + f*strideL, Winv + f*(n_pad/64)*4096, Ks + f*strideK, sumsq + f*q_pad):
 * every GEMM launch covers all factors.  The hyper-parameter sweep uses it for y' K^-1 y = |L^-1 y|^2 (exact_gp.py:199). */
int gpmpc_trsm_sumsq_batched(const double* L, int n_pad, long lda, long strideL, const double* Winv, double* Ks,
                             int q_pad, long ldk, long strideK, double* sumsq, int batch, void* stream);

/* Linv <- L^{-1} (n_pad x n_pad, lower triangular, row-major) from the factor and the inverted diagonal blocks that
 * gpmpc_potrf returns.  Formed once per fit; lets the horizon pipeline evaluate the predictive variances of
 * exact_gp.py:259-264 (v = solve_triangular(L, K*^T); sum v^2) as one triangular GEMM launch, see gpmpc_linv_sumsq. */
int gpmpc_tri_inverse(const double* L, int n_pad, long lda, const double* Winv, double* Linv, long ldi, void* stream);

/* sumsq[q] = | Linv Ks[q,:]^T |^2 for the q_pad rows of Ks (q_pad x n_pad row-major, NOT modified).
 * work: (n_pad/64) x q_pad doubles of scratch (per-tile partial sums, reduced in a fixed order).
 * one_cta_per_sm != 0: keep one CTA resident per SM (same result) so that a kernel on another stream -- the ADMM
 * solve, which does not read the variances -- can share the SMs. */
int gpmpc_linv_sumsq(const double* Linv, int n_pad, long ldi, const double* Ks, int q_pad, long ldk, double* work,
                     double* sumsq, int one_cta_per_sm, void* stream);

/* The triangular GEMM of gpmpc_linv_sumsq alone: work[t * q_pad + q], t < n_pad / 64, per-tile partial sums that the
 * caller reduces in the order t = 0, 1, ... (gpmpc_solve_batch_3dof does so inside its variance / covariance kernel). */
int gpmpc_linv_sumsq_parts(const double* Linv, int n_pad, long ldi, const double* Ks, int q_pad, long ldk, double* work,
                           void* stream);
/* Same, told the number of training points n <= n_pad (0: unknown): the 8-row DMMA strips that consist of padding rows
 * n..n_pad-1 only (identity rows of Linv against zero columns of Ks) are skipped.  Both forms also skip, inside each diagonal
 * block, the 16-column chunks that lie entirely above a strip (Linv is exactly zero there): same sums, fewer DMMAs
 * (branches around whole strips: a predicated-off DMMA still occupies the tensor pipe). */
int gpmpc_linv_sumsq_parts_n(const double* Linv, int n, int n_pad, long ldi, const double* Ks, int q_pad, long ldk,
                             double* work, void* stream);

/* Small (M <= 144) lower-triangular solve against many columns, optional sumsq[c] += sign * sum_i x_ic^2:
 * sparse_gp.py:190 (A = L_uu^{-1} K_uf), :293-299 (v, w and their squared norms). */
int gpmpc_trsm_small(const double* L, int M, long ldl, double* B, long ldb, int ncols, int b_colmajor, double* sumsq,
                     double sumsq_sign, void* stream);

/* mean[q, col0+p] = (sum_j Ks[q,j] alpha[p,j]) * ystd[p] + ymean[p]  (exact_gp.py:240-245, sparse_gp.py:281-284). */
int gpmpc_predict_mean(const double* Ks, long ldk, int Q, int n, const double* alpha, int n_pad, int P,
                       const double* ymean, const double* ystd, double* mean, long ldm, int col0, void* stream);

/* var[q, col0+p] = max(prior - sumsq[q] (+ plus[q]), floor) * ystd[p]^2  (exact_gp.py:259-263, sparse_gp.py:299-300). */
int gpmpc_predict_var(const double* sumsq, const double* plus, int Q, int P, double prior, double floor_,
                      const double* ystd, double* var, long ldv, int col0, void* stream);

/* C = alpha * A * B^T + beta * C on FP64 tensor cores (M,N multiples of 128, K of 16): sparse_gp.py:203,226. */
int gpmpc_dgemm_nt(const double* A, long lda, const double* B, long ldb, double* C, long ldc, int M, int N, int K,
                   double alpha, double beta, int lower_only, void* stream);

/* Feature extraction, features.py extract()/extract_batch(): X (B x ldx), U (B x ldu) -> Z (B x ldz). */
int gpmpc_features(int which, const double* X, long ldx, const double* U, long ldu, int B, double* Z, long ldz,
                   void* stream);


/* ---- sparse (inducing-point) GP glue: sparse_gp.py:150-253 ------------------------------------------------------- */

/* A[i,i] += value for i < n (K_uu + jitter I at sparse_gp.py:186, B + jitter I at :227). */
int gpmpc_add_diag(double* A, int n, long lda, double value, void* stream);

/* FITC diagonal (sparse_gp.py:194-208): Lam[i] = max(prior - qff[i] + noise, floor); As[m,i] = A[m,i] / sqrt(Lam[i]);
 * Ylam[p,i] = Yn[p,i] / Lam[i].  A, As: M x n_pad row-major; Yn, Ylam: P x n_pad; padding columns get 1 / 0 / 0. */
int gpmpc_fitc_scale(const double* A, long lda, int M, int n, int n_pad, const double* qff, double prior, double noise,
                     double floor_, const double* Yn, int P, double* Lam, double* As, long ldas, double* Ylam,
                     void* stream);

/* FITC log marginal likelihood (method 0, sparse_gp.py:212-218) / VFE bound (method 1, :234-251), one value per output.
 * Kuf: M_pad x n_pad; Kuu, LB, Luu: M_pad x M_pad; Yn: P x n_pad; alpha, c: P x M_pad; Lam, qff: n_pad. */
int gpmpc_sparse_lml(int method, int M, int n, int n_pad, int M_pad, const double* Kuf, const double* Kuu,
                     const double* LB, const double* Luu, const double* Yn, const double* alpha, const double* c,
                     const double* Lam, const double* qff, double sv, double noise, int P, double* lml, void* stream);


/* ---- MPC: dynamics, horizon rollout / propagation, QP assembly, ADMM ---------------------------------------------- */

/* OSQP status codes (osqp constants; the reference tests result.info.status, osqp_rti.py:532). */
#define GPMPC_QP_SOLVED 1
#define GPMPC_QP_SOLVED_INACCURATE 2
#define GPMPC_QP_MAX_ITER -2
#define GPMPC_QP_PRIMAL_INFEASIBLE -3
#define GPMPC_QP_PRIMAL_INFEASIBLE_INACCURATE 3
#define GPMPC_QP_DUAL_INFEASIBLE -4
#define GPMPC_QP_DUAL_INFEASIBLE_INACCURATE 4
#define GPMPC_QP_UNSOLVED -10

/* Solver settings: the keyword arguments of osqp.OSQP.setup at osqp_rti.py:465-478 plus the OSQP 0.6 defaults the
 * reference leaves untouched.  adaptive_rho_interval is iteration based (OSQP's default is timing based). */
typedef struct GpmpcAdmmSettings {
  double rho, sigma, alpha;
  double eps_abs, eps_rel, eps_prim_inf, eps_dual_inf;
  double adaptive_rho_tolerance;
  int max_iter, scaling, check_every;
  int adaptive_rho, adaptive_rho_interval;
  int warm_x, warm_y;      /* use x / y buffers as warm start (osqp_rti.py:522-524) */
  int rho_from_input;      /* start from rho[b] when > 0 (OSQP keeps its adapted rho between solves) */
  int commit_y_solved_only; /* write y back only when status == SOLVED (GP-MPC SQP loop keeps the last good duals) */
  int kernel_variant;     /* 0: the fastest kernel measured for the shape: csrc/admm3.cu (light iterations + cyclic reduction)
                             for nx 14 at N 30, csrc/admm_stage.cu (stage-parallel) otherwise; 3: force csrc/admm3.cu (nx 7:
                             N 20, nx 14: N 30); 2: force csrc/admm_stage.cu; 1: force the first-generation one-chain kernel */
  int bound_stride;       /* 0: lb / ub (n) are shared by all QPs (the reference's fixed box, osqp_rti.py:197-200);
                             > 0: QP b reads lb + b * bound_stride, ub + b * bound_stride (per-scenario tightened boxes) */
} GpmpcAdmmSettings;

/* 3-DoF Euler model parameters (nominal_mpc.py:579-605, rocket_3dof.py:33-65). */
typedef struct GpmpcDyn3 {
  double alpha;   /* 1 / (I_sp g0) */
  double g[3];    /* gravity vector g_I */
  double dt;
} GpmpcDyn3;

/* Exact-GP model handle for the fused horizon kernels: P outputs sharing ONE kernel / factor (the
 * MultiOutputExactGP default, exact_gp.py:464-474). All device pointers. */
typedef struct GpmpcExactModel {
  int kind, n, n_pad, D, DP, P;
  double signal_var;
  const double* lengthscales; /* D */
  const double* Xs;           /* n_pad x DP scaled training inputs */
  const double* sq;           /* n_pad */
  const double* alpha;        /* P x n_pad */
  const double* ymean;        /* P */
  const double* ystd;         /* P */
  const double* L;            /* n_pad x n_pad */
  const double* Winv;         /* n_pad/64 x 64 x 64 */
} GpmpcExactModel;

/* Batched ADMM solve of B stage-structured QPs (see csrc/admm.cu): replaces osqp.OSQP.setup/update/warm_start/solve,
 * osqp_rti.py:464-478,496,517,524,527.  AB: B x N x nx x (nx+nu) row-major stage blocks [A_k | B_k]; beq: B x nx(N+1)
 * right-hand sides [x_init ; b_0 .. b_{N-1}]; q: B x n; Pdiag, lb, ub: n (shared by all scenarios; +-1e30 = infinite).
 * x (B x n) and y (B x m): warm start in, solution out.  rho (B): in (when rho_from_input) / out.
 * active (B ints or NULL): scenarios with 0 are skipped. */
int gpmpc_admm_solve(int nx, int nu, int N, int B, const double* AB, const double* beq, const double* q,
                     const double* Pdiag, const double* lb, const double* ub, const GpmpcAdmmSettings* settings,
                     const int* active, double* x, double* y, double* rho, int* status, int* iters, double* obj,
                     double* pri_res, double* dua_res, void* stream);

/* Allocation-free variant of gpmpc_admm_solve for (nx, nu) = (7, 3), second-generation (stage) kernel.  It has no counterpart
 * in the reference: it was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  It calls no cudaMalloc / cudaFree / allocating library
 * routine: the per-QP scaling scratch (n + m doubles, 561 at N = 20) lives in a static device array of
 * GPMPC_NOALLOC_MAX_DOUBLES doubles (4.6 MB, held by every GPU that loads the library until exit), so B x (n + m) beyond that
 * -- more than 1024 QPs at N = 20 -- is rejected with GPMPC_ERR_ARG and a message.  Same arguments, results, status codes
 * and iteration counts as gpmpc_admm_solve with kernel_variant 0 / 2.  All calls share the one array: calls that overlap on
 * the GPU (different streams or host threads) race; calls ordered on one stream do not.  Nothing in it needs clearing between
 * calls (the kernel writes every entry before reading it). */
#define GPMPC_NOALLOC_MAX_DOUBLES 574464
int gpmpc_admm_solve_noalloc(int N, int B, const double* AB, const double* beq, const double* q, const double* Pdiag,
                             const double* lb, const double* ub, const GpmpcAdmmSettings* settings, const int* active,
                             double* x, double* y, double* rho, int* status, int* iters, double* obj, double* pri_res,
                             double* dua_res, void* stream);

/* GP-augmented horizon rollout for B scenarios (gp_mpc.py:277-281 + _predict_with_gp :139-175, 7-state model):
 * X[b,0]=x0[b]; X[b,k+1] = step(X[b,k],U[b,k]) + [0,0,0,0, d_mean*dt].  Also writes the cross-covariance rows
 * Ks[(b*N+k), :] (consumed by gpmpc_trsm_sumsq for the variances) and dmean (B x N x 3).  model==NULL: nominal. */
int gpmpc_rollout_3dof(const GpmpcExactModel* model, const GpmpcDyn3* dyn, int B, int N, const double* x0,
                       const double* U, double* X, double* dmean, double* Ks, long ldk, const int* active,
                       void* stream);

/* GP posterior means at given trajectory points (X[b,k], U[b,k]), k < N: the affine terms c_k of one SQP iteration
 * (gp_mpc.py:309-317: gp.predict at every stage of the current iterate).  dmean: B x N x 3; Ks as in rollout. */
int gpmpc_gp_mean_points_3dof(const GpmpcExactModel* model, int B, int N, const double* X, const double* U,
                              double* dmean, double* Ks, long ldk, const int* active, void* stream);

/* Discrete Jacobians [A_d | B_d] (npts x 7 x 10) at npts points (X: npts x 7, U: npts x 3): the dynamics.linearize
 * protocol consumed at osqp_rti.py:295,340, gp_mpc.py:305, uncertainty_prop.py:147.  lin_mode as in qp_assemble. */
int gpmpc_linearize_3dof(const GpmpcDyn3* dyn, int npts, int lin_mode, const double* X, const double* U, double* AB,
                         void* stream);

/* z = [x_0,u_0,...,x_N] from (X, U): OSQPRTIMPC._solution_to_vector, osqp_rti.py:601-614 (primal warm start). */
int gpmpc_pack_traj_3dof(int B, int N, const double* X, const double* U, double* z, const int* active, void* stream);

/* Per-scenario reset of the batched controller state for respawned Monte Carlo runs (a new GPMPC per run in the reference,
 * monte_carlo.py:606-616): where mask[b] != 0 the ADMM duals y[b] (m) <- 0, rho[b] <- 0 (= take the configured default),
 * has_warm[b] <- 0 (may be NULL: the next solve then starts from the hover guess) and, when U_warm and hover0 are given,
 * U_warm[b] <- hover (first control component hover0[b], others 0). */
int gpmpc_reset_scenarios(int B, int N, int nu, int m, const int* mask, const double* hover0, double* U_warm, double* y,
                          double* rho, int* has_warm, void* stream);

/* Unscented propagation (uncertainty_prop.py:179-264), 7-state model, one stage at a time:
 * gpmpc_ut_sigma_3dof: SP[b] (15 x 7) = x_k, x_k + columns, x_k - columns of chol(scale * Sigma_k + 1e-10 I), and the stage
 * control repeated 15 times in Usp (the layout gpmpc_rollout_3dof takes with N = 1); info[0] = b + 1 of a scenario whose
 * matrix is not positive definite (np.linalg.cholesky raises).  gpmpc_ut_combine_3dof: x_{k+1} = sum w_m image_i,
 * Sigma_{k+1} = sum w_c (image_i - x_{k+1})(...)' + dt^2 diag(var) in rows 4:7, from the rollout's images Xn
 * ((B*15) x 2 x 7); n_inert = the reference's extra states (14 - 7), whose sigma points coincide with the centre. */
int gpmpc_ut_sigma_3dof(int B, double scale, const double* xk, const double* Sk, const double* U, long ldu, double* SP,
                        double* Usp, int* info, void* stream);
int gpmpc_ut_combine_3dof(int B, int n_inert, double wm0, double wc0, double wi, const double* Xn, const double* var,
                          double dt, double* xnext, double* Snext, void* stream);

/* Particle stage of the Monte Carlo propagation (uncertainty_prop.py:295-315), 7-state model: for scenario b and particle i
 * particles[b,i] = Xn[b*S+i][1] (the N = 1 rollout image: step + GP mean * dt) with sqrt(var[b*S+i]) * xi[b,i] * dt added to the
 * velocity (the draw from the GP posterior, :305); mean[b] = sum_i / S (:311), cov[b] = diff' diff / (S - 1) (:312-313).
 * One warp per scenario.  Xn: (B*S, 2, 7); var: (B*S, 3); xi: (B, S, 3) standard-normal draws; particles: (B*S, 7). */
int gpmpc_mc_combine_3dof(int B, int S, const double* Xn, const double* var, const double* xi, double dt, double* particles,
                          double* mean, double* cov, void* stream);

/* Linear covariance propagation (uncertainty_prop.py:117-177, 7-state): Sigma_{k+1} = A_k Sigma_k A_k' + diag(var)dt^2
 * on the velocity block, A_k = analytic discrete Jacobian at (X[b,k],U[b,k]).  covs: B x (N+1) x 7 x 7. */
int gpmpc_propagate_cov_3dof(const GpmpcDyn3* dyn, int B, int N, const double* X, const double* U, const double* dvar,
                             double sigma0, double* covs, const int* active, void* stream);

/* QP assembly in the reference layout (osqp_rti.py:203-372): stage blocks AB (entries with |.|<=1e-10 dropped as at
 * :301,309), beq, q.  lin_mode 0 = analytic Jacobians (FastRTI3DoF._linearize :656-710), 1 = finite differences
 * (OSQPRTIMPC._linearize :374-401).  rhs_mode 0 = reference RTI rows (l=u=+c_k, :338-344), 1 = GP-MPC deviation form
 * (gp_mpc.py:402-411).  c_gp (B x N x 3 GP means, may be NULL) enters as the affine term [0,0,0,0,d*dt]. */
int gpmpc_qp_assemble_3dof(const GpmpcDyn3* dyn, int B, int N, int lin_mode, int rhs_mode, const double* X_lin,
                           const double* U_lin, const double* x_init, const double* x_ref, long ldxref,
                           const double* c_gp, const double* Qdiag, const double* Qfdiag, double* AB, double* beq,
                           double* q, const int* active, void* stream);

/* Unpack QP solutions z -> (X, U), SQP convergence test (gp_mpc.py:336-345), cost, fallback handling and the shifted
 * warm start (osqp_rti.py:532-552).  See csrc/mpc.cu for the exact semantics of `mode`. */
int gpmpc_mpc_update_3dof(int B, int N, int mode, const double* z, const int* qp_status, const double* x_ref,
                          long ldxref, const double* Qdiag, const double* Qfdiag, const double* Rdiag, double* X,
                          double* U, double* X_prev, double* U_prev, double* cost, int* converged, int* active,
                          int* n_iter, void* stream);

/* ---- convex thrust / glide-slope rows inside the ADMM QP (second half) -----------------------------------
 *
 * The reference's sub-problem (gp_mpc.py:417-429) constrains every stage k < N by  T_min^2 <= |u_k|^2 <= T_max^2  and the
 * glide-slope cone  r_x^2 tan^2(gamma) >= r_y^2 + r_z^2  and hands them to IPOPT.  For the OSQP-form QP they become
 * polyhedral INNER approximations (every point the QP accepts satisfies the original constraints):
 *   - |u| >= T_min (non-convex)  ->  u_x >= T_min: the lower box bound of the axial component, set by the host;
 *   - |u| <= T_max  ->  u_x +- b u_y <= T_max,  u_x +- b u_z <= T_max  with  b = sqrt(2 (T_max - T_min) / (T_max + T_min)):
 *     a square pyramid with its apex at T_max on the axial direction, inside the ball for u_x >= T_min;
 *   - the cone at the NEXT stage, sqrt(y^2 + z^2) <= alt tan(gamma): the pyramid over the triangle inscribed in the circle
 *     with one vertex in the radial direction e of the nominal point,  n_i . (y, z) <= alt tan(gamma) / 2  for the normals
 *     n_i = R(+60 deg) e, R(-60 deg) e, -e  (e = (1, 0) when the nominal point is on the axis).  The host leaves the glide
 *     rows of stage 1 unbounded: x_1 = x_0 + dt v_0 whatever the controls, so they are either redundant or make the QP
 *     infeasible by the solver tolerance of the previous step.
 * To keep the stage structure the ADMM kernels factorise (block tridiagonal, no extra row type), the 7 row values travel as
 * 7 extra "states":  xi_{k+1} = G_k [x_k; u_k] - g_k  with  -inf < xi <= (T_max x 4, 0 x 3)  as box bounds, i.e. the QP is
 * solved in the stage form (n_x, n_u) = (14, 3) with stage variables [x (7); xi (7); u (3)] and zero cost on xi;
 * xi_0 = 0 is fixed by the initial-state rows and unbounded.  n = 17 N + 14, m = 14 (N + 1) + n. */
typedef struct GpmpcCone3 {
  int enabled;
  double T_max;        /* constraints.py:39 */
  double facet_b;      /* b above */
  double half_tan_gs;  /* tan(gamma_gs) / 2, constraints.py:50 */
} GpmpcCone3;

/* (AB, beq, q) of gpmpc_qp_assemble_3dof + the nominal trajectory -> the augmented stage QP (AB_c: B x N x 14 x 17,
 * beq_c: B x 14 (N + 1), q_c: B x (17 N + 14)) and its primal warm start z_c (may be NULL). */
int gpmpc_qp_augment_cone_3dof(const GpmpcCone3* cone, int B, int N, const double* AB, const double* beq, const double* q,
                               const double* X_nom, const double* U_nom, double* AB_c, double* beq_c, double* q_c,
                               double* z_c, const int* active, void* stream);
/* z_c (B x (17 N + 14)) -> z (B x (10 N + 7)): the [x; u] part of every stage. */
int gpmpc_qp_extract_cone_3dof(int B, int N, const double* z_c, double* z, const int* active, void* stream);

/* ---- fused per-control-step orchestration (SURVEY 8b proposal `gpmpc_solve_batch`) ---------------------------------------
 *
 * One call issues the whole device sequence of `GPMPC.solve` (gp_mpc.py:229-369, 7-state specialisation) for B scenarios:
 *   prologue (reference trajectory from the target, initial controls = warm start | U_ref | hover, gp_mpc.py:266-275)
 *   -> GP-augmented rollout (:277-281) -> GP variances + covariance chain (:284-289, uncertainty_prop.py:117-177)
 *   -> max_sqp_iter x [ GP means at the iterate (:309-317) -> QP assembly + primal warm start (osqp_rti.py:203-372,601-614)
 *                       -> ADMM (osqp_rti.py:464-527) -> unpack / step-size test / cost (:336-345)
 *                       -> re-propagation of the uncertainty (:347-353) ]
 *   -> warm start for the next call (:355-357).
 * 8 kernel launches for max_sqp_iter = 1, no host synchronisation, no allocation: capturable in a CUDA graph.
 * Every pointer is a device pointer owned by the caller and must stay valid until the stream has run the call. */
typedef struct GpmpcSolve3 {
  int B, N;
  int max_sqp_iter;        /* gp_mpc.py:293 (10); 1 = real-time iteration */
  int use_gp_mean;         /* GPMPCConfig.use_gp_mean (needs model) */
  int use_gp_uncertainty;  /* GPMPCConfig.use_gp_uncertainty (needs model, Linv or L/Winv, Ks, covs) */
  int repropagate;         /* gp_mpc.py:347-353 */
  int target_mode;         /* 0: x_target (B x 7) given; 1: derived per scenario from the current state as the Monte Carlo
                              loop does for `.solve` controllers (monte_carlo.py:497-499: x, v = 0, alt = max(0.5, alt - 2)) */
  int has_model;
  GpmpcDyn3 dyn;
  double g0;               /* hover thrust m g0 along axis 0 (nominal_mpc.py:651-657) */
  GpmpcAdmmSettings admm;
  GpmpcExactModel model;
  const double* Linv;      /* n_pad x n_pad explicit inverse factor (gpmpc_tri_inverse) or NULL -> blocked in-place solve */
  long ldi;
  double tighten_kappa;    /* > 0: per-stage box tightening (gpmpc_tighten_bounds) with lb_t / ub_t */
  double tighten_max_variance;
  /* inputs */
  const double* x0;        /* B x 7 */
  const double* x_target;  /* B x 7 (target_mode 0) */
  const double* X_ref;     /* optional B x (N+1) x 7 reference trajectory (overrides the target), row stride ld_xref_in (0: shared) */
  long ld_xref_in;
  const double* U_ref;     /* optional (N x 3), stride ld_uref per scenario (0: shared) */
  long ld_uref;
  const int* active_in;    /* optional B */
  /* shared QP data (n = 10 N + 7) */
  const double *Pdiag, *lb, *ub, *Qd, *Qfd, *Rd;
  /* per-scenario state and scratch */
  double *x_ref;           /* B x (N+1) x 7 */
  double *X, *U;           /* B x (N+1) x 7, B x N x 3: iterate in / solution out */
  double *Xs, *dmean, *dmean_s;
  double *Ks; long ldk; int q_pad;
  double *var_work, *sumsq, *dvar, *covs;
  double *lb_t, *ub_t;
  double *AB, *beq, *q;
  double *cost; int *converged, *active, *n_iter, *qp_iters;
  double *z, *y, *rho, *obj, *pri, *dua; int *qp_status, *qp_it;
  double *U_warm, *X_warm; int* has_warm;   /* per-scenario warm start; has_warm[b] = 0 -> hover / U_ref */
  /* optional convex rows (cone.enabled): the QP is solved in the augmented (14, 3) stage form; y_c (B x (31 N + 28)) takes the
     place of y as the dual state that persists between control steps; Pdiag_c / lb_c / ub_c: shared, length 17 N + 14 */
  GpmpcCone3 cone;
  double *AB_c, *beq_c, *q_c, *z_c, *y_c;
  const double *Pdiag_c, *lb_c, *ub_c;
  /* optional phase timing: events[i] (cudaEvent_t) is recorded after phase event_phase[i] (host arrays written by the call) */
  void** events; int* event_phase; int event_cap; int* event_count;
} GpmpcSolve3;

#define GPMPC_PHASE_BEGIN 0
#define GPMPC_PHASE_PROLOGUE 1
#define GPMPC_PHASE_ROLLOUT 2
#define GPMPC_PHASE_GP_VARIANCE 3
#define GPMPC_PHASE_COV 4
#define GPMPC_PHASE_GP_MEAN 5
#define GPMPC_PHASE_ASSEMBLE 6
#define GPMPC_PHASE_ADMM 7
#define GPMPC_PHASE_UPDATE 8
#define GPMPC_PHASE_MC 9

int gpmpc_solve_batch_3dof(const GpmpcSolve3* plan, int* n_launches, void* stream);

/* Closed-loop Monte Carlo advance after a solve (experiments/monte_carlo.py:455-549), one thread per scenario, all on the
 * device: [fail_policy 1: a scenario whose solve did not succeed ends as DIVERGENCE without a plant step (:507-510, the
 * branch taken for controllers with `.solve`); 0: the flag is only counted (as for `.step` controllers, :492-494)]
 * -> plant step (+ drag residual) -> t_step += 1 -> TIMEOUT when t_step reaches max_steps (the for/else at :539-542, the last
 * state is not classified) else the termination rules :460-488 + check_landing :54-104.  respawn != 0: a finished
 * scenario restarts from x_init[b] with a fresh controller (has_warm = 0, duals 0, rho unset) so the batch stays full.
 * classify_only != 0: no solve / plant step, only the termination rules on the current state (the reference applies them
 * to x0 before the first control). */
typedef struct GpmpcMcStep3 {
  int B, N, m;             /* m = rows of the QP (length of y per scenario) */
  GpmpcDyn3 dyn;
  double drag_cda;
  int max_steps, fail_policy, respawn, classify_only;
  double* x;               /* B x 7 in / out */
  const double* u;         /* B x ldu: first control of the solution */
  long ldu;
  const int* success;      /* B: the solve's success flag (converged) or NULL */
  const double *m0, *x_init;
  int *t_step, *active, *outcome, *fail_steps;
  double* x_final;
  unsigned long long* counts;   /* 8 counters: [1..6] runs ended per outcome code, [0] solves consumed, [7] their ADMM iterations */
  int* has_warm; double *y, *rho;   /* controller state cleared on respawn (may be NULL) */
  const int* qp_iters;          /* B: ADMM iterations of this step's solve (may be NULL) */
} GpmpcMcStep3;
int gpmpc_mc_advance_3dof(const GpmpcMcStep3* args, void* stream);

/* Predictive safety filter on the first control (src/safety/safety_filter.py:184-361, the branch that runs without CasADi):
 * per scenario  (1) `_check_safety` :237-279 -- immediate constraints (:315-346: thrust magnitude in [T_min, T_max], glide
 * slope r_x^2 tan^2(gamma) >= r_y^2 + r_z^2 above 0.1 m; the 7-state model has identity attitude, so the tilt test always
 * passes), N-step prediction with u first and the backup law u_eq - K (x - x_eq) afterwards (:281-313,
 * backup_controller.py:231-252), terminal Lyapunov value (x_N - x_eq)' P (x_N - x_eq) against alpha (invariant_sets.py:178-184),
 * path constraints;  (2) when unsafe, `_solve_gradient_qp` :483-531 -- <= n_iters steps of  u -= lr (dV/du + 2 (u - u_nom))
 * with forward-difference gradients (eps 1e-4) through the same prediction, stopped when V <= alpha * backup_margin, each
 * followed by the projection on the thrust annulus;  (3) the check of the filtered control.
 * flags[b]: bit 0 = control was modified, bit 1 = the returned control is safe.  U_nom row stride ldu (u0 = U_opt[:, 0]). */
typedef struct GpmpcSafety3 {
  GpmpcDyn3 dyn;          /* dt = the filter's step (SafetyFilterConfig.dt) */
  int N, n_iters;         /* SafetyFilterConfig.N (10); 50 */
  int has_u_limits;       /* saturate the backup law to [u_min, u_max] (BackupControllerConfig.u_min / u_max) */
  int reserved;
  double lr, alpha, backup_margin, T_min, T_max, tan_gamma;
  double K[21];           /* 3 x 7 backup gain, row-major */
  double P[49];           /* 7 x 7 Lyapunov matrix */
  double x_eq[7], u_eq[3], u_min[3], u_max[3];
} GpmpcSafety3;
int gpmpc_safety_filter_3dof(const GpmpcSafety3* setup, int B, const double* X, const double* U_nom, long ldu, double* U_safe,
                             int* flags, double* margin, double* backup_value, const int* active, void* stream);

/* sizeof / offsetof probes for the binding tests: which = 0 GpmpcSolve3, 1 GpmpcMcStep3, 2 GpmpcAdmmSettings,
 * 3 GpmpcExactModel, 4 GpmpcSafety3; field < 0 -> sizeof, else the offset of a sentinel field (see csrc/mpc.cu). */
long gpmpc_struct_layout(int which, int field);

/* CUDA event pool for the phase timers (plain cudaEvent_t behind void*). */
int gpmpc_events_create(int n, void** events_out);
int gpmpc_events_destroy(int n, void** events);
int gpmpc_events_elapsed_ms(void* a, void* b, float* ms_out);   /* synchronises on b */

/* ---- 6-DoF quaternion model (BASELINE configs[3]): 14 states [m, r_I(3), v_I(3), q_BI(4), w_B(3)], 3 controls ----------- */

/* Parameters of the continuous model stated at nominal_mpc.py:163-203 (defaults rocket_6dof.py:35-89). */
typedef struct GpmpcDyn6 {
  double alpha;      /* 1 / (I_sp g0) */
  double g[3];       /* g_I */
  double J[9];       /* J_B row-major */
  double Jinv[9];    /* J_B^{-1} row-major */
  double rT[3];      /* r_T_B */
  double dt;
  int substeps;      /* 1: Euler; 4: the reference's "rk4" (= 4 Euler sub-steps, nominal_mpc.py:205-236) */
  int normalize_quat; /* renormalise q after the step (rocket_6dof.py:370-387) */
} GpmpcDyn6;

/* Xn[i] = step(X[i], U[i]) for npts points (dynamics.step protocol, rocket_6dof.py:370-376). */
int gpmpc_step_6dof(const GpmpcDyn6* dyn, int npts, const double* X, const double* U, double* Xn, void* stream);

/* Forward-difference discrete Jacobians [A | B] (npts x 14 x 17), eps = 1e-6: osqp_rti.py:374-401. */
int gpmpc_linearize_6dof(const GpmpcDyn6* dyn, int npts, const double* X, const double* U, double* AB, void* stream);

/* GP-augmented horizon rollout, 14-state _predict_with_gp (gp_mpc.py:139-175,277-281):
 * X[b,k+1] = step(X[b,k],U[b,k]) + dt [0..,d_v,0..,d_w].  gp_v / gp_w: posterior-mean models of the translational
 * (13 features, features.py:194-258) and rotational (12 features, :305-354) residual GPs -- for a sparse GP the
 * "training inputs" of the struct are the scaled inducing points and alpha its M-vector (sparse_gp.py:281), L unused.
 * Either may be NULL (nominal).  dmean: B x N x 6 or NULL. */
int gpmpc_rollout_6dof(const GpmpcExactModel* gp_v, const GpmpcExactModel* gp_w, const GpmpcDyn6* dyn, int B, int N,
                       const double* x0, const double* U, double* X, double* dmean, const int* active, void* stream);

/* QP assembly in the OSQP-RTI layout for n_x = 14 (osqp_rti.py:203-372; cost fallback branch :172-175): stage blocks
 * AB (B x N x 14 x 17, FD Jacobians with |a| <= 1e-10 dropped), beq (B x 14(N+1)) = [x_init ; -(X_lin[k+1] - A x - B u +
 * dt c_gp)] (gp_mpc.py:402-411), q (B x n) = -Q x_ref.  x_ref: (N+1) x 14 shared (ldxref = 0) or per scenario. */
int gpmpc_qp_assemble_6dof(const GpmpcDyn6* dyn, int B, int N, const double* X_lin, const double* U_lin,
                           const double* x_init, const double* x_ref, long ldxref, const double* c_gp, const double* Qd,
                           const double* Qfd, double* AB, double* beq, double* q, const int* active, void* stream);

/* Per-scenario, per-stage box tightening from the propagated covariances: for state i of stage k >= 1,
 * lb' = lb + kappa sqrt(Sigma_k[i][i]), ub' = ub - kappa sqrt(Sigma_k[i][i]) (finite bounds only; a box that would invert
 * collapses to its midpoint; variances above max_variance leave the stage untouched, gp_mpc.py:206-208).  Controls and
 * stage 0 keep the shared box.  lb, ub: n = N (nx + nu) + nx shared entries; lb_out, ub_out: B x n.  This is the per-stage
 * use of Sigma_k that gp_mpc.py:177-215,414 prepares and then applies at k = 0 only. */
int gpmpc_tighten_bounds(int nx, int nu, int B, int N, const double* lb, const double* ub, const double* covs,
                         double kappa, double max_variance, double* lb_out, double* ub_out, const int* active,
                         void* stream);

/* Tube half-widths e_0 = 0, e_{k+1} = |A_k| e_k + w_k for B trajectories (uncertainty_prop.py:435-470 with
 * w = d_max dt in the velocity / rate rows; safety/tube_mpc.py:118-183 with w = n_sigma * GP standard deviations).
 * A as in gpmpc_propagate_cov; w: (B,N,nx); tubes: (B,N+1,nx). */
int gpmpc_propagate_tube(int nx, int B, int N, const double* A, long ld_row, long stage_stride, const double* w,
                         double* tubes, const int* active, void* stream);

/* Linear covariance chain Sigma_0 = sigma0 I, Sigma_{k+1} = A_k Sigma_k A_k' + diag(qdiag[b,k]) for any nx <= 16:
 * uncertainty_prop.py:117-177 (the 14-state propagator GPMPC.solve calls at gp_mpc.py:283-289,347-351).  A: entry (i,j) of
 * the stage-(b,k) Jacobian at A[(b*N+k)*stage_stride + i*ld_row + j] (e.g. the [A|B] blocks of gpmpc_linearize_6dof with
 * ld_row 17, stage_stride 14*17); qdiag: (B,N,nx) = GP variances * dt^2 in the velocity / rate rows; covs: (B,N+1,nx,nx). */
int gpmpc_propagate_cov(int nx, int B, int N, const double* A, long ld_row, long stage_stride, const double* qdiag,
                        double sigma0, double* covs, const int* active, void* stream);

/* z = [x_0,u_0,...,x_N] <-> (X, U) for any state dimension (osqp_rti.py:601-631). */
int gpmpc_pack_traj(int nx, int nu, int B, int N, const double* X, const double* U, double* z, void* stream);
int gpmpc_unpack_traj(int nx, int nu, int B, int N, const double* z, double* X, double* U, void* stream);

/* Post-QP bookkeeping for any state dimension: unpack z into (X, U), SQP step-size test max|dz| < 1e-4 (gp_mpc.py:336-345,
 * mode 1; converged scenarios are cleared in `active`) or RTI acceptance (osqp_rti.py:532-552, mode 0), stage cost,
 * optional shifted warm start (X_prev, U_prev).  Unsolved QPs leave (X, U) untouched and report cost = inf. */
int gpmpc_mpc_update(int nx, int nu, int B, int N, int mode, const double* z, const int* qp_status, const double* x_ref,
                     long ldxref, const double* Qd, const double* Qfd, const double* Rd, double* X, double* U,
                     double* X_prev, double* U_prev, double* cost, int* converged, int* active, int* n_iter,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GPMPC_B200_H */
