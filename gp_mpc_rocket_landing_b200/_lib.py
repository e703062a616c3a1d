"""ctypes binding of libgpmpc_b200.so (include/gpmpc_b200.h).

The library is the product: there is no CPU fallback.  Importing this module fails loudly when the shared
library has not been built (``python -m gp_mpc_rocket_landing_b200.build``), and every compute entry point
fails loudly when no CUDA device is present.  torch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgpmpc_b200.so")

c_dp = C.c_void_p  # device pointers travel as integers
c_int, c_long, c_double = C.c_int, C.c_long, C.c_double


class AdmmSettings(C.Structure):
    """GpmpcAdmmSettings: osqp.OSQP.setup keywords (osqp_rti.py:465-478) + untouched OSQP 0.6 defaults."""
    _fields_ = [
        ("rho", c_double), ("sigma", c_double), ("alpha", c_double),
        ("eps_abs", c_double), ("eps_rel", c_double), ("eps_prim_inf", c_double), ("eps_dual_inf", c_double),
        ("adaptive_rho_tolerance", c_double),
        ("max_iter", c_int), ("scaling", c_int), ("check_every", c_int),
        ("adaptive_rho", c_int), ("adaptive_rho_interval", c_int),
        ("warm_x", c_int), ("warm_y", c_int), ("rho_from_input", c_int),
        ("commit_y_solved_only", c_int), ("kernel_variant", c_int), ("bound_stride", c_int),
    ]

    @classmethod
    def reference_defaults(cls, **over):
        s = cls(rho=0.1, sigma=1e-6, alpha=1.6, eps_abs=1e-4, eps_rel=1e-4, eps_prim_inf=1e-4, eps_dual_inf=1e-4,
                adaptive_rho_tolerance=5.0, max_iter=50, scaling=3, check_every=25, adaptive_rho=1,
                adaptive_rho_interval=25, warm_x=0, warm_y=0, rho_from_input=0, commit_y_solved_only=0, kernel_variant=0,
                bound_stride=0)
        for k, v in over.items():
            setattr(s, k, v)
        return s


class Dyn3(C.Structure):
    """GpmpcDyn3: Euler 3-DoF parameters (nominal_mpc.py:579-605, rocket_3dof.py:33-65)."""
    _fields_ = [("alpha", c_double), ("g", c_double * 3), ("dt", c_double)]


class Dyn6(C.Structure):
    """GpmpcDyn6: 6-DoF quaternion model parameters (nominal_mpc.py:163-203, rocket_6dof.py:35-89)."""
    _fields_ = [("alpha", c_double), ("g", c_double * 3), ("J", c_double * 9), ("Jinv", c_double * 9),
                ("rT", c_double * 3), ("dt", c_double), ("substeps", c_int), ("normalize_quat", c_int)]


class ExactModel(C.Structure):
    """GpmpcExactModel: P outputs sharing one kernel / Cholesky factor (exact_gp.py:464-474)."""
    _fields_ = [
        ("kind", c_int), ("n", c_int), ("n_pad", c_int), ("D", c_int), ("DP", c_int), ("P", c_int),
        ("signal_var", c_double),
        ("lengthscales", c_dp), ("Xs", c_dp), ("sq", c_dp), ("alpha", c_dp), ("ymean", c_dp), ("ystd", c_dp),
        ("L", c_dp), ("Winv", c_dp),
    ]


class Cone3(C.Structure):
    """GpmpcCone3: polyhedral thrust / glide-slope rows (gp_mpc.py:417-429 as inner approximations)."""
    _fields_ = [("enabled", c_int), ("T_max", c_double), ("facet_b", c_double), ("half_tan_gs", c_double)]


class Solve3(C.Structure):
    """GpmpcSolve3: the plan of one fused batched solve (gpmpc_solve_batch_3dof); field order = include/gpmpc_b200.h."""
    _fields_ = [
        ("B", c_int), ("N", c_int), ("max_sqp_iter", c_int), ("use_gp_mean", c_int), ("use_gp_uncertainty", c_int),
        ("repropagate", c_int), ("target_mode", c_int), ("has_model", c_int),
        ("dyn", Dyn3), ("g0", c_double), ("admm", AdmmSettings), ("model", ExactModel),
        ("Linv", c_dp), ("ldi", c_long), ("tighten_kappa", c_double), ("tighten_max_variance", c_double),
        ("x0", c_dp), ("x_target", c_dp), ("X_ref", c_dp), ("ld_xref_in", c_long), ("U_ref", c_dp), ("ld_uref", c_long),
        ("active_in", c_dp),
        ("Pdiag", c_dp), ("lb", c_dp), ("ub", c_dp), ("Qd", c_dp), ("Qfd", c_dp), ("Rd", c_dp),
        ("x_ref", c_dp), ("X", c_dp), ("U", c_dp), ("Xs", c_dp), ("dmean", c_dp), ("dmean_s", c_dp),
        ("Ks", c_dp), ("ldk", c_long), ("q_pad", c_int),
        ("var_work", c_dp), ("sumsq", c_dp), ("dvar", c_dp), ("covs", c_dp), ("lb_t", c_dp), ("ub_t", c_dp),
        ("AB", c_dp), ("beq", c_dp), ("q", c_dp),
        ("cost", c_dp), ("converged", c_dp), ("active", c_dp), ("n_iter", c_dp), ("qp_iters", c_dp),
        ("z", c_dp), ("y", c_dp), ("rho", c_dp), ("obj", c_dp), ("pri", c_dp), ("dua", c_dp), ("qp_status", c_dp),
        ("qp_it", c_dp),
        ("U_warm", c_dp), ("X_warm", c_dp), ("has_warm", c_dp),
        ("cone", Cone3), ("AB_c", c_dp), ("beq_c", c_dp), ("q_c", c_dp), ("z_c", c_dp), ("y_c", c_dp),
        ("Pdiag_c", c_dp), ("lb_c", c_dp), ("ub_c", c_dp),
        ("events", C.POINTER(C.c_void_p)), ("event_phase", C.POINTER(c_int)), ("event_cap", c_int),
        ("event_count", C.POINTER(c_int)),
    ]


class McStep3(C.Structure):
    """GpmpcMcStep3: arguments of gpmpc_mc_advance_3dof."""
    _fields_ = [
        ("B", c_int), ("N", c_int), ("m", c_int), ("dyn", Dyn3), ("drag_cda", c_double),
        ("max_steps", c_int), ("fail_policy", c_int), ("respawn", c_int), ("classify_only", c_int),
        ("x", c_dp), ("u", c_dp), ("ldu", c_long), ("success", c_dp), ("m0", c_dp), ("x_init", c_dp),
        ("t_step", c_dp), ("active", c_dp), ("outcome", c_dp), ("fail_steps", c_dp), ("x_final", c_dp), ("counts", c_dp),
        ("has_warm", c_dp), ("y", c_dp), ("rho", c_dp), ("qp_iters", c_dp),
    ]


class Safety3(C.Structure):
    """GpmpcSafety3: set-up of the predictive safety filter (gpmpc_safety_filter_3dof)."""
    _fields_ = [
        ("dyn", Dyn3), ("N", c_int), ("n_iters", c_int), ("has_u_limits", c_int), ("reserved", c_int),
        ("lr", c_double), ("alpha", c_double), ("backup_margin", c_double), ("T_min", c_double), ("T_max", c_double),
        ("tan_gamma", c_double),
        ("K", c_double * 21), ("P", c_double * 49), ("x_eq", c_double * 7), ("u_eq", c_double * 3),
        ("u_min", c_double * 3), ("u_max", c_double * 3),
    ]


PHASE_NAMES = ("begin", "prologue", "rollout", "gp_variance", "cov", "gp_mean", "assemble", "admm", "update", "mc")


_PROTOS = {
    "gpmpc_version": (c_int, []),
    "gpmpc_last_error": (C.c_char_p, []),
    "gpmpc_scale_inputs": (c_int, [c_dp, c_int, c_int, c_long, c_dp, c_dp, c_int, c_int, c_dp, c_dp]),
    "gpmpc_gram": (c_int, [c_int, c_dp, c_dp, c_int, c_int, c_dp, c_dp, c_int, c_int, c_int, c_double, c_double,
                           c_int, c_int, c_dp, c_long, c_dp]),
    "gpmpc_gram_grad_lengthscale": (c_int, [c_dp, c_dp, c_int, c_int, c_dp, c_dp, c_int, c_int, c_int, c_double,
                                            c_int, c_dp, c_long, c_dp]),
    "gpmpc_potrf": (c_int, [c_dp, c_int, c_long, c_long, c_int, c_dp, c_dp, c_dp]),
    "gpmpc_potrs": (c_int, [c_dp, c_int, c_long, c_long, c_dp, c_dp, c_int, c_int, c_dp]),
    "gpmpc_trsm_right": (c_int, [c_dp, c_int, c_long, c_dp, c_dp, c_int, c_long, c_dp]),
    "gpmpc_trsm_right_batched": (c_int, [c_dp, c_int, c_long, c_long, c_dp, c_dp, c_int, c_long, c_long, c_int, c_dp]),
    "gpmpc_trinv64_blocks": (c_int, [c_dp, c_int, c_long, c_dp, c_int, c_int, c_dp]),
    "gpmpc_normalize_targets": (c_int, [c_dp, c_int, c_int, c_long, c_int, c_int, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_lml": (c_int, [c_dp, c_int, c_int, c_long, c_long, c_dp, c_dp, c_int, c_int, c_dp, c_dp]),
    "gpmpc_lml_grad_se_work_doubles": (C.c_size_t, [c_int, c_int, c_int]),
    "gpmpc_lml_grad_se": (c_int, [c_dp, c_int, c_int, c_int, c_int, c_double, c_double, c_dp, c_long, c_dp, c_long, c_int,
                                  c_dp, c_dp, c_dp]),
    "gpmpc_trsm_sumsq": (c_int, [c_dp, c_int, c_long, c_dp, c_dp, c_int, c_long, c_dp, c_dp]),
    "gpmpc_reset_scenarios": (c_int, [c_int, c_int, c_int, c_int, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_ut_sigma_3dof": (c_int, [c_int, c_double, c_dp, c_dp, c_dp, c_long, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_ut_combine_3dof": (c_int, [c_int, c_int, c_double, c_double, c_double, c_dp, c_dp, c_double, c_dp, c_dp, c_dp]),
    "gpmpc_tighten_bounds": (c_int, [c_int, c_int, c_int, c_int, c_dp, c_dp, c_dp, c_double, c_double, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_propagate_tube": (c_int, [c_int, c_int, c_int, c_dp, c_long, c_long, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_propagate_cov": (c_int, [c_int, c_int, c_int, c_dp, c_long, c_long, c_dp, c_double, c_dp, c_dp, c_dp]),
    "gpmpc_step_6dof": (c_int, [C.POINTER(Dyn6), c_int, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_linearize_6dof": (c_int, [C.POINTER(Dyn6), c_int, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_rollout_6dof": (c_int, [C.POINTER(ExactModel), C.POINTER(ExactModel), C.POINTER(Dyn6), c_int, c_int, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_qp_assemble_6dof": (c_int, [C.POINTER(Dyn6), c_int, c_int, c_dp, c_dp, c_dp, c_dp, c_long, c_dp, c_dp, c_dp, c_dp,
                                       c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_mpc_update": (c_int, [c_int, c_int, c_int, c_int, c_int, c_dp, c_dp, c_dp, c_long, c_dp, c_dp, c_dp, c_dp, c_dp,
                                 c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_pack_traj": (c_int, [c_int, c_int, c_int, c_int, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_unpack_traj": (c_int, [c_int, c_int, c_int, c_int, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_tri_inverse": (c_int, [c_dp, c_int, c_long, c_dp, c_dp, c_long, c_dp]),
    "gpmpc_linv_sumsq": (c_int, [c_dp, c_int, c_long, c_dp, c_int, c_long, c_dp, c_dp, c_int, c_dp]),
    "gpmpc_trsm_sumsq_batched": (c_int, [c_dp, c_int, c_long, c_long, c_dp, c_dp, c_int, c_long, c_long, c_dp, c_int,
                                         c_dp]),
    "gpmpc_trsm_small": (c_int, [c_dp, c_int, c_long, c_dp, c_long, c_int, c_int, c_dp, c_double, c_dp]),
    "gpmpc_predict_mean": (c_int, [c_dp, c_long, c_int, c_int, c_dp, c_int, c_int, c_dp, c_dp, c_dp, c_long, c_int,
                                   c_dp]),
    "gpmpc_predict_var": (c_int, [c_dp, c_dp, c_int, c_int, c_double, c_double, c_dp, c_dp, c_long, c_int, c_dp]),
    "gpmpc_dgemm_nt": (c_int, [c_dp, c_long, c_dp, c_long, c_dp, c_long, c_int, c_int, c_int, c_double, c_double,
                               c_int, c_dp]),
    "gpmpc_features": (c_int, [c_int, c_dp, c_long, c_dp, c_long, c_int, c_dp, c_long, c_dp]),
    "gpmpc_add_diag": (c_int, [c_dp, c_int, c_long, c_double, c_dp]),
    "gpmpc_fitc_scale": (c_int, [c_dp, c_long, c_int, c_int, c_int, c_dp, c_double, c_double, c_double, c_dp, c_int,
                                 c_dp, c_dp, c_long, c_dp, c_dp]),
    "gpmpc_sparse_lml": (c_int, [c_int, c_int, c_int, c_int, c_int, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp,
                                 c_dp, c_double, c_double, c_int, c_dp, c_dp]),
    "gpmpc_admm_solve": (c_int, [c_int, c_int, c_int, c_int, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp,
                                 C.POINTER(AdmmSettings), c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_admm_solve_noalloc": (c_int, [c_int, c_int, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp,
                                         C.POINTER(AdmmSettings), c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_rollout_3dof": (c_int, [C.POINTER(ExactModel), C.POINTER(Dyn3), c_int, c_int, c_dp, c_dp, c_dp, c_dp,
                                   c_dp, c_long, c_dp, c_dp]),
    "gpmpc_gp_mean_points_3dof": (c_int, [C.POINTER(ExactModel), c_int, c_int, c_dp, c_dp, c_dp, c_dp, c_long, c_dp,
                                          c_dp]),
    "gpmpc_linearize_3dof": (c_int, [C.POINTER(Dyn3), c_int, c_int, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_pack_traj_3dof": (c_int, [c_int, c_int, c_dp, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_propagate_cov_3dof": (c_int, [C.POINTER(Dyn3), c_int, c_int, c_dp, c_dp, c_dp, c_double, c_dp, c_dp,
                                         c_dp]),
    "gpmpc_qp_assemble_3dof": (c_int, [C.POINTER(Dyn3), c_int, c_int, c_int, c_int, c_dp, c_dp, c_dp, c_dp, c_long,
                                       c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_linv_sumsq_parts": (c_int, [c_dp, c_int, c_long, c_dp, c_int, c_long, c_dp, c_dp]),
    "gpmpc_linv_sumsq_parts_n": (c_int, [c_dp, c_int, c_int, c_long, c_dp, c_int, c_long, c_dp, c_dp]),
    "gpmpc_solve_batch_3dof": (c_int, [C.POINTER(Solve3), C.POINTER(c_int), c_dp]),
    "gpmpc_qp_augment_cone_3dof": (c_int, [C.POINTER(Cone3), c_int, c_int, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp,
                                           c_dp, c_dp, c_dp]),
    "gpmpc_qp_extract_cone_3dof": (c_int, [c_int, c_int, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_mc_advance_3dof": (c_int, [C.POINTER(McStep3), c_dp]),
    "gpmpc_mc_combine_3dof": (c_int, [c_int, c_int, c_dp, c_dp, c_dp, c_double, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_safety_filter_3dof": (c_int, [C.POINTER(Safety3), c_int, c_dp, c_dp, c_long, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp]),
    "gpmpc_struct_layout": (c_long, [c_int, c_int]),
    "gpmpc_events_create": (c_int, [c_int, C.POINTER(C.c_void_p)]),
    "gpmpc_events_destroy": (c_int, [c_int, C.POINTER(C.c_void_p)]),
    "gpmpc_events_elapsed_ms": (c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_float)]),
    "gpmpc_mpc_update_3dof": (c_int, [c_int, c_int, c_int, c_dp, c_dp, c_dp, c_long, c_dp, c_dp, c_dp, c_dp, c_dp,
                                      c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp]),
}

EXPORTED_SYMBOLS = tuple(_PROTOS)


class GpmpcError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m gp_mpc_rocket_landing_b200.build` "
            "(the CUDA library is the product; there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise GpmpcError(f"{what} failed ({rc}): {lib.gpmpc_last_error().decode()}")


def require_cuda():
    import torch
    if not torch.cuda.is_available():
        raise GpmpcError("gp_mpc_rocket_landing_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch


def ptr(t):
    """Device pointer of a CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "expected a contiguous CUDA tensor"
    return t.data_ptr()


def stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


def call(name: str, *args) -> None:
    check(getattr(lib, name)(*args), name)
