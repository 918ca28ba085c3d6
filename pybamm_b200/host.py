"""ctypes mirror of include/pbgpu.h and the host-side helpers shared by the solver front ends.

The time-grid canonicalisation follows IDAKLUSolverGroup::solve
(packages/pybammsolvers/src/pybammsolvers/idaklu_source/IDAKLUSolverGroup.cpp:25-43): t_eval sorted +
unique; t_interp sorted + unique with the t_eval values removed.
"""
from __future__ import annotations

import ctypes as C

import numpy as np


class Options(C.Structure):
    _fields_ = [
        ("rtol", C.c_double),
        ("max_order_bdf", C.c_int),
        ("max_num_steps", C.c_longlong),
        ("max_error_test_failures", C.c_int),
        ("max_nonlinear_iterations", C.c_int),
        ("max_convergence_failures", C.c_int),
        ("nonlinear_convergence_coefficient", C.c_double),
        ("nonlinear_convergence_coefficient_ic", C.c_double),
        ("newton_step_tol", C.c_double),
        ("max_num_iterations_ic", C.c_int),
        ("max_linesearch_backtracks_ic", C.c_int),
        ("linear_solution_scaling", C.c_int),
        ("calc_ic", C.c_int),
        ("dt_init", C.c_double),
        ("dt_max", C.c_double),
        ("num_steps_no_progress", C.c_int),
        ("t_no_progress", C.c_double),
    ]


class ModelDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int), ("n_states", C.c_int), ("n_diff", C.c_int), ("n_inputs", C.c_int),
        ("n_events", C.c_int), ("nnz_jac", C.c_int), ("nnz_lu", C.c_int), ("n_vars", C.c_int),
        ("team_size", C.c_int), ("smem_bytes", C.c_int), ("n_tmp", C.c_int), ("lu_levels", C.c_int * 3),
        ("flops", C.c_longlong * 6), ("scratch_doubles_per_team", C.c_longlong), ("name", C.c_char * 64),
    ]


class Result(C.Structure):
    _fields_ = [
        ("n_max", C.c_int), ("n_ret", C.c_int), ("t", C.c_void_p), ("y", C.c_void_p), ("n_t", C.c_void_p),
        ("flag", C.c_void_p), ("y_term", C.c_void_p), ("stats", C.c_void_p), ("yp", C.c_void_p),
        ("event_idx", C.c_void_p), ("t_event", C.c_void_p), ("kernel_ms", C.c_float), ("total_ms", C.c_float),
    ]


class ObserveArgs(C.Structure):
    _fields_ = [
        ("n_sys", C.c_int), ("n_max", C.c_int), ("n_obs", C.c_int), ("n_t", C.c_void_p), ("t", C.c_void_p), ("y", C.c_void_p),
        ("yp", C.c_void_p), ("inputs", C.c_void_p), ("t_obs", C.c_void_p), ("out", C.c_void_p), ("scratch", C.c_void_p),
    ]


class Schedule(C.Structure):
    _fields_ = [
        ("n_steps", C.c_int), ("n_over", C.c_int), ("over_slot", C.c_void_p), ("over_val", C.c_void_p),
        ("duration", C.c_void_p), ("n_pts", C.c_void_p), ("stop_on_final_time", C.c_void_p), ("experiment_event", C.c_int),
    ]


class LaunchArgs(C.Structure):
    _fields_ = [
        ("n_sys", C.c_int), ("n_eval", C.c_int), ("n_interp", C.c_int), ("n_max", C.c_int), ("save_vars", C.c_int),
        ("n_ret", C.c_int), ("t_eval", C.c_void_p), ("t_interp", C.c_void_p), ("inputs", C.c_void_p),
        ("y0", C.c_void_p), ("yp0", C.c_void_p), ("atol", C.c_void_p), ("t_out", C.c_void_p), ("y_out", C.c_void_p),
        ("n_out", C.c_void_p), ("flag", C.c_void_p), ("y_term", C.c_void_p), ("stats", C.c_void_p),
        ("yp_out", C.c_void_p), ("event_idx", C.c_void_p), ("t_event", C.c_void_p),
        ("queue", C.c_void_p), ("scratch", C.c_void_p), ("order", C.c_void_p), ("opt", Options),
        ("sched", Schedule), ("t_start", C.c_double), ("step_end", C.c_void_p), ("step_term", C.c_void_p),
    ]


# IDAKLUSolver option names -> defaults (idaklu_solver.py:286-325) for the fields the device path uses
OPTION_DEFAULTS = dict(
    max_order_bdf=5, max_num_steps=100000, max_error_test_failures=10, max_nonlinear_iterations=40,
    max_convergence_failures=100, nonlinear_convergence_coefficient=0.33,
    nonlinear_convergence_coefficient_ic=0.0033, newton_step_tol=1e-4, max_num_iterations_ic=100,
    max_linesearch_backtracks_ic=5, linear_solution_scaling=True, calc_ic=True, dt_init=0.0, dt_max=0.0,
    num_steps_no_progress=0, t_no_progress=0.0,
)


def default_options(rtol=1e-4, **overrides) -> Options:
    o = Options()
    o.rtol = float(rtol)
    vals = dict(OPTION_DEFAULTS)
    for k, v in overrides.items():
        if k in vals:
            vals[k] = v
    for k, v in vals.items():
        cur = getattr(o, k)
        setattr(o, k, type(cur)(v))
    return o


def canonical_time_grids(t_eval, t_interp):
    t_eval = np.unique(np.asarray(t_eval, dtype=np.float64).reshape(-1))
    if t_interp is None:
        t_interp = np.zeros(0)
    t_interp = np.unique(np.asarray(t_interp, dtype=np.float64).reshape(-1))
    t_interp = np.ascontiguousarray(np.setdiff1d(t_interp, t_eval))
    if len(t_eval) < 2:
        raise ValueError("t_eval must have at least 2 entries")
    if len(t_interp):
        if t_interp[0] < t_eval[0]:
            raise ValueError(f"t_interp values must be greater than the smallest t_eval value: {t_eval[0]:f}")
        if t_interp[-1] > t_eval[-1]:
            raise ValueError(f"t_interp values must be less than the greatest t_eval value: {t_eval[-1]:f}")
    return np.ascontiguousarray(t_eval), t_interp
