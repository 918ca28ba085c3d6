"""``IDAKLUSolver``: the pybamm-side drop-in.

A ``pybamm.BaseSolver`` subclass with the constructor, options, ``solve(model, t_eval, inputs=[{...}, ...])`` contract and
result objects of the reference ``pybamm.IDAKLUSolver`` (packages/pybamm/src/pybamm/solvers/idaklu_solver.py:62-1013),
whose native back end is the B200 batched solver of this package instead of pybammsolvers' IDAS + KLU:

=====================================  =====================================================================================
reference                               here
=====================================  =====================================================================================
``set_up`` (idaklu_solver.py:366-643)   no CasADi: the discretised model is lowered to a scalar DAG (``bundle_from_pybamm``),
                                        differentiated, compiled to an sm_100a model image (``build.build_model``) and loaded
                                        through the C ABI (``include/pbgpu.h``)
``_set_consistent_initialization``      nothing on the host: y0(p), the algebraic IC Newton and ydot0 run inside the kernel
(:892-965), ``_rhs_dot_...`` (:967)     (``pbm_ic``, ``pb_consistent_init``); ``model.y0_list`` is still filled for BaseSolver
``_integrate`` (:665-718)               one ``pbgpu_solve`` call for the whole list of input sets
``_post_process_solution`` (:734-894)   same ``pybamm.Solution(t, y.T, model, inputs, t_event, y_event, termination, all_yps=...,
                                        all_t_evals=...)`` construction, ``closest_event_idx`` from the kernel
=====================================  =====================================================================================

``BaseSolver.solve`` / ``step`` themselves are inherited untouched.  Importing this module needs pybamm; nothing else in
the package does.  There is no CPU fallback: without a usable GPU ``solve`` raises ``pybamm.SolverError``.
"""
from __future__ import annotations

import itertools
import warnings

import numpy as np
import pybamm

from . import build
from .bundle import bundle_from_pybamm
from .solver import IDA_ROOT_RETURN, BatchedIDASolver, SolverError, sundials_error_message

_UNSET = object()
_REBUILDABLE_STATE_KEYS = ("_setup", "_model_set_up")


def _flatten_inputs(inputs_dict):
    """``{name: value}`` -> 1-D float array in dict-key order (idaklu_solver.py:19-23)."""
    if not inputs_dict:
        return np.zeros(0)
    return np.concatenate([np.asarray(v, dtype=np.float64).reshape(-1) for v in inputs_dict.values()])


class IDAKLUSolver(pybamm.BaseSolver):
    """Batched variable-order BDF (IDA-style) DAE solver on B200 GPUs behind ``pybamm.IDAKLUSolver``'s interface.

    Parameters are those of the reference (idaklu_solver.py:240-252): ``rtol``, ``atol``, ``root_method``, ``root_tol``,
    ``extrap_tol``, ``on_extrapolation``, ``output_variables``, ``on_failure``, ``options`` (same names and defaults,
    :286-325; the ones that select SUNDIALS linear solvers / threading are accepted and have no effect),
    ``store_first_last``.  Extra, keyword-only: ``devices`` (CUDA ordinals the list of input sets is sharded over,
    default ``[0]``).
    """

    def __init__(self, rtol=1e-4, atol=1e-6, root_method=_UNSET, root_tol=1e-6, extrap_tol=None, on_extrapolation=None,
                 output_variables=None, on_failure=None, options=None, store_first_last=False, *, devices=None,
                 _backend_factory=None):
        self.output_variables = [] if output_variables is None else output_variables
        self._options = self._combine_options(options)
        # the algebraic initial conditions are solved inside the kernel (calc_ic); with calc_ic off the reference falls back
        # to a Python root finder (idaklu_solver.py:258-262), which needs CasADi: not available on this path
        if root_method is _UNSET:
            root_method = None
        if root_method is not None:
            raise pybamm.SolverError("the B200 IDAKLUSolver initialises algebraic states inside the kernel: root_method must be None")
        super().__init__(method="ida", rtol=rtol, atol=atol, root_method=None, root_tol=root_tol, extrap_tol=extrap_tol,
                         output_variables=output_variables, on_extrapolation=on_extrapolation, on_failure=on_failure,
                         store_first_last=store_first_last)
        self.name = "IDA KLU solver (B200 batched)"
        self._supports_interp = True
        self._supports_t_eval_discontinuities = True
        self._devices = list(devices) if devices is not None else [0]
        self._backend_factory = _backend_factory
        self._setup = None
        pybamm.citations.register("Hindmarsh2000")
        pybamm.citations.register("Hindmarsh2005")

    # -- options (idaklu_solver.py:279-363) -----------------------------------------------------------------------------
    def _combine_options(self, user_options):
        user_options = user_options or {}
        default_options = {
            "print_stats": False, "compile": False, "jacobian": "sparse", "preconditioner": "BBDP",
            "precon_half_bandwidth": 5, "precon_half_bandwidth_keep": 5, "num_threads": 1,
            "num_solvers": user_options.get("num_threads", 1), "linear_solver": "SUNLinSol_KLU",
            "linsol_max_iterations": 5, "epsilon_linear_tolerance": 0.05, "increment_factor": 1.0,
            "linear_solution_scaling": True, "silence_sundials_errors": False, "max_order_bdf": 5,
            "max_num_steps": 100000, "dt_init": 0.0, "dt_min": 0.0, "dt_max": 0.0, "max_error_test_failures": 10,
            "max_nonlinear_iterations": 40, "max_convergence_failures": 100, "nonlinear_convergence_coefficient": 0.33,
            "suppress_algebraic_error": False, "hermite_interpolation": True, "hermite_reduction_factor": 1.0,
            "nonlinear_convergence_coefficient_ic": 0.0033, "max_num_steps_ic": 50, "max_num_jacobians_ic": 40,
            "max_num_iterations_ic": 100, "max_linesearch_backtracks_ic": 5, "linesearch_off_ic": False,
            "init_all_y_ic": False, "calc_ic": True, "newton_step_tol": 1e-4, "newton_mode": "auto",
            "num_steps_no_progress": 0, "t_no_progress": 0.0,
        }
        unknown = [k for k in user_options if k not in default_options and k != "max_saved_steps"]
        if unknown:
            raise pybamm.SolverError(f"unknown IDAKLU option(s): {unknown}")
        options = default_options | user_options
        if options["hermite_reduction_factor"] > 1.0:
            raise pybamm.SolverError("hermite_reduction_factor > 1 (knot reduction) is not implemented on the B200 path")
        if options["hermite_reduction_factor"] < 1.0:
            raise pybamm.SolverError("hermite_reduction_factor must be >= 1.0.")
        if not isinstance(options["compile"], bool):
            raise pybamm.SolverError("compile must be a bool")
        if options["num_solvers"] > options["num_threads"]:
            raise pybamm.SolverError("Number of solvers must be less than or equal to the number of threads")
        return options

    @property
    def options(self):
        return self._options

    @property
    def _internal_initialisation(self) -> bool:
        return bool(self._options["calc_ic"])

    def _check_atol_type(self, atol, model):
        if isinstance(atol, float):
            return np.full(model.len_rhs_and_alg, atol)
        if isinstance(atol, np.ndarray):
            return atol
        raise pybamm.SolverError("Absolute tolerances must be a numpy array or float")

    # -- pickling drops the native state (idaklu_solver.py:645-659) ----------------------------------------------------
    def __getstate__(self):
        state = self.__dict__.copy()
        for key in _REBUILDABLE_STATE_KEYS:
            state.pop(key, None)
        return state

    def __setstate__(self, d):
        self.__dict__.update(d)
        self._model_set_up = {}
        self._setup = None

    # -- set-up: lower + compile instead of building CasADi functions --------------------------------------------------
    def set_up(self, model, inputs=None, t_eval=None, ics_only=False):
        if isinstance(inputs, dict):
            inputs = [inputs]
        inputs = inputs or [{}]
        pybamm.logger.info("Start solver set-up")
        self._check_and_prepare_model_inplace(model)
        if getattr(model, "calculate_sensitivities", None):
            raise pybamm.SolverError("forward sensitivities are not implemented on the B200 IDAKLUSolver")
        model.calculate_sensitivities = []
        if not model.is_standard_form_dae:
            raise pybamm.SolverError("the B200 IDAKLUSolver needs a standard-form DAE (identity mass block)")
        if ics_only and self._setup is not None and self._setup["model"] is model:
            # new initial conditions for the same equations (BaseSolver.solve, base_solver.py:985-990): only y0(p) changes,
            # and it is recompiled with the rest; the residual / Jacobian lowering is cached by the bundle digest
            pass

        self._set_up_discontinuities(model, t_eval)
        missing = [v for v in self.output_variables if v not in model.variables]
        if missing:
            raise pybamm.SolverError(f"output variable(s) {missing} are not variables of model '{model.name}'")
        bundle = bundle_from_pybamm(model, inputs=inputs[0], variables=list(self.output_variables), name="pb_model")
        bundle.name = self._bundle_name(model, bundle)
        # the static pivot order of the sparse LU is chosen from sample matrices at the first input set (codegen.sample_matrices)
        bundle.meta["nominal_inputs"] = [float(x) for x in bundle.flatten_inputs([inputs[0]])[0]] if bundle.n_inputs else []
        atol = self._check_atol_type(getattr(model, "atol", self.atol), model)
        opts = {k: v for k, v in self._options.items() if k in BatchedIDASolver.FRONT_END_OPTIONS or k in _DEVICE_OPTIONS}
        if self._backend_factory is not None:
            backend = self._backend_factory(bundle, rtol=self.rtol, atol=atol, options=opts, output_variables=list(self.output_variables))
        else:
            try:
                backend = BatchedIDASolver(bundle, rtol=self.rtol, atol=atol, options=opts,
                                           output_variables=list(self.output_variables), devices=self._devices,
                                           on_failure="ignore", image_path=build.build_model(bundle))
                backend._ensure()
            except SolverError as e:
                raise pybamm.SolverError(str(e)) from None
        self._setup = {"model": model, "bundle": bundle, "backend": backend, "num_of_events": bundle.n_events,
                       "event_names": list(bundle.event_names)}
        # attributes BaseSolver.solve reads after set_up
        model.terminate_events_eval = [self._event_callable(bundle, k) for k in range(bundle.n_events)]
        model.interpolant_extrapolation_events_eval = []
        self._set_initial_conditions(model, 0.0, inputs)
        self.computed_var_fcns = {}
        pybamm.logger.info("Finish solver set-up")

    @staticmethod
    def _bundle_name(model, bundle):
        """Image name: the model's name plus a digest of the lowered equations, so that two models called alike do not
        share a compiled image and an unchanged model finds its image again (build.build_model caches by source digest)."""
        import hashlib

        h = hashlib.sha256()
        for k, v in sorted(bundle.graph.to_arrays().items()):
            h.update(k.encode())
            h.update(np.ascontiguousarray(v).tobytes())
        for a in (bundle.rhs_alg, bundle.ic, bundle.events, bundle.jac_nodes):
            h.update(np.ascontiguousarray(a).tobytes())
        h.update(repr((bundle.input_names, bundle.input_sizes, sorted(bundle.variables))).encode())
        stem = "".join(c if c.isalnum() else "_" for c in str(model.name))[:24].strip("_").lower()
        return f"{stem}_{h.hexdigest()[:10]}"

    def _set_up_discontinuities(self, model, t_eval):
        """The time-discontinuity part of BaseSolver._set_up_events (base_solver.py:560-625): constant ``t < X`` /
        ``t % X`` breakpoints become extra t_eval stops, the rest become DISCONTINUITY events."""
        from pybamm.expression_tree.binary_operators import _Heaviside

        tf = np.max(t_eval) if t_eval is not None else None
        t_discon_symbols = []
        for symbol in itertools.chain(model.concatenated_rhs.pre_order(), model.concatenated_algebraic.pre_order()):
            if isinstance(symbol, _Heaviside):
                if symbol.right == pybamm.t:
                    expr = symbol.left
                elif symbol.left == pybamm.t:
                    expr = symbol.right
                else:
                    continue
                if expr.is_constant():
                    t_discon_symbols.append((symbol, expr, None))
                else:
                    model.events.append(pybamm.Event(str(symbol), expr, pybamm.EventType.DISCONTINUITY))
            elif isinstance(symbol, pybamm.Modulo) and symbol.left == pybamm.t:
                expr = symbol.right
                num_events = 200 if tf is None else (tf // expr.value)
                if expr.is_constant():
                    t_discon_symbols.append((symbol, expr, num_events))
                else:
                    for i in np.arange(num_events):
                        model.events.append(pybamm.Event(str(symbol), expr * pybamm.Scalar(i + 1), pybamm.EventType.DISCONTINUITY))
        model.t_discon_constant_symbols = t_discon_symbols
        model.discontinuity_events_eval = [e for e in model.events if e.event_type == pybamm.EventType.DISCONTINUITY]

    @staticmethod
    def _event_callable(bundle, k):
        node = [int(bundle.events[k])]

        def event(t, y, inputs):
            p = _flatten_inputs(inputs) if isinstance(inputs, dict) else np.asarray(inputs, dtype=np.float64).reshape(-1)
            return float(bundle.graph.evaluate(node, float(t), np.asarray(y, dtype=np.float64).reshape(-1), p)[0])

        return event

    # -- initial conditions: one vectorised evaluation of y0(p) for BaseSolver's bookkeeping ---------------------------
    def _set_initial_conditions(self, model, time, inputs):
        bundle = self._setup["bundle"]
        P = self._stack_inputs(inputs, bundle)
        if bundle.n_inputs and len(inputs) > 1:
            y0 = bundle.graph.evaluate(bundle.ic, float(time), None, P)
        else:
            y0 = np.tile(bundle.graph.evaluate(bundle.ic, float(time), None, P[0] if len(P) else None), (len(inputs), 1))
        self._setup["y0_block"] = y0
        model.y0_list = [y0[i].reshape(-1, 1) for i in range(len(inputs))]
        model.y0S_list = None
        self._setup["y0_ids"] = [id(a) for a in model.y0_list]

    @staticmethod
    def _stack_inputs(inputs_list, bundle):
        if not bundle.n_inputs:
            return np.zeros((len(inputs_list), 0))
        try:
            return bundle.flatten_inputs(list(inputs_list))
        except ValueError as e:
            raise pybamm.SolverError(str(e)) from None

    def _set_consistent_initialization(self, model, time, inputs_list):
        # y0(p), the algebraic IC Newton and ydot0 = f_d(t0, y0) run inside the kernel (pbm_ic, pb_consistent_init):
        # nothing to do on the host; model.y0_list keeps the un-initialised guess, as with calc_ic in the reference
        return

    def _check_event_violation_on_initialisation(self, *args, **kwargs):
        if self._internal_initialisation:
            return
        return self._check_event_violation(*args, **kwargs)

    def _check_event_violation_post_solve(self, *args, **kwargs):
        if not self._internal_initialisation:
            return
        return self._check_event_violation(*args, **kwargs)

    def _check_event_violation(self, t_eval, model, y0, inputs_dict):
        evs = model.terminate_events_eval
        if not evs:
            return
        vals = np.array([ev(t_eval[0], y0, inputs_dict) for ev in evs])
        if vals.min() <= 0:
            names = [self._setup["event_names"][i] for i in np.where(vals <= 0)[0]]
            raise pybamm.SolverError(f"Events {names} are non-positive at initial conditions with inputs {inputs_dict}")

    # -- the batched integration ------------------------------------------------------------------------------------------
    def _integrate(self, model, t_eval, inputs_list=None, t_interp=None, nproc=None):
        inputs_list = inputs_list or [{}]
        setup = self._setup
        bundle, backend = setup["bundle"], setup["backend"]
        P = self._stack_inputs(inputs_list, bundle)
        # y0: the kernel evaluates y0(p) itself unless BaseSolver replaced the states (a later time segment)
        y0 = None
        if [id(a) for a in model.y0_list] != setup.get("y0_ids"):
            y0 = np.vstack([np.asarray(a, dtype=np.float64).reshape(1, -1) for a in model.y0_list])
        timer = pybamm.Timer()
        try:
            r = backend.solve_arrays(np.asarray(t_eval, dtype=np.float64), P,
                                     t_interp=None if t_interp is None or len(t_interp) == 0 else np.asarray(t_interp, dtype=np.float64),
                                     y0=y0)
        except SolverError as e:
            raise pybamm.SolverError(str(e)) from None
        integration_time = timer.time()
        return self._post_process_batch(r, model, integration_time, inputs_list, np.asarray(t_eval, dtype=np.float64))

    def _post_process_batch(self, r, model, integration_time, inputs_list, t_eval):
        """_post_process_solution (idaklu_solver.py:734-894) for every system of the batch."""
        n_states = model.len_rhs_and_alg
        outputs_only = bool(self.output_variables)
        n_t_all = np.minimum(r["n_t"], r["n_max"]).tolist()
        flags, ev_idx = r["flag"].tolist(), r["event_idx"].tolist()
        t_all, y_all, yp_all, yterm = r["t"], r["y"], r["yp"], r["y_term"]
        sol_options = {"compile": self._options["compile"]}
        sols = []
        for i, inputs_dict in enumerate(inputs_list):
            n, flag = n_t_all[i], flags[i]
            t = t_all[i, :n]
            if outputs_only:
                y_out_T = np.zeros((0, n * n_states))
                y_event = yterm[i]
            else:
                y_out_T = y_all[i, :n].T          # the transpose *view* of the time-major block (idaklu_solver.py:747)
                y_event = y_all[i, n - 1]
            if n == 1 and flag == IDA_ROOT_RETURN:
                self._check_event_violation_post_solve(t_eval, model, y_event, inputs_dict)
            if flag == IDA_ROOT_RETURN:
                termination = "event"
            elif flag >= 0:
                termination = "final time"
            else:
                termination = "failure"
                msg = sundials_error_message(flag)
                if self._on_failure == "warn":
                    warnings.warn(msg + ", returning a partial solution.", stacklevel=2)
                elif self._on_failure == "error":
                    raise pybamm.SolverError(msg)
            yp = yp_all[i, :n].T if yp_all is not None else None
            te = t_eval
            if t[-1] != te[-1]:
                te = te[: np.searchsorted(te, t[-1]) + 1].copy()
                te[-1] = t[-1]
            newsol = pybamm.Solution(t, y_out_T, model, inputs_dict, np.array([t[-1]]), np.asarray(y_event)[:, np.newaxis],
                                     termination, all_sensitivities={}, all_yps=yp, all_t_evals=te,
                                     variables_returned=outputs_only, options=sol_options)
            if flag == IDA_ROOT_RETURN and ev_idx[i] >= 0:
                newsol.closest_event_idx = int(ev_idx[i])
            newsol.integration_time = integration_time
            if outputs_only:
                self._attach_output_variables(newsol, model, y_all[i, :n])
            sols.append(newsol)
        return sols

    def _attach_output_variables(self, newsol, model, data):
        """Variables computed in the solver (idaklu_solver.py:845-893): one ProcessedVariableComputed per name."""
        bundle = self._setup["bundle"]
        off = 0
        for var in bundle.variables:
            n = len(bundle.variables[var])
            if var in self.output_variables:
                sym = model.get_processed_variable_or_event(var)
                newsol._variables[var] = pybamm.ProcessedVariableComputed([sym], [_dense_shape_function(n)], [data[:, off:off + n]],
                                                                          newsol, time_indep=False)
            off += n


# option names the device integrator reads (pybamm_b200.host.OPTION_DEFAULTS)
from .host import OPTION_DEFAULTS as _OPT  # noqa: E402

_DEVICE_OPTIONS = set(_OPT)


def _dense_shape_function(n):
    """ProcessedVariableComputed only asks its ``base_variables_casadi`` entry for the sparsity of the variable
    (processed_variable_computed.py:_unroll_nnz); the compiled output variables are dense, so a real ``casadi.Function``
    with a dense n x 1 output is all it needs (built only where CasADi is installed)."""
    try:
        import casadi

        t, y, p = casadi.MX.sym("t"), casadi.MX.sym("y"), casadi.MX.sym("p")
        f = casadi.Function("var", [t, y, p], [casadi.MX.zeros(n, 1) + t])
        if hasattr(f, "sparsity_out"):
            return f
    except Exception:
        pass
    return None
