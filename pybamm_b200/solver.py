"""Host front end of the B200 batched DAE solver (no pybamm import needed).

``BatchedIDASolver`` mirrors the data contract of ``IDAKLUSolver._integrate`` /
``IDAKLUSolverGroup.solve`` (packages/pybamm/src/pybamm/solvers/idaklu_solver.py:665-718,
packages/pybammsolvers/src/pybammsolvers/idaklu_source/IDAKLUSolverGroup.cpp:6-170): one blocking
call per batch, inputs ``(n_sys, n_p)`` ordered by sorted input name, results per system as
``t (n_t,)``, ``y (n_t, n_ret)`` time-major, ``flag``, ``y_term``.  It reaches CUDA through the C ABI
of ``include/pbgpu.h`` with ctypes.  There is no CPU fallback: constructing the solver without a
usable GPU (or without the compiled libraries) raises ``SolverError``.

``pybamm_b200.idaklu_solver.IDAKLUSolver`` (a ``pybamm.BaseSolver`` subclass, importable where pybamm is)
drives this class from ``solve(model, t_eval, inputs=[...])`` and returns ``pybamm.Solution`` objects.
"""
from __future__ import annotations

import collections.abc
import ctypes as C
import operator
import os
import warnings

import numpy as np

from . import build
from .bundle import ModelBundle
from .host import ModelDesc, Options, Result, canonical_time_grids, default_options, OPTION_DEFAULTS

IDA_ROOT_RETURN = 2

# flag -> message table of the reference (idaklu_source/sundials_error_handler.hpp:11-61), the codes this path can return
IDA_MESSAGES = {
    -1: "IDA_TOO_MUCH_WORK: The solver took mxstep internal steps but could not reach tout.",
    -2: "IDA_TOO_MUCH_ACC: The solver could not satisfy the accuracy demanded by the user for some internal step.",
    -3: "IDA_ERR_FAIL: Error test failures occurred too many times during one internal time step or minimum step size was reached.",
    -4: "IDA_CONV_FAIL: Convergence test failures occurred too many times during one internal time step or minimum step size was reached.",
    -6: "IDA_LSETUP_FAIL: The linear solver's setup function failed in an unrecoverable manner.",
    -22: "IDA_ILL_INPUT: One of the function inputs is illegal.",
}

STAT_NAMES = ("nsteps", "nrevals", "nlinsetups", "netfails", "nniters", "nncfails", "njevals", "nge")


class SolverError(Exception):
    """Same role as pybamm.SolverError (raised for argument errors and on_failure='error')."""


def sundials_error_message(flag: int) -> str:
    return IDA_MESSAGES.get(int(flag), f"IDA failure (flag: {int(flag)})")


class Solution:
    """Minimal stand-in for pybamm.Solution holding one system's result.

    Field meaning follows the arguments ``_post_process_solution`` passes to ``pybamm.Solution``
    (idaklu_solver.py:809-822): ``t``, ``y`` with shape (n_states, n_t) (the transpose view of the
    time-major block), ``t_event``, ``y_event``, ``termination``.
    """

    __slots__ = ("t", "_rows", "all_inputs", "flag", "y_event", "_stats", "_variables", "closest_event_idx", "termination",
                 "_yp_rows")

    def __init__(self, t, y_rows, inputs, flag, y_term, stats, variables, event_names, closest_event_idx):
        # cheap on purpose: a batch builds ten thousand of these; everything derivable is a property
        self.t = t
        self._rows = y_rows
        self.all_inputs = [inputs]
        self.flag = flag
        self.y_event = y_term
        self._stats = stats
        self._variables = variables
        self.closest_event_idx = closest_event_idx
        self._yp_rows = None
        if flag == IDA_ROOT_RETURN:
            self.termination = "event: " + (event_names[closest_event_idx] if closest_event_idx is not None else "?")
        elif flag >= 0:
            self.termination = "final time"
        else:
            self.termination = "failure"

    @property
    def t_event(self):
        return np.array([self.t[-1]]) if len(self.t) else np.array([])

    @property
    def yp(self):
        """dy/dt at the stored points, shape (n_states, n_t) (all_yps of idaklu_solver.py:819), or None."""
        return None if self._yp_rows is None else self._yp_rows.T

    @property
    def stats(self):
        return dict(zip(STAT_NAMES, (int(s) for s in self._stats))) if self._stats is not None else {}

    @property
    def y(self):
        if self._variables is not None:
            raise KeyError("full state was not stored (output_variables mode); use solution[name]")
        return self._rows.T

    def __getitem__(self, name):
        if self._variables is None:
            raise KeyError(f"variable '{name}' was not requested through output_variables")
        sl = self._variables[name]
        return self._rows[:, sl]


class SolutionBatch(collections.abc.Sequence):
    """What ``solve(inputs=[...])`` returns: a read-only sequence of ``Solution`` objects backed by the batch arrays.

    Behaves like the list the reference returns (``len``, indexing, slicing, iteration), but a ``Solution`` -- a handful of
    views into the arrays of the whole batch -- is only made when it is asked for, and then kept: with ten thousand input
    sets the eager loop was 35 ms of Python, a tenth of an 8-GPU solve.  ``arrays`` is the dict
    of ``solve_arrays`` for callers that want the batch as a whole."""

    __slots__ = ("arrays", "_inputs", "_variables", "_names", "_n_t", "_flags", "_closest", "_made")

    def __init__(self, arrays, inputs_list, variables, event_names):
        self.arrays = arrays
        self._inputs = inputs_list
        self._variables = variables
        self._names = event_names
        self._n_t = np.minimum(arrays["n_t"], arrays["n_max"]).tolist()
        self._flags = arrays["flag"].tolist()
        self._closest = arrays["event_idx"].tolist()   # found by the kernel at the root (argmin |g|, idaklu_solver.py:826-834)
        self._made = {}

    def __len__(self):
        return len(self._inputs)

    def _make(self, i):
        sol = self._made.get(i)
        if sol is None:
            r, n, flag = self.arrays, self._n_t[i], self._flags[i]
            idx = self._closest[i] if (flag == IDA_ROOT_RETURN and self._closest[i] >= 0) else None
            sol = Solution(r["t"][i, :n], r["y"][i, :n], self._inputs[i], flag, r["y_term"][i],
                           r["stats"][i] if r["stats"] is not None else None, self._variables, self._names, idx)
            if r["yp"] is not None:
                sol._yp_rows = r["yp"][i, :n]
            self._made[i] = sol
        return sol

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self._make(k) for k in range(*i.indices(len(self)))]
        i = operator.index(i)
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError("solution index out of range")
        return self._make(i)

    def __iter__(self):
        return (self._make(i) for i in range(len(self)))


def _pinned_empty(shape, dtype):
    """Page-locked host array (through torch) so that H2D/D2H copies run at full PCIe speed."""
    try:
        import torch

        if torch.cuda.is_available():
            tdt = {np.float64: torch.float64, np.int32: torch.int32, np.int64: torch.int64}[np.dtype(dtype).type]
            t = torch.empty(tuple(int(s) for s in shape), dtype=tdt, pin_memory=True)
            return t.numpy(), t
    except Exception:
        pass
    return np.empty(shape, dtype=dtype), None


class _Lib:
    _inst = None

    @classmethod
    def get(cls):
        if cls._inst is None:
            path = os.path.join(build.BUILD, "libpbgpu.so")
            if not os.path.exists(path):
                try:
                    path = build.build_api()
                except Exception as e:
                    raise SolverError(f"libpbgpu.so is not built and cannot be built here: {e}") from None
            try:
                lib = C.CDLL(path)
            except OSError as e:
                raise SolverError(f"cannot load {path}: {e}") from None
            lib.pbgpu_last_error.restype = C.c_char_p
            vp = C.c_void_p
            lib.pbgpu_model_load.argtypes = [C.c_char_p, C.POINTER(vp)]
            lib.pbgpu_model_describe.argtypes = [vp, C.POINTER(ModelDesc)]
            lib.pbgpu_model_unload.argtypes = [vp]
            lib.pbgpu_solver_create.argtypes = [vp, vp, C.POINTER(Options), C.c_int, C.POINTER(vp)]
            lib.pbgpu_solver_destroy.argtypes = [vp]
            lib.pbgpu_result_capacity.argtypes = [C.c_int, C.c_int]
            lib.pbgpu_solve.argtypes = [vp, C.c_int, vp, C.c_int, vp, C.c_int, vp, vp, vp, C.c_int, C.POINTER(Result)]
            lib.pbgpu_solve_device.argtypes = [vp, C.c_int, vp, C.c_int, vp, C.c_int, vp, vp, vp, C.c_int,
                                               C.POINTER(Result), vp, C.c_int]
            lib.pbgpu_solve_device_noalloc.argtypes = [vp, C.c_int, vp, C.c_int, vp, C.c_int, vp, vp, vp, C.c_int,
                                               C.POINTER(Result), vp, C.c_int]
            lib.pbgpu_model_eval.argtypes = [vp, C.c_int, C.c_int, vp, vp, C.c_double, C.c_double, vp, vp, C.c_int]
            lib.pbgpu_solve_experiment.argtypes = [vp, C.c_int, C.c_double, vp, vp, vp, C.c_int, C.POINTER(Result), vp, vp]
            lib.pbgpu_observe.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, C.c_int, vp]
            lib.pbgpu_observe_device.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, C.c_int, vp, vp, C.c_int]
            cls._inst = lib
        return cls._inst


def _err(lib):
    return lib.pbgpu_last_error().decode(errors="replace")


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class BatchedIDASolver:
    """Batched IDA-style BDF solver on one B200.

    Parameters mirror ``pybamm.IDAKLUSolver`` (idaklu_solver.py:240-252): ``rtol``, ``atol`` (float or
    (n_states,) array), ``options`` (same option names; options that select other linear solvers or
    threading are accepted and ignored), ``output_variables`` (names compiled into the bundle),
    ``on_failure`` in {"warn", "error", "ignore"}.
    """

    IGNORED_OPTIONS = {"print_stats", "compile", "jacobian", "preconditioner", "precon_half_bandwidth",
                       "precon_half_bandwidth_keep", "num_threads", "num_solvers", "linear_solver",
                       "linsol_max_iterations", "epsilon_linear_tolerance", "increment_factor",
                       "silence_sundials_errors", "suppress_algebraic_error",
                       "hermite_reduction_factor", "max_num_steps_ic", "max_num_jacobians_ic", "linesearch_off_ic",
                       "init_all_y_ic", "newton_mode", "dt_min"}
    # options of this front end itself: hermite_interpolation (store dy/dt with the adaptive steps, idaklu_solver.py:318),
    # max_saved_steps (first guess of the adaptive-step capacity; the block is re-sized and the affected rows re-solved
    # when a system takes more steps: the reference grows its arrays instead, IDAKLUSolverOpenMP.inl:608-612)
    FRONT_END_OPTIONS = {"hermite_interpolation": True, "max_saved_steps": 4096, "repivot": True}

    def __init__(self, model, rtol=1e-4, atol=1e-6, options=None, output_variables=None, device=0,
                 on_failure="warn", image_path=None, devices=None, reuse_buffers=False):
        self.bundle: ModelBundle = model if isinstance(model, ModelBundle) else build.load_bundle(model)
        self.rtol = float(rtol)
        self.atol = atol
        self.on_failure = on_failure or "warn"
        if self.on_failure not in ("warn", "error", "ignore"):
            raise SolverError("on_failure must be 'warn', 'error' or 'ignore'")
        self.output_variables = list(output_variables or [])
        for v in self.output_variables:
            if v not in self.bundle.variables:
                raise SolverError(f"output variable '{v}' is not compiled into model bundle '{self.bundle.name}'")
        options = dict(options or {})
        unknown = [k for k in options if k not in OPTION_DEFAULTS and k not in self.IGNORED_OPTIONS
                   and k not in self.FRONT_END_OPTIONS]
        if unknown:
            raise SolverError(f"unknown IDAKLU option(s): {unknown}")
        if options.get("num_solvers", 1) > options.get("num_threads", options.get("num_solvers", 1)):
            raise SolverError("Number of solvers must be less than or equal to the number of threads")
        if int(options.get("num_steps_no_progress", 0)) > 32:
            raise SolverError("num_steps_no_progress is limited to 32 on this path")
        self.options = options
        # devices: CUDA ordinals the batch is sharded over inside this process (one solver handle and one host
        # thread per device, no collective: SURVEY.md section 8e).  Default: the single ordinal ``device``.
        self.devices = [int(d) for d in devices] if devices is not None else [int(device)]
        if not self.devices:
            raise SolverError("devices must name at least one CUDA ordinal")
        self.device = self.devices[0]
        self.reuse_buffers = bool(reuse_buffers)
        self._image_path = image_path
        self._lib = None
        self._model = None
        self._solver = None
        self._solvers = []
        self.desc = None
        self.last_kernel_ms = None
        self.last_total_ms = None
        from .shard import CostModel

        self._cost_model = CostModel()
        # rows whose integration failed are re-solved once on an image whose static pivot order is also valid for the
        # matrices of the failed systems (see _repivot); "repivot": False in the options switches it off
        self.repivot = bool(options.get("repivot", True))
        self._fallback = None
        self._is_fallback = False

    # -- native handles ---------------------------------------------------------------------
    def _ensure(self):
        if self._solver is not None:
            return
        lib = _Lib.get()
        path = self._image_path or os.path.join(build.BUILD, f"pbgpu_model_{self.bundle.name}.so")
        if not os.path.exists(path):
            try:
                path = build.build_model(self.bundle)
            except Exception as e:
                raise SolverError(f"model image for '{self.bundle.name}' is not built and cannot be built here: {e}") from None
        mh = C.c_void_p()
        if lib.pbgpu_model_load(path.encode(), C.byref(mh)):
            raise SolverError(_err(lib))
        desc = ModelDesc()
        lib.pbgpu_model_describe(mh, C.byref(desc))
        if desc.n_states != self.bundle.n_states or desc.n_inputs != self.bundle.n_inputs:
            raise SolverError("model image does not match the bundle (rebuild with pybamm_b200.build.build_model)")
        atol = np.ascontiguousarray(np.broadcast_to(np.asarray(self.atol, dtype=np.float64), (desc.n_states,)).copy())
        opt = default_options(rtol=self.rtol, **self.options)
        handles = []
        for dev in self.devices:
            sh = C.c_void_p()
            if lib.pbgpu_solver_create(mh, _ptr(atol), C.byref(opt), dev, C.byref(sh)):
                msg = _err(lib)
                for h in handles:
                    lib.pbgpu_solver_destroy(h)
                raise SolverError(msg)
            handles.append(sh)
        self._lib, self._model, self._solver, self._solvers, self.desc = lib, mh, handles[0], handles, desc

    def close(self):
        if getattr(self, "_fallback", None) is not None:
            self._fallback.close()
            self._fallback = None
        if self._solver is not None:
            for h in self._solvers:
                self._lib.pbgpu_solver_destroy(h)
            self._lib.pbgpu_model_unload(self._model)
            self._solver = self._model = None
            self._solvers = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # pickling drops native state and rebuilds lazily (idaklu_solver.py:645-659)
    def __getstate__(self):
        d = dict(self.__dict__)
        d["_lib"] = d["_model"] = d["_solver"] = d["desc"] = None
        d["_solvers"] = []
        d["_pinned"] = {}
        return d

    def __setstate__(self, d):
        self.__dict__.update(d)

    # -- batch solve --------------------------------------------------------------------------
    def solve_arrays(self, t_eval, inputs, t_interp=None, y0=None, output_mode=None, want_stats=True, yp0=None,
                     want_yp=None):
        """Array-level entry: returns dict(t, y, yp, n_t, flag, y_term, stats, event_idx, t_event) of padded arrays.

        ``want_yp``: also return dy/dt at the stored points (full-state mode only).  Default: the reference's rule
        ``hermite_interpolation and adaptive steps and not output_variables`` (save_hermite, IDAKLUSolverOpenMP.inl:447-457)."""
        self._ensure()
        lib, D = self._lib, self.desc
        try:
            t_eval, t_interp = canonical_time_grids(t_eval, t_interp)
        except ValueError as e:
            raise SolverError(str(e)) from None
        inputs = np.asarray(inputs, dtype=np.float64)
        if inputs.ndim != 2:
            raise SolverError(f"inputs has wrong number of dimensions. Expected 2 but got {inputs.ndim}")
        n_sys = inputs.shape[0]
        if inputs.shape[1] != D.n_inputs:
            raise SolverError(f"inputs has wrong number of cols. Expected {D.n_inputs} but got {inputs.shape[1]}")
        inputs = np.ascontiguousarray(inputs)

        def state_block(a, what):
            if a is None:
                return None
            a = np.asarray(a, dtype=np.float64)
            if a.ndim != 2:
                raise SolverError(f"{what} has wrong number of dimensions. Expected 2 but got {a.ndim}")
            if a.shape[1] != D.n_states:
                raise SolverError(f"{what} has wrong number of cols. Expected {D.n_states} but got {a.shape[1]}")
            if a.shape[0] != n_sys:
                raise SolverError(f"{what} has wrong number of rows. Expected {n_sys} but got {a.shape[0]}")
            return np.ascontiguousarray(a)

        y0, yp0 = state_block(y0, "y0"), state_block(yp0, "yp0")
        if output_mode is None:
            output_mode = 1 if self.output_variables else 0
        n_ret = D.n_vars if output_mode == 1 else D.n_states
        adaptive = len(t_interp) == 0
        if want_yp is None:
            want_yp = bool(self.options.get("hermite_interpolation", True)) and adaptive
        want_yp = bool(want_yp) and output_mode == 0
        if not adaptive:
            n_max = lib.pbgpu_result_capacity(len(t_eval), len(t_interp))
        else:
            n_max = max(int(self.options.get("max_saved_steps", 4096)), len(t_eval) + 2)
        r = self._solve_batch(t_eval, t_interp, inputs, y0, yp0, output_mode, n_ret, n_max, want_stats, want_yp)
        over = np.nonzero(r["n_t"] > n_max)[0]
        if len(over):
            # adaptive-step mode: some systems took more steps than the block holds.  The kernel counted them all, so one
            # more pass over those rows with an exact capacity completes them.
            n_big = int(r["n_t"][over].max())
            sub = self._solve_batch(t_eval, t_interp, np.ascontiguousarray(inputs[over]),
                                    None if y0 is None else np.ascontiguousarray(y0[over]),
                                    None if yp0 is None else np.ascontiguousarray(yp0[over]), output_mode, n_ret, n_big,
                                    want_stats, want_yp)

            def grow(a):
                if a is None:
                    return None
                b = np.zeros((a.shape[0], n_big) + a.shape[2:], dtype=a.dtype)
                b[:, :a.shape[1]] = a
                return b

            for key in ("t", "y", "yp"):
                if r[key] is not None:
                    big = grow(np.asarray(r[key]))
                    big[over] = sub[key]
                    r[key] = big
            for key in ("n_t", "flag", "y_term", "stats", "event_idx", "t_event"):
                if r[key] is not None:
                    r[key] = np.array(r[key])
                    r[key][over] = sub[key]
            r["n_max"] = n_big
            r["_keep"] = None
        bad = np.nonzero(np.asarray(r["flag"]) < 0)[0]
        if len(bad) and self.repivot and not self._is_fallback and r["y_term"] is not None:
            self._repivot(r, bad, t_eval, t_interp, inputs, y0, yp0, output_mode, want_stats, want_yp)
        if want_stats and r["stats"] is not None:
            self._cost_model.fit(inputs, r["stats"][:, 0])
        return r

    def _repivot(self, r, bad, t_eval, t_interp, inputs, y0, yp0, output_mode, want_stats, want_yp):
        """Slow path for rows whose integration failed (flag < 0).

        The pivot order of the sparse LU is fixed per image from sample matrices taken around the initial state; KLU in
        the reference keeps its order between refactorisations too, but factorises afresh when a refactorisation turns out
        ill-conditioned.  The counterpart here: J - cj M of the failed systems at the states where they failed (evaluated
        on the device) joins the samples, a second image is generated and compiled for that order (a minute, once; cached
        by content), and the failed rows are solved again from the start on it.  Rows that fail again keep their flag."""
        import hashlib

        bad = bad[:16]                                    # a handful of failure states is enough to choose an order
        yb, pb_ = np.ascontiguousarray(r["y_term"][bad]), np.ascontiguousarray(inputs[bad])
        tb = np.asarray(r["t_event"])[bad] if r.get("t_event") is not None else np.zeros(len(bad))
        if not np.all(np.isfinite(yb)):
            return
        extra = []
        for cj in (1.0, 1e-3, 1e3):
            vals = np.vstack([self.eval("jac", yb[k:k + 1], pb_[k:k + 1], float(tb[k]), cj)[0] for k in range(len(bad))])
            extra += [v for v in vals if np.all(np.isfinite(v))]
        if not extra:
            return
        if self._fallback is None:
            tag = "_rp" + hashlib.sha256(np.round(np.asarray(extra), 6).tobytes()).hexdigest()[:10]
            try:
                img = build.build_model(self.bundle, tag=tag, extra_samples=extra)
            except Exception as e:                        # no order is valid for old and new samples alike
                warnings.warn(f"re-pivoting failed: {e}", stacklevel=3)
                return
            fb = BatchedIDASolver(self.bundle, rtol=self.rtol, atol=self.atol, options=dict(self.options),
                                  output_variables=self.output_variables, device=self.device, on_failure="ignore",
                                  image_path=img)
            fb._is_fallback = True
            self._fallback = fb
        rows = np.nonzero(np.asarray(r["flag"]) < 0)[0]
        sub = self._fallback.solve_arrays(t_eval, np.ascontiguousarray(inputs[rows]), t_interp=t_interp if len(t_interp) else None,
                                          y0=None if y0 is None else y0[rows], output_mode=output_mode, want_stats=want_stats,
                                          yp0=None if yp0 is None else yp0[rows], want_yp=want_yp)
        better = np.asarray(sub["flag"]) >= 0
        if not better.any():
            return
        n_keep = min(r["n_max"], sub["n_max"])
        for key in ("t", "y", "yp"):
            if r[key] is not None and sub[key] is not None:
                r[key] = np.array(r[key])
                r[key][rows[better], :n_keep] = np.asarray(sub[key])[better, :n_keep]
        for key in ("n_t", "flag", "y_term", "stats", "event_idx", "t_event"):
            if r[key] is not None and sub[key] is not None:
                r[key] = np.array(r[key])
                r[key][rows[better]] = np.asarray(sub[key])[better]
        r["n_t"][rows[better]] = np.minimum(r["n_t"][rows[better]], n_keep) if sub["n_max"] != r["n_max"] else r["n_t"][rows[better]]
        r["_keep"] = None
        r["repivoted_rows"] = rows[better]

    def _solve_batch(self, t_eval, t_interp, inputs, y0, yp0, output_mode, n_ret, n_max, want_stats, want_yp):
        n_sys = inputs.shape[0]
        if len(self._solvers) == 1 or n_sys < 2 * len(self._solvers):
            r = self._solve_shard(0, t_eval, t_interp, inputs, y0, yp0, output_mode, n_ret, n_max, want_stats, want_yp)
            self.last_kernel_ms, self.last_total_ms = r.pop("kernel_ms"), r.pop("total_ms")
            return r
        # shard over the devices: one host thread per device (ctypes releases the GIL during pbgpu_solve)
        import threading

        from .shard import cost_proxy, partition

        shards = partition(n_sys, len(self._solvers), cost_proxy(inputs, self._cost_model))
        D = self.desc
        # batch-order result arrays; every worker scatters its own rows into them as soon as its device is done (numpy
        # releases the GIL for these copies, so the 8 scatters run side by side instead of 40 ms of serial gathers)
        out = dict(t=np.empty((n_sys, n_max)), y=np.empty((n_sys, n_max, n_ret)), n_t=np.empty(n_sys, np.int32),
                   flag=np.empty(n_sys, np.int32), y_term=np.empty((n_sys, D.n_states)), event_idx=np.empty(n_sys, np.int32),
                   t_event=np.empty(n_sys))
        out["yp"] = np.empty((n_sys, n_max, D.n_states)) if want_yp else None
        out["stats"] = np.empty((n_sys, 8), np.int64) if want_stats else None
        times, errors = [None] * len(shards), []

        def work(k):
            try:
                idx = shards[k]
                if not len(idx):
                    times[k] = (0.0, 0.0)
                    return
                part = self._solve_shard(k, t_eval, t_interp, np.ascontiguousarray(inputs[idx]),
                                         None if y0 is None else np.ascontiguousarray(y0[idx]),
                                         None if yp0 is None else np.ascontiguousarray(yp0[idx]), output_mode, n_ret,
                                         n_max, want_stats, want_yp)
                for key, dst in out.items():
                    if dst is not None:
                        dst[idx] = part[key]
                times[k] = (part["kernel_ms"], part["total_ms"])
            except Exception as e:  # re-raised on the calling thread
                errors.append(e)

        threads = [threading.Thread(target=work, args=(k,)) for k in range(len(shards))]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        if errors:
            raise errors[0]
        self.last_kernel_ms = max(t_[0] for t_ in times)
        self.last_total_ms = max(t_[1] for t_ in times)
        out["n_max"] = n_max
        out["_keep"] = None
        return out

    def _solve_shard(self, k, t_eval, t_interp, inputs, y0, yp0, output_mode, n_ret, n_max, want_stats, want_yp):
        """One pbgpu_solve call on device slot k with host buffers (pinned result block, reused across calls)."""
        lib, D = self._lib, self.desc
        n_sys = inputs.shape[0]
        if self.reuse_buffers:
            # the returned arrays alias this cache: they are valid until the next solve on this solver
            key = (k, n_sys, n_max, n_ret)
            cache = self.__dict__.setdefault("_pinned", {})
            if key not in cache:
                for old in [q for q in cache if q[0] == k]:
                    del cache[old]
                cache[key] = (_pinned_empty((n_sys, n_max), np.float64), _pinned_empty((n_sys, n_max, n_ret), np.float64))
            (t_out, k1), (y_out, k2) = cache[key]
        else:
            (t_out, k1), (y_out, k2) = _pinned_empty((n_sys, n_max), np.float64), _pinned_empty((n_sys, n_max, n_ret), np.float64)
        yp_out, k3 = _pinned_empty((n_sys, n_max, D.n_states), np.float64) if want_yp else (None, None)
        n_t = np.zeros(n_sys, np.int32)
        flag = np.zeros(n_sys, np.int32)
        event_idx = np.zeros(n_sys, np.int32)
        t_event = np.zeros(n_sys)
        y_term = np.zeros((n_sys, D.n_states))
        stats = np.zeros((n_sys, 8), np.int64) if want_stats else None
        res = Result()
        res.n_max, res.n_ret = n_max, n_ret
        res.t, res.y, res.n_t, res.flag = _ptr(t_out), _ptr(y_out), _ptr(n_t), _ptr(flag)
        res.y_term, res.stats, res.yp = _ptr(y_term), _ptr(stats), _ptr(yp_out)
        res.event_idx, res.t_event = _ptr(event_idx), _ptr(t_event)
        # expensive systems first: the batch ends when its slowest system does (results stay in row order)
        from .shard import launch_order

        order = np.ascontiguousarray(launch_order(inputs, self._cost_model), dtype=np.int32)
        if lib.pbgpu_solver_set_launch_order(self._solvers[k], _ptr(order), n_sys):
            raise SolverError(_err(lib))
        rc = lib.pbgpu_solve(self._solvers[k], n_sys, _ptr(t_eval), len(t_eval), _ptr(t_interp) if len(t_interp) else None,
                             len(t_interp), _ptr(y0), _ptr(yp0), _ptr(inputs), output_mode, C.byref(res))
        if rc:
            raise SolverError(_err(lib))
        return dict(t=t_out, y=y_out, yp=yp_out, n_t=n_t, flag=flag, y_term=y_term, stats=stats, event_idx=event_idx,
                    t_event=t_event, n_max=n_max, _keep=(k1, k2, k3), kernel_ms=float(res.kernel_ms),
                    total_ms=float(res.total_ms))

    def solve(self, t_eval, inputs=None, t_interp=None, y0=None):
        """solve(t_eval, inputs=[{...}, ...]) -> SolutionBatch, a list-like sequence of Solution (a single Solution for a single dict)."""
        single = isinstance(inputs, dict) or inputs is None
        inputs_list = [inputs or {}] if single else list(inputs)
        try:
            p = self.bundle.flatten_inputs(inputs_list)
        except ValueError as e:
            raise SolverError(str(e)) from None
        r = self.solve_arrays(t_eval, p, t_interp=t_interp, y0=y0)
        variables = None
        if self.output_variables:
            variables, off = {}, 0
            for name in self.bundle.variables:
                n = len(self.bundle.variables[name])
                variables[name] = slice(off, off + n)
                off += n
        names = self.bundle.event_names
        # failures and events that are already non-positive at t0 are reported for the whole batch at once, like the loop did
        flags = r["flag"]
        if np.any(flags < 0):
            msg = sundials_error_message(int(flags[np.argmax(flags < 0)]))
            if self.on_failure == "error":
                raise SolverError(msg)
            if self.on_failure == "warn":
                for f in flags[flags < 0].tolist():
                    warnings.warn(sundials_error_message(f) + ", returning a partial solution.", stacklevel=2)
        if np.any((r["n_t"] == 1) & (flags == IDA_ROOT_RETURN)):
            raise SolverError(f"Events {[names[k] for k in range(self.bundle.n_events)]} are non-positive at initial conditions")
        sols = SolutionBatch(r, inputs_list, variables, names)
        return sols[0] if single else sols

    # -- experiment schedules on the device (pybamm_b200.experiment; simulation.py:991-1167) -------------------------------
    def solve_experiment(self, schedule, inputs, t0=0.0, y0=None, output_mode=None, want_stats=True):
        """Run an ``ExperimentSchedule`` for every row of ``inputs`` (n_sys, n_inputs; the columns the schedule overrides are
        ignored).  Returns the dict of ``solve_arrays`` plus ``step_end`` / ``step_term`` (n_sys, n_steps): stored points up
        to and including each step, and how each step ended (experiment.STEP_* or a negative IDA flag).
        flag: 1 every step ran, 0 a default step duration ran out, 2 a termination event other than the experiment's, < 0 failure."""
        from .host import Schedule

        self._ensure()
        lib, D = self._lib, self.desc
        inputs = np.ascontiguousarray(np.asarray(inputs, dtype=np.float64))
        if inputs.ndim != 2 or inputs.shape[1] != D.n_inputs:
            raise SolverError(f"inputs has wrong shape. Expected (n_sys, {D.n_inputs}) but got {inputs.shape}")
        n_sys = inputs.shape[0]
        try:
            slots, ev = schedule.slots(self.bundle), schedule.event_index(self.bundle)
        except ValueError as e:
            raise SolverError(str(e)) from None
        if output_mode is None:
            output_mode = 1 if self.output_variables else 0
        n_ret = D.n_vars if output_mode == 1 else D.n_states
        n_max = schedule.capacity
        q = Schedule()
        q.n_steps, q.n_over, q.experiment_event = schedule.n_steps, len(slots), ev
        q.over_slot, q.over_val, q.duration = _ptr(slots), _ptr(schedule.over_val), _ptr(schedule.duration)
        q.n_pts, q.stop_on_final_time = _ptr(schedule.n_pts), _ptr(schedule.stop_on_final_time)
        if y0 is not None:
            y0 = np.ascontiguousarray(np.asarray(y0, dtype=np.float64).reshape(n_sys, D.n_states))
        (t_out, k1), (y_out, k2) = _pinned_empty((n_sys, n_max), np.float64), _pinned_empty((n_sys, n_max, n_ret), np.float64)
        n_t, flag = np.zeros(n_sys, np.int32), np.zeros(n_sys, np.int32)
        event_idx, t_event = np.zeros(n_sys, np.int32), np.zeros(n_sys)
        y_term = np.zeros((n_sys, D.n_states))
        stats = np.zeros((n_sys, 8), np.int64) if want_stats else None
        step_end = np.zeros((n_sys, schedule.n_steps), np.int32)
        step_term = np.zeros((n_sys, schedule.n_steps), np.int32)
        res = Result()
        res.n_max, res.n_ret = n_max, n_ret
        res.t, res.y, res.n_t, res.flag = _ptr(t_out), _ptr(y_out), _ptr(n_t), _ptr(flag)
        res.y_term, res.stats, res.event_idx, res.t_event = _ptr(y_term), _ptr(stats), _ptr(event_idx), _ptr(t_event)
        if lib.pbgpu_solver_set_launch_order(self._solver, None, 0):
            raise SolverError(_err(lib))
        if lib.pbgpu_solve_experiment(self._solver, n_sys, float(t0), C.byref(q), _ptr(y0), _ptr(inputs), output_mode, C.byref(res),
                                      _ptr(step_end), _ptr(step_term)):
            raise SolverError(_err(lib))
        self.last_kernel_ms, self.last_total_ms = float(res.kernel_ms), float(res.total_ms)
        return dict(t=t_out, y=y_out, yp=None, n_t=n_t, flag=flag, y_term=y_term, stats=stats, event_idx=event_idx,
                    t_event=t_event, n_max=n_max, step_end=step_end, step_term=step_term, _keep=(k1, k2))

    # -- batched variable observation (idaklu.observe / observe_hermite_interp, observe.cpp:319-354) ----------------------
    def observe(self, result, inputs, t_obs=None):
        """Evaluate the bundle's compiled variables for every system of a full-state result of ``solve_arrays``.

        ``t_obs=None``: at the stored points, returns ``(n_sys, n_max, n_vars)`` (rows beyond ``n_t`` are zero);
        otherwise at the sorted times ``t_obs`` by cubic Hermite interpolation of ``y`` from ``(y, yp)`` (needs the ``yp``
        block: adaptive steps with ``hermite_interpolation``), returns ``(n_sys, len(t_obs), n_vars)``.
        The columns follow ``bundle.variables`` in order."""
        self._ensure()
        D = self.desc
        if D.n_vars < 1:
            raise SolverError(f"model bundle '{self.bundle.name}' has no compiled variables to observe")
        t = np.ascontiguousarray(result["t"], dtype=np.float64)
        y = np.ascontiguousarray(result["y"], dtype=np.float64)
        if y.ndim != 3 or y.shape[2] != D.n_states:
            raise SolverError("observe needs the full state (solve without output_variables)")
        n_sys, n_max = t.shape
        n_t = np.ascontiguousarray(np.minimum(result["n_t"], n_max), dtype=np.int32)
        p = np.ascontiguousarray(np.asarray(inputs, dtype=np.float64).reshape(n_sys, -1))
        if p.shape[1] != D.n_inputs:
            raise SolverError(f"inputs has wrong number of cols. Expected {D.n_inputs} but got {p.shape[1]}")
        yp = None
        if t_obs is not None:
            if result.get("yp") is None:
                raise SolverError("Hermite observation needs yp: solve with adaptive steps and hermite_interpolation")
            yp = np.ascontiguousarray(result["yp"], dtype=np.float64)
            t_obs = np.ascontiguousarray(np.asarray(t_obs, dtype=np.float64).reshape(-1))
            if np.any(np.diff(t_obs) < 0):
                raise SolverError("t_obs must be sorted")
        rows = len(t_obs) if t_obs is not None else n_max
        out = np.zeros((n_sys, rows, D.n_vars))
        if self._lib.pbgpu_observe(self._solver, n_sys, n_max, _ptr(n_t), _ptr(t), _ptr(y), _ptr(yp), _ptr(p), _ptr(t_obs),
                                   0 if t_obs is None else len(t_obs), _ptr(out)):
            raise SolverError(_err(self._lib))
        return out

    # -- device evaluation of the generated functions (parity tests, event identification) ---------
    def eval(self, which, y, p, t=0.0, cj=0.0, y2=None):
        self._ensure()
        D = self.desc
        idx = {"res": 0, "jac": 1, "events": 2, "ic": 3, "vars": 4, "solve": 5, "cycles": 6, "solve_loop": 7}[which]
        out_len = {"res": D.n_states, "jac": D.nnz_jac, "events": max(D.n_events, 1), "ic": D.n_states,
                   "vars": max(D.n_vars, 1), "solve": D.n_states, "cycles": 128, "solve_loop": 8}[which]
        p = np.ascontiguousarray(np.atleast_2d(np.asarray(p, dtype=np.float64)))
        n_sys = p.shape[0]
        if y is not None:
            y = np.ascontiguousarray(np.atleast_2d(np.asarray(y, dtype=np.float64)))
        if y2 is not None:
            y2 = np.ascontiguousarray(np.atleast_2d(np.asarray(y2, dtype=np.float64)))
        t_arr = np.broadcast_to(np.asarray(t, dtype=np.float64), (n_sys,))
        out = np.zeros((n_sys, out_len))
        if which == "events" and not np.all(t_arr == t_arr[0]):
            # one launch for the whole batch: per-system times travel in column 0 of the y2 block (which = 8)
            tcol = np.zeros((n_sys, D.n_states))
            tcol[:, 0] = t_arr
            if self._lib.pbgpu_model_eval(self._solver, 8, n_sys, _ptr(y), _ptr(p), 0.0, float(cj), _ptr(tcol), _ptr(out), out_len):
                raise SolverError(_err(self._lib))
            return out
        if np.all(t_arr == t_arr[0]):
            if self._lib.pbgpu_model_eval(self._solver, idx, n_sys, _ptr(y), _ptr(p), float(t_arr[0]), float(cj),
                                          _ptr(y2), _ptr(out), out_len):
                raise SolverError(_err(self._lib))
        else:
            for i in range(n_sys):
                oi = np.zeros((1, out_len))
                if self._lib.pbgpu_model_eval(self._solver, idx, 1, _ptr(np.ascontiguousarray(y[i:i + 1])) if y is not None else None,
                                              _ptr(np.ascontiguousarray(p[i:i + 1])), float(t_arr[i]), float(cj),
                                              _ptr(np.ascontiguousarray(y2[i:i + 1])) if y2 is not None else None,
                                              _ptr(oi), out_len):
                    raise SolverError(_err(self._lib))
                out[i] = oi[0]
        return out
