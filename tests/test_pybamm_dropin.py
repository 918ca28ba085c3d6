"""The pybamm-side drop-in (pybamm_b200/idaklu_solver.py) on the CPU tier.

Runs only where the reference checkout is mounted (this container): imports the *reference* pybamm through
tools/pybamm_env.py, builds a model exactly as a user would, and drives ``IDAKLUSolver.solve(model, t_eval, inputs=[...])``.
The native back end is replaced by the host emulation of the device integrator (tests/emu: the same pbgpu_ida.cuh and
generated CUDA-dialect code, compiled for the CPU), so what is tested here is everything the subclass adds: lowering at
set_up, BaseSolver's solve() flow, batching, result wrapping into pybamm.Solution.  The GPU tier runs the same class on
the real back end from a frozen bundle (tests/test_gpu.py).
"""
import numpy as np
import pytest

import tools.pybamm_env as env
from tests import util

pytestmark = pytest.mark.skipif(not env.available(), reason="reference pybamm checkout not mounted")


class EmuBackend:
    """solve_arrays() of BatchedIDASolver on top of tests/emu (tests only)."""

    FRONT_END_OPTIONS = {}

    def __init__(self, bundle, rtol, atol, options, output_variables):
        from tests.emu.emu import Emu, build_emu

        self.bundle, self.rtol, self.atol, self.options = bundle, rtol, atol, dict(options)
        self.output_variables = output_variables
        self.emu = Emu(build_emu(bundle))

    def solve_arrays(self, t_eval, inputs, t_interp=None, y0=None):
        from pybamm_b200.host import OPTION_DEFAULTS

        adaptive = t_interp is None or len(np.setdiff1d(t_interp, t_eval)) == 0
        save_vars = bool(self.output_variables)
        want_yp = adaptive and not save_vars and bool(self.options.get("hermite_interpolation", True))
        r = self.emu.solve(t_eval, inputs if inputs.shape[1] else np.zeros((len(inputs), 0)), t_interp=t_interp, rtol=self.rtol,
                           atol=self.atol, y0=y0, save_vars=save_vars, want_yp=want_yp,
                           options={k: v for k, v in self.options.items() if k in OPTION_DEFAULTS})
        n = len(r["t"])
        n_max = max(len(t) for t in r["t"])
        n_ret = r["y"][0].shape[1]
        t = np.zeros((n, n_max))
        y = np.zeros((n, n_max, n_ret))
        yp = np.zeros((n, n_max, self.bundle.n_states)) if want_yp else None
        for i in range(n):
            k = len(r["t"][i])
            t[i, :k], y[i, :k] = r["t"][i], r["y"][i]
            if want_yp:
                yp[i, :k] = r["yp"][i]
        return dict(t=t, y=y, yp=yp, n_t=np.array([len(x) for x in r["t"]], np.int32), flag=r["flag"], y_term=r["y_term"],
                    stats=r["stats"], event_idx=r["event_idx"], t_event=r["t_event"], n_max=n_max)


def _solver(**kw):
    from pybamm_b200.idaklu_solver import IDAKLUSolver

    return IDAKLUSolver(_backend_factory=EmuBackend, **kw)


@pytest.fixture(scope="module")
def spm():
    from tools.build_models import build

    pybamm, model, names = build("spm")
    return pybamm, model, names


def _inputs(n, seed=3):
    rng = np.random.default_rng(seed)
    return [{"Current function [A]": float(rng.uniform(2.5, 7.5)),
             "Negative particle diffusivity [m2.s-1]": float(10 ** rng.uniform(-14, -13))} for _ in range(n)]


def test_is_a_base_solver_with_the_reference_surface():
    pybamm = env.load()
    s = _solver(rtol=1e-6, atol=1e-8, options={"max_order_bdf": 4, "num_threads": 2})
    assert isinstance(s, pybamm.BaseSolver) and s.supports_interp and s.supports_t_eval_discontinuities
    assert s.options["max_order_bdf"] == 4 and s.options["num_solvers"] == 2 and s.options["linear_solver"] == "SUNLinSol_KLU"
    with pytest.raises(pybamm.SolverError):
        _solver(options={"num_threads": 1, "num_solvers": 2})
    with pytest.raises(pybamm.SolverError):
        _solver(options={"no_such_option": 1})
    with pytest.raises(pybamm.SolverError):
        _solver(options={"hermite_reduction_factor": 0.5})


def test_solve_model_t_eval_inputs_returns_pybamm_solutions(spm):
    pybamm, model, names = spm
    solver = _solver(rtol=1e-8, atol=1e-8)
    inputs = _inputs(3)
    t_interp = np.linspace(0, 3600, 61)
    sols = solver.solve(model, [0, 3700], inputs=inputs, t_interp=t_interp)
    assert isinstance(sols, list) and len(sols) == 3 and all(isinstance(s, pybamm.Solution) for s in sols)
    b, o = util.bundle("spm"), util.oracle("spm")
    P = np.array([[d[n] for n in b.input_names] for d in inputs])
    ref = o.solve([0, 3700], P, t_interp=t_interp, rtol=1e-8, atol=1e-8)
    volt = model.get_processed_variable_or_event("Voltage [V]")
    for i, s in enumerate(sols):
        assert s.y.shape == (model.len_rhs_and_alg, len(s.t))
        np.testing.assert_allclose(s.t[:-1], ref["t"][i][:-1], rtol=0, atol=0)
        assert abs(s.t[-1] - ref["t"][i][-1]) < 1e-3
        np.testing.assert_allclose(s.y[:, :-1], ref["y"][i][:-1].T, rtol=0, atol=1e-7)
        if ref["flag"][i] == 2:
            assert s.termination == "event: Minimum voltage [V]"
            assert s.t_event[0] == s.t[-1] and s.y_event.shape == (model.len_rhs_and_alg, 1)
            v_end = float(np.asarray(volt.evaluate(s.t[-1], s.y[:, -1], inputs=inputs[i])).reshape(-1)[0])
            assert abs(v_end - 2.5) < 1e-6
        else:
            assert s.termination == "final time" and s.t[-1] == 3700.0
        assert s.all_inputs[0].keys() == inputs[i].keys()
        assert s.integration_time is not None and s.solve_time is not None and s.set_up_time is not None
    # a single dict gives a single Solution (base_solver.py:1123-1127); the solver is bound to its model (:970-977)
    one = solver.solve(model, [0, 3700], inputs=inputs[0], t_interp=t_interp)
    assert isinstance(one, pybamm.Solution)
    np.testing.assert_array_equal(one.y, sols[0].y)


def test_adaptive_steps_carry_hermite_data(spm):
    pybamm, model, _ = spm
    solver = _solver(rtol=1e-6, atol=1e-8)
    s = solver.solve(model, [0, 600], inputs=_inputs(1)[0])
    assert s.all_yps is not None and s.all_yps[0].shape == s.y.shape     # hermite_interpolation=True, adaptive steps
    assert s.t[0] == 0 and s.t[-1] == 600 and len(s.t) > 10
    # dy/dt of the interpolant satisfies the model's own rhs (the reference's Symbol.evaluate) at a stored step
    k = len(s.t) // 2
    f = np.asarray(model.concatenated_rhs.evaluate(s.t[k], s.y[:, k], inputs=_inputs(1)[0])).reshape(-1)
    np.testing.assert_allclose(s.all_yps[0][:, k], f, rtol=2e-2, atol=1e-5 * np.max(np.abs(f)))   # rtol = 1e-6 run
    fixed = solver.solve(model, [0, 600], inputs=_inputs(1)[0], t_interp=np.linspace(0, 600, 7))
    np.testing.assert_array_equal(fixed.t, np.linspace(0, 600, 7))


def test_missing_input_and_second_model_are_errors(spm):
    pybamm, model, _ = spm
    solver = _solver()
    with pytest.raises(pybamm.SolverError, match="No value provided for input"):
        solver.solve(model, [0, 10], inputs=[{"Current function [A]": 5.0}])
    solver.solve(model, [0, 10], inputs=_inputs(1))
    other = pybamm.BaseModel("other")
    v = pybamm.Variable("v")
    other.rhs = {v: -v}
    other.initial_conditions = {v: 1}
    with pytest.raises(RuntimeError, match="already been initialised"):
        solver.solve(other, [0, 1])


def test_known_answer_models_of_the_reference_suite():
    # packages/pybamm/tests/unit/test_solvers/test_idaklu_solver.py:92-129 (DAE + event at t = 2), :131-170 (batch of inputs)
    pybamm = env.load()
    model = pybamm.BaseModel()
    u, v = pybamm.Variable("u"), pybamm.Variable("v")
    model.rhs = {u: 0.1 * v}
    model.algebraic = {v: 1 - v}
    model.initial_conditions = {u: 0, v: 1}
    model.events = [pybamm.Event("1", 0.2 - u), pybamm.Event("2", v)]
    model.variables = {"u": u}
    pybamm.Discretisation().process_model(model)
    sol = _solver(rtol=1e-8, atol=1e-8).solve(model, np.linspace(0, 3, 100))
    assert sol.termination == "event: 1"
    np.testing.assert_allclose(sol.t[-1], 2.0, atol=1e-6)
    np.testing.assert_allclose(sol.y[0, -1], 0.2, atol=1e-6)
    np.testing.assert_allclose(sol.y[0], 0.1 * sol.t, atol=1e-6)

    m2 = pybamm.BaseModel()
    var = pybamm.Variable("var")
    m2.rhs = {var: -pybamm.InputParameter("rate") * var}
    m2.initial_conditions = {var: 2}
    pybamm.Discretisation().process_model(m2)
    t_interp = np.linspace(0, 1, 10)
    rates = np.arange(1, 9) * 0.1
    sols = _solver(rtol=1e-8, atol=1e-8).solve(m2, [0, 1], inputs=[{"rate": r} for r in rates], t_interp=t_interp)
    for r, s in zip(rates, sols):
        np.testing.assert_array_equal(s.t, t_interp)
        np.testing.assert_allclose(s.y[0], 2 * np.exp(-r * t_interp), atol=1e-6)


def test_pickle_drops_the_native_state(spm):
    import pickle

    pybamm, model, _ = spm
    solver = _solver(rtol=1e-6, atol=1e-8)
    solver.solve(model, [0, 10], inputs=_inputs(1))
    clone = pickle.loads(pickle.dumps(solver.__getstate__()))
    assert "_setup" not in clone and "_model_set_up" not in clone
