"""GPU tier: the CUDA path, called through the C ABI (ctypes -> libpbgpu.so -> model image), against
the reference's golden vectors, its known-answer tests, the CPU oracle and size-independent properties.

Run on the B200 box:  python -m pytest tests -m gpu
"""
import ctypes as C

import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from pybamm_b200 import BatchedIDASolver, SolverError
from tests import util

pytestmark = pytest.mark.gpu

_solvers = {}


def gpu(name, rtol=1e-8, atol=1e-8, out=True, **kw):
    key = (name, rtol, atol, out, repr(sorted(kw.items())))
    if key not in _solvers:
        b = util.bundle(name)
        ov = [next(iter(b.variables))] if (out and b.variables) else None
        _solvers[key] = BatchedIDASolver(b, rtol=rtol, atol=atol, output_variables=ov, **kw)
    return _solvers[key]


def _close(got, ref, rtol=1e-12):
    scale = max(1.0, float(np.max(np.abs(ref))))
    np.testing.assert_allclose(got, ref, rtol=0, atol=rtol * scale)


# ---- generated device functions vs the reference's own evaluation ---------------------------------------
@pytest.mark.parametrize("name", util.BATTERY + util.LARGE)
def test_device_functions_match_reference_evaluation(name):
    z = np.load(f"{util.GOLDEN}/{name}_eval.npz")
    b, s = util.bundle(name), gpu(name)
    diag = ((b.jac_rows == b.jac_cols) & (b.jac_rows < b.n_diff)).astype(float)
    for k in range(len(z["t"])):
        t, y, p = float(z["t"][k]), z["y"][k], z["p"][k]
        _close(s.eval("res", y, p, t)[0], z["rhs_alg"][k])
        _close(s.eval("events", y, p, t)[0][:b.n_events], z["events"][k])
        _close(s.eval("vars", y, p, t)[0][:1], z["voltage"][k])
        _close(s.eval("ic", None, p)[0], z["ic"][k])
        _close(s.eval("jac", y, p, t, 0.37)[0], z["jac_vals"][k] - 0.37 * diag, 1e-11)
    # batched call == per-system calls (bitwise)
    allres = s.eval("res", z["y"], z["p"], 100.0)
    for k in range(len(z["t"])):
        np.testing.assert_array_equal(allres[k], s.eval("res", z["y"][k], z["p"][k], 100.0)[0])


@pytest.mark.parametrize("name", util.BATTERY + util.LARGE)
def test_device_sparse_lu_matches_scipy(name):
    z = np.load(f"{util.GOLDEN}/{name}_eval.npz")
    b, s = util.bundle(name), gpu(name)
    rng = np.random.default_rng(3)
    diag = ((b.jac_rows == b.jac_cols) & (b.jac_rows < b.n_diff)).astype(float)
    for cj in (1e-3, 1.0, 250.0):
        rhs = rng.standard_normal((len(z["t"]), b.n_states))
        x = s.eval("solve", z["y"], z["p"], 100.0, cj, rhs)
        for k in range(len(z["t"])):
            jv = s.eval("jac", z["y"][k], z["p"][k], 100.0, cj)[0]
            A = sp.csc_matrix((jv, (b.jac_rows, b.jac_cols)), shape=(b.n_states,) * 2)
            xr = spla.spsolve(A, rhs[k])
            assert np.max(np.abs(x[k] - xr)) <= 1e-9 * np.max(np.abs(xr))


# ---- the reference's known-answer tests through the CUDA path -------------------------------------------
def test_kat_roberts():
    s = gpu("kat_roberts", 1e-4, 1e-6, out=False)
    sol = s.solve([0, 3])
    assert sol.termination == "event: 1"
    np.testing.assert_allclose(sol.t[-1], 2.0, rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(sol.y[0, -1], 0.2, rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(sol.y[1, :], np.ones(sol.t.shape), rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(sol.y[0, :], 0.1 * sol.t, rtol=1e-7, atol=1e-6)
    np.testing.assert_array_equal(sol.y_event, sol.y[:, -1])


def test_kat_multiple_inputs():
    s = gpu("kat_decay", 1e-5, 1e-5, out=False, options={"num_threads": 8, "num_solvers": 2})
    t_interp = np.linspace(0, 1, 10)
    inputs = [{"rate": 0.01 * (i + 1)} for i in range(8)]
    sols = s.solve([0, 1], inputs=inputs, t_interp=t_interp)
    for d, sol in zip(inputs, sols):
        assert sol.all_inputs == [d] and sol.termination == "final time"
        np.testing.assert_array_equal(sol.t, t_interp)
        np.testing.assert_allclose(sol.y[0], 2 * np.exp(-d["rate"] * sol.t), atol=1e-4, rtol=1e-4)


def test_kat_growth_events():
    t_interp = np.linspace(0, 1, 100)
    sol = gpu("kat_growth", out=False).solve([0, 1], t_interp=t_interp)
    np.testing.assert_array_equal(sol.t, t_interp)
    np.testing.assert_allclose(sol.y[0], np.exp(0.1 * sol.t), rtol=1e-6, atol=1e-5)
    sol = gpu("kat_growth_ev", out=False).solve([0, 1], t_interp=t_interp)
    assert len(sol.t) < len(t_interp) and sol.termination == "event: an event"
    np.testing.assert_allclose(sol.y[0], np.exp(0.1 * sol.t), rtol=1e-6, atol=1e-5)
    np.testing.assert_allclose(sol.t[-1], np.log(1.01) / 0.1, rtol=0, atol=1e-6)
    sol = gpu("kat_growth_dae", out=False).solve([0, 5])
    assert sol.termination == "event: var2 = 2.5"   # 2 e^{0.1 t} = 2.5 at t = 2.23 comes first
    assert np.all(sol.y[0, :-1] < 1.5) and np.all(sol.y[-1, :-1] < 2.5)
    np.testing.assert_equal(sol.t_event[0], sol.t[-1])
    np.testing.assert_allclose(sol.y[0], np.exp(0.1 * sol.t), rtol=1e-6, atol=1e-5)
    np.testing.assert_allclose(sol.y[-1], 2 * np.exp(0.1 * sol.t), rtol=1e-6, atol=1e-5)


def test_kat_scalar_and_vector_inputs():
    s = gpu("kat_inputs", 1e-4, 1e-6, out=False)
    t_interp = np.linspace(0, 3, 100)
    sol = s.solve([0, 3], inputs={"a": 0.1, "b": np.array([[0.2], [0.3]])}, t_interp=t_interp)
    np.testing.assert_allclose(sol.y[3], np.ones(sol.t.shape), rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(sol.y[0], 0.1 * sol.t, rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(sol.y[1:3], np.array([[0.2], [0.3]]) * sol.t, rtol=1e-7, atol=1e-6)
    with pytest.raises(SolverError, match="No value provided for input"):
        s.solve([0, 3], inputs={"a": 0.1})


def test_event_at_initial_condition_raises():
    # idaklu_solver.py:754-755, base_solver.py:1215-1242
    s = gpu("kat_decay", out=False)
    b = util.bundle("kat_growth_ev")
    s2 = gpu("kat_growth_ev", out=False)
    y0 = np.array([[1.5]])   # event 1.01 - var is already negative
    with pytest.raises(SolverError, match="non-positive at initial conditions"):
        s2.solve([0, 1], y0=y0)
    assert s is not None and b.n_events == 1


# ---- battery models: golden solutions, oracle parity -------------------------------------------------------
@pytest.mark.parametrize("name", util.BATTERY)
def test_golden_solutions(name):
    z = np.load(f"{util.GOLDEN}/{name}_solve.npz")
    r = gpu(name).solve_arrays(z["t_eval"], z["p"], t_interp=z["t_interp"])
    for i in range(len(z["p"])):
        n, m = int(r["n_t"][i]), int(z["ref_n"][i])
        for tag in ("ref", "truth"):
            dv, dt = util.compare_voltage(r["t"][i, :n], r["y"][i, :n, 0], z[f"{tag}_t"][i, :m], z[f"{tag}_v"][i, :m],
                                          r["flag"][i], z[f"{tag}_flag"][i])
            assert dv < util.V_TOL and dt < util.T_EVENT_TOL, (name, i, tag, dv, dt)


@pytest.mark.parametrize("name,n", [("spm", 96), ("spme", 64), ("dfn", 32)])
def test_parity_against_oracle_seeded_batch(name, n):
    b, o = util.bundle(name), util.oracle(name)
    p = util.sample_inputs(b, n, np.random.default_rng(4242))
    r = gpu(name).solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    q = o.solve(util.T_EVAL, p, t_interp=util.T_INTERP, rtol=1e-8, atol=1e-8, save_vars=True, n_threads=16)
    worst = [0.0, 0.0]
    for i in range(n):
        k = int(r["n_t"][i])
        dv, dt = util.compare_voltage(r["t"][i, :k], r["y"][i, :k, 0], q["t"][i], q["y"][i][:, 0], r["flag"][i], q["flag"][i])
        worst = [max(worst[0], dv), max(worst[1], dt)]
    assert worst[0] < util.V_TOL and worst[1] < util.T_EVENT_TOL, worst


def test_full_state_output_matches_oracle():
    b, o = util.bundle("dfn"), util.oracle("dfn")
    p = util.sample_inputs(b, 4, np.random.default_rng(99))
    s = gpu("dfn", out=False)
    r = s.solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    q = o.solve(util.T_EVAL, p, t_interp=util.T_INTERP, rtol=1e-8, atol=1e-8, n_threads=4)
    for i in range(4):
        n = int(r["n_t"][i])
        assert n == len(q["t"][i]) and r["y"].shape[2] == b.n_states
        m = n - 1 if r["flag"][i] == 2 else n
        np.testing.assert_allclose(r["y"][i, :m], q["y"][i][:m], rtol=0, atol=2e-6)
        np.testing.assert_allclose(r["y_term"][i], r["y"][i, n - 1], rtol=0, atol=1e-12)


def test_batch_equals_independent_and_is_order_invariant():
    b = util.bundle("spme")
    s = gpu("spme")
    p = util.sample_inputs(b, 300, np.random.default_rng(11))
    r = s.solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    t, y, n = r["t"].copy(), r["y"].copy(), r["n_t"].copy()
    for i in (0, 17, 299):
        one = s.solve_arrays(util.T_EVAL, p[i:i + 1], t_interp=util.T_INTERP)
        np.testing.assert_array_equal(one["t"][0, :n[i]], t[i, :n[i]])
        np.testing.assert_array_equal(one["y"][0, :n[i]], y[i, :n[i]])
    perm = np.random.default_rng(1).permutation(len(p))
    r2 = s.solve_arrays(util.T_EVAL, p[perm], t_interp=util.T_INTERP)
    np.testing.assert_array_equal(r2["n_t"], n[perm])
    for k in range(0, len(p), 7):
        np.testing.assert_array_equal(r2["y"][k, :n[perm[k]]], y[perm[k], :n[perm[k]]])


def test_interior_breakpoints_and_adaptive_storage():
    b, o = util.bundle("dfn"), util.oracle("dfn")
    p = np.array([b.meta["nominal_inputs"]])
    ti = np.linspace(0, 1200, 25)
    r = gpu("dfn").solve_arrays([0.0, 600.0, 1200.0], p, t_interp=ti)
    q = o.solve([0.0, 600.0, 1200.0], p, t_interp=ti, rtol=1e-8, atol=1e-8, save_vars=True)
    n = int(r["n_t"][0])
    np.testing.assert_array_equal(r["t"][0, :n], ti)
    assert np.max(np.abs(r["y"][0, :n, 0] - q["y"][0][:, 0])) < util.V_TOL
    a = gpu("spm").solve_arrays([0.0, 600.0], np.array([util.bundle("spm").meta["nominal_inputs"]]))   # adaptive steps stored
    na = int(a["n_t"][0])
    ta = a["t"][0, :na]
    assert ta[0] == 0.0 and ta[-1] == 600.0 and na > 10 and np.all(np.diff(ta) > 0)


def test_failures_are_flags_and_argument_errors_raise():
    b = util.bundle("spm")
    p = util.sample_inputs(b, 3, np.random.default_rng(2))
    s = BatchedIDASolver(b, rtol=1e-8, atol=1e-8, options={"max_num_steps": 20}, on_failure="ignore")
    r = s.solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    assert np.all(r["flag"] == -1) and np.all(r["n_t"] >= 1) and np.all(r["n_t"] < len(util.T_INTERP))
    sols = s.solve(util.T_EVAL, inputs=[dict(zip(b.input_names, row)) for row in p], t_interp=util.T_INTERP)
    assert all(x.termination == "failure" for x in sols)
    s_err = BatchedIDASolver(b, rtol=1e-8, atol=1e-8, options={"max_num_steps": 20}, on_failure="error")
    with pytest.raises(SolverError, match="IDA_TOO_MUCH_WORK"):
        s_err.solve(util.T_EVAL, inputs=dict(zip(b.input_names, p[0])), t_interp=util.T_INTERP)
    g = gpu("spm")
    with pytest.raises(SolverError, match="wrong number of cols. Expected 2 but got 3"):
        g.solve_arrays(util.T_EVAL, np.zeros((2, 3)))
    with pytest.raises(SolverError, match="at least 2"):
        g.solve_arrays([0.0], p)
    with pytest.raises(SolverError, match="less than the greatest"):
        g.solve_arrays([0.0, 1.0], p, t_interp=[2.0])


def test_device_pointer_entry_point():
    import torch

    b = util.bundle("spme")
    s = gpu("spme")
    s._ensure()
    from pybamm_b200.host import Result

    n = 40
    p = util.sample_inputs(b, n, np.random.default_rng(8))
    ref = s.solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    dev = torch.device("cuda", 0)
    te = torch.tensor(util.T_EVAL, dtype=torch.float64, device=dev)
    ti = torch.tensor(util.T_INTERP[1:], dtype=torch.float64, device=dev)   # t_eval values removed (canonical form)
    pin = torch.tensor(p, dtype=torch.float64, device=dev)
    n_max = s._lib.pbgpu_result_capacity(2, len(ti))
    t_out = torch.zeros((n, n_max), dtype=torch.float64, device=dev)
    y_out = torch.zeros((n, n_max, 1), dtype=torch.float64, device=dev)
    n_t = torch.zeros(n, dtype=torch.int32, device=dev)
    flag = torch.zeros(n, dtype=torch.int32, device=dev)
    res = Result()
    res.n_max, res.n_ret = n_max, 1
    res.t, res.y, res.n_t, res.flag = t_out.data_ptr(), y_out.data_ptr(), n_t.data_ptr(), flag.data_ptr()
    rc = s._lib.pbgpu_solve_device(s._solver, n, C.c_void_p(te.data_ptr()), 2, C.c_void_p(ti.data_ptr()), len(ti), None, None,
                                   C.c_void_p(pin.data_ptr()), 1, C.byref(res), None, 1)
    assert rc == 0 and res.kernel_ms > 0
    np.testing.assert_array_equal(n_t.cpu().numpy(), ref["n_t"])
    for i in range(n):
        k = int(ref["n_t"][i])
        np.testing.assert_array_equal(y_out[i, :k, 0].cpu().numpy(), ref["y"][i, :k, 0])


def test_batched_event_identification_matches_per_system_evaluation():
    """solve(inputs=[{...}] * n) names the event that ended every system from ONE device launch (which = 8, per-system
    times); it must agree with the per-system evaluation of the event functions (idaklu_solver.py:826-834)."""
    b = util.bundle("spme")
    s = gpu("spme")
    n = 24
    p = util.sample_inputs(b, n, np.random.default_rng(11))
    dicts = [dict(zip(b.input_names, row.tolist())) for row in p]
    sols = s.solve(util.T_EVAL, inputs=dicts, t_interp=util.T_INTERP)
    r = s.solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    t_end = np.array([r["t"][i, r["n_t"][i] - 1] for i in range(n)])
    batched = s.eval("events", r["y_term"], p, t_end)
    single = np.vstack([s.eval("events", r["y_term"][i:i + 1], p[i:i + 1], float(t_end[i])) for i in range(n)])
    np.testing.assert_array_equal(batched, single)
    assert np.any(r["flag"] == 2)
    for i, sol in enumerate(sols):
        want = f"event: {b.event_names[int(np.argmin(np.abs(single[i])))]}" if r["flag"][i] == 2 else "final time"
        assert sol.termination == want, (i, sol.termination, want)


# ---- BASELINE.json sizes: size-independent properties ---------------------------------------------------------
def test_config1_spm_1000_sets_properties():
    b = util.bundle("spm")
    p = util.sample_inputs(b, 1000, np.random.default_rng(20260))
    r = gpu("spm").solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    assert set(np.unique(r["flag"]).tolist()) <= {1, 2}
    cur = p[:, b.input_names.index("Current function [A]")]
    for i in range(0, 1000, 9):
        n = int(r["n_t"][i])
        v = r["y"][i, :n, 0]
        assert np.all(np.diff(v) < 0)                  # constant-current discharge: V falls monotonically
        if r["flag"][i] == 2:
            assert abs(v[-1] - 2.5) < 1e-6 and r["t"][i, n - 1] < 3700.0
        else:
            assert n == len(util.T_INTERP) + 1 and r["t"][i, n - 1] == 3700.0
    # discharged capacity at the end = I * t_end / 3600 (state 0 is the discharge capacity [A.h])
    full = gpu("spm", out=False).solve_arrays(util.T_EVAL, p[:64], t_interp=util.T_INTERP)
    for i in range(64):
        n = int(full["n_t"][i])
        assert abs(full["y"][i, n - 1, 0] - cur[i] * full["t"][i, n - 1] / 3600.0) < 1e-6


def test_config2_spme_10k_sets_all_end_on_the_voltage_cutoff():
    b = util.bundle("spme")
    p = util.sample_inputs(b, 10000, np.random.default_rng(20260))
    s = gpu("spme")
    t_eval = [0.0, 8000.0]
    ti = np.linspace(0, 7990, 800)
    r = s.solve_arrays(t_eval, p, t_interp=ti)
    assert np.all(r["flag"] == 2)
    rows = np.arange(len(p))
    v_end = r["y"][rows, r["n_t"] - 1, 0]
    assert np.max(np.abs(v_end - 3.0)) < 1e-6
    t_end = r["t"][rows, r["n_t"] - 1]
    cur = p[:, b.input_names.index("Current function [A]")]
    # capacity to the 3.0 V cut-off is bounded by the cell capacity and falls with the current
    cap = cur * t_end / 3600.0
    assert 3.0 < cap.min() and cap.max() < 5.3
    ev = s.eval("events", r["y_term"], p, t_end)
    assert np.max(np.abs(ev[:, 0])) < 1e-6 and np.all(ev[:, 1] > 0)


def test_config3_dfn_properties_and_linearity_of_capacity():
    b = util.bundle("dfn")
    p = util.sample_inputs(b, 592, np.random.default_rng(20260))   # 4 waves of 148 SMs
    r = gpu("dfn").solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    assert set(np.unique(r["flag"]).tolist()) <= {1, 2}
    rows = np.arange(len(p))
    v_end = r["y"][rows, r["n_t"] - 1, 0]
    ev = r["flag"] == 2
    assert np.max(np.abs(v_end[ev] - 2.5)) < 1e-6
    assert np.all(r["n_t"][~ev] == len(util.T_INTERP) + 1)
    st = r["stats"]
    assert np.all(st[:, 0] > 100) and np.all(st[:, 1] >= st[:, 0]) and np.all(st[:, 2] == st[:, 6])


def test_multi_device_sharding_matches_single_device():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    b = util.bundle("spme")
    p = util.sample_inputs(b, 257, np.random.default_rng(5))
    one = gpu("spme").solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    t1, y1, n1 = one["t"].copy(), one["y"].copy(), one["n_t"].copy()
    two = BatchedIDASolver(b, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"], devices=[0, 1])
    r = two.solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    np.testing.assert_array_equal(r["n_t"], n1)
    for i in range(len(p)):
        np.testing.assert_array_equal(r["y"][i, :n1[i]], y1[i, :n1[i]])
        np.testing.assert_array_equal(r["t"][i, :n1[i]], t1[i, :n1[i]])


@pytest.mark.gpu
def test_adaptive_steps_beyond_the_first_block_are_re_solved_not_truncated():
    # a system with more steps than max_saved_steps used to come back short with a success flag
    from pybamm_b200 import BatchedIDASolver

    b = util.bundle("kat_decay")
    s = BatchedIDASolver(b, rtol=1e-8, atol=1e-10, options={"max_saved_steps": 8})
    big = BatchedIDASolver(b, rtol=1e-8, atol=1e-10)
    p = np.array([[0.3], [0.05], [1.0]])
    r, q = s.solve_arrays([0.0, 10.0], p), big.solve_arrays([0.0, 10.0], p)
    np.testing.assert_array_equal(r["n_t"], q["n_t"])
    assert int(r["n_t"].max()) > 8 and r["n_max"] == int(r["n_t"].max())
    for i in range(3):
        k = int(q["n_t"][i])
        np.testing.assert_array_equal(r["t"][i, :k], q["t"][i, :k])
        np.testing.assert_array_equal(r["y"][i, :k], q["y"][i, :k])
        assert r["t"][i, k - 1] == 10.0
    sols = s.solve([0.0, 10.0], inputs=[{"rate": 0.3}])
    assert sols[0].t[-1] == 10.0 and sols[0].termination == "final time"


@pytest.mark.gpu
def test_hermite_data_event_index_and_event_time_from_the_device():
    from pybamm_b200 import BatchedIDASolver

    s = BatchedIDASolver(util.bundle("kat_decay"), rtol=1e-9, atol=1e-11)
    sol = s.solve([0.0, 10.0], inputs=[{"rate": 0.3}])[0]
    assert sol.yp is not None and sol.yp.shape == sol.y.shape           # adaptive steps + full state: hermite_interpolation
    np.testing.assert_allclose(sol.yp[0], -0.3 * sol.y[0], rtol=1e-4, atol=1e-9)
    assert s.solve([0.0, 10.0], inputs=[{"rate": 0.3}], t_interp=np.linspace(0, 10, 11))[0].yp is None
    e = util.emu("kat_decay").solve([0.0, 10.0], np.array([[0.3]]), rtol=1e-9, atol=1e-11, want_yp=True)
    np.testing.assert_allclose(sol.t, e["t"][0], rtol=1e-6, atol=1e-12)       # same step sequence (FMA round-off apart)
    rb = BatchedIDASolver(util.bundle("kat_roberts"), rtol=1e-8, atol=1e-8)
    r = rb.solve_arrays([0.0, 5.0], np.zeros((2, 0)), t_interp=np.linspace(0, 5, 11))
    assert list(r["flag"]) == [2, 2] and list(r["event_idx"]) == [0, 0]
    np.testing.assert_allclose(r["t_event"], 2.0, atol=1e-6)
    st = BatchedIDASolver(util.bundle("kat_decay"), rtol=1e-8, atol=1e-10, on_failure="ignore",
                          options={"num_steps_no_progress": 5, "t_no_progress": 1e9})
    assert st.solve_arrays([0.0, 10.0], np.array([[0.3]]))["flag"][0] == -3


# ---- BASELINE configs at batch size: GPU vs oracle@1e-8, GPU vs oracle@1e-11 ("truth"), oracle@1e-8 vs truth, per system ------
@pytest.mark.parametrize("name,n", [("spm", 1000), ("spme", 1024), ("dfn", 1024)])
def test_parity_on_the_baseline_batches(name, n):
    """The first n of bench.py's seeded input sets (configs 1-3), the whole t_interp grid including the last point before a
    cut-off.  Bars: same termination and grid; event times within 1e-3 s; per system the voltage is
    within 1e-6 V of the reference algorithm's own solution at rtol = atol = 1e-8 or, where those tolerances do not deliver
    1e-6 V themselves (the knee before the cut-off), no further from the 1e-11 truth than twice the reference algorithm's
    own error there.  bench.py applies the same rule to its sample and exits non-zero when it fails."""
    import os

    import bench

    b = util.bundle(name)
    p = bench.sample_inputs(b, n, np.random.default_rng(20260))
    r = gpu(name).solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    nt = np.asarray(r["n_t"])
    t_end = np.array([r["t"][i, int(nt[i]) - 1] for i in range(n)])
    par, _ = bench.parity_block(b, p, nt, np.asarray(r["flag"]), np.asarray(r["y"])[:, :, 0], t_end, n, os.cpu_count() or 1)
    print(name, {k: v for k, v in par.items() if k != "bars"})
    assert par["same_termination"]
    assert par["max_abs_dt_end_gpu_vs_oracle"] <= util.T_EVENT_TOL and par["max_abs_dt_end_gpu_vs_truth"] <= util.T_EVENT_TOL
    assert par["systems_failing_the_bar"] == 0, par
    if name != "dfn":
        # the ODE models meet the plain bar everywhere, knee included
        assert par["max_abs_dV_gpu_vs_oracle_1e-8"] <= util.V_TOL and par["max_abs_dV_gpu_vs_truth_1e-11"] <= util.V_TOL
    # the GPU's error against the truth is of the size of the reference algorithm's own (same tolerances, same method)
    assert par["median_abs_dV_gpu_vs_truth"] <= 3.0 * par["median_abs_dV_oracle_vs_truth"] + 1e-9


def test_failed_static_pivot_is_repaired_by_the_repivoted_image():
    """kat_pivot_flip: the pivot the order was chosen for at t = 0 is exactly zero from t = 2 ln 2 on.  Without the slow path
    the system fails there; with it (default) the failed row is solved again on an image whose order is also valid for the
    matrices at the failure state, and lands on the analytic solution x = 1 / (1 + max(2 u - 1, 0)), u = exp(-t / 2)."""
    b = util.bundle("kat_pivot_flip")
    t_i = np.linspace(0.0, 4.0, 41)
    plain = BatchedIDASolver(b, rtol=1e-8, atol=1e-10, output_variables=["x"], options={"repivot": False}, on_failure="ignore")
    r0 = plain.solve_arrays([0.0, 4.0], np.zeros((3, 0)), t_interp=t_i)
    assert np.all(r0["flag"] < 0) and np.all(r0["t_event"] < 2.0 * np.log(2.0) + 0.11)
    s = BatchedIDASolver(b, rtol=1e-8, atol=1e-10, output_variables=["x"], on_failure="ignore")
    r = s.solve_arrays([0.0, 4.0], np.zeros((3, 0)), t_interp=t_i)
    assert np.all(r["flag"] == 1), r["flag"]
    assert sorted(r["repivoted_rows"].tolist()) == [0, 1, 2]
    n = int(r["n_t"][0])
    t = r["t"][0, :n]
    np.testing.assert_allclose(r["y"][0, :n, 0], 1.0 / (1.0 + np.maximum(2.0 * np.exp(-0.5 * t) - 1.0, 0.0)), rtol=0, atol=2e-6)
    ref = util.oracle("kat_pivot_flip").solve([0.0, 4.0], np.zeros((1, 0)), t_interp=t_i, rtol=1e-8, atol=1e-10, save_vars=True)
    np.testing.assert_allclose(r["y"][0, :n, 0], ref["y"][0][:, 0], rtol=0, atol=1e-6)


def test_config4_thermal_dfn_batch_against_the_oracle():
    """BASELINE config 4: DFN + lumped thermal on 40-point meshes (3 523 states; one 512-thread team per SM, L+U values in shared
    memory, everything streamed in L2), a C-rate sweep of two waves of teams.  Same bars as configs 1-3.  The checker here is the
    oracle's static sparse LU (the pivoting banded LU would be a dense 3523 x 3523 factorisation per system)."""
    import os

    import bench

    b = util.bundle("dfn_thermal40")
    n = 296
    p = bench.sample_inputs(b, n, np.random.default_rng(20264))
    r = gpu("dfn_thermal40").solve_arrays(util.T_EVAL, p, t_interp=util.T_INTERP)
    nt = np.asarray(r["n_t"])
    assert set(np.unique(r["flag"]).tolist()) <= {1, 2}
    t_end = np.array([r["t"][i, int(nt[i]) - 1] for i in range(n)])
    par, _ = bench.parity_block(b, p, nt, np.asarray(r["flag"]), np.asarray(r["y"])[:, :, 0], t_end, n, os.cpu_count() or 1)
    print({k: v for k, v in par.items() if k != "bars"})
    assert par["same_termination"]
    assert par["max_abs_dt_end_gpu_vs_oracle"] <= util.T_EVENT_TOL and par["max_abs_dt_end_gpu_vs_truth"] <= util.T_EVENT_TOL
    assert par["systems_failing_the_bar"] == 0, par
    # a larger current heats the cell more: the final temperature state grows with the current among full discharges
    cur = p[:, 0]
    done = np.nonzero(r["flag"] == 2)[0]
    temp = r["y_term"][done, b.n_diff - 1]
    order = np.argsort(cur[done])
    assert np.all(np.diff(temp[order]) > 0)


@pytest.mark.parametrize("name", ["spm", "spme"])
def test_cuda_path_matches_scipy_radau_anchor(name):
    # an integrator that is neither IDA nor ours (tests/test_anchors.py, tools/make_scipy_anchors.py)
    from tests.test_anchors import anchor_inputs, check_against_anchor

    b = util.bundle(name)
    r = gpu(name).solve_arrays(util.T_EVAL, anchor_inputs(b), t_interp=util.T_INTERP)
    n = int(r["n_t"][0])
    check_against_anchor(name, r["t"][0, n - 1], r["flag"][0], r["t"][0, :n], r["y"][0, :n, 0])


def test_batched_observation_on_the_device():
    # idaklu.observe / observe_hermite_interp (observe.cpp:319-354) for a whole batch in one launch, against the host
    # emulation of the same routine and the numpy evaluation of the bundle
    b = util.bundle("spme")
    s = BatchedIDASolver(b, rtol=1e-6, atol=1e-8)
    p = util.sample_inputs(b, 64, np.random.default_rng(21))
    r = s.solve_arrays([0.0, 900.0], p)                       # adaptive steps: y and yp
    assert r["yp"] is not None
    obs = s.observe(r, p)
    for i in (0, 31, 63):
        k = int(r["n_t"][i]) - 1
        ref = b.eval_variable("Voltage [V]", r["t"][i, k], r["y"][i, k], p[i])[0]
        assert abs(obs[i, k, 0] - ref) < 1e-11
    t_obs = np.array([0.0, 10.0, 450.0, 899.9, 900.0, 901.0])
    got = s.observe(r, p, t_obs=t_obs)
    ref = util.emu("spme").observe(r["t"], r["y"], np.minimum(r["n_t"], r["n_max"]), p, yp=r["yp"], t_obs=t_obs)
    np.testing.assert_allclose(got, ref, rtol=0, atol=1e-11)
    fixed = BatchedIDASolver(b, rtol=1e-6, atol=1e-8, output_variables=["Voltage [V]"]).solve_arrays([0.0, 900.0], p, t_interp=t_obs[1:4])
    np.testing.assert_allclose(got[:, 1:4, 0], fixed["y"][:, 1:4, 0], rtol=0, atol=1e-4)
    with pytest.raises(SolverError, match="yp"):
        s.observe(s.solve_arrays([0.0, 900.0], p, t_interp=np.linspace(0, 900, 10)), p, t_obs=t_obs)


# ---- experiment schedules on the device (8f-2, BASELINE config 5) ---------------------------------------------------------
def test_experiment_toy_closed_forms_on_the_device():
    from tests.test_experiment import closed_form, cccv_schedule
    from pybamm_b200.experiment import STEP_EXPERIMENT_EVENT, STEP_FINAL_TIME

    b = util.bundle("kat_cccv")
    s = BatchedIDASolver(b, rtol=1e-9, atol=1e-11)
    Rs = np.linspace(0.05, 0.25, 33)
    p = np.zeros((len(Rs), 5))
    p[:, b.input_names.index("R")] = Rs
    r = s.solve_experiment(cccv_schedule(2), p)
    assert np.all(r["flag"] == 1)
    for i, R in enumerate(Rs):
        dts, qs = closed_form(R)
        np.testing.assert_array_equal(r["step_term"][i][:4], [STEP_EXPERIMENT_EVENT, STEP_FINAL_TIME, STEP_EXPERIMENT_EVENT, STEP_EXPERIMENT_EVENT])
        t_end = 0.0
        for k in range(4):
            j = r["step_end"][i, k] - 1
            t_end += dts[k]
            assert abs(r["t"][i, j] - t_end) < 1e-6 and abs(r["y"][i, j, 0] - qs[k]) < 1e-7


def test_dfn_cccv_three_cycles_batch_against_the_reference_loop():
    """BASELINE config 5 at test size: DFN unified-experiment model, 3 x (discharge 1C to 2.5 V / rest 1 h / charge 0.3C to
    4.2 V / hold 4.2 V to C/50) for a batch of input sets in ONE launch; two of them against the reference's step loop restated
    around the CPU oracle (tests/util.oracle_experiment)."""
    from pybamm_b200.experiment import STEP_EXPERIMENT_EVENT, STEP_FINAL_TIME, ExperimentSchedule
    from tests.util import oracle_experiment

    b = util.bundle("dfn_cccv")
    sched = ExperimentSchedule.from_meta(b.meta["experiment"])
    rng = np.random.default_rng(20265)
    n = 64
    p = np.tile(np.array(b.meta["nominal_inputs"]), (n, 1))
    for nm, lo, hi in (("Negative particle diffusivity [m2.s-1]", -14, -13), ("Positive particle diffusivity [m2.s-1]", -15, -14)):
        p[:, b.input_names.index(nm)] = 10 ** rng.uniform(lo, hi, n)
    p[:, b.input_names.index("Initial concentration in electrolyte [mol.m-3]")] = rng.uniform(900, 1100, n)
    s = BatchedIDASolver(b, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]", "Current [A]"])
    r = s.solve_experiment(sched, p)
    assert np.all(r["flag"] == 1), np.unique(r["flag"], return_counts=True)
    expect = [STEP_EXPERIMENT_EVENT, STEP_FINAL_TIME, STEP_EXPERIMENT_EVENT, STEP_EXPERIMENT_EVENT] * 3
    for i in range(n):
        np.testing.assert_array_equal(r["step_term"][i], expect)
        ends = r["step_end"][i] - 1
        np.testing.assert_allclose(r["y"][i, ends[0::4], 0], 2.5, atol=1e-6)         # every discharge ends at the cut-off
        np.testing.assert_allclose(r["y"][i, ends[2::4], 0], 4.2, atol=1e-6)         # every charge at 4.2 V
        np.testing.assert_allclose(r["y"][i, ends[3::4], 1], -0.1, atol=1e-6)        # every hold at C/50
    ref = oracle_experiment("dfn_cccv", sched, p[:2], rtol=1e-8, atol=1e-8)
    for i in range(2):
        np.testing.assert_array_equal(r["step_end"][i], ref[i]["step_end"])
        k = int(r["n_t"][i])
        # the hold steps end where |dI/dt| ~ 1e-6 A/s: at rtol = atol = 1e-8 their event times are defined to ~1e-2 s only,
        # and every later step inherits the shift
        np.testing.assert_allclose(r["t"][i, :k], ref[i]["t"], rtol=0, atol=5e-2)
        ends = set(ref[i]["step_end"] - 1)
        inner = np.array([j for j in range(k) if j not in ends])
        assert np.max(np.abs(r["y"][i, inner, 0] - ref[i]["v"][inner, 0])) < 5e-6      # voltage over 3 cycles (errors of event times add up)
