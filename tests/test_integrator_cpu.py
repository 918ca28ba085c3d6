"""Integrator parity on the CPU tier.

* the oracle reproduces its committed golden solutions (tests/golden/<model>_solve.npz,
  tools/make_golden_solutions.py) and those agree with the 1e-11 "truth" run inside the parity bars;
* the host emulation of the *device* integrator (same pbgpu_ida.cuh + CUDA-dialect model code, static sparse LU)
  matches the oracle (independent C restatement, banded LU with partial pivoting) inside the bars.
"""
import numpy as np
import pytest

from tests import util


def _golden(name):
    return np.load(f"{util.GOLDEN}/{name}_solve.npz")


@pytest.mark.parametrize("name", util.BATTERY)
def test_oracle_reproduces_golden(name):
    z, o = _golden(name), util.oracle(name)
    n = 4 if name == "dfn" else len(z["p"])
    r = o.solve(z["t_eval"], z["p"][:n], t_interp=z["t_interp"], rtol=1e-8, atol=1e-8, save_vars=True, n_threads=4)
    for i in range(n):
        m = int(z["ref_n"][i])
        dv, dt = util.compare_voltage(r["t"][i], r["y"][i][:, 0], z["ref_t"][i, :m], z["ref_v"][i, :m], r["flag"][i], z["ref_flag"][i])
        assert dv < 1e-9 and dt < 1e-7   # same code, same inputs: only libm / compiler differences
        np.testing.assert_array_equal(r["stats"][i][:7], z["ref_stats"][i][:7])


@pytest.mark.parametrize("name", util.BATTERY)
def test_golden_is_converged(name):
    z = _golden(name)
    for i in range(len(z["p"])):
        m = int(z["ref_n"][i])
        dv, dt = util.compare_voltage(z["ref_t"][i, :m], z["ref_v"][i, :m], z["truth_t"][i, :m], z["truth_v"][i, :m],
                                      z["ref_flag"][i], z["truth_flag"][i])
        assert dv < 0.5 * util.V_TOL and dt < 0.5 * util.T_EVENT_TOL


@pytest.mark.parametrize("name", util.BATTERY)
def test_device_integrator_emulation_matches_golden(name):
    z, e = _golden(name), util.emu(name)
    n = 3 if name == "dfn" else len(z["p"])
    r = e.solve(z["t_eval"], z["p"][:n], t_interp=z["t_interp"], rtol=1e-8, atol=1e-8, save_vars=True)
    for i in range(n):
        m = int(z["ref_n"][i])
        dv, dt = util.compare_voltage(r["t"][i], r["y"][i][:, 0], z["ref_t"][i, :m], z["ref_v"][i, :m], r["flag"][i], z["ref_flag"][i])
        assert dv < util.V_TOL and dt < util.T_EVENT_TOL
        # step counts of the two implementations stay comparable (at 1e-8 they are round-off sensitive: the LU
        # differs, and so do error-test failures near the noise floor; both show the same spread across inputs)
        assert 0.5 <= int(r["stats"][i][0]) / int(z["ref_stats"][i][0]) <= 2.0


def test_full_state_output_and_y_term():
    b, e, o = util.bundle("spm"), util.emu("spm"), util.oracle("spm")
    p = util.sample_inputs(b, 3, np.random.default_rng(5))
    r = e.solve(util.T_EVAL, p, t_interp=util.T_INTERP, rtol=1e-8, atol=1e-8)
    q = o.solve(util.T_EVAL, p, t_interp=util.T_INTERP, rtol=1e-8, atol=1e-8)
    for i in range(3):
        assert r["y"][i].shape == (len(r["t"][i]), b.n_states)
        np.testing.assert_allclose(r["y"][i], q["y"][i], rtol=0, atol=1e-9)
        np.testing.assert_allclose(r["y_term"][i], r["y"][i][-1], rtol=0, atol=1e-12)


def test_batch_equals_independent_solves():
    # packages/pybamm/tests/integration/test_solvers/test_idaklu.py:244-280
    b, e = util.bundle("spme"), util.emu("spme")
    p = util.sample_inputs(b, 4, np.random.default_rng(11))
    r = e.solve(util.T_EVAL, p, t_interp=util.T_INTERP, rtol=1e-6, atol=1e-8, save_vars=True)
    for i in range(4):
        s = e.solve(util.T_EVAL, p[i:i + 1], t_interp=util.T_INTERP, rtol=1e-6, atol=1e-8, save_vars=True)
        np.testing.assert_array_equal(s["t"][0], r["t"][i])
        np.testing.assert_array_equal(s["y"][0], r["y"][i])


def test_interior_breakpoint_reinitialises():
    # IDAKLUSolverOpenMP.inl:618-645: interior t_eval values are stop times followed by a re-initialisation
    b, e, o = util.bundle("dfn"), util.emu("dfn"), util.oracle("dfn")
    p = np.array([b.meta["nominal_inputs"]])
    t_eval = [0.0, 600.0, 1200.0]
    ti = np.linspace(0, 1200, 25)
    r = e.solve(t_eval, p, t_interp=ti, rtol=1e-8, atol=1e-8, save_vars=True)
    q = o.solve(t_eval, p, t_interp=ti, rtol=1e-8, atol=1e-8, save_vars=True)
    np.testing.assert_array_equal(r["t"][0], ti)
    dv, _ = util.compare_voltage(r["t"][0], r["y"][0][:, 0], q["t"][0], q["y"][0][:, 0], r["flag"][0], q["flag"][0])
    assert dv < util.V_TOL
    one = e.solve([0.0, 1200.0], p, t_interp=ti, rtol=1e-8, atol=1e-8, save_vars=True)
    assert np.max(np.abs(one["y"][0][:, 0] - r["y"][0][:, 0])) < util.V_TOL


def test_failure_is_a_flag_not_an_exception():
    # IDAKLUSolverOpenMP.inl:401-406,443: integration failures end the loop with flag < 0 and the partial solution
    b, e = util.bundle("spm"), util.emu("spm")
    p = util.sample_inputs(b, 1, np.random.default_rng(2))
    r = e.solve(util.T_EVAL, p, t_interp=util.T_INTERP, rtol=1e-8, atol=1e-8, options={"max_num_steps": 20})
    assert r["flag"][0] == -1 and 1 <= len(r["t"][0]) < len(util.T_INTERP)


def test_adaptive_block_overflow_is_counted_not_silently_truncated():
    # ExtendAdaptiveArrays (IDAKLUSolverOpenMP.inl:608-612) grows without bound; the fixed block counts every point
    # so that the host can re-solve with the exact capacity (solver.solve_arrays)
    e = util.emu("kat_decay")
    full = e.solve([0.0, 10.0], np.array([[0.3]]), rtol=1e-8, atol=1e-10)
    n_full = int(full["n_out"][0])
    assert n_full > 20 and len(full["t"][0]) == n_full
    small = e.solve([0.0, 10.0], np.array([[0.3]]), rtol=1e-8, atol=1e-10, n_max=10)
    assert int(small["n_out"][0]) == n_full           # counted past the capacity
    np.testing.assert_array_equal(small["t"][0], full["t"][0][:10])
    assert small["flag"][0] == full["flag"][0] and small["t_event"][0] == full["t"][0][-1]


def test_hermite_derivatives_are_stored_with_the_adaptive_steps():
    # save_hermite (IDAKLUSolverOpenMP.inl:432-435, 447-457): yp = IDAGetDky(t, 1) at every stored step
    e = util.emu("kat_decay")
    rate = 0.3
    r = e.solve([0.0, 10.0], np.array([[rate]]), rtol=1e-9, atol=1e-11, want_yp=True)
    t, y, yp = r["t"][0], r["y"][0][:, 0], r["yp"][0][:, 0]
    np.testing.assert_allclose(y, 2.0 * np.exp(-rate * t), rtol=1e-6)
    # the derivative of the interpolating polynomial satisfies the ODE to integration accuracy
    np.testing.assert_allclose(yp, -rate * y, rtol=1e-4, atol=1e-9)


def test_event_index_and_time_come_back_from_the_kernel():
    # closest_event_idx (idaklu_solver.py:826-834): roberts ends on event 0 ("0.2 - u") at t = 2
    e = util.emu("kat_roberts")
    r = e.solve([0.0, 5.0], np.zeros((1, 0)), t_interp=np.linspace(0, 5, 11), rtol=1e-8, atol=1e-8)
    assert r["flag"][0] == 2 and r["event_idx"][0] == 0
    assert abs(r["t_event"][0] - 2.0) < 1e-6 and r["t_event"][0] == r["t"][0][-1]
    d = util.emu("kat_decay").solve([0.0, 1.0], np.array([[0.1]]), rtol=1e-6, atol=1e-8)
    assert d["flag"][0] in (0, 1) and d["event_idx"][0] == -1


def test_no_progress_guard_ends_a_stalled_integration():
    # NoProgressGuard.hpp:12-66: fewer than t_no_progress seconds over the last num_steps_no_progress steps -> IDA_ERR_FAIL
    e = util.emu("kat_decay")
    ok = e.solve([0.0, 10.0], np.array([[0.3]]), rtol=1e-8, atol=1e-10,
                 options={"num_steps_no_progress": 5, "t_no_progress": 1e-9})
    assert ok["flag"][0] in (0, 1)
    stalled = e.solve([0.0, 10.0], np.array([[0.3]]), rtol=1e-8, atol=1e-10,
                      options={"num_steps_no_progress": 5, "t_no_progress": 1e9})
    assert stalled["flag"][0] == -3 and stalled["t"][0][-1] < 10.0
    off = e.solve([0.0, 10.0], np.array([[0.3]]), rtol=1e-8, atol=1e-10,
                  options={"num_steps_no_progress": 0, "t_no_progress": 1e9})
    assert off["flag"][0] in (0, 1)


def test_yp0_is_used_only_without_calc_ic():
    # SetupInitialState (IDAKLUSolverOpenMP.inl:514-563): with calc_ic the derivatives are recomputed, without it the
    # caller's ydot0 is what IDA starts from
    e = util.emu("kat_decay")
    p = np.array([[0.3]])
    base = e.solve([0.0, 1.0], p, rtol=1e-8, atol=1e-10)
    with_ic = e.solve([0.0, 1.0], p, rtol=1e-8, atol=1e-10, yp0=np.array([[123.0]]))
    np.testing.assert_array_equal(with_ic["t"][0], base["t"][0])
    good = e.solve([0.0, 1.0], p, rtol=1e-8, atol=1e-10, yp0=np.array([[-0.6]]), options={"calc_ic": False})
    np.testing.assert_allclose(good["y"][0][-1, 0], 2.0 * np.exp(-0.3), rtol=1e-6)
    bad = e.solve([0.0, 1.0], p, rtol=1e-8, atol=1e-10, yp0=np.array([[123.0]]), options={"calc_ic": False})
    assert not np.array_equal(bad["t"][0], base["t"][0])      # the wrong derivative changes the first step


def _pad(r, key, n_max, width):
    out = np.zeros((len(r["t"]), n_max) + ((width,) if width else ()))
    for i, a in enumerate(r[key]):
        out[i, :len(a)] = a
    return out


def test_device_observation_matches_the_reference_semantics():
    # observe.cpp:25-64 (Hermite knots), :100-166 (interval search, extrapolation with the last interval), :319-354
    b, e = util.bundle("spm"), util.emu("spm")
    p = util.sample_inputs(b, 2, np.random.default_rng(8))
    r = e.solve([0.0, 900.0], p, rtol=1e-6, atol=1e-8, want_yp=True)          # adaptive steps, y and y' stored
    n_t = np.array([len(t) for t in r["t"]], np.int32)
    n_max = int(n_t.max())
    t, y, yp = _pad(r, "t", n_max, 0), _pad(r, "y", n_max, b.n_states), _pad(r, "yp", n_max, b.n_states)
    # (1) at the stored points: var(t_i, y_i, p)
    obs = e.observe(t, y, n_t, p)
    for i in range(2):
        ref = np.array([b.eval_variable("Voltage [V]", t[i, k], y[i, k], p[i])[0] for k in range(n_t[i])])
        np.testing.assert_allclose(obs[i, :n_t[i], 0], ref, rtol=0, atol=1e-12)
    # (2) Hermite interpolation at new times, including one beyond the last stored point
    t_obs = np.array([0.0, 1.0, 123.4, 450.0, 899.0, 900.0, 905.0])
    got = e.observe(t, y, n_t, p, yp=yp, t_obs=t_obs)
    for i in range(2):
        ti, yi, ypi = t[i, :n_t[i]], y[i, :n_t[i]], yp[i, :n_t[i]]
        for k, to in enumerate(t_obs):
            j = min(max(int(np.searchsorted(ti, to, side="left")) - 1, 0), len(ti) - 2)
            if ti[j + 1] < to:
                j = len(ti) - 2
            H, h = ti[j + 1] - ti[j], to - ti[j]
            c = 3 * (yi[j + 1] - yi[j]) / H ** 2 - (2 * ypi[j] + ypi[j + 1]) / H
            d = 2 * (yi[j] - yi[j + 1]) / H ** 3 + (ypi[j] + ypi[j + 1]) / H ** 2
            yh = yi[j] + ypi[j] * h + c * h * h + d * h ** 3
            ref = b.eval_variable("Voltage [V]", to, yh, p[i])[0]
            assert abs(got[i, k, 0] - ref) < 1e-10, (i, k, got[i, k, 0], ref)
    # the interpolant is as accurate as the integration: against a fixed-grid solve at the same times
    fixed = e.solve([0.0, 900.0], p, t_interp=t_obs[1:5], rtol=1e-6, atol=1e-8, save_vars=True)
    for i in range(2):
        np.testing.assert_allclose(got[i, 1:5, 0], fixed["y"][i][1:5, 0], rtol=0, atol=5e-5)


def test_failed_ic_newton_gets_a_second_pass_with_a_real_line_search():
    # IDAKLUSolverOpenMP.inl:838-850: when the Newton of NonlinearSolver fails, IDACalcIC takes over
    e = util.emu("kat_hard_ic")
    r = e.solve([0.0, 1.0], np.zeros((1, 0)), t_interp=np.linspace(0, 1, 5), rtol=1e-8, atol=1e-10)
    assert r["flag"][0] in (0, 1)
    np.testing.assert_allclose(r["y"][0][:, 1], 1.0, atol=1e-8)          # the algebraic state found its root
    np.testing.assert_allclose(r["y"][0][:, 0], np.exp(-r["t"][0]), rtol=1e-6)
    one = e.solve([0.0, 1.0], np.zeros((1, 0)), t_interp=np.linspace(0, 1, 5), rtol=1e-8, atol=1e-10,
                  options={"max_linesearch_backtracks_ic": 30})
    np.testing.assert_allclose(one["y"][0], r["y"][0], atol=1e-9)
