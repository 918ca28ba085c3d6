"""Known-answer tests of the reference, restated on the scalar DAG (tests/kat_models.py), run through
the CPU oracle and the host emulation of the device integrator (CPU tier).  The GPU tier runs the same
cases through the C ABI (tests/test_gpu.py).  Tolerances are the reference tests' own.
"""
import numpy as np
import pytest

from tests import util

SOLVERS = ("oracle", "emu")


def _solve(kind, name, t_eval, inputs, t_interp=None, rtol=1e-4, atol=1e-6):
    h = util.oracle(name) if kind == "oracle" else util.emu(name)
    inputs = np.atleast_2d(np.asarray(inputs, dtype=float))
    if inputs.size == 0:
        inputs = np.zeros((1, 0))
    if kind == "oracle":
        return h.solve(t_eval, inputs, t_interp=t_interp, rtol=rtol, atol=atol)
    return h.solve(t_eval, inputs if inputs.shape[1] else np.zeros((inputs.shape[0], 1)), t_interp=t_interp, rtol=rtol, atol=atol)


@pytest.mark.parametrize("kind", SOLVERS)
def test_roberts_event_at_two(kind):
    # test_idaklu_solver.py:92-129 (default tolerances, adaptive steps stored)
    r = _solve(kind, "kat_roberts", [0, 3], np.zeros((1, 0)))
    t, y = r["t"][0], r["y"][0]
    assert r["flag"][0] == 2
    np.testing.assert_allclose(t[-1], 2.0, rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(y[-1, 0], 0.2, rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(y[:, 1], np.ones_like(t), rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(y[:, 0], 0.1 * t, rtol=1e-7, atol=1e-6)


@pytest.mark.parametrize("kind", SOLVERS)
def test_multiple_inputs_decay(kind):
    # test_idaklu_solver.py:131-170
    t_interp = np.linspace(0, 1, 10)
    rates = 0.01 * (np.arange(8) + 1)
    r = _solve(kind, "kat_decay", [0, 1], rates[:, None], t_interp=t_interp, rtol=1e-5, atol=1e-5)
    for i, rate in enumerate(rates):
        np.testing.assert_array_equal(r["t"][i], t_interp)
        np.testing.assert_allclose(r["y"][i][:, 0], 2 * np.exp(-rate * t_interp), atol=1e-4, rtol=1e-4)
        assert r["flag"][i] in (0, 1)


@pytest.mark.parametrize("kind", SOLVERS)
def test_exponential_decay_tight(kind):
    # pybammsolvers tests/test_integration.py:118-138,203-234 (exp decay, 1e-5 at rtol=atol=1e-8)
    t_interp = np.linspace(0, 1, 50)
    r = _solve(kind, "kat_decay", [0, 1], [[1.0]], t_interp=t_interp, rtol=1e-8, atol=1e-8)
    np.testing.assert_allclose(r["y"][0][:, 0], 2 * np.exp(-t_interp), rtol=1e-5, atol=1e-8)


@pytest.mark.parametrize("kind", SOLVERS)
def test_growth_with_and_without_event(kind):
    # test_idaklu_solver.py:172-231
    t_interp = np.linspace(0, 1, 100)
    r = _solve(kind, "kat_growth", [0, 1], np.zeros((1, 0)), t_interp=t_interp, rtol=1e-8, atol=1e-8)
    np.testing.assert_array_equal(r["t"][0], t_interp)
    np.testing.assert_allclose(r["y"][0][:, 0], np.exp(0.1 * t_interp), rtol=1e-6, atol=1e-5)
    r = _solve(kind, "kat_growth_ev", [0, 1], np.zeros((1, 0)), t_interp=t_interp, rtol=1e-8, atol=1e-8)
    t = r["t"][0]
    assert r["flag"][0] == 2 and len(t) < len(t_interp)
    np.testing.assert_allclose(r["y"][0][:, 0], np.exp(0.1 * t), rtol=1e-6, atol=1e-5)
    np.testing.assert_allclose(t[-1], np.log(1.01) / 0.1, rtol=0, atol=1e-6)


@pytest.mark.parametrize("kind", SOLVERS)
def test_growth_dae_two_events(kind):
    # test_idaklu_solver.py:233-275
    r = _solve(kind, "kat_growth_dae", [0, 5], np.zeros((1, 0)), rtol=1e-8, atol=1e-8)
    t, y = r["t"][0], r["y"][0]
    assert r["flag"][0] == 2
    assert np.all(y[:-1, 0] < 1.5) and np.all(y[:-1, -1] < 2.5)
    np.testing.assert_array_equal(r["y_term"][0], y[-1])
    np.testing.assert_allclose(y[:, 0], np.exp(0.1 * t), rtol=1e-6, atol=1e-5)
    np.testing.assert_allclose(y[:, -1], 2 * np.exp(0.1 * t), rtol=1e-6, atol=1e-5)
    np.testing.assert_allclose(t[-1], np.log(1.25) / 0.1, rtol=0, atol=1e-5)


@pytest.mark.parametrize("kind", SOLVERS)
def test_scalar_and_vector_inputs(kind):
    # test_idaklu_solver.py:277-331
    t_interp = np.linspace(0, 3, 100)
    r = _solve(kind, "kat_inputs", [0, 3], [[0.1, 0.2, 0.3]], t_interp=t_interp)
    t, y = r["t"][0], r["y"][0]
    np.testing.assert_allclose(y[:, 3], np.ones_like(t), rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(y[:, 0], 0.1 * t, rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(y[:, 1], 0.2 * t, rtol=1e-7, atol=1e-6)
    np.testing.assert_allclose(y[:, 2], 0.3 * t, rtol=1e-7, atol=1e-6)


@pytest.mark.parametrize("kind", SOLVERS)
def test_time_grid_semantics(kind):
    # packages/pybamm/tests/integration/test_solvers/test_idaklu.py:123-184: t_eval breakpoints are always stored,
    # t_interp points in between, duplicates removed, adaptive steps when t_interp is empty
    t_eval = [0.0, 0.4, 1.0]
    t_interp = np.array([0.1, 0.4, 0.7])
    r = _solve(kind, "kat_decay", t_eval, [[0.5]], t_interp=t_interp, rtol=1e-6, atol=1e-8)
    np.testing.assert_array_equal(r["t"][0], [0.0, 0.1, 0.4, 0.7, 1.0])
    np.testing.assert_allclose(r["y"][0][:, 0], 2 * np.exp(-0.5 * r["t"][0]), rtol=1e-5)
    r = _solve(kind, "kat_decay", t_eval, [[0.5]], rtol=1e-6, atol=1e-8)
    t = r["t"][0]
    assert t[0] == 0.0 and t[-1] == 1.0 and 0.4 in t and len(t) > 5 and np.all(np.diff(t) > 0)
