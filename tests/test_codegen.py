"""Code generator parity (CPU tier): the lowered models reproduce the reference's own evaluation.

tests/golden/<model>_eval.npz were produced by the *reference* pybamm (Symbol.evaluate and the symbolic
pybamm.Jacobian, tools/freeze_models.py) at seeded (t, y, p) points.  Checked here:
  scalar DAG (ir.Graph.evaluate)  ==  golden
  generated C dialect (the oracle's model functions)  ==  golden
  generated CUDA dialect compiled for the host (tests/emu), incl. the scatter into the static LU slots  ==  golden
  static sparse LU tables  ==  scipy.sparse.linalg.spsolve
"""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from pybamm_b200 import codegen, program, symbolic_lu
from tests import util

RTOL = 1e-12  # relative to the vector's scale: different summation order than numpy's sparse mat-vecs


def _golden(name):
    return np.load(f"{util.GOLDEN}/{name}_eval.npz")


def _close(got, ref, rtol=RTOL):
    scale = max(1.0, float(np.max(np.abs(ref))))
    np.testing.assert_allclose(got, ref, rtol=0, atol=rtol * scale)


@pytest.mark.parametrize("name", util.BATTERY + util.LARGE)
def test_graph_matches_reference_evaluation(name):
    z, b = _golden(name), util.bundle(name)
    assert list(z["input_names"]) == b.input_names
    np.testing.assert_array_equal(z["jac_rows"], b.jac_rows)
    np.testing.assert_array_equal(z["jac_cols"], b.jac_cols)
    for k in range(len(z["t"])):
        t, y, p = float(z["t"][k]), z["y"][k], z["p"][k]
        _close(b.eval_rhs_alg(t, y, p), z["rhs_alg"][k])
        _close(b.eval_events(t, y, p), z["events"][k])
        _close(b.eval_variable("Voltage [V]", t, y, p), z["voltage"][k])
        _close(b.eval_ic(p), z["ic"][k])
        _close(b.graph.evaluate(b.jac_nodes, t, y, p), z["jac_vals"][k], 1e-11)


@pytest.mark.parametrize("name", util.BATTERY + util.LARGE)
def test_c_dialect_matches_reference_evaluation(name):
    z, o = _golden(name), util.oracle(name)
    b = util.bundle(name)
    diag = ((b.jac_rows == b.jac_cols) & (b.jac_rows < b.n_diff)).astype(float)
    for k in range(len(z["t"])):
        t, y, p = float(z["t"][k]), z["y"][k], z["p"][k]
        _close(o.eval("res", y, p, t), z["rhs_alg"][k])
        _close(o.eval("events", y, p, t)[:b.n_events], z["events"][k])
        _close(o.eval("vars", y, p, t)[:1], z["voltage"][k])
        _close(o.eval("ic", None, p), z["ic"][k])
        cj = 0.37
        _close(o.eval("jac", y, p, t, cj), z["jac_vals"][k] - cj * diag, 1e-11)


@pytest.mark.parametrize("name", util.BATTERY)
def test_cuda_dialect_matches_reference_evaluation(name):
    z, e = _golden(name), util.emu(name)
    b = util.bundle(name)
    diag = ((b.jac_rows == b.jac_cols) & (b.jac_rows < b.n_diff)).astype(float)
    for k in range(len(z["t"])):
        t, y, p = float(z["t"][k]), z["y"][k], z["p"][k]
        _close(e.eval("res", y, p, t)[0], z["rhs_alg"][k])
        _close(e.eval("events", y, p, t)[0][:b.n_events], z["events"][k])
        _close(e.eval("vars", y, p, t)[0][:1], z["voltage"][k])
        _close(e.eval("ic", None, p)[0], z["ic"][k])
        cj = 0.37
        _close(e.eval("jac", y, p, t, cj)[0], z["jac_vals"][k] - cj * diag, 1e-11)


@pytest.mark.parametrize("name", util.BATTERY)
def test_static_lu_solves_like_scipy(name):
    z, e, b = _golden(name), util.emu(name), util.bundle(name)
    rng = np.random.default_rng(3)
    diag = ((b.jac_rows == b.jac_cols) & (b.jac_rows < b.n_diff)).astype(float)
    for k in (0, 3):
        for cj in (1e-3, 1.0, 250.0):
            rhs = rng.standard_normal(b.n_states)
            x, rc = e.eval("solve", z["y"][k], z["p"][k], float(z["t"][k]), cj, rhs)
            assert rc == 0
            A = sp.csc_matrix((z["jac_vals"][k] - cj * diag, (b.jac_rows, b.jac_cols)), shape=(b.n_states,) * 2)
            xr = spla.spsolve(A, rhs)
            assert np.max(np.abs(x - xr)) <= 1e-9 * np.max(np.abs(xr))


def test_symbolic_lu_tables_numpy_emulation():
    b = util.bundle("spme")
    samples = codegen.sample_matrices(b)
    s = symbolic_lu.analyse(b.n_states, b.jac_rows, b.jac_cols, samples)
    for v in samples:
        assert symbolic_lu._validate(s, b.jac_rows, b.jac_cols, np.asarray(v)) < 1e-10
    # every structural entry has a slot, fill slots are disjoint from them
    assert len(set(s.a_slot.tolist())) == len(b.jac_rows)
    assert s.nnz_lu >= len(b.jac_rows)


def test_program_interpreter_equals_graph():
    """The re-rolled staged templates compute exactly what the scalar DAG does (bitwise: same op order)."""
    b = util.bundle("spm")
    z = _golden("spm")
    prog = program.build_program(b.graph, b.rhs_alg, "res")
    y, p, t = z["y"][2], z["p"][2], float(z["t"][2])
    got = program.run_program(prog, t, y, p)
    ref = b.eval_rhs_alg(t, y, p)
    np.testing.assert_array_equal(got, ref)
    ev = program.build_program(b.graph, b.events, "events")
    np.testing.assert_array_equal(program.run_program(ev, t, y, p), b.eval_events(t, y, p))


def test_gradients_against_finite_differences():
    b = util.bundle("spme")
    z = _golden("spme")
    y, p, t = z["y"][1], z["p"][1], float(z["t"][1])
    J = b.eval_jac(t, y, p).toarray()
    rng = np.random.default_rng(0)
    for j in rng.choice(b.n_states, 12, replace=False):
        h = 1e-6 * max(1.0, abs(y[j]))
        yp, ym = y.copy(), y.copy()
        yp[j] += h
        ym[j] -= h
        fd = (b.eval_rhs_alg(t, yp, p) - b.eval_rhs_alg(t, ym, p)) / (2 * h)
        np.testing.assert_allclose(J[:, j], fd, rtol=2e-5, atol=2e-6 * max(1.0, np.max(np.abs(fd))))
