"""Small known-answer models of the reference's own IDAKLU tests, built directly on the scalar DAG
(no pybamm needed), so that the oracle, the host emulation and the CUDA path can all be pinned
against the analytic answers those tests assert.

Each builder cites the reference test it restates
(packages/pybamm/tests/unit/test_solvers/test_idaklu_solver.py, packages/pybammsolvers/tests/test_integration.py).
"""
from __future__ import annotations

import numpy as np

from pybamm_b200 import ir
from pybamm_b200.bundle import ModelBundle


def _finish(b: ModelBundle, name, rhs_alg, ic, n_diff, events=(), event_names=(), variables=None,
            input_names=(), input_sizes=(), nominal=()):
    b.name = name
    b.n_states = len(rhs_alg)
    b.n_diff = n_diff
    b.rhs_alg = np.asarray(rhs_alg, np.int64)
    b.ic = np.asarray(ic, np.int64)
    b.events = np.asarray(list(events), np.int64)
    b.event_names = list(event_names)
    b.input_names = list(input_names)
    b.input_sizes = list(input_sizes)
    b.graph.n_states = b.n_states
    for k, v in (variables or {}).items():
        b.variables[k] = np.asarray(v, np.int64)
    b.meta["nominal_inputs"] = [float(x) for x in nominal]
    return b.finalize()


def roberts():
    """test_idaklu_solver.py:92-129: u' = 0.1 v, 0 = 1 - v, events 0.2 - u and v; event at t = 2, u = 0.2."""
    b = ModelBundle()
    g = b.graph
    u, v = g.state(0), g.state(1)
    rhs = [g.mul(g.const(0.1), v), g.sub(g.const(1.0), v)]
    ic = [g.const(0.0), g.const(1.0)]
    ev = [g.sub(g.const(0.2), u), v]
    return _finish(b, "kat_roberts", rhs, ic, 1, ev, ["1", "2"], {"u": [u]})


def decay():
    """test_idaklu_solver.py:131-170 (and test_integration.py:118-138): var' = -rate var, var(0) = 2."""
    b = ModelBundle()
    g = b.graph
    var, rate = g.state(0), g.input(0)
    rhs = [g.neg(g.mul(rate, var))]
    return _finish(b, "kat_decay", rhs, [g.const(2.0)], 1, variables={"var": [var]},
                   input_names=["rate"], input_sizes=[1], nominal=[0.05])


def growth(event_at=None):
    """test_idaklu_solver.py:172-231: var' = 0.1 var, var(0) = 1; optional event event_at - var."""
    b = ModelBundle()
    g = b.graph
    var = g.state(0)
    rhs = [g.mul(g.const(0.1), var)]
    ev, names = [], []
    if event_at is not None:
        ev, names = [g.sub(g.const(float(event_at)), var)], ["an event"]
    name = "kat_growth" if event_at is None else "kat_growth_ev"
    return _finish(b, name, rhs, [g.const(1.0)], 1, ev, names, {"var": [var]})


def growth_dae(n=12):
    """test_idaklu_solver.py:233-275: var1' = 0.1 var1 (n nodes), 0 = 2 var1 - var2, events min(1.5 - var1), min(2.5 - var2)."""
    b = ModelBundle()
    g = b.graph
    v1 = [g.state(i) for i in range(n)]
    v2 = [g.state(n + i) for i in range(n)]
    rhs = [g.mul(g.const(0.1), x) for x in v1] + [g.sub(g.mul(g.const(2.0), a), c) for a, c in zip(v1, v2)]
    ic = [g.const(1.0)] * n + [g.const(2.0)] * n

    def gmin(nodes):
        acc = nodes[0]
        for x in nodes[1:]:
            acc = g.binary(ir.MIN, acc, x)
        return acc

    ev = [gmin([g.sub(g.const(1.5), x) for x in v1]), gmin([g.sub(g.const(2.5), x) for x in v2])]
    return _finish(b, "kat_growth_dae", rhs, ic, n, ev, ["var1 = 1.5", "var2 = 2.5"], {"var1": v1})


def input_params():
    """test_idaklu_solver.py:277-331: u1' = a v, u2' = b[0], u3' = b[1], 0 = 1 - v (scalar + vector inputs)."""
    b = ModelBundle()
    g = b.graph
    v = g.state(3)
    a, b0, b1 = g.input(0), g.input(1), g.input(2)
    rhs = [g.mul(a, v), b0, b1, g.sub(g.const(1.0), v)]
    ic = [g.const(0.0)] * 3 + [g.const(1.0)]
    return _finish(b, "kat_inputs", rhs, ic, 3, variables={"u1": [g.state(0)]},
                   input_names=["a", "b"], input_sizes=[1, 2], nominal=[0.1, 0.2, 0.3])


def cccv_toy():
    """A constant-current / constant-voltage toy with the structure of the unified experiment model
    (simulation.py:269-416): one differential state q, one algebraic "current variable" i whose constraint is selected by the
    input ``mode`` (1: i = value, 2: v = value with v = q - R i), and a combined experiment event selected the same way
    (mode 1: sgn (v - thr), mode 2: |i| - thr).  q' = -i, q(0) = 1.  Every step has a closed form (tests/test_experiment.py)."""
    b = ModelBundle()
    g = b.graph
    q, i = g.state(0), g.state(1)
    R, mode, sgn, thr, value = (g.input(k) for k in range(5))     # sorted names: R, mode, sgn, thr, value
    v = g.sub(q, g.mul(R, i))
    is_cc = g.binary(ir.LE, mode, g.const(1.5))
    rhs = [g.neg(i), g.select(is_cc, g.sub(i, value), g.sub(v, value))]
    ic = [g.const(1.0), g.const(0.0)]
    ev_exp = g.select(is_cc, g.mul(sgn, g.sub(v, thr)), g.sub(g.unary(ir.ABS, i), thr))
    ev = [g.sub(g.const(10.0), q), ev_exp]
    return _finish(b, "kat_cccv", rhs, ic, 1, ev, ["q max", "Combined termination [experiment]"], {"v": [v], "i": [i]},
                   input_names=["R", "mode", "sgn", "thr", "value"], input_sizes=[1] * 5, nominal=[0.1, 1.0, 1.0, 0.5, 1.0])


def hard_ic():
    """An algebraic initial condition a plain Newton cannot reach: u' = -u, 0 = atan(10 (v - 1)), v(0) = 5.  From the flat
    part of the arctangent the full step overshoots by two orders of magnitude; five halvings are not enough, the
    IDACalcIC-like second pass (pb_consistent_init) is."""
    b = ModelBundle()
    g = b.graph
    u, v = g.state(0), g.state(1)
    rhs = [g.neg(u), g.unary(ir.ATAN, g.mul(g.const(10.0), g.sub(v, g.const(1.0))))]
    return _finish(b, "kat_hard_ic", rhs, [g.const(1.0), g.const(5.0)], 1, variables={"v": [v]})


def pivot_flip():
    """A static pivot that dies on the way: u' = -u / 2, 0 = a x + z - 1, 0 = x - z with a = max(2 u - 1, 0) and u(0) = 1,
    so x = z = 1 / (1 + a).  On the natural diagonal the pivot of x is a: one at t = 0, where the pivot order is chosen,
    exactly zero from t = 2 ln 2 on (a pivot that merely passes through zero is stepped over).  KLU's refactorisation would
    notice the condition number and factorise afresh; here the failed system is re-solved on an image whose order is also
    valid for its matrices (BatchedIDASolver._repivot)."""
    b = ModelBundle()
    g = b.graph
    u, x, z = g.state(0), g.state(1), g.state(2)
    a = g.binary(ir.MAX, g.sub(g.mul(g.const(2.0), u), g.const(1.0)), g.const(0.0))
    rhs = [g.mul(g.const(-0.5), u), g.sub(g.add(g.mul(a, x), z), g.const(1.0)), g.sub(x, z)]
    return _finish(b, "kat_pivot_flip", rhs, [g.const(1.0), g.const(0.5), g.const(0.5)], 1, variables={"x": [x]})


ALL = {
    "kat_pivot_flip": pivot_flip,
    "kat_hard_ic": hard_ic,
    "kat_cccv": cccv_toy,
    "kat_roberts": roberts,
    "kat_decay": decay,
    "kat_growth": lambda: growth(None),
    "kat_growth_ev": lambda: growth(1.01),
    "kat_growth_dae": growth_dae,
    "kat_inputs": input_params,
}
