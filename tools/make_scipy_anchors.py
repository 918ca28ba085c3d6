"""Independent anchors for the ODE models (SPM, SPMe): scipy's Radau and BDF integrators -- neither is IDA, neither shares a
line with oracle/ida_oracle.c or the device integrator -- driven by the numpy evaluation of the frozen bundles' scalar DAG
(ModelBundle.graph.evaluate: no generated C / CUDA code involved either).  BASELINE.md section 5 lists the survey's values for
SPM (3567.860565 s at I = 5 A, D_n = 3.3e-14, 2.5 V cut-off), reproduced here; the SPMe bundle of BASELINE config 2 carries the
3.0 V cut-off, so its anchor is computed for that bundle.

Writes tests/golden/scipy_anchors.json.  Usage: python -m tools.make_scipy_anchors
"""
from __future__ import annotations

import json
import os

import numpy as np
import scipy.sparse as sp
from scipy.integrate import solve_ivp

from tests import util

P_NOM = {"Current function [A]": 5.0, "Negative particle diffusivity [m2.s-1]": 3.3e-14}
T_SAMPLE = [600.0, 1800.0, 3000.0]


def run(name, method, tol):
    b = util.bundle(name)
    assert b.n_alg == 0
    p = np.array([P_NOM[n] for n in b.input_names])
    y0 = b.eval_ic(p)
    f = lambda t, y: b.eval_rhs_alg(t, y, p)          # noqa: E731
    jac = lambda t, y: sp.csc_matrix(b.eval_jac(t, y, p))   # noqa: E731
    k_min = b.event_names.index("Minimum voltage [V]")

    def ev(t, y):
        return float(b.eval_events(t, y, p)[k_min])

    ev.terminal, ev.direction = True, -1
    s = solve_ivp(f, (0.0, 3700.0), y0, method=method, rtol=tol, atol=tol, jac=jac, events=ev, dense_output=True)
    t_event = float(s.t_events[0][0]) if len(s.t_events[0]) else None
    v = [float(b.eval_variable("Voltage [V]", t, s.sol(t), p)[0]) for t in T_SAMPLE]
    return {"t_event": t_event, "steps": int(len(s.t)), "v": v}


if __name__ == "__main__":
    out = {"inputs": P_NOM, "t_sample": T_SAMPLE, "models": {}}
    for name in ("spm", "spme"):
        out["models"][name] = {}
        for method, tol in (("Radau", 1e-10), ("BDF", 1e-10)):
            r = run(name, method, tol)
            out["models"][name][method] = r
            print(name, method, r, flush=True)
    path = os.path.join(util.GOLDEN, "scipy_anchors.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print("->", path)
