"""Freeze the BASELINE.json model configurations into ModelBundles and golden vectors.

Runs only where /root/reference is mounted (this container): imports the *reference*
pybamm (through tools/pybamm_env.py stubs), builds + discretises each model, lowers it
(pybamm_b200.bundle) and writes

  pybamm_b200/models/<config>.npz       the frozen bundle (travels to the GPU box)
  tests/golden/<config>_eval.npz        golden vectors produced by the reference's own
                                        Symbol.evaluate / pybamm.Jacobian at seeded points:
                                        rhs_alg, events, voltage, initial conditions and the
                                        sparse Jacobian (dense-evaluated at its structural pattern)

Usage: python -m tools.freeze_models [config ...]
"""
from __future__ import annotations

import os
import sys

import numpy as np
import scipy.sparse as sp

from tools.build_models import CONFIGS, EXPERIMENTS, build, build_experiment, nominal_inputs
from pybamm_b200.bundle import bundle_from_pybamm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def freeze(config: str, n_points=6, seed=20260):
    schedule = None
    if config in EXPERIMENTS:
        pybamm, model, names, schedule = build_experiment(config)
        b = bundle_from_pybamm(model, inputs=names, variables=("Voltage [V]", "Current [A]"), name=config)
        b.meta["experiment"] = schedule.to_meta()
    else:
        pybamm, model, names = build(config)
        b = bundle_from_pybamm(model, inputs=names, name=config)
    nom = nominal_inputs(config)
    b.meta["nominal_inputs"] = [float(nom[n]) for n in b.input_names]
    b.meta["config"] = config
    os.makedirs(os.path.join(ROOT, "pybamm_b200", "models"), exist_ok=True)
    b.save(os.path.join(ROOT, "pybamm_b200", "models", f"{config}.npz"))

    # golden vectors from the reference's own evaluation
    rng = np.random.default_rng(seed)
    rhs_alg = pybamm.NumpyConcatenation(model.concatenated_rhs, model.concatenated_algebraic) \
        if model.concatenated_algebraic.shape[0] else model.concatenated_rhs
    y_sym = pybamm.StateVector(slice(0, b.n_states))
    jac_sym = rhs_alg.jac(y_sym)
    events = [ev.expression for ev in model.events if "TERMINATION" in str(ev.event_type)]
    volt = model.get_processed_variable_or_event("Voltage [V]")
    P, Y, T, F, G, V, JV, IC = [], [], [], [], [], [], [], []
    p_nom = np.array(b.meta["nominal_inputs"])
    for k in range(n_points):
        p = p_nom * (1.0 + (0.2 * rng.uniform(-1, 1, len(p_nom)) if k else 0.0))
        if schedule is not None:      # the schedule's own inputs are selectors: take them from its rows, never perturbed
            row = schedule.over_val[k % schedule.n_steps]
            for nm, v in zip(schedule.over_names, row):
                p[b.input_names.index(nm)] = v
        inputs = {n: float(v) for n, v in zip(b.input_names, p)}
        y0 = model.concatenated_initial_conditions.evaluate(0.0, None, None, inputs).reshape(-1)
        noise = rng.standard_normal(len(y0))
        for amp in ((0.003,) if schedule is not None else (0.03, 0.01, 0.003, 0.001)):      # an unphysical state overflows sinh
            y = y0 * (1.0 + (amp * noise if k else 0.0))
            if np.all(np.isfinite(rhs_alg.evaluate(float(100.0 * k), y, None, inputs))):
                break
        t = float(100.0 * k)
        f = rhs_alg.evaluate(t, y, None, inputs).reshape(-1)
        g = np.array([float(np.asarray(e.evaluate(t, y, None, inputs)).reshape(-1)[0]) for e in events])
        v = np.asarray(volt.evaluate(t, y, None, inputs)).reshape(-1)
        J = sp.csc_matrix(jac_sym.evaluate(t, y, None, inputs))
        jv = np.asarray(J[b.jac_rows, b.jac_cols]).reshape(-1)
        # every non-zero of the reference Jacobian must lie inside the bundle's pattern
        Jp = sp.csc_matrix((np.ones(len(b.jac_rows)), (b.jac_rows, b.jac_cols)), shape=J.shape)
        outside = J - J.multiply(Jp)
        assert abs(outside).max() == 0.0, "reference Jacobian has entries outside the bundle pattern"
        P.append(p); Y.append(y); T.append(t); F.append(f); G.append(g); V.append(v); JV.append(jv); IC.append(y0)
    os.makedirs(os.path.join(ROOT, "tests", "golden"), exist_ok=True)
    np.savez_compressed(
        os.path.join(ROOT, "tests", "golden", f"{config}_eval.npz"),
        p=np.array(P), y=np.array(Y), t=np.array(T), rhs_alg=np.array(F), events=np.array(G),
        voltage=np.array(V), jac_vals=np.array(JV), ic=np.array(IC),
        jac_rows=b.jac_rows, jac_cols=b.jac_cols, input_names=np.array(b.input_names),
    )
    return b


if __name__ == "__main__":
    cfgs = sys.argv[1:] or ["spm", "spme", "dfn"]
    for c in cfgs:
        b = freeze(c)
        print(c, "N", b.n_states, "nnz", len(b.jac_nodes), "graph nodes", len(b.graph))
