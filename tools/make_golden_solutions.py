"""Generate tests/golden/<model>_solve.npz: voltage curves and end times of seeded input sets.

The reference IDAKLU cannot run in this environment, so the fixtures come from the
CPU oracle (oracle/ida_oracle.c) at the north-star tolerances rtol = atol = 1e-8 ("ref") and, as a
tolerance-refinement anchor, at rtol = atol = 1e-11 ("truth").  Both must agree to well inside the
parity bars (1e-6 V, 1e-3 s) for the fixture to be written.

Usage: python -m tools.make_golden_solutions [model ...]
"""
from __future__ import annotations

import os
import sys

import numpy as np

from tests import util

N_SETS = 8


def make(name):
    b, o = util.bundle(name), util.oracle(name)
    rng = np.random.default_rng(20260)
    p = util.sample_inputs(b, N_SETS, rng)
    out = {"p": p, "t_eval": np.asarray(util.T_EVAL), "t_interp": util.T_INTERP}
    for tag, tol in (("ref", 1e-8), ("truth", 1e-11)):
        r = o.solve(util.T_EVAL, p, t_interp=util.T_INTERP, rtol=tol, atol=tol, save_vars=True, n_threads=8)
        n_max = max(len(t) for t in r["t"])
        T = np.full((N_SETS, n_max), np.nan)
        V = np.full((N_SETS, n_max), np.nan)
        for i in range(N_SETS):
            T[i, :len(r["t"][i])] = r["t"][i]
            V[i, :len(r["t"][i])] = r["y"][i][:, 0]
        out[f"{tag}_t"], out[f"{tag}_v"] = T, V
        out[f"{tag}_n"], out[f"{tag}_flag"] = r["n_out"].copy(), r["flag"].copy()
        out[f"{tag}_stats"] = r["stats"].copy()
    worst_v = worst_t = 0.0
    for i in range(N_SETS):
        n = int(out["ref_n"][i])
        dv, dt = util.compare_voltage(out["ref_t"][i, :n], out["ref_v"][i, :n], out["truth_t"][i, :n], out["truth_v"][i, :n],
                                      out["ref_flag"][i], out["truth_flag"][i])
        worst_v, worst_t = max(worst_v, dv), max(worst_t, dt)
    print(f"{name}: 1e-8 vs 1e-11: max|dV| {worst_v:.2e} V, max|dt_end| {worst_t:.2e} s; flags {out['ref_flag']}")
    assert worst_v < 0.5 * util.V_TOL and worst_t < 0.5 * util.T_EVENT_TOL
    np.savez_compressed(os.path.join(util.GOLDEN, f"{name}_solve.npz"), **out)


if __name__ == "__main__":
    for m in sys.argv[1:] or util.BATTERY:
        make(m)
