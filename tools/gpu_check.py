"""First-light GPU check: device evaluation and integration parity against the CPU oracle + timings.
Run on a GPU box:  python -m tools.gpu_check [models...]"""
from __future__ import annotations

import sys
import time

import numpy as np

from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b
from oracle.oracle import Oracle


def sample_inputs(bundle, n, rng):
    cols = []
    for name in bundle.input_names:
        if name == "Current function [A]":
            cols.append(rng.uniform(2.5, 7.5, n))
        elif name == "Negative particle diffusivity [m2.s-1]":
            cols.append(10 ** rng.uniform(-14, -13, n))
        elif name == "Positive particle diffusivity [m2.s-1]":
            cols.append(10 ** rng.uniform(-15, -14, n))
        elif name == "Initial concentration in electrolyte [mol.m-3]":
            cols.append(rng.uniform(900, 1100, n))
        else:
            raise KeyError(name)
    return np.column_stack(cols)


def main(models):
    t_eval = [0.0, 3700.0]
    t_interp = np.linspace(0, 3600, 361)
    for name in models:
        bundle = b.load_bundle(name)
        rng = np.random.default_rng(20260)
        o = Oracle(b.build_oracle(bundle))
        o.set_rcm(bundle.jac_rows, bundle.jac_cols)
        s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"])
        s._ensure()
        print(f"== {name}: team {s.desc.team_size} smem {s.desc.smem_bytes} B nnz_lu {s.desc.nnz_lu}", flush=True)
        n = 16
        p = sample_inputs(bundle, n, rng)
        y = np.array([o.eval("ic", None, p[i]) for i in range(n)]) * (1 + 0.01 * rng.standard_normal((n, bundle.n_states)))
        for w in ("res", "jac", "events", "vars"):
            ref = np.array([o.eval(w, y[i], p[i], 10.0, 0.5) for i in range(n)])
            got = s.eval(w, y, p, 10.0, 0.5)[:, :ref.shape[1]]
            print(f"   eval {w}: max abs diff {np.max(np.abs(ref - got)):.2e} (scale {np.max(np.abs(ref)):.2e})", flush=True)
        import scipy.sparse as sp, scipy.sparse.linalg as spla
        rhs = rng.standard_normal((n, bundle.n_states))
        x = s.eval("solve", y, p, 10.0, 0.5, rhs)
        jv = o.eval("jac", y[0], p[0], 10.0, 0.5)
        A = sp.csc_matrix((jv, (bundle.jac_rows, bundle.jac_cols)), shape=(bundle.n_states,) * 2)
        xr = spla.spsolve(A, rhs[0])
        print(f"   lu solve rel err {np.max(np.abs(x[0] - xr)) / np.max(np.abs(xr)):.2e}", flush=True)
        # integration parity, small batch
        t0 = time.time()
        r = s.solve_arrays(t_eval, p, t_interp=t_interp)
        print(f"   solve x{n}: kernel {s.last_kernel_ms:.1f} ms total {s.last_total_ms:.1f} ms wall {time.time() - t0:.2f}s flags {np.unique(r['flag'])}", flush=True)
        ref = o.solve(t_eval, p, t_interp=t_interp, rtol=1e-8, atol=1e-8, save_vars=True, n_threads=8)
        wv = wt = 0.0
        for i in range(n):
            nt = int(r["n_t"][i])
            if nt != len(ref["t"][i]) or r["flag"][i] != ref["flag"][i]:
                print("   MISMATCH sys", i, nt, len(ref["t"][i]), r["flag"][i], ref["flag"][i])
                continue
            wt = max(wt, abs(r["t"][i, nt - 1] - ref["t"][i][-1]))
            wv = max(wv, float(np.max(np.abs(r["y"][i, :nt - 1, 0] - ref["y"][i][:nt - 1, 0]))))
        print(f"   parity vs oracle: max|dV| {wv:.2e} V  max|dt_end| {wt:.2e} s", flush=True)
        print("   stats gpu[0]", r["stats"][0].tolist(), "oracle[0]", ref["stats"][0].tolist(), flush=True)
        for nb in (1000, 10000):
            pb = sample_inputs(bundle, nb, rng)
            t0 = time.time()
            rr = s.solve_arrays(t_eval, pb, t_interp=t_interp)
            wall = time.time() - t0
            fl, cnt = np.unique(rr["flag"], return_counts=True)
            print(f"   batch {nb}: kernel {s.last_kernel_ms:.1f} ms ({nb / s.last_kernel_ms * 1e3:.0f} sims/s) total {s.last_total_ms:.1f} ms wall {wall:.2f}s "
                  f"flags {dict(zip(fl.tolist(), cnt.tolist()))} mean steps {rr['stats'][:, 0].mean():.0f}", flush=True)
        s.close()


if __name__ == "__main__":
    main(sys.argv[1:] or ["spm", "spme", "dfn"])
