"""BASELINE config 4 (DFN + lumped thermal, 40-point meshes, 3 523 states) on the GPU against the CPU oracle.
Run on a GPU box:  python -m tools.gpu_thermal [n_parity] [n_batch]"""
import sys
import time

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from oracle.oracle import Oracle
from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b

NAME = "dfn_thermal40"
n_par = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n_batch = int(sys.argv[2]) if len(sys.argv) > 2 else 592
bundle = b.load_bundle(NAME)
rng = np.random.default_rng(20260)
o = Oracle(b.build_oracle(bundle, fast=True))
o.set_sparse(*b.lu_order(bundle))
s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"])
s._ensure()
print(f"== {NAME}: team {s.desc.team_size} smem {s.desc.smem_bytes} B nnz_lu {s.desc.nnz_lu}", flush=True)
p = rng.uniform(2.5, 7.5, (n_par, 1))
y = np.array([o.eval("ic", None, p[i]) for i in range(n_par)]) * (1 + 0.003 * rng.standard_normal((n_par, bundle.n_states)))
for w in ("res", "jac", "events", "vars"):
    ref = np.array([o.eval(w, y[i], p[i], 10.0, 0.5) for i in range(n_par)])
    got = s.eval(w, y, p, 10.0, 0.5)[:, :ref.shape[1]]
    print(f"   eval {w}: max abs diff {np.max(np.abs(ref - got)):.2e} (scale {np.max(np.abs(ref)):.2e})", flush=True)
rhs = rng.standard_normal((n_par, bundle.n_states))
x = s.eval("solve", y, p, 10.0, 0.5, rhs)
jv = o.eval("jac", y[0], p[0], 10.0, 0.5)
A = sp.csc_matrix((jv, (bundle.jac_rows, bundle.jac_cols)), shape=(bundle.n_states,) * 2)
xr = spla.spsolve(A, rhs[0])
print(f"   lu solve rel err {np.max(np.abs(x[0] - xr)) / np.max(np.abs(xr)):.2e}", flush=True)
te, ti = [0.0, 3700.0], np.linspace(0, 3600, 361)
t0 = time.time()
r = s.solve_arrays(te, p, t_interp=ti)
print(f"   solve x{n_par}: kernel {s.last_kernel_ms:.1f} ms wall {time.time() - t0:.2f}s flags {np.unique(r['flag'])} steps {r['stats'][:, 0].tolist()}", flush=True)
t0 = time.time()
ref = o.solve(te, p, t_interp=ti, rtol=1e-8, atol=1e-8, save_vars=True, n_threads=16, lu="sparse")
dt_cpu = time.time() - t0
print(f"   oracle x{n_par}: {dt_cpu:.1f} s on 16 threads ({n_par / dt_cpu:.2f} sims/s), steps {[int(q[0]) for q in ref['stats']]}", flush=True)
wv = wt = 0.0
for i in range(n_par):
    nt = int(r["n_t"][i])
    if nt != len(ref["t"][i]) or r["flag"][i] != ref["flag"][i]:
        print("   MISMATCH sys", i, nt, len(ref["t"][i]), r["flag"][i], ref["flag"][i])
        continue
    wt = max(wt, abs(r["t"][i, nt - 1] - ref["t"][i][-1]))
    wv = max(wv, float(np.max(np.abs(r["y"][i, :nt - 1, 0] - ref["y"][i][:nt - 1, 0]))))
print(f"   parity vs oracle: max|dV| {wv:.2e} V  max|dt_end| {wt:.2e} s", flush=True)
pb = rng.uniform(2.5, 7.5, (n_batch, 1))
for rep in range(2):
    rr = s.solve_arrays(te, pb, t_interp=ti)
    fl, cnt = np.unique(rr["flag"], return_counts=True)
    print(f"   batch {n_batch}: kernel {s.last_kernel_ms:.1f} ms ({n_batch / s.last_kernel_ms * 1e3:.1f} sims/s) flags {dict(zip(fl.tolist(), cnt.tolist()))} "
          f"mean steps {rr['stats'][:, 0].mean():.0f}", flush=True)
