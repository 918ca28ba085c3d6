"""Per-system parity table of the systems that are off: python -m tools.gpu_parity_debug <model> <n>"""
import os
import sys

import numpy as np

import bench
from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b

name, n = sys.argv[1], int(sys.argv[2])
bundle = b.load_bundle(name)
p = bench.sample_inputs(bundle, 10000, np.random.default_rng(20260))[:n]      # bench.py's rank-0 batch
s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"])
r = s.solve_arrays(bench.T_EVAL, p, t_interp=bench.T_INTERP)
cores = os.cpu_count()
o = bench.cpu_oracle(bundle, fast=False)
ref, _ = bench.run_cpu(o, p, cores, "band")
refs, _ = bench.run_cpu(o, p, cores, "sparse")
tru, _ = bench.run_cpu(o, p, cores, "sparse", tol=1e-11)
tru10, _ = bench.run_cpu(o, p, cores, "sparse", tol=1e-10)
print("truth flags", np.unique(tru["flag"], return_counts=True))
for i in range(n):
    k = int(r["n_t"][i])
    ks = [k, len(ref["t"][i]), len(refs["t"][i]), len(tru["t"][i]), len(tru10["t"][i])]
    fl = [int(r["flag"][i]), int(ref["flag"][i]), int(refs["flag"][i]), int(tru["flag"][i]), int(tru10["flag"][i])]
    m = min(ks) - 1
    if m < 1:
        print(i, "short", ks, fl)
        continue
    g, a, a2, t, t10 = r["y"][i, :m, 0], ref["y"][i][:m, 0], refs["y"][i][:m, 0], tru["y"][i][:m, 0], tru10["y"][i][:m, 0]
    e = [np.max(np.abs(g - a)), np.max(np.abs(g - t)), np.max(np.abs(a - t)), np.max(np.abs(a2 - t)), np.max(np.abs(t10 - t))]
    if len(set(ks)) > 1 or len(set(fl)) > 1 or max(e[:4]) > 1e-6:
        where = int(np.argmax(np.abs(g - t)))
        print(f"sys {i:4d} n_t {ks} flags {fl} I={p[i, 0]:.3f}  |g-ref| {e[0]:.2e} |g-tru| {e[1]:.2e} |ref-tru| {e[2]:.2e} |refsparse-tru| {e[3]:.2e} "
              f"|tru10-tru| {e[4]:.2e} at point {where}/{m} V={t[where]:.4f}  t_end gpu {r['t'][i, k - 1]:.6f} ref {ref['t'][i][-1]:.6f} tru {tru['t'][i][-1]:.6f}")
