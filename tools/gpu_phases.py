"""Per-phase cycle breakdown of the integration kernel (profiling build, -DPB_PROFILE):
python -m tools.gpu_phases <model> [n_sys]   -- run on a GPU box."""
import sys

import numpy as np

from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b
from tools.gpu_check import sample_inputs

name = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 148
bundle = b.load_bundle(name)
img = b.build_model(bundle, profile=True)
p = sample_inputs(bundle, n, np.random.default_rng(20260))
ref = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"])
r0 = ref.solve_arrays([0.0, 3700.0], p, t_interp=np.linspace(0, 3600, 361))
s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"], image_path=img)
r = s.solve_arrays([0.0, 3700.0], p, t_interp=np.linspace(0, 3600, 361))
c = r["stats"].astype(float).mean(axis=0)
st = r0["stats"].astype(float).mean(axis=0)
res, slv, jac, fac, nls, step, ev, out = c
tot = step + ev + out
print(f"{name} x{n}: kernel {s.last_kernel_ms:.1f} ms (plain build {ref.last_kernel_ms:.1f} ms); mean per system:")
print(f"  counters: steps {st[0]:.0f} res {st[1]:.0f} setups {st[2]:.0f} etf {st[3]:.0f} nni {st[4]:.0f}")
rows = [("residual", res, st[1]), ("tri solves", slv, st[4]), ("jacobian", jac, st[6]), ("factor", fac, st[2]),
        ("newton vector work", nls - res - slv - jac - fac, st[4]), ("step control", step - nls, st[0]),
        ("events/rootfind", ev, st[0]), ("output points", out, 362)]
for nm, cyc, cnt in rows:
    print(f"  {nm:20s} {cyc / 1e6:9.2f} Mcyc  {100 * cyc / tot:5.1f} %   {cyc / max(cnt, 1):9.0f} cyc/call")
print(f"  {'total (instrumented)':20s} {tot / 1e6:9.2f} Mcyc")
