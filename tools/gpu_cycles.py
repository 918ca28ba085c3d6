"""clock64() cycles per call of the integrator's building blocks, measured in place by the eval kernel (which = 6):
python -m tools.gpu_cycles <model> [n_sys] [image.so]   -- n_sys = 1: one team alone on an SM; 296: two per SM."""
import sys

import numpy as np

from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b
from tools.gpu_check import sample_inputs
from oracle.oracle import Oracle

name = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1
img = sys.argv[3] if len(sys.argv) > 3 else None
bundle = b.load_bundle(name)
rng = np.random.default_rng(20260)
p = sample_inputs(bundle, n, rng)
o = Oracle(b.build_oracle(bundle))
y = np.array([o.eval("ic", None, p[i]) for i in range(n)]) * (1 + 0.01 * rng.standard_normal((n, bundle.n_states)))
s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"], image_path=img)
c = s.eval("cycles", y, p, 10.0, 0.5, rng.standard_normal((n, bundle.n_states)))
m = c.mean(axis=0)
print(f"{name} x{n} team {s.desc.team_size}: cycles/call  res {m[0]:.0f}  lu_solve {m[1]:.0f}  jac {m[2]:.0f}  factor {m[3]:.0f}  ewt_set {m[4]:.0f}  wrms {m[5]:.0f}  events {m[6]:.0f}  vars {m[7]:.0f}")
print("  lu_solve trace (gather, records..., scatter):", " ".join(f"{v:.0f}" for v in m[8:] if v > 0))
