"""ncu target: a kernel that only runs triangular solves.  python -m tools.gpu_solve_loop <model> [reps]"""
import sys

import numpy as np

from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b
from tools.gpu_check import sample_inputs
from oracle.oracle import Oracle

name = sys.argv[1]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
bundle = b.load_bundle(name)
rng = np.random.default_rng(20260)
p = sample_inputs(bundle, 1, rng)
o = Oracle(b.build_oracle(bundle))
y = np.array([o.eval("ic", None, p[0])])
s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"])
c = s.eval("solve_loop", y, p, 10.0, float(reps), rng.standard_normal((1, bundle.n_states)))
print(f"{name}: {c[0, 0]:.0f} cycles per triangular solve ({reps} repetitions)")
