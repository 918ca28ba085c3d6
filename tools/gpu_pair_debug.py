"""Join statistics of the two-teams-per-CTA kernel (image built with -DPB_PAIR=1 -DPB_PAIR_DEBUG): python -m tools.gpu_pair_debug <image.so> [n]"""
import sys

import numpy as np

from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b
from tools.gpu_check import sample_inputs

bundle = b.load_bundle("dfn")
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2960
p = sample_inputs(bundle, n, np.random.default_rng(20260))
s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"], image_path=sys.argv[1], on_failure="ignore")
for rep in range(2):
    r = s.solve_arrays([0.0, 3700.0], p, t_interp=np.linspace(0, 3600, 361))
st = r["stats"].astype(float).mean(axis=0)
print(f"kernel {s.last_kernel_ms:.1f} ms  {n / s.last_kernel_ms * 1e3:.0f} sims/s; per system: residuals solo {st[6]:.0f} ({st[4] / max(st[6], 1):.0f} cycles each) "
      f"joint {st[7]:.0f} ({st[5] / max(st[7], 1):.0f} cycles each)")
