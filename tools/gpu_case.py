"""Run one batch of one model on the GPU (used under ncu):  python -m tools.gpu_case <model> <n_sys> [reps]"""
import sys
import numpy as np
from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b
from tools.gpu_check import sample_inputs

name, n = sys.argv[1], int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
bundle = b.load_bundle(name)
rng = np.random.default_rng(20260)
p = sample_inputs(bundle, n, rng)
s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"])
for _ in range(reps):
    r = s.solve_arrays([0.0, 3700.0], p, t_interp=np.linspace(0, 3600, 361))
    print(name, n, "kernel ms", s.last_kernel_ms, "sims/s", n / s.last_kernel_ms * 1e3, "flags", np.unique(r["flag"]), "steps", r["stats"][:, 0].mean())
