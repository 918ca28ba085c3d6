"""Sub-phases of the triangular solves (profiling build -DPB_PROFILE=2): python -m tools.gpu_phases2 <model> [n_sys]"""
import sys
import numpy as np
from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b
from tools.gpu_check import sample_inputs

name = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 148
bundle = b.load_bundle(name)
img = b.build_model(bundle, profile=2)
p = sample_inputs(bundle, n, np.random.default_rng(20260))
s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"], image_path=img)
r = s.solve_arrays([0.0, 3700.0], p, t_interp=np.linspace(0, 3600, 361))
c = r["stats"].astype(float).mean(axis=0)
print(f"{name} x{n} kernel {s.last_kernel_ms:.1f} ms; cycles per system (M): res {c[0]/1e6:.2f} solve {c[1]/1e6:.2f} [gather {c[5]/1e6:.2f} levels {c[6]/1e6:.2f} scatter {c[7]/1e6:.2f}] jac {c[2]/1e6:.2f} factor {c[3]/1e6:.2f} nls {c[4]/1e6:.2f}")
