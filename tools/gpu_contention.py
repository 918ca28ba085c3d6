"""How much do co-running teams slow each other?  The SAME system n times (every team does identical work, in phase at
first, drifting apart never: identical control flow), then n different systems of similar length (out of phase):
python -m tools.gpu_contention <model> [image.so]"""
import sys
import numpy as np
from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b
from tools.gpu_check import sample_inputs

name = sys.argv[1]
img = sys.argv[2] if len(sys.argv) > 2 else None
bundle = b.load_bundle(name)
P = sample_inputs(bundle, 2000, np.random.default_rng(20260))
s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"], image_path=img)
ti = np.linspace(0, 3600, 361)
r = s.solve_arrays([0.0, 3700.0], P, t_interp=ti)
steps = np.asarray(r["stats"])[:, 0]
for n in (1, 37, 74, 148, 296):
    p = np.repeat(P[:1], n, axis=0)
    for rep in range(2):
        s.solve_arrays([0.0, 3700.0], p, t_interp=ti)
    print(f"same system x{n:4d}: kernel {s.last_kernel_ms:8.2f} ms", flush=True)
if steps is not None:
    order = np.argsort(np.abs(steps - np.median(steps)))[:296]      # 296 systems of nearly the same length
    for n in (1, 37, 74, 148, 296):
        p = P[order[:n]]
        for rep in range(2):
            s.solve_arrays([0.0, 3700.0], p, t_interp=ti)
        print(f"similar systems x{n:4d}: kernel {s.last_kernel_ms:8.2f} ms  (steps {steps[order[:n]].min():.0f}..{steps[order[:n]].max():.0f})", flush=True)
s.close()
