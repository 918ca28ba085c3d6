"""Time prebuilt image variants: python -m tools.gpu_variants <model> <n_sys> <image.so> [<image.so> ...]"""
import sys
import numpy as np
from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b
from tools.gpu_check import sample_inputs

name, n = sys.argv[1], int(sys.argv[2])
bundle = b.load_bundle(name)
p = sample_inputs(bundle, n, np.random.default_rng(20260))
for img in sys.argv[3:]:
    s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"], image_path=img)
    for rep in range(2):
        r = s.solve_arrays([0.0, 3700.0], p, t_interp=np.linspace(0, 3600, 361))
    print(f"{img.split('/')[-1]:40s} team {s.desc.team_size} smem {s.desc.smem_bytes}: kernel {s.last_kernel_ms:.1f} ms  {n / s.last_kernel_ms * 1e3:.0f} sims/s  flags {np.unique(r['flag'])}", flush=True)
    s.close()
