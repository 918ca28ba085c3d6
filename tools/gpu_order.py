"""Effect of the launch order on the persistent kernel's tail: python -m tools.gpu_order <model> <n_sys>"""
import sys
import numpy as np
from pybamm_b200 import BatchedIDASolver
from pybamm_b200 import build as b
from pybamm_b200.shard import cost_proxy
from tools.gpu_check import sample_inputs

name, n = sys.argv[1], int(sys.argv[2])
bundle = b.load_bundle(name)
p = sample_inputs(bundle, n, np.random.default_rng(20260))
s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"])
c = cost_proxy(p, bundle.input_names)
for label, order in (("natural", np.arange(n)), ("descending 1/I", np.argsort(-c)), ("ascending 1/I", np.argsort(c))):
    for rep in range(2):
        r = s.solve_arrays([0.0, 3700.0], p[order], t_interp=np.linspace(0, 3600, 361), want_stats=True)
    st = r["stats"][:, 0]
    print(f"{label:16s} kernel {s.last_kernel_ms:8.1f} ms  {n / s.last_kernel_ms * 1e3:7.0f} sims/s   steps: mean {st.mean():.0f} min {st.min()} max {st.max()}"
          f"  corr(steps, 1/I) {np.corrcoef(st, c[order])[0, 1]:.2f}", flush=True)
