"""Effect of the launch order on the persistent kernel's tail: python -m tools.gpu_order <model> <n_sys>
natural order, the cost proxy of shard.launch_order, and the bound: descending MEASURED step count of a first run."""
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
ti = np.linspace(0, 3600, 361)
r = s.solve_arrays([0.0, 3700.0], p, t_interp=ti, want_stats=True)
work = r["stats"][:, 1] + 8.0 * r["stats"][:, 2]      # residuals + 8 x set-ups: roughly proportional to time
c = cost_proxy(p, s._cost_model)      # fitted on the step counts of the solve above
print("inputs:", bundle.input_names)
X = np.column_stack([np.ones(n)] + [p[:, k] for k in range(p.shape[1])] + [np.log(np.abs(p[:, k]) + 1e-300) for k in range(p.shape[1])])
coef, *_ = np.linalg.lstsq(X, work, rcond=None)
fit = X @ coef
print("corr(work, proxy) %.2f   corr(work, linear+log fit) %.2f" % (np.corrcoef(work, c)[0, 1], np.corrcoef(work, fit)[0, 1]))
for label, order in (("natural", np.arange(n)), ("proxy", np.argsort(-c)), ("fit", np.argsort(-fit)), ("measured work", np.argsort(-work))):
    for rep in range(2):
        r = s.solve_arrays([0.0, 3700.0], p[order], t_interp=ti, want_stats=True)
    print(f"{label:16s} kernel {s.last_kernel_ms:8.1f} ms  {n / s.last_kernel_ms * 1e3:7.0f} sims/s", flush=True)
