"""A/B measurements of build variants of one model image.

Build here (no GPU needed):   python -m tools.gpu_variants build <model>
Measure on a GPU box:         python -m tools.gpu_variants run <model> [n_sys]

Every variant is a (tag, environment of the code generator, extra -D defines) triple; the images are built in-tree
(pybamm_b200/_build/pbgpu_model_<model><tag>.so) so that they travel to the GPU box with the snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys

import numpy as np

VARIANTS = [
    # tag, env of the code generator, -D defines, build_model keyword arguments
    ("", {}, (), {}),
    ("_m8", {}, (), {"min_teams": 8}),
    ("_m16", {}, (), {"min_teams": 16}),
    ("_m24", {}, (), {"min_teams": 24}),
    ("_m32", {}, (), {"min_teams": 32}),
]


def build(model):
    procs = []
    for tag, env, defs, kw in VARIANTS:
        code = (f"from pybamm_b200 import build as b; bundle = b.load_bundle('{model}'); "
                f"print(b.build_model(bundle, tag='{tag}', defines={tuple(defs)!r}, **{kw!r}))")
        e = dict(os.environ)
        e.update(env)
        procs.append((tag, subprocess.Popen([sys.executable, "-c", code], env=e)))
    for tag, p in procs:
        if p.wait():
            raise SystemExit(f"variant {tag!r} failed to build")


def run(model, n_sys):
    from pybamm_b200 import BatchedIDASolver
    from pybamm_b200 import build as b
    from tools.gpu_check import sample_inputs

    bundle = b.load_bundle(model)
    rng = np.random.default_rng(20260)
    P = sample_inputs(bundle, n_sys, rng)
    ti = np.linspace(0, 3600, 361)
    y = np.asarray(bundle.graph.evaluate(bundle.ic, 0.0, None, P[0]))[None, :] * (1 + 0.01 * rng.standard_normal((296, bundle.n_states)))
    rhs = rng.standard_normal((296, bundle.n_states))
    base_v = None
    for tag, env, defs, kw in VARIANTS:
        img = os.path.join(b.BUILD, f"pbgpu_model_{model}{tag}.so")
        if not os.path.exists(img):
            print(f"{tag or '(default)':12s} missing image {img}")
            continue
        s = BatchedIDASolver(bundle, rtol=1e-8, atol=1e-8, output_variables=["Voltage [V]"], image_path=img, on_failure="ignore")
        best = 1e30
        for rep in range(3):
            r = s.solve_arrays([0.0, 3700.0], P, t_interp=ti)
            best = min(best, s.last_kernel_ms)
        v = np.array(r["y"][:, :, 0])
        dv = 0.0 if base_v is None else float(np.nanmax(np.abs(np.where(np.arange(v.shape[1])[None, :] < (r["n_t"] - 1)[:, None], v - base_v, 0.0))))
        if base_v is None:
            base_v = v
        c1 = s.eval("cycles", y[:1], P[:1], 10.0, 0.5, rhs[:1]).mean(axis=0)
        c2 = s.eval("cycles", y, np.repeat(P[:1], 296, axis=0), 10.0, 0.5, rhs).mean(axis=0)
        print(f"{tag or '(default)':12s} team {s.desc.team_size} smem {s.desc.smem_bytes:6d}  {n_sys} systems: {best:8.1f} ms  {n_sys / best * 1e3:7.0f} sims/s  "
              f"flags ok {bool(np.all(r['flag'] > 0))}  max|dV| vs first {dv:.2e}  mean steps {r['stats'][:, 0].mean():.1f}\n"
              f"{'':12s} cycles alone : res {c1[0]:7.0f} solve {c1[1]:7.0f} jac {c1[2]:7.0f} factor {c1[3]:7.0f}\n"
              f"{'':12s} cycles x296  : res {c2[0]:7.0f} solve {c2[1]:7.0f} jac {c2[2]:7.0f} factor {c2[3]:7.0f}", flush=True)
        s.close()


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build(sys.argv[2])
    else:
        run(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 2960)
