"""Shared helpers of the test tiers: cached bundles, oracle / emulation handles, seeded input samplers."""
from __future__ import annotations

import functools
import os

import numpy as np

from pybamm_b200 import build as pbuild

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
BATTERY = ("spm", "spme", "dfn")
LARGE = ("dfn_thermal40",)     # BASELINE config 4: DFN + lumped thermal, 40-point meshes, 3 523 states
T_EVAL = [0.0, 3700.0]
T_INTERP = np.linspace(0.0, 3600.0, 361)

# parity bars of BASELINE.json's north_star (rtol = atol = 1e-8)
V_TOL = 1e-6       # V, max-abs terminal voltage over t_eval
T_EVENT_TOL = 1e-3  # s


@functools.lru_cache(maxsize=None)
def bundle(name):
    if name.startswith("kat_"):
        from tests import kat_models

        return kat_models.ALL[name]()
    return pbuild.load_bundle(name)


@functools.lru_cache(maxsize=None)
def oracle(name):
    from oracle.oracle import Oracle

    b = bundle(name)
    o = Oracle(pbuild.build_oracle(b))
    o.set_rcm(b.jac_rows, b.jac_cols)
    return o


@functools.lru_cache(maxsize=None)
def emu(name):
    from tests.emu.emu import Emu, build_emu

    return Emu(build_emu(bundle(name)))


def sample_inputs(b, n, rng):
    """Seeded input sets of SURVEY.md section 8d (C1-C3)."""
    cols = []
    for name in b.input_names:
        if name == "Current function [A]":
            cols.append(rng.uniform(2.5, 7.5, n))
        elif name == "Negative particle diffusivity [m2.s-1]":
            cols.append(10 ** rng.uniform(-14, -13, n))
        elif name == "Positive particle diffusivity [m2.s-1]":
            cols.append(10 ** rng.uniform(-15, -14, n))
        elif name == "Initial concentration in electrolyte [mol.m-3]":
            cols.append(rng.uniform(900, 1100, n))
        else:
            raise KeyError(name)
    return np.column_stack(cols) if cols else np.zeros((n, 0))


def compare_voltage(res_t, res_v, ref_t, ref_v, flag, ref_flag):
    """(max |dV| over the common grid excluding a terminal event point, |dt_end|)."""
    assert int(flag) == int(ref_flag), (flag, ref_flag)
    assert len(res_t) == len(ref_t), (len(res_t), len(ref_t))
    n = len(res_t)
    m = n - 1 if int(flag) == 2 else n
    np.testing.assert_array_equal(res_t[:m], ref_t[:m])
    dv = float(np.max(np.abs(np.asarray(res_v)[:m] - np.asarray(ref_v)[:m]))) if m else 0.0
    return dv, abs(float(res_t[-1]) - float(ref_t[-1]))


def oracle_experiment(name, schedule, p, **kw):
    """The reference's experiment loop restated around the CPU oracle (oracle/experiment_loop.py)."""
    from oracle.experiment_loop import oracle_experiment as run

    return run(bundle(name), oracle(name), schedule, p, **kw)
