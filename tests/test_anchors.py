"""Independent pins of the integrators ("the oracle is unpinned ... and shares an author with the device
integrator"): scipy's Radau and BDF -- other methods, other code -- on the ODE models, driven by the numpy evaluation of the
frozen bundles (tests/golden/scipy_anchors.json, tools/make_scipy_anchors.py).  The SPM value reproduces the survey's own
anchor of BASELINE.md section 5 (3567.860565 s), which was computed with pybamm's EvaluatorPython.

The oracle and the host emulation of the device integrator are held to it here; tests/test_gpu.py holds the CUDA path to it.
"""
import json

import numpy as np
import pytest

from tests import util

ANCH = json.load(open(f"{util.GOLDEN}/scipy_anchors.json"))


def anchor_inputs(b):
    return np.array([[ANCH["inputs"][n] for n in b.input_names]])


def check_against_anchor(name, t_end, flag, t, v):
    a = ANCH["models"][name]
    assert abs(a["Radau"]["t_event"] - a["BDF"]["t_event"]) < 1e-5         # the two scipy methods agree between themselves
    assert int(flag) == 2
    assert abs(t_end - a["Radau"]["t_event"]) < util.T_EVENT_TOL, (t_end, a["Radau"]["t_event"])
    for ts, vs in zip(ANCH["t_sample"], a["Radau"]["v"]):
        k = int(np.argmin(np.abs(t - ts)))
        assert t[k] == ts
        assert abs(v[k] - vs) < util.V_TOL, (ts, v[k], vs)


def test_survey_anchor_is_reproduced():
    # SPM, I = 5 A, D_n = 3.3e-14, 2.5 V: Radau 1e-8 -> 3567.860565387 s
    assert abs(ANCH["models"]["spm"]["Radau"]["t_event"] - 3567.860565387) < 1e-5


@pytest.mark.parametrize("name", ["spm", "spme"])
@pytest.mark.parametrize("lu", ["band", "sparse"])
def test_oracle_matches_scipy_radau(name, lu):
    from pybamm_b200 import build as pbuild

    b, o = util.bundle(name), util.oracle(name)
    if lu == "sparse":
        pbuild.generate(b, "cuda")
        o.set_sparse(*pbuild.lu_order(b))
    r = o.solve(util.T_EVAL, anchor_inputs(b), t_interp=util.T_INTERP, rtol=1e-8, atol=1e-8, save_vars=True, lu=lu)
    check_against_anchor(name, r["t"][0][-1], r["flag"][0], r["t"][0], r["y"][0][:, 0])


@pytest.mark.parametrize("name", ["spm", "spme"])
def test_device_integrator_emulation_matches_scipy_radau(name):
    b, e = util.bundle(name), util.emu(name)
    r = e.solve(util.T_EVAL, anchor_inputs(b), t_interp=util.T_INTERP, rtol=1e-8, atol=1e-8, save_vars=True)
    check_against_anchor(name, r["t"][0][-1], r["flag"][0], r["t"][0], r["y"][0][:, 0])
