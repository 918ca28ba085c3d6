"""Shared helpers of the test tiers: cached bundles, oracle / emulation handles, seeded input samplers."""
from __future__ import annotations

import functools
import os

import numpy as np

from pybamm_b200 import build as pbuild

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
BATTERY = ("spm", "spme", "dfn")
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


def oracle_experiment(name, schedule, p, rtol=1e-8, atol=1e-8, save_vars=True, n_steps=None):
    """The reference's experiment loop, restated in Python around the CPU oracle: one oracle solve per step and per system,
    exactly the sequence Simulation.solve -> solver.step makes (simulation.py:991-1167, base_solver.py:1334-1588): the step's
    inputs, t_eval = [nextafter(t_start), t_start + duration], t_interp = t_start + linspace(0, duration, n_pts), the previous
    step's last state as y0, consistent initialisation inside the solve, an event that is already exceeded at the start of a
    step skips it when it is the experiment's and ends the run otherwise.
    Returns per system dict(t, y, step_end, step_term, flag) like the device path."""
    from pybamm_b200.experiment import (STEP_EXPERIMENT_EVENT, STEP_FINAL_TIME, STEP_NOT_RUN, STEP_OTHER_EVENT, STEP_SKIPPED)

    b, o = bundle(name), oracle(name)
    slots, ev_exp = schedule.slots(b), schedule.event_index(b)
    ns = schedule.n_steps if n_steps is None else n_steps
    out = []
    for row in np.atleast_2d(p):
        q = np.array(row, dtype=np.float64)
        t_cur, y_cur = 0.0, None
        ts, ys, step_end, step_term, flag = [], [], [0] * ns, [STEP_NOT_RUN] * ns, 1
        for k in range(ns):
            q[slots] = schedule.over_val[k]
            dur, npts = float(schedule.duration[k]), int(max(schedule.n_pts[k], 2))
            t0 = t_cur if t_cur == 0.0 else float(np.nextafter(t_cur, np.inf))
            grid = t_cur + np.linspace(0.0, dur, npts)
            r = o.solve([t0, t_cur + dur], q[None, :], t_interp=grid[1:-1], rtol=rtol, atol=atol,
                        y0=None if y_cur is None else y_cur[None, :], save_vars=False)
            t, y, fl = r["t"][0], r["y"][0], int(r["flag"][0])
            if fl == 2 and len(t) == 1:                     # event already exceeded after the consistent initialisation
                g = b.eval_events(t[0], y[0], q)
                bad = np.nonzero(g <= 0)[0]
                if len(bad) and np.all(bad == ev_exp):
                    step_term[k], step_end[k] = STEP_SKIPPED, sum(len(x) for x in ts)
                    continue
                step_term[k], step_end[k], flag = STEP_OTHER_EVENT, sum(len(x) for x in ts), 2
                break
            ts.append(t)
            ys.append(y)
            step_end[k] = sum(len(x) for x in ts)
            t_cur, y_cur = float(t[-1]), y[-1].copy()
            if fl < 0:
                step_term[k], flag = fl, fl
                break
            if fl == 2:
                g = np.abs(b.eval_events(t[-1], y[-1], q))
                if int(np.argmin(g)) == ev_exp:
                    step_term[k] = STEP_EXPERIMENT_EVENT
                    continue
                step_term[k], flag = STEP_OTHER_EVENT, 2
                break
            step_term[k] = STEP_FINAL_TIME
            if schedule.stop_on_final_time[k]:
                flag = 0
                break
        t_all = np.concatenate(ts) if ts else np.zeros(0)
        y_all = np.concatenate(ys) if ys else np.zeros((0, b.n_states))
        v = None
        if save_vars and len(t_all):
            v = np.array([np.concatenate([b.eval_variable(nm, t_all[i], y_all[i], q) for nm in b.variables]) for i in range(len(t_all))])
        out.append(dict(t=t_all, y=y_all, v=v, step_end=np.array(step_end), step_term=np.array(step_term), flag=flag))
    return out
