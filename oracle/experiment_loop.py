"""The reference's experiment loop around the CPU oracle.

TEST / BASELINE INFRASTRUCTURE ONLY (like everything under oracle/): imported by tests/ and by bench.py's CPU legs, never
by the product package.
"""
from __future__ import annotations

import numpy as np


def oracle_experiment(b, o, schedule, p, rtol=1e-8, atol=1e-8, save_vars=True, n_steps=None, lu="band"):
    """The reference's experiment loop, restated in Python around the CPU oracle: one oracle solve per step and per system,
    exactly the sequence Simulation.solve -> solver.step makes (simulation.py:991-1167, base_solver.py:1334-1588): the step's
    inputs, t_eval = [nextafter(t_start), t_start + duration], t_interp = t_start + linspace(0, duration, n_pts), the previous
    step's last state as y0, consistent initialisation inside the solve, an event that is already exceeded at the start of a
    step skips it when it is the experiment's and ends the run otherwise.
    Returns per system dict(t, y, step_end, step_term, flag) like the device path."""
    from pybamm_b200.experiment import (STEP_EXPERIMENT_EVENT, STEP_FINAL_TIME, STEP_NOT_RUN, STEP_OTHER_EVENT, STEP_SKIPPED)

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
                        y0=None if y_cur is None else y_cur[None, :], save_vars=False, lu=lu)
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
