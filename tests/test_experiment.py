"""Experiment schedules run inside the integrator: the device step loop (pb_solve_experiment, here
through the host emulation of the same source) against closed forms and against the oracle driven step by step the way
Simulation.solve drives solver.step (simulation.py:991-1167, base_solver.py:1334-1588)."""
import numpy as np
import pytest

from pybamm_b200.experiment import (STEP_EXPERIMENT_EVENT, STEP_FINAL_TIME, STEP_NOT_RUN, STEP_OTHER_EVENT, STEP_SKIPPED,
                                    ExperimentSchedule)
from tests import util


def cccv_schedule(cycles=1):
    """Discharge at 1 until v = 0.5 / rest for 1 / charge at 0.5 until v = 0.9 / hold v = 0.9 until |i| = 0.05 (kat_cccv)."""
    #            mode sgn  thr    value
    rows = [[1.0, 1.0, 0.5, 1.0],
            [1.0, 1.0, -10.0, 0.0],
            [1.0, -1.0, 0.9, -0.5],
            [2.0, 1.0, 0.05, 0.9]] * cycles
    return ExperimentSchedule(["mode", "sgn", "thr", "value"], np.array(rows), [5.0, 1.0, 5.0, 5.0] * cycles,
                              [51, 11, 51, 51] * cycles, [1, 0, 1, 1] * cycles, "Combined termination [experiment]",
                              ["discharge", "rest", "charge", "hold"] * cycles)


def closed_form(R):
    """End times (relative to the step start) and end charges of the four steps of one cycle."""
    tA = 0.5 - R                       # q = 1 - t, v = q - R = 0.5
    qA = 1.0 - tA
    tC = (0.9 - 0.5 * R - qA) / 0.5    # q = qA + t / 2, v = q + R / 2 = 0.9
    qC = qA + 0.5 * tC
    tD = R * np.log((0.9 - qC) / (0.05 * R))     # q - 0.9 = (qC - 0.9) e^{-t / R},  |i| = (0.9 - q) / R = 0.05
    qD = 0.9 - 0.05 * R
    return (tA, 1.0, tC, tD), (qA, qA, qC, qD)


def test_cccv_toy_against_closed_forms():
    b, e = util.bundle("kat_cccv"), util.emu("kat_cccv")
    sched = cccv_schedule()
    Rs = np.array([0.1, 0.2, 0.05])
    p = np.zeros((3, 5))
    p[:, b.input_names.index("R")] = Rs
    r = e.solve_experiment(b, sched, p, rtol=1e-9, atol=1e-11)
    assert list(r["flag"]) == [1, 1, 1]
    for s, R in enumerate(Rs):
        dts, qs = closed_form(R)
        np.testing.assert_array_equal(r["step_term"][s], [STEP_EXPERIMENT_EVENT, STEP_FINAL_TIME, STEP_EXPERIMENT_EVENT, STEP_EXPERIMENT_EVENT])
        t, y = r["t"][s], r["y"][s]
        t_end = 0.0
        for k in range(4):
            j = r["step_end"][s, k] - 1
            t_end += dts[k]
            assert abs(t[j] - t_end) < 1e-6, (s, k, t[j], t_end)
            assert abs(y[j, 0] - qs[k]) < 1e-7, (s, k, y[j, 0], qs[k])
        # the algebraic current jumps at every step boundary and is consistent right after it
        j0 = r["step_end"][s, 0]                  # first point of the rest step
        assert t[j0] == np.nextafter(t[j0 - 1], np.inf) and abs(y[j0, 1]) < 1e-9 and abs(y[j0 - 1, 1] - 1.0) < 1e-9
        # output grid of a step: t_start + j * duration / (n_pts - 1) up to the event
        np.testing.assert_allclose(t[1:4], [0.1, 0.2, 0.3], rtol=0, atol=1e-15)
        assert np.all(np.diff(t) > 0)


def test_step_skipped_when_its_event_is_exceeded_at_the_start_and_safety_event_ends_the_run():
    b, e = util.bundle("kat_cccv"), util.emu("kat_cccv")
    p = np.zeros((1, 5))
    p[0, b.input_names.index("R")] = 0.1
    # "discharge until v = 2": v(0) = 0.9 < 2 is already past the threshold (sgn (v - thr) < 0) -> skipped, then a rest
    rows = [[1.0, 1.0, 2.0, 1.0], [1.0, 1.0, -10.0, 0.0]]
    sk = ExperimentSchedule(["mode", "sgn", "thr", "value"], np.array(rows), [5.0, 1.0], [6, 3], [0, 0])
    r = e.solve_experiment(b, sk, p, rtol=1e-8, atol=1e-10)
    np.testing.assert_array_equal(r["step_term"][0], [STEP_SKIPPED, STEP_FINAL_TIME])
    assert r["flag"][0] == 1 and r["step_end"][0, 0] == 0 and r["t"][0][0] == 0.0 and r["t"][0][-1] == 1.0
    # charging at 1 with no experiment threshold in reach runs into the safety event q = 10
    rows = [[1.0, 1.0, -100.0, -1.0], [1.0, 1.0, -10.0, 0.0]]
    sf = ExperimentSchedule(["mode", "sgn", "thr", "value"], np.array(rows), [20.0, 1.0], [5, 3], [0, 0])
    r = e.solve_experiment(b, sf, p, rtol=1e-8, atol=1e-10)
    assert r["flag"][0] == 2 and r["event_idx"][0] == 0
    np.testing.assert_array_equal(r["step_term"][0], [STEP_OTHER_EVENT, STEP_NOT_RUN])
    assert abs(r["t"][0][-1] - 9.0) < 1e-6
    # a default duration that runs out makes the experiment infeasible: flag 0, later steps not run
    rows = [[1.0, 1.0, -10.0, 0.0], [1.0, 1.0, -10.0, 0.0]]
    inf = ExperimentSchedule(["mode", "sgn", "thr", "value"], np.array(rows), [1.0, 1.0], [3, 3], [1, 0])
    r = e.solve_experiment(b, inf, p, rtol=1e-8, atol=1e-10)
    assert r["flag"][0] == 0
    np.testing.assert_array_equal(r["step_term"][0], [STEP_FINAL_TIME, STEP_NOT_RUN])


def test_schedule_meta_round_trip():
    s = cccv_schedule(2)
    t = ExperimentSchedule.from_meta(s.to_meta())
    np.testing.assert_array_equal(s.over_val, t.over_val)
    assert t.n_steps == 8 and t.capacity == s.capacity == 2 * (52 + 12 + 52 + 52)
    with pytest.raises(ValueError):
        ExperimentSchedule(["a"], np.zeros((2, 2)), [1, 1], [2, 2], [0, 0])


def test_toy_device_loop_equals_the_reference_loop_around_the_oracle():
    from tests.util import oracle_experiment

    b, e = util.bundle("kat_cccv"), util.emu("kat_cccv")
    sched = cccv_schedule(2)
    p = np.zeros((2, 5))
    p[:, b.input_names.index("R")] = [0.1, 0.15]
    r = e.solve_experiment(b, sched, p, rtol=1e-8, atol=1e-10)
    ref = oracle_experiment("kat_cccv", sched, p, rtol=1e-8, atol=1e-10)
    for s in range(2):
        np.testing.assert_array_equal(r["step_term"][s], ref[s]["step_term"])
        np.testing.assert_array_equal(r["step_end"][s], ref[s]["step_end"])
        assert r["flag"][s] == ref[s]["flag"] and len(r["t"][s]) == len(ref[s]["t"])
        ends = set(ref[s]["step_end"] - 1)
        inner = np.array([i for i in range(len(ref[s]["t"])) if i not in ends])
        k0 = int(ref[s]["step_end"][0]) - 1
        np.testing.assert_array_equal(r["t"][s][:k0], ref[s]["t"][:k0])               # first step's grid: bit-identical times
        np.testing.assert_allclose(r["t"][s], ref[s]["t"], rtol=0, atol=1e-6)          # event times
        np.testing.assert_allclose(r["y"][s][inner], ref[s]["y"][inner], rtol=0, atol=1e-7)


def test_dfn_cccv_first_cycle_device_loop_vs_reference_loop():
    # BASELINE config 5 (one cycle of it on the CPU tier): DFN unified-experiment model, discharge / rest / charge / hold
    from tests.util import oracle_experiment

    b, e = util.bundle("dfn_cccv"), util.emu("dfn_cccv")
    sched = ExperimentSchedule.from_meta(b.meta["experiment"])
    assert sched.n_steps == 12 and sched.over_names == ["Ambient temperature [K]", "Experiment step index", "Experiment step value"]
    one = ExperimentSchedule(sched.over_names, sched.over_val[:4], sched.duration[:4], sched.n_pts[:4], sched.stop_on_final_time[:4],
                             sched.experiment_event, sched.step_names[:4])
    p = np.array([b.meta["nominal_inputs"]])
    r = e.solve_experiment(b, one, p, rtol=1e-6, atol=1e-6, save_vars=True)
    ref = oracle_experiment("dfn_cccv", one, p, rtol=1e-6, atol=1e-6)[0]
    np.testing.assert_array_equal(r["step_term"][0], ref["step_term"])
    np.testing.assert_array_equal(r["step_term"][0], [STEP_EXPERIMENT_EVENT, STEP_FINAL_TIME, STEP_EXPERIMENT_EVENT, STEP_EXPERIMENT_EVENT])
    np.testing.assert_array_equal(r["step_end"][0], ref["step_end"])
    np.testing.assert_allclose(r["t"][0], ref["t"], rtol=0, atol=0.5)                  # event times at rtol 1e-6: within half a second
    ends = set(ref["step_end"] - 1)
    inner = np.array([i for i in range(len(ref["t"])) if i not in ends])
    np.testing.assert_allclose(r["y"][0][inner, 0], ref["v"][inner, 0], rtol=0, atol=2e-4)   # voltage
    np.testing.assert_allclose(r["y"][0][inner, 1], ref["v"][inner, 1], rtol=0, atol=2e-3)   # current
    v_end = [r["y"][0][k - 1, 0] for k in r["step_end"][0]]
    assert abs(v_end[0] - 2.5) < 1e-5 and abs(v_end[2] - 4.2) < 1e-5 and abs(r["y"][0][r["step_end"][0][3] - 1, 1] + 0.1) < 1e-4
