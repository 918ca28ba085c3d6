"""Experiment schedules for the batched solver (BASELINE config 5).

The reference runs an ``Experiment`` as a Python loop of ``solver.step`` calls, one per step and per input set
(packages/pybamm/src/pybamm/simulation/simulation.py:991-1167), and refuses a list of input sets (:832-835).  For the
*unified* experiment model (:269-416) every step is the same DAE with three inputs switched -- ``Experiment step index``,
``Experiment step value``, ``Ambient temperature [K]`` -- so the whole loop can run inside the kernel
(``pb_solve_experiment`` in csrc/pbgpu_ida.cuh).  ``ExperimentSchedule`` is the data that loop needs, free of pybamm
objects so that it travels with a frozen bundle; ``schedule_from_simulation`` builds it where pybamm is importable.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

STEP_NOT_RUN, STEP_FINAL_TIME, STEP_EXPERIMENT_EVENT, STEP_SKIPPED, STEP_OTHER_EVENT = 0, 1, 2, 3, 4
STEP_TERMINATIONS = {STEP_NOT_RUN: "not run", STEP_FINAL_TIME: "final time", STEP_EXPERIMENT_EVENT: "event: [experiment]",
                     STEP_SKIPPED: "skipped: event exceeded in initial conditions", STEP_OTHER_EVENT: "event"}


@dataclass
class ExperimentSchedule:
    over_names: list                      # input names the schedule sets per step
    over_val: np.ndarray                  # (n_steps, n_over)
    duration: np.ndarray                  # (n_steps,) seconds
    n_pts: np.ndarray                     # (n_steps,) points of linspace(0, duration, n_pts) (BaseStep.default_time_vector)
    stop_on_final_time: np.ndarray        # (n_steps,) bool: the duration is a default, reaching it means "infeasible"
    experiment_event: str | None = "Combined termination [experiment]"
    step_names: list = field(default_factory=list)

    def __post_init__(self):
        self.over_val = np.ascontiguousarray(np.atleast_2d(np.asarray(self.over_val, dtype=np.float64)))
        self.duration = np.ascontiguousarray(self.duration, dtype=np.float64).reshape(-1)
        self.n_pts = np.ascontiguousarray(self.n_pts, dtype=np.int32).reshape(-1)
        self.stop_on_final_time = np.ascontiguousarray(self.stop_on_final_time, dtype=np.uint8).reshape(-1)
        n = len(self.duration)
        if self.over_val.shape != (n, len(self.over_names)) or len(self.n_pts) != n or len(self.stop_on_final_time) != n:
            raise ValueError("inconsistent schedule arrays")

    @property
    def n_steps(self):
        return len(self.duration)

    @property
    def capacity(self):
        """Output points per system a result block must hold (pbgpu_schedule_capacity)."""
        return int(np.sum(np.maximum(self.n_pts, 2) + 1))

    def slots(self, bundle):
        """Columns of the bundle's flattened inputs the schedule overrides."""
        layout = bundle.input_layout()
        out = []
        for n in self.over_names:
            if n not in layout or layout[n][1] != 1:
                raise ValueError(f"schedule input '{n}' is not a scalar input of model bundle '{bundle.name}'")
            out.append(layout[n][0])
        return np.ascontiguousarray(out, dtype=np.int32)

    def event_index(self, bundle):
        if self.experiment_event is None:
            return -1
        if self.experiment_event not in bundle.event_names:
            raise ValueError(f"experiment event '{self.experiment_event}' is not a termination event of '{bundle.name}'")
        return bundle.event_names.index(self.experiment_event)

    def to_meta(self):
        return dict(over_names=list(self.over_names), over_val=self.over_val.tolist(), duration=self.duration.tolist(),
                    n_pts=self.n_pts.tolist(), stop_on_final_time=self.stop_on_final_time.tolist(),
                    experiment_event=self.experiment_event, step_names=list(self.step_names))

    @classmethod
    def from_meta(cls, m):
        return cls(m["over_names"], np.asarray(m["over_val"]), m["duration"], m["n_pts"], m["stop_on_final_time"],
                   m.get("experiment_event"), list(m.get("step_names", [])))


def schedule_from_simulation(sim) -> ExperimentSchedule:
    """Schedule of a ``pybamm.Simulation`` built with ``experiment_model_mode="unified"`` (after ``build_for_experiment``):
    what ``_build_experiment_step_inputs`` (simulation.py:386-416) and ``BaseStep.setup_timestepping``
    (experiment/step/base_step.py:355-421) would hand to ``solver.step`` for every step."""
    import pybamm

    if not getattr(sim, "_experiment_uses_unified_model", False):
        raise ValueError("the simulation does not use the unified experiment model (experiment_model_mode='unified')")
    pv = sim._parameter_values
    names = [sim._AMBIENT_TEMPERATURE_INPUT, sim._STEP_INDEX_INPUT]
    if sim._experiment_uses_value_input:
        names.append(sim._STEP_VALUE_INPUT)
    rows, dur, npts, stop, labels = [], [], [], [], []
    for step, idx in zip(sim.experiment.steps, sim._experiment_step_indices):
        if step.end_time is not None or step.next_start_time is not None or step.is_drive_cycle:
            raise NotImplementedError("start/end times and drive cycles are not lowered to a device schedule")
        ambient = step.temperature or pv[sim._AMBIENT_TEMPERATURE_INPUT]
        if isinstance(ambient, pybamm.Scalar):
            ambient = ambient.value
        row = [float(ambient), float(idx)]
        if sim._experiment_uses_value_input:
            target = step.control_target_value(pv)
            row.append(float(target) if target is not None else float("nan"))
        rows.append(row)
        tf = float(step.duration)
        period = step.period if step.period is not None else step.default_period()
        dur.append(tf)
        npts.append(max(int(round(abs(tf) / period)) + 1, 2))
        stop.append(bool(step.uses_default_duration))
        labels.append(str(step))
    return ExperimentSchedule(names, np.array(rows), dur, npts, stop, sim._COMBINED_TERMINATION_EVENT, labels)
