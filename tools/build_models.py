"""Build the BASELINE.json configurations with the *reference* pybamm (dev tooling).

Runs only where /root/reference is mounted (this container).  Produces discretised
pybamm models; ``tools/freeze_models.py`` lowers them to ModelBundles that travel.
Inputs follow SURVEY.md section 8d.
"""
from __future__ import annotations

import tools.pybamm_env as env

INPUTS_C1 = ["Current function [A]", "Negative particle diffusivity [m2.s-1]"]
INPUTS_C3 = [
    "Current function [A]",
    "Negative particle diffusivity [m2.s-1]",
    "Positive particle diffusivity [m2.s-1]",
    "Initial concentration in electrolyte [mol.m-3]",
]

CONFIGS = {
    # name: (model class name, options, var_pts override, input names, extra parameter updates)
    "spm": ("SPM", None, None, INPUTS_C1, {}),
    "spme": ("SPMe", None, None, INPUTS_C1, {"Lower voltage cut-off [V]": 3.0}),
    "dfn": ("DFN", None, None, INPUTS_C3, {}),
    "dfn_thermal40": ("DFN", {"thermal": "lumped"}, 40, ["Current function [A]"], {}),
}


INPUTS_C5 = INPUTS_C3[1:]          # config 5: the inputs of config 3 minus the current (the experiment sets it)
CCCV_CYCLE = ("Discharge at 1C until 2.5 V", "Rest for 1 hour", "Charge at 0.3C until 4.2 V", "Hold at 4.2 V until C/50")
EXPERIMENTS = {"dfn_cccv": ("DFN", [CCCV_CYCLE] * 3, INPUTS_C5)}


def build_experiment(config: str):
    """BASELINE config 5: (pybamm, discretised *unified* experiment model, input names, ExperimentSchedule).
    Simulation(..., experiment_model_mode="unified").build_for_experiment() (simulation.py:269-416)."""
    pybamm = env.load()
    from pybamm.input.parameters.lithium_ion.Chen2020 import get_parameter_values

    from pybamm_b200.experiment import schedule_from_simulation

    cls, cycles, input_names = EXPERIMENTS[config]
    model = getattr(pybamm.lithium_ion, cls)()
    pv = pybamm.ParameterValues(get_parameter_values())
    pv.update({n: "[input]" for n in input_names})
    sim = pybamm.Simulation(model, parameter_values=pv, experiment=pybamm.Experiment(cycles), experiment_model_mode="unified")
    sim.build_for_experiment()
    built = sim._built_experiment_model
    return pybamm, built, sorted(ip.name for ip in built.input_parameters), schedule_from_simulation(sim)


def build(config: str):
    """Return (pybamm, discretised model, input names) for a named config."""
    if config in EXPERIMENTS:
        return build_experiment(config)[:3]
    pybamm = env.load()
    from pybamm.input.parameters.lithium_ion.Chen2020 import get_parameter_values

    cls, options, npts, input_names, updates = CONFIGS[config]
    model = getattr(pybamm.lithium_ion, cls)(options) if options else getattr(pybamm.lithium_ion, cls)()
    pv = pybamm.ParameterValues(get_parameter_values())
    pv.update(updates)
    pv.update({n: "[input]" for n in input_names})
    geometry = model.default_geometry
    pv.process_model(model)
    pv.process_geometry(geometry)
    var_pts = model.default_var_pts
    if npts:
        var_pts = {k: npts for k in var_pts}
    mesh = pybamm.Mesh(geometry, model.default_submesh_types, var_pts)
    disc = pybamm.Discretisation(mesh, model.default_spatial_methods)
    disc.process_model(model)
    return pybamm, model, sorted(input_names)


def nominal_inputs(config: str) -> dict:
    vals = {
        "Current function [A]": 5.0,
        "Negative particle diffusivity [m2.s-1]": 3.3e-14,
        "Positive particle diffusivity [m2.s-1]": 4e-15,
        "Initial concentration in electrolyte [mol.m-3]": 1000.0,
    }
    vals.update({"Ambient temperature [K]": 298.15, "Experiment step index": 1.0, "Experiment step value": 5.0})
    if config in EXPERIMENTS:
        return {n: vals[n] for n in list(EXPERIMENTS[config][2]) + ["Ambient temperature [K]", "Experiment step index", "Experiment step value"]}
    return {n: vals[n] for n in CONFIGS[config][3]}
