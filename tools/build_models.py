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


def build(config: str):
    """Return (pybamm, discretised model, input names) for a named config."""
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
    return {n: vals[n] for n in CONFIGS[config][3]}
