"""pybamm_b200 -- B200-native batched DAE solver behind pybamm.IDAKLUSolver's API.

Only what the hot path needs: the lowering of a discretised pybamm model to a
scalar DAG (``frontend``, ``ir``, ``bundle``), the re-rolling code generator (``program``, ``codegen``,
``symbolic_lu``), the build drivers (``build``), the ctypes mirror of the C ABI (``host``), the
stand-alone front end (``solver``) and the pybamm drop-in (``idaklu_solver``).
"""
from .bundle import ModelBundle, bundle_from_pybamm  # noqa: F401
from .solver import BatchedIDASolver, Solution, SolutionBatch, SolverError  # noqa: F401

__version__ = "0.1.0"
