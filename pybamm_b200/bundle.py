"""ModelBundle: the frozen, self-contained description of one discretised model.

It holds what ``IDAKLUSolver.set_up`` hands to the native solver in the reference
(packages/pybamm/src/pybamm/solvers/idaklu_solver.py:402-628): the residual
``rhs_algebraic``, the sparse ``J - cj*M`` (values + CSC pattern), the mass matrix
ids, the termination-event root functions, the initial conditions ``y0(p)`` and
optional output-variable functions -- but as node ids into one scalar DAG
(``ir.Graph``) instead of serialized CasADi functions.

A bundle can be saved to / loaded from a single ``.npz`` file, so models lowered
where pybamm is importable can be integrated on machines where it is not.
"""
from __future__ import annotations

import json

import numpy as np

from . import ir
from .frontend import Lowerer


class ModelBundle:
    def __init__(self):
        self.name = "model"
        self.graph: ir.Graph = ir.Graph()
        self.n_states = 0
        self.n_diff = 0                       # len_rhs; states [n_diff, n_states) are algebraic
        self.input_names: list[str] = []      # sorted, as base_solver.py:1824
        self.input_sizes: list[int] = []
        self.rhs_alg = np.zeros(0, np.int64)  # (N,) node ids of f(t, y, p)
        self.ic = np.zeros(0, np.int64)       # (N,) node ids of y0(p)
        self.events = np.zeros(0, np.int64)   # (n_events,) termination events g(t, y, p)
        self.event_names: list[str] = []
        self.variables: dict[str, np.ndarray] = {}   # name -> node ids (flattened)
        # Jacobian of rhs_alg wrt y in COO, sorted column-major (CSC order like
        # jac_times_cjmass_{colptrs,rowvals}); the mass-matrix diagonal of the
        # differential rows is always part of the pattern.
        self.jac_rows = np.zeros(0, np.int32)
        self.jac_cols = np.zeros(0, np.int32)
        self.jac_nodes = np.zeros(0, np.int64)
        self.meta: dict = {}

    # ------------------------------------------------------------------
    @property
    def n_inputs(self):
        return int(sum(self.input_sizes))

    @property
    def n_alg(self):
        return self.n_states - self.n_diff

    @property
    def n_events(self):
        return len(self.events)

    def input_layout(self):
        out, off = {}, 0
        for n, s in zip(self.input_names, self.input_sizes):
            out[n] = (off, s)
            off += s
        return out

    def flatten_inputs(self, inputs_list) -> np.ndarray:
        """list of {name: value} -> (n_sys, n_inputs) float64, columns in bundle order."""
        if isinstance(inputs_list, dict):
            inputs_list = [inputs_list]
        out = np.empty((len(inputs_list), self.n_inputs))
        layout = self.input_layout()
        names = list(self.input_names)
        if all(size == 1 for _, size in layout.values()):
            # scalar inputs (the usual case): one pass over the dicts, no per-entry numpy calls
            try:
                out[:, [layout[n][0] for n in names]] = [[d[n] for n in names] for d in inputs_list]
                return out
            except KeyError:
                pass              # fall through to the loop below for the reference's error message
            except (TypeError, ValueError):
                pass              # size-1 arrays etc.: the general path handles them
        for i, d in enumerate(inputs_list):
            missing = [n for n in names if n not in d]
            if missing:
                raise ValueError(
                    "No value provided for input: [" + ", ".join(f"'{m}'" for m in missing) + "]"
                )
            for n, (off, size) in layout.items():
                out[i, off:off + size] = np.asarray(d[n], dtype=np.float64).reshape(-1)
        return out

    def finalize(self):
        """Differentiate, add the mass diagonal to the pattern, compact the graph."""
        g = self.graph
        grads = g.gradients(self.rhs_alg, wrt="state")
        entries = {}
        for r, gr in enumerate(grads):
            for c, node in gr.items():
                entries[(r, c)] = node
        for i in range(self.n_diff):
            entries.setdefault((i, i), g.const(0.0))
        keys = sorted(entries.keys(), key=lambda rc: (rc[1], rc[0]))
        self.jac_rows = np.array([k[0] for k in keys], dtype=np.int32)
        self.jac_cols = np.array([k[1] for k in keys], dtype=np.int32)
        self.jac_nodes = np.array([entries[k] for k in keys], dtype=np.int64)
        self._compact()
        return self

    def _compact(self):
        var_names = list(self.variables)
        lists = [self.rhs_alg, self.ic, self.events, self.jac_nodes] + [self.variables[n] for n in var_names]
        g, new = self.graph.compact(lists)
        g.n_states = self.n_states
        g.n_inputs = self.n_inputs
        self.graph = g
        self.rhs_alg = np.asarray(new[0], np.int64)
        self.ic = np.asarray(new[1], np.int64)
        self.events = np.asarray(new[2], np.int64)
        self.jac_nodes = np.asarray(new[3], np.int64)
        for k, n in enumerate(var_names):
            self.variables[n] = np.asarray(new[4 + k], np.int64)

    # -- numpy evaluation helpers (tests / oracle callbacks) -----------------
    def eval_rhs_alg(self, t, y, p):
        return self.graph.evaluate(self.rhs_alg, t, y, p)

    def eval_ic(self, p):
        return self.graph.evaluate(self.ic, 0.0, None, p)

    def eval_events(self, t, y, p):
        return self.graph.evaluate(self.events, t, y, p)

    def eval_variable(self, name, t, y, p):
        return self.graph.evaluate(self.variables[name], t, y, p)

    def eval_jac(self, t, y, p):
        """scipy CSC matrix of d rhs_alg / d y (single system)."""
        import scipy.sparse as sp

        vals = self.graph.evaluate(self.jac_nodes, t, y, p)
        return sp.csc_matrix((vals, (self.jac_rows, self.jac_cols)), shape=(self.n_states, self.n_states))

    # -- persistence --------------------------------------------------------
    def save(self, path):
        arrs = {"g_" + k: v for k, v in self.graph.to_arrays().items()}
        arrs.update(
            rhs_alg=self.rhs_alg, ic=self.ic, events=self.events,
            jac_rows=self.jac_rows, jac_cols=self.jac_cols, jac_nodes=self.jac_nodes,
        )
        var_names = list(self.variables)
        for k, n in enumerate(var_names):
            arrs[f"var_{k}"] = self.variables[n]
        meta = dict(
            name=self.name, n_states=self.n_states, n_diff=self.n_diff,
            input_names=self.input_names, input_sizes=self.input_sizes,
            event_names=self.event_names, variable_names=var_names, meta=self.meta,
        )
        arrs["meta_json"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
        np.savez_compressed(path, **arrs)

    @classmethod
    def load(cls, path):
        z = np.load(path)
        meta = json.loads(bytes(z["meta_json"]).decode())
        b = cls()
        b.graph = ir.Graph.from_arrays({k[2:]: z[k] for k in z.files if k.startswith("g_")})
        b.name = meta["name"]
        b.n_states = meta["n_states"]
        b.n_diff = meta["n_diff"]
        b.input_names = meta["input_names"]
        b.input_sizes = meta["input_sizes"]
        b.event_names = meta["event_names"]
        b.meta = meta.get("meta", {})
        b.rhs_alg = z["rhs_alg"].astype(np.int64)
        b.ic = z["ic"].astype(np.int64)
        b.events = z["events"].astype(np.int64)
        b.jac_rows = z["jac_rows"].astype(np.int32)
        b.jac_cols = z["jac_cols"].astype(np.int32)
        b.jac_nodes = z["jac_nodes"].astype(np.int64)
        for k, n in enumerate(meta["variable_names"]):
            b.variables[n] = z[f"var_{k}"].astype(np.int64)
        b.graph.n_states = b.n_states
        b.graph.n_inputs = b.n_inputs
        return b


def bundle_from_pybamm(model, inputs=None, variables=("Voltage [V]",), name=None) -> ModelBundle:
    """Lower a *discretised* pybamm model.

    ``inputs``: example input dict (or list of names) fixing the input layout;
    defaults to the model's own input parameters, scalar each.  Input slots are
    ordered by sorted name (base_solver.py:1824).  Only TERMINATION events are
    kept, in model order (base_solver.py:560-680 drops SWITCH events for IDAKLU).
    """
    b = ModelBundle()
    b.name = name or getattr(model, "name", "model")
    sizes = {}
    for ipar in model.input_parameters:
        sizes[ipar.name] = int(getattr(ipar, "_expected_size", 1) or 1)
    if inputs is not None:
        if isinstance(inputs, dict):
            for k, v in inputs.items():
                sizes[k] = int(np.asarray(v).size)
        else:
            for k in inputs:
                sizes.setdefault(k, 1)
    b.input_names = sorted(sizes)
    b.input_sizes = [sizes[n] for n in b.input_names]
    b.n_diff = int(model.concatenated_rhs.shape[0]) if model.concatenated_rhs is not None else 0
    n_alg = int(model.concatenated_algebraic.shape[0])
    b.n_states = b.n_diff + n_alg

    # standard-form check: M = I on the differential block, 0 on the algebraic
    # block (discretisation.py:695-751, idaklu_solver.py:451-457)
    M = model.mass_matrix.entries
    import scipy.sparse as sp

    Mref = sp.diags(np.r_[np.ones(b.n_diff), np.zeros(n_alg)]).tocsr()
    if abs(sp.csr_matrix(M) - Mref).max() != 0.0:
        raise NotImplementedError("only standard-form DAEs (identity mass block) are supported")

    low = Lowerer(b.graph, b.input_layout())
    parts = []
    if b.n_diff:
        parts.append(low.lower_vector(model.concatenated_rhs))
    if n_alg:
        parts.append(low.lower_vector(model.concatenated_algebraic))
    b.rhs_alg = np.concatenate(parts)
    assert len(b.rhs_alg) == b.n_states
    b.ic = low.lower_vector(model.concatenated_initial_conditions)
    assert len(b.ic) == b.n_states

    ev_nodes, ev_names = [], []
    for ev in model.events:
        if "TERMINATION" in str(ev.event_type):
            v = low.lower_vector(ev.expression)
            if len(v) != 1:
                raise NotImplementedError(f"event '{ev.name}' is not scalar")
            ev_nodes.append(int(v[0]))
            ev_names.append(ev.name)
    b.events = np.asarray(ev_nodes, np.int64)
    b.event_names = ev_names

    for vname in variables or ():
        sym = model.get_processed_variable_or_event(vname)
        b.variables[vname] = low.lower_vector(sym)

    try:
        b.meta["y_slices"] = {
            k.name: [[s.start, s.stop] for s in v] for k, v in model.y_slices.items()
        }
    except Exception:
        pass
    return b.finalize()
