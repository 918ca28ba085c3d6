"""Lower a discretised pybamm ``Symbol`` tree to the scalar DAG of ``ir.py``.

This replaces ``Symbol.to_casadi`` + ``BaseSolver.process``
(packages/pybamm/src/pybamm/solvers/base_solver.py:1830-2002) for the hot path.
The node semantics follow ``Symbol.evaluate`` and the in-tree python lowering
(packages/pybamm/src/pybamm/expression_tree/operations/evaluate_python.py:141-391).

The module never imports pybamm: it dispatches on class *names* along the MRO,
so it works with any object that quacks like the reference expression tree.
Only trees that are already discretised (no Variables, no spatial operators,
no Broadcasts) are accepted; anything else fails loudly.
"""
from __future__ import annotations

import numbers

import numpy as np
import scipy.sparse as sp

from . import ir


class LoweringError(NotImplementedError):
    pass


def _mro_names(obj):
    return [c.__name__ for c in type(obj).__mro__]


class Const:
    """A constant value: python float, dense ndarray (2-D) or scipy sparse matrix."""

    __slots__ = ("v",)

    def __init__(self, v):
        if sp.issparse(v):
            v = v.tocsr()
        elif isinstance(v, numbers.Number):
            v = np.array([[float(v)]])
        else:
            v = np.asarray(v, dtype=np.float64)
            if v.ndim == 0:
                v = v.reshape(1, 1)
            elif v.ndim == 1:
                v = v.reshape(-1, 1)
        self.v = v

    @property
    def shape(self):
        return self.v.shape


_FUNCTION_BY_NAME = {
    "exp": ir.EXP, "log": ir.LOG, "sqrt": ir.SQRT, "tanh": ir.TANH, "sinh": ir.SINH,
    "cosh": ir.COSH, "sin": ir.SIN, "cos": ir.COS, "arcsinh": ir.ASINH, "arctan": ir.ATAN,
    "erf": ir.ERF, "erfc": ir.ERFC, "log10": ir.LOG10, "abs": ir.ABS, "absolute": ir.ABS,
    "sign": ir.SIGN, "floor": ir.FLOOR, "ceil": ir.CEIL,
}
_SPECIFIC_FUNCTION = {
    "Exp": ir.EXP, "Log": ir.LOG, "Sqrt": ir.SQRT, "Tanh": ir.TANH, "Sinh": ir.SINH,
    "Cosh": ir.COSH, "Sin": ir.SIN, "Cos": ir.COS, "Arcsinh": ir.ASINH, "Arctan": ir.ATAN,
    "Erf": ir.ERF,
}
_BINARY = {
    "Addition": ir.ADD, "Subtraction": ir.SUB, "Multiplication": ir.MUL, "Division": ir.DIV,
    "Power": ir.POW, "Minimum": ir.MIN, "Maximum": ir.MAX, "Inner": ir.MUL,
    "EqualHeaviside": ir.LE, "NotEqualHeaviside": ir.LT, "Modulo": ir.MOD, "Hypot": ir.HYPOT,
}
_UNARY = {
    "Negate": ir.NEG, "AbsoluteValue": ir.ABS, "Sign": ir.SIGN, "Floor": ir.FLOOR,
    "Ceiling": ir.CEIL,
}


class Lowerer:
    """Lower Symbols into ``graph``. ``input_layout``: {name: (offset, size)}."""

    def __init__(self, graph: ir.Graph, input_layout: dict):
        self.g = graph
        self.input_layout = input_layout
        self._memo: dict = {}

    # -- helpers ------------------------------------------------------------
    def ids(self, v) -> np.ndarray:
        """Value -> 2-D int64 array of node ids."""
        if isinstance(v, Const):
            a = v.v
            if sp.issparse(a):
                a = a.toarray()
            out = np.empty(a.shape, dtype=np.int64)
            flat = out.reshape(-1)
            for i, x in enumerate(a.reshape(-1)):
                flat[i] = self.g.const(x)
            return out
        return v

    def lower_vector(self, sym) -> np.ndarray:
        """Lower to a flat vector of node ids (column vector expected)."""
        v = self.ids(self.lower(sym))
        if v.ndim == 2 and v.shape[1] != 1 and v.shape[0] == 1:
            v = v.T
        if v.ndim == 2 and v.shape[1] != 1:
            raise LoweringError(f"expected a column vector, got shape {v.shape}")
        return v.reshape(-1)

    # -- main dispatch --------------------------------------------------------
    def lower(self, sym):
        key = getattr(sym, "id", None)
        if key is not None and key in self._memo:
            return self._memo[key]
        out = self._lower(sym)
        if key is not None:
            self._memo[key] = out
        return out

    def _lower(self, sym):
        names = _mro_names(sym)
        g = self.g
        # constants first (Scalar, Vector, Matrix, or any constant subtree)
        if "StateVector" not in names and "InputParameter" not in names and "Time" not in names:
            try:
                is_const = sym.is_constant()
            except Exception:
                is_const = False
            if is_const:
                return Const(sym.evaluate())
        if "StateVectorDot" in names:
            raise LoweringError("StateVectorDot (explicit y') is not supported on this path")
        if "StateVector" in names:
            idx = np.argwhere(np.asarray(sym.evaluation_array)).reshape(-1)
            return np.array([g.state(i) for i in idx], dtype=np.int64).reshape(-1, 1)
        if "InputParameter" in names:
            if sym.name not in self.input_layout:
                raise LoweringError(f"input parameter '{sym.name}' has no slot in the input layout")
            off, size = self.input_layout[sym.name]
            return np.array([g.input(off + k) for k in range(size)], dtype=np.int64).reshape(-1, 1)
        if "Time" in names:
            return np.array([[g.time()]], dtype=np.int64)
        if "Conditional" in names:
            return self._conditional(sym)
        if "MatrixMultiplication" in names:
            return self._matmul(self.lower(sym.children[0]), self.lower(sym.children[1]))
        if "BinaryOperator" in names:
            for nm in names:
                if nm in _BINARY:
                    return self._elementwise2(_BINARY[nm], self.lower(sym.children[0]), self.lower(sym.children[1]))
            raise LoweringError(f"binary operator {names[0]}")
        if "Index" in names:
            v = self.ids(self.lower(sym.children[0]))
            s = sym.slice
            if isinstance(s, slice):
                return v[s, :]
            return v[int(s):int(s) + 1, :]
        if "Transpose" in names:
            return self.ids(self.lower(sym.children[0])).T
        if "NotConstant" in names:
            return self.lower(sym.children[0])
        if "UnaryOperator" in names:
            for nm in names:
                if nm in _UNARY:
                    return self._elementwise1(_UNARY[nm], self.lower(sym.children[0]))
            raise LoweringError(f"unary operator {names[0]} (tree must be discretised first)")
        if "Function" in names:
            return self._function(sym, names)
        if "DomainConcatenation" in names:
            return self._domain_concatenation(sym)
        if "NumpyConcatenation" in names or "SparseStack" in names:
            parts = [self.ids(self.lower(c)) for c in sym.children]
            return np.vstack(parts)
        raise LoweringError(f"cannot lower symbol of type {names[0]} ('{getattr(sym, 'name', '?')}')")

    # -- elementwise ------------------------------------------------------------
    def _elementwise1(self, op, a):
        a = self.ids(a)
        out = np.empty(a.shape, dtype=np.int64)
        of, af = out.reshape(-1), a.reshape(-1)
        for i in range(af.size):
            of[i] = self.g.unary(op, int(af[i]))
        return out

    def _elementwise2(self, op, a, b):
        if isinstance(a, Const) and isinstance(b, Const):  # pragma: no cover - folded upstream
            a = self.ids(a)
        a, b = self.ids(a), self.ids(b)
        try:
            shape = np.broadcast_shapes(a.shape, b.shape)
        except ValueError as e:
            raise LoweringError(f"cannot broadcast {a.shape} with {b.shape}") from e
        ab = np.broadcast_to(a, shape).reshape(-1)
        bb = np.broadcast_to(b, shape).reshape(-1)
        out = np.empty(shape, dtype=np.int64)
        of = out.reshape(-1)
        g = self.g
        for i in range(of.size):
            of[i] = g.binary(op, int(ab[i]), int(bb[i]))
        return out

    def _matmul(self, a, b):
        g = self.g
        if isinstance(a, Const) and isinstance(b, Const):  # pragma: no cover
            return Const(a.v @ b.v)
        if isinstance(a, Const):
            M = a.v if sp.issparse(a.v) else sp.csr_matrix(a.v)
            M = M.tocsr()
            M.sort_indices()
            v = self.ids(b)
            if M.shape[1] != v.shape[0]:
                if M.shape == (1, 1):
                    return self._elementwise2(ir.MUL, a, v)
                raise LoweringError(f"matmul shape mismatch {M.shape} @ {v.shape}")
            out = np.empty((M.shape[0], v.shape[1]), dtype=np.int64)
            indptr, indices, data = M.indptr, M.indices, M.data
            for j in range(v.shape[1]):
                col = v[:, j]
                for i in range(M.shape[0]):
                    acc = g.const(0.0)
                    for k in range(indptr[i], indptr[i + 1]):
                        acc = g.add(acc, g.mul(g.const(data[k]), int(col[indices[k]])))
                    out[i, j] = acc
            return out
        if isinstance(b, Const):
            # (A @ B) = (B^T @ A^T)^T
            bt = Const(b.v.T)
            return self._matmul(bt, self.ids(a).T).T
        A, B = self.ids(a), self.ids(b)
        if A.shape[1] != B.shape[0]:
            raise LoweringError(f"matmul shape mismatch {A.shape} @ {B.shape}")
        out = np.empty((A.shape[0], B.shape[1]), dtype=np.int64)
        for i in range(A.shape[0]):
            for j in range(B.shape[1]):
                out[i, j] = g.sum(g.mul(int(A[i, k]), int(B[k, j])) for k in range(A.shape[1]))
        return out

    # -- functions ----------------------------------------------------------------
    def _function(self, sym, names):
        g = self.g
        kids = [self.lower(c) for c in sym.children]
        if "RegPower" in names:
            # x * (x^2 + delta^2)^((a-1)/2) * scale^a with x = base/scale  (functions.py:937-941)
            base, expo, scale = (self.ids(k) for k in kids)
            d2 = Const(float(sym.delta) ** 2)
            x = self._elementwise2(ir.DIV, base, scale)
            x2 = self._elementwise2(ir.ADD, self._elementwise1(ir.SQR, x), d2)
            e = self._elementwise2(ir.DIV, self._elementwise2(ir.SUB, expo, Const(1.0)), Const(2.0))
            r = self._elementwise2(ir.MUL, x, self._elementwise2(ir.POW, x2, e))
            return self._elementwise2(ir.MUL, r, self._elementwise2(ir.POW, scale, expo))
        if "Arcsinh2" in names:
            # arcsinh(a / (sgn(b) hypot(b, eps))), sgn(0) = +1  (functions.py:412-431)
            a, b = (self.ids(k) for k in kids)
            eps = Const(float(sym.eps))
            ge0 = self._elementwise2(ir.LE, Const(0.0), b)
            sgn = self._elementwise2(ir.SUB, self._elementwise2(ir.MUL, Const(2.0), ge0), Const(1.0))
            beff = self._elementwise2(ir.MUL, sgn, self._elementwise2(ir.HYPOT, b, eps))
            return self._elementwise1(ir.ASINH, self._elementwise2(ir.DIV, a, beff))
        for nm in names:
            if nm in _SPECIFIC_FUNCTION:
                return self._elementwise1(_SPECIFIC_FUNCTION[nm], kids[0])
        if "Max" in names or "Min" in names:
            v = self.ids(kids[0]).reshape(-1)
            op = ir.MAX if "Max" in names else ir.MIN
            acc = int(v[0])
            for x in v[1:]:
                acc = g.binary(op, acc, int(x))
            return np.array([[acc]], dtype=np.int64)
        fn = getattr(sym, "function", None)
        fname = getattr(fn, "__name__", None)
        if isinstance(fn, np.ufunc) and fname in _FUNCTION_BY_NAME and len(kids) == 1:
            return self._elementwise1(_FUNCTION_BY_NAME[fname], kids[0])
        if isinstance(fn, np.ufunc) and fname in ("minimum", "maximum", "power", "hypot") and len(kids) == 2:
            op = {"minimum": ir.MIN, "maximum": ir.MAX, "power": ir.POW, "hypot": ir.HYPOT}[fname]
            return self._elementwise2(op, kids[0], kids[1])
        raise LoweringError(f"function '{getattr(sym, 'name', fname)}' ({names[0]}) has no device lowering")

    def _conditional(self, sym):
        # branch k (1-based) is taken when k-0.5 < selector < k+0.5, else 0
        # (evaluate_python.py:163-171)
        sel = self.lower(sym.children[0])
        out = Const(0.0)
        branches = [self.lower(c) for c in sym.children[1:]]
        for k in range(len(branches), 0, -1):
            lo = self._elementwise2(ir.LT, Const(k - 0.5), sel)
            hi = self._elementwise2(ir.LT, sel, Const(k + 0.5))
            cond = self._elementwise2(ir.MUL, lo, hi)
            out = self._select(cond, branches[k - 1], out)
        return out

    def _select(self, c, a, b):
        c, a, b = self.ids(c), self.ids(a), self.ids(b)
        shape = np.broadcast_shapes(c.shape, a.shape, b.shape)
        cb = np.broadcast_to(c, shape).reshape(-1)
        ab = np.broadcast_to(a, shape).reshape(-1)
        bb = np.broadcast_to(b, shape).reshape(-1)
        out = np.empty(shape, dtype=np.int64)
        of = out.reshape(-1)
        for i in range(of.size):
            of[i] = self.g.select(int(cb[i]), int(ab[i]), int(bb[i]))
        return out

    def _domain_concatenation(self, sym):
        # ordering rule of evaluate_python.py:341-367
        kids = [self.ids(self.lower(c)) for c in sym.children]
        if len(kids) == 1 and sym.secondary_dimensions_npts == 1:
            return kids[0]
        # Children pieces land where ``_slices[domain][i]`` says
        # (concatenations.py DomainConcatenation._concatenation_evaluate); the
        # slices tile the output, so sorting the pieces by start reproduces it.
        slice_starts = []
        vecs = []
        for i in range(sym.secondary_dimensions_npts):
            for kid, slices in zip(kids, sym._children_slices):
                for child_dom, child_slice in slices.items():
                    slice_starts.append(sym._slices[child_dom][i].start)
                    vecs.append(kid[child_slice[i].start:child_slice[i].stop, :])
        order = sorted(range(len(vecs)), key=lambda k: slice_starts[k])
        return np.vstack([vecs[k] for k in order])
