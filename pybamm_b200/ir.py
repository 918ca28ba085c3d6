"""Scalar expression DAG: the intermediate representation of a discretised model.

Role in the reference: CasADi's SX/MX graphs built by ``BaseSolver.process``
(packages/pybamm/src/pybamm/solvers/base_solver.py:1830-2002) and evaluated by
CasADi's virtual machine (idaklu_source/Expressions/Casadi/CasadiFunctions.cpp:96-105).
Here the discretised pybamm ``Symbol`` tree is lowered (``frontend.py``) to a
hash-consed scalar DAG on which we do constant folding, algebraic simplification
and sparse forward-mode differentiation (the Jacobian ``J = d f / d y`` that the
reference obtains from ``casadi.jacobian``, base_model.py:568-610).  The DAG is
then re-rolled into team-parallel templates (``program.py``) and emitted as
sm_100a device code (``codegen_cuda.py``) or plain C (``codegen_c.py``).

A node is (op, a0, a1, a2, val).  Nodes are immutable and unique, so node ids
double as value numbers (CSE is free).
"""
from __future__ import annotations

import math

import numpy as np

# ---------------------------------------------------------------------------
# op codes
# ---------------------------------------------------------------------------
CONST, STATE, INPUT, TIME, CJ = 0, 1, 2, 3, 4   # CJ: the BDF leading coefficient cj (Jacobian of F wrt y' scale)
ADD, SUB, MUL, DIV, POW, MIN, MAX = 10, 11, 12, 13, 14, 15, 16
LT, LE, MOD, HYPOT = 17, 18, 19, 20
NEG, ABS, SIGN, EXP, LOG, SQRT, TANH, SINH, COSH = 30, 31, 32, 33, 34, 35, 36, 37, 38
SIN, COS, ASINH, ATAN, ERF, FLOOR, CEIL, RECIP, SQR = 39, 40, 41, 42, 43, 44, 45, 46, 47
LOG10, ERFC = 48, 49
SELECT = 60  # SELECT(c, a, b) = c != 0 ? a : b

LEAF_OPS = (CONST, STATE, INPUT, TIME, CJ)
BINARY_OPS = (ADD, SUB, MUL, DIV, POW, MIN, MAX, LT, LE, MOD, HYPOT)
UNARY_OPS = (NEG, ABS, SIGN, EXP, LOG, SQRT, TANH, SINH, COSH, SIN, COS, ASINH,
             ATAN, ERF, FLOOR, CEIL, RECIP, SQR, LOG10, ERFC)
COMMUTATIVE = (ADD, MUL, MIN, MAX, HYPOT)

OP_NAMES = {
    CONST: "const", STATE: "state", INPUT: "input", TIME: "time", CJ: "cj",
    ADD: "add", SUB: "sub", MUL: "mul", DIV: "div", POW: "pow", MIN: "min", MAX: "max",
    LT: "lt", LE: "le", MOD: "mod", HYPOT: "hypot",
    NEG: "neg", ABS: "abs", SIGN: "sign", EXP: "exp", LOG: "log", SQRT: "sqrt",
    TANH: "tanh", SINH: "sinh", COSH: "cosh", SIN: "sin", COS: "cos", ASINH: "asinh",
    ATAN: "atan", ERF: "erf", FLOOR: "floor", CEIL: "ceil", RECIP: "recip", SQR: "sqr",
    LOG10: "log10", ERFC: "erfc", SELECT: "select",
}

# Rough cost model in FP64 "DFMA equivalents" on sm_100 (used for the
# materialisation decision in program.py and for the work model of DESIGN.md;
# transcendental lengths were read off cuobjdump -sass of libdevice calls).
OP_COST = {
    CONST: 0, STATE: 0, INPUT: 0, TIME: 0, CJ: 0,
    ADD: 1, SUB: 1, MUL: 1, DIV: 10, POW: 60, MIN: 1, MAX: 1, LT: 1, LE: 1, MOD: 20,
    HYPOT: 12, NEG: 0, ABS: 0, SIGN: 2, EXP: 20, LOG: 25, SQRT: 8, TANH: 30, SINH: 35,
    COSH: 30, SIN: 30, COS: 30, ASINH: 45, ATAN: 35, ERF: 35, FLOOR: 1, CEIL: 1,
    RECIP: 8, SQR: 1, LOG10: 27, ERFC: 45, SELECT: 1,
}


def _np_sign(x):
    return np.sign(x)


_UNARY_NUMPY = {
    NEG: np.negative, ABS: np.abs, SIGN: _np_sign, EXP: np.exp, LOG: np.log,
    SQRT: np.sqrt, TANH: np.tanh, SINH: np.sinh, COSH: np.cosh, SIN: np.sin,
    COS: np.cos, ASINH: np.arcsinh, ATAN: np.arctan, FLOOR: np.floor, CEIL: np.ceil,
    RECIP: lambda x: 1.0 / x, SQR: lambda x: x * x, LOG10: np.log10,
}


def _erf(x):
    from scipy.special import erf

    return erf(x)


def _erfc(x):
    from scipy.special import erfc

    return erfc(x)


_UNARY_NUMPY[ERF] = _erf
_UNARY_NUMPY[ERFC] = _erfc

_BINARY_NUMPY = {
    ADD: np.add, SUB: np.subtract, MUL: np.multiply, DIV: np.divide, POW: np.power,
    MIN: np.minimum, MAX: np.maximum,
    LT: lambda a, b: (a < b).astype(np.float64) if isinstance(a < b, np.ndarray) else float(a < b),
    LE: lambda a, b: (a <= b).astype(np.float64) if isinstance(a <= b, np.ndarray) else float(a <= b),
    MOD: np.mod, HYPOT: np.hypot,
}


class Graph:
    """Hash-consed scalar DAG with on-the-fly simplification."""

    def __init__(self):
        self.op: list[int] = []
        self.a0: list[int] = []
        self.a1: list[int] = []
        self.a2: list[int] = []
        self.val: list[float] = []
        self._cse: dict = {}
        self.n_states = 0
        self.n_inputs = 0
        self._zero = self.const(0.0)
        self._one = self.const(1.0)

    # -- construction -----------------------------------------------------
    def __len__(self):
        return len(self.op)

    def _new(self, op, a0=-1, a1=-1, a2=-1, val=0.0):
        key = (op, a0, a1, a2, val if op in LEAF_OPS else 0.0)
        if op == CONST:
            # distinguish -0.0 / nan bit patterns
            key = (op, np.float64(val).tobytes())
        nid = self._cse.get(key)
        if nid is not None:
            return nid
        nid = len(self.op)
        self.op.append(op)
        self.a0.append(a0)
        self.a1.append(a1)
        self.a2.append(a2)
        self.val.append(float(val))
        self._cse[key] = nid
        return nid

    def const(self, v) -> int:
        v = float(v)
        if v == 0.0:
            v = 0.0  # collapse -0.0
        return self._new(CONST, val=v)

    def state(self, i) -> int:
        i = int(i)
        self.n_states = max(self.n_states, i + 1)
        return self._new(STATE, val=i)

    def input(self, j) -> int:
        j = int(j)
        self.n_inputs = max(self.n_inputs, j + 1)
        return self._new(INPUT, val=j)

    def time(self) -> int:
        return self._new(TIME)

    def cj(self) -> int:
        return self._new(CJ)

    def is_const(self, n) -> bool:
        return self.op[n] == CONST

    def cval(self, n) -> float:
        return self.val[n]

    # -- unary ------------------------------------------------------------
    def unary(self, op, a) -> int:
        if self.op[a] == CONST:
            with np.errstate(all="ignore"):
                return self.const(float(_UNARY_NUMPY[op](np.float64(self.val[a]))))
        if op == NEG:
            if self.op[a] == NEG:
                return self.a0[a]
            if self.op[a] == SUB:
                return self.binary(SUB, self.a1[a], self.a0[a])
        if op == ABS and self.op[a] in (ABS, SQR, EXP, COSH):
            return a
        if op == SQR and self.op[a] == NEG:
            return self.unary(SQR, self.a0[a])
        if op == SQR and self.op[a] == SQRT:
            pass  # keep: sqrt(x)^2 != x in floating point
        return self._new(op, a)

    def neg(self, a):
        return self.unary(NEG, a)

    # -- binary -----------------------------------------------------------
    def binary(self, op, a, b) -> int:
        opa, opb = self.op[a], self.op[b]
        if opa == CONST and opb == CONST:
            with np.errstate(all="ignore"):
                r = _BINARY_NUMPY[op](np.float64(self.val[a]), np.float64(self.val[b]))
            return self.const(float(r))
        ca = self.val[a] if opa == CONST else None
        cb = self.val[b] if opb == CONST else None
        if op == ADD:
            if ca == 0.0:
                return b
            if cb == 0.0:
                return a
            if opb == NEG:
                return self.binary(SUB, a, self.a0[b])
            if opa == NEG:
                return self.binary(SUB, b, self.a0[a])
        elif op == SUB:
            if cb == 0.0:
                return a
            if ca == 0.0:
                return self.neg(b)
            if a == b:
                return self._zero
            if opb == NEG:
                return self.binary(ADD, a, self.a0[b])
        elif op == MUL:
            if ca == 0.0 or cb == 0.0:
                return self._zero
            if ca == 1.0:
                return b
            if cb == 1.0:
                return a
            if ca == -1.0:
                return self.neg(b)
            if cb == -1.0:
                return self.neg(a)
            if a == b:
                return self.unary(SQR, a)
            if opa == NEG and opb == NEG:
                return self.binary(MUL, self.a0[a], self.a0[b])
            if opa == NEG and opb == CONST:
                return self.binary(MUL, self.a0[a], self.const(-cb))
            if opb == NEG and opa == CONST:
                return self.binary(MUL, self.a0[b], self.const(-ca))
            # const * (const * x) -> (const*const) * x
            if opa == CONST and opb == MUL:
                b0, b1 = self.a0[b], self.a1[b]
                if self.op[b0] == CONST:
                    return self.binary(MUL, self.const(ca * self.val[b0]), b1)
                if self.op[b1] == CONST:
                    return self.binary(MUL, self.const(ca * self.val[b1]), b0)
            if opb == CONST and opa == MUL:
                return self.binary(MUL, b, a)
        elif op == DIV:
            if ca == 0.0:
                return self._zero
            if cb == 1.0:
                return a
            if cb == -1.0:
                return self.neg(a)
            if cb is not None and cb != 0.0 and _is_pow2(cb):
                return self.binary(MUL, a, self.const(1.0 / cb))  # exact
            if ca == 1.0:
                return self.unary(RECIP, b)
        elif op == POW:
            if cb is not None:
                if cb == 0.0:
                    return self._one
                if cb == 1.0:
                    return a
                if cb == 2.0:
                    return self.unary(SQR, a)
                if cb == 0.5:
                    return self.unary(SQRT, a)
                if cb == -1.0:
                    return self.unary(RECIP, a)
                if cb == 3.0:
                    return self.binary(MUL, self.unary(SQR, a), a)
                if cb == 4.0:
                    return self.unary(SQR, self.unary(SQR, a))
                if cb == -2.0:
                    return self.unary(RECIP, self.unary(SQR, a))
                if cb == 1.5:
                    return self.binary(MUL, a, self.unary(SQRT, a))
                if cb == -0.5:
                    return self.unary(RECIP, self.unary(SQRT, a))
            if ca == 1.0:
                return self._one
        elif op in (MIN, MAX):
            if a == b:
                return a
        if op in COMMUTATIVE and a > b:
            a, b = b, a
        return self._new(op, a, b)

    def add(self, a, b):
        return self.binary(ADD, a, b)

    def sub(self, a, b):
        return self.binary(SUB, a, b)

    def mul(self, a, b):
        return self.binary(MUL, a, b)

    def div(self, a, b):
        return self.binary(DIV, a, b)

    def pow(self, a, b):
        return self.binary(POW, a, b)

    def select(self, c, a, b) -> int:
        if self.op[c] == CONST:
            return a if self.val[c] != 0.0 else b
        if a == b:
            return a
        return self._new(SELECT, c, a, b)

    def sum(self, terms) -> int:
        """Left-to-right sum (matches a CSR row dot product order)."""
        acc = self._zero
        for t in terms:
            acc = self.add(acc, t)
        return acc

    # -- analysis ---------------------------------------------------------
    def args(self, n):
        op = self.op[n]
        if op in LEAF_OPS:
            return ()
        if op == SELECT:
            return (self.a0[n], self.a1[n], self.a2[n])
        if op in UNARY_OPS:
            return (self.a0[n],)
        return (self.a0[n], self.a1[n])

    def reachable(self, roots) -> np.ndarray:
        """Sorted ids of all nodes reachable from roots (ids are topological)."""
        seen = np.zeros(len(self.op), dtype=bool)
        stack = [int(r) for r in roots]
        while stack:
            n = stack.pop()
            if seen[n]:
                continue
            seen[n] = True
            for a in self.args(n):
                if not seen[a]:
                    stack.append(a)
        return np.nonzero(seen)[0]

    def depends(self, roots):
        """(set of state indices, set of input indices, uses_time) of roots."""
        ids = self.reachable(roots)
        st, inp, tm = set(), set(), False
        for n in ids:
            o = self.op[n]
            if o == STATE:
                st.add(int(self.val[n]))
            elif o == INPUT:
                inp.add(int(self.val[n]))
            elif o == TIME:
                tm = True
        return st, inp, tm

    # -- numpy evaluation (test/oracle helper; vectorised over a batch) ----
    def evaluate(self, roots, t=0.0, y=None, p=None, cj=0.0):
        """Evaluate roots. y: (N,) or (B,N); p: (P,) or (B,P). Returns (len(roots),) or (B,len(roots))."""
        roots = np.asarray(roots, dtype=np.int64).reshape(-1)
        y = None if y is None else np.asarray(y, dtype=np.float64)
        p = None if p is None else np.asarray(p, dtype=np.float64)
        batched = (y is not None and y.ndim == 2) or (p is not None and p.ndim == 2) or np.ndim(t) == 1
        ids = self.reachable(roots)
        vals = {}
        with np.errstate(all="ignore"):
            for n in ids:
                o = self.op[n]
                if o == CONST:
                    v = np.float64(self.val[n])
                elif o == STATE:
                    v = y[..., int(self.val[n])]
                elif o == INPUT:
                    v = p[..., int(self.val[n])]
                elif o == TIME:
                    v = np.asarray(t, dtype=np.float64)
                elif o == CJ:
                    v = np.float64(cj)
                elif o == SELECT:
                    v = np.where(vals[self.a0[n]] != 0.0, vals[self.a1[n]], vals[self.a2[n]])
                elif o in _UNARY_NUMPY:
                    v = _UNARY_NUMPY[o](vals[self.a0[n]])
                else:
                    v = _BINARY_NUMPY[o](vals[self.a0[n]], vals[self.a1[n]])
                vals[n] = v
        if batched:
            B = 1
            for arr in (y, p):
                if arr is not None and arr.ndim == 2:
                    B = arr.shape[0]
            if np.ndim(t) == 1:
                B = len(t)
            out = np.empty((B, len(roots)))
            for k, r in enumerate(roots):
                out[:, k] = vals[int(r)]
            return out
        return np.array([float(vals[int(r)]) for r in roots])

    # -- differentiation --------------------------------------------------
    def gradients(self, roots, wrt="state"):
        """Sparse forward-mode derivatives.

        Returns a list (one per root) of dicts {index: node id of d root / d x_index}
        for x = states (wrt='state') or inputs (wrt='input').  Structural zeros are
        absent; entries that simplify to the constant 0 are dropped.
        """
        leaf = STATE if wrt == "state" else INPUT
        ids = self.reachable(roots)
        grad: dict[int, dict[int, int]] = {}
        for n in ids:
            n = int(n)
            o = self.op[n]
            if o in LEAF_OPS:
                grad[n] = {int(self.val[n]): self._one} if o == leaf else {}
                continue
            a = self.a0[n]
            ga = grad[a]
            if o in UNARY_OPS:
                if not ga or o in (SIGN, FLOOR, CEIL):
                    grad[n] = {}
                    continue
                d = self._unary_derivative(o, a, n)
                grad[n] = self._scale(ga, d)
                continue
            if o == SELECT:
                b, c = self.a1[n], self.a2[n]
                gb, gc = grad[b], grad[c]
                out = {}
                for k in set(gb) | set(gc):
                    out[k] = self.select(a, gb.get(k, self._zero), gc.get(k, self._zero))
                grad[n] = self._clean(out)
                continue
            b = self.a1[n]
            gb = grad[b]
            if not ga and not gb:
                grad[n] = {}
                continue
            if o == ADD:
                grad[n] = self._combine(ga, gb, None, None)
            elif o == SUB:
                grad[n] = self._combine(ga, self._scale(gb, self.const(-1.0)), None, None)
            elif o == MUL:
                grad[n] = self._combine(ga, gb, b, a)
            elif o == DIV:
                # d(a/b) = a'/b - (a/b) b'/b
                inv = self.unary(RECIP, b)
                grad[n] = self._combine(ga, gb, inv, self.neg(self.mul(n, inv)))
            elif o == POW:
                da = db = None
                if ga:
                    if self.op[b] == CONST:
                        e = self.val[b]
                        da = self.mul(b, self.pow(a, self.const(e - 1.0)))
                    else:
                        da = self.mul(b, self.pow(a, self.sub(b, self._one)))
                if gb:
                    db = self.mul(n, self.unary(LOG, a))
                grad[n] = self._combine(ga if ga else {}, gb if gb else {}, da, db)
            elif o == MIN:
                # reference convention (binary_operators.py:923-929): ties go left
                w = self.binary(LE, a, b)
                grad[n] = self._combine(ga, gb, w, self.sub(self._one, w))
            elif o == MAX:
                w = self.binary(LE, b, a)
                grad[n] = self._combine(ga, gb, w, self.sub(self._one, w))
            elif o == HYPOT:
                inv = self.unary(RECIP, n)
                grad[n] = self._combine(ga, gb, self.mul(a, inv), self.mul(b, inv))
            elif o == MOD:
                grad[n] = self._combine(ga, gb, None, self.neg(self.unary(FLOOR, self.div(a, b))))
            elif o in (LT, LE):
                grad[n] = {}
            else:  # pragma: no cover
                raise NotImplementedError(OP_NAMES[o])
        return [grad[int(r)] for r in roots]

    def _clean(self, g):
        return {k: v for k, v in g.items() if not (self.op[v] == CONST and self.val[v] == 0.0)}

    def _scale(self, g, d):
        if d is None:
            return dict(g)
        return self._clean({k: self.mul(d, v) for k, v in g.items()})

    def _combine(self, ga, gb, da, db):
        """da*ga + db*gb (None scale = 1)."""
        out = {}
        for k, v in ga.items():
            out[k] = v if da is None else self.mul(da, v)
        for k, v in gb.items():
            w = v if db is None else self.mul(db, v)
            out[k] = self.add(out[k], w) if k in out else w
        return self._clean(out)

    def _unary_derivative(self, o, a, n):
        """d op(a) / d a as a node (n = id of op(a))."""
        if o == NEG:
            return self.const(-1.0)
        if o == ABS:
            return self.unary(SIGN, a)
        if o == EXP:
            return n
        if o == LOG:
            return self.unary(RECIP, a)
        if o == LOG10:
            return self.unary(RECIP, self.mul(self.const(math.log(10.0)), a))
        if o == SQRT:
            return self.unary(RECIP, self.mul(self.const(2.0), n))
        if o == TANH:
            return self.unary(RECIP, self.unary(SQR, self.unary(COSH, a)))
        if o == SINH:
            return self.unary(COSH, a)
        if o == COSH:
            return self.unary(SINH, a)
        if o == SIN:
            return self.unary(COS, a)
        if o == COS:
            return self.neg(self.unary(SIN, a))
        if o == ASINH:
            return self.unary(RECIP, self.unary(SQRT, self.add(self.unary(SQR, a), self._one)))
        if o == ATAN:
            return self.unary(RECIP, self.add(self.unary(SQR, a), self._one))
        if o == ERF:
            return self.mul(self.const(2.0 / math.sqrt(math.pi)), self.unary(EXP, self.neg(self.unary(SQR, a))))
        if o == ERFC:
            return self.mul(self.const(-2.0 / math.sqrt(math.pi)), self.unary(EXP, self.neg(self.unary(SQR, a))))
        if o == RECIP:
            return self.neg(self.unary(SQR, n))
        if o == SQR:
            return self.mul(self.const(2.0), a)
        raise NotImplementedError(OP_NAMES[o])  # pragma: no cover

    # -- (de)serialisation -------------------------------------------------
    def to_arrays(self):
        return {
            "op": np.asarray(self.op, dtype=np.int16),
            "a0": np.asarray(self.a0, dtype=np.int32),
            "a1": np.asarray(self.a1, dtype=np.int32),
            "a2": np.asarray(self.a2, dtype=np.int32),
            "val": np.asarray(self.val, dtype=np.float64),
        }

    @classmethod
    def from_arrays(cls, d):
        g = cls.__new__(cls)
        g.op = [int(x) for x in d["op"]]
        g.a0 = [int(x) for x in d["a0"]]
        g.a1 = [int(x) for x in d["a1"]]
        g.a2 = [int(x) for x in d["a2"]]
        g.val = [float(x) for x in d["val"]]
        g._cse = {}
        g.n_states = 0
        g.n_inputs = 0
        for n, o in enumerate(g.op):
            if o == CONST:
                key = (o, np.float64(g.val[n]).tobytes())
            else:
                key = (o, g.a0[n], g.a1[n], g.a2[n], g.val[n] if o in LEAF_OPS else 0.0)
            g._cse[key] = n
            if o == STATE:
                g.n_states = max(g.n_states, int(g.val[n]) + 1)
            elif o == INPUT:
                g.n_inputs = max(g.n_inputs, int(g.val[n]) + 1)
        g._zero = g.const(0.0)
        g._one = g.const(1.0)
        return g

    def compact(self, roots_lists):
        """Return (new_graph, remapped roots lists) keeping only reachable nodes."""
        allroots = [int(r) for rl in roots_lists for r in rl]
        ids = self.reachable(allroots)
        remap = {}
        g = Graph()
        for n in ids:
            n = int(n)
            o = self.op[n]
            if o == CONST:
                m = g.const(self.val[n])
            elif o == STATE:
                m = g.state(self.val[n])
            elif o == INPUT:
                m = g.input(self.val[n])
            elif o == TIME:
                m = g.time()
            elif o == CJ:
                m = g.cj()
            elif o == SELECT:
                m = g._new(o, remap[self.a0[n]], remap[self.a1[n]], remap[self.a2[n]])
            elif o in UNARY_OPS:
                m = g._new(o, remap[self.a0[n]])
            else:
                m = g._new(o, remap[self.a0[n]], remap[self.a1[n]])
            remap[n] = m
        g.n_states = max(g.n_states, self.n_states)
        g.n_inputs = max(g.n_inputs, self.n_inputs)
        return g, [[remap[int(r)] for r in rl] for rl in roots_lists]


def _is_pow2(x: float) -> bool:
    m, _ = math.frexp(abs(x))
    return m == 0.5
