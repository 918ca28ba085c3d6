"""Re-roll a scalar DAG into staged, team-parallel *templates*.

The discretised battery models are a few hundred vector operations over mesh
domains; after scalarisation (``frontend.py``) that structure is still there as
thousands of scalar roots that are *the same expression up to leaf indices and
constants* (one per mesh node).  This pass recovers it:

* every root (a residual row, a Jacobian entry, an event, ...) is reduced to its
  local expression DAG down to leaves (states, inputs, constants, time, cj) or
  *materialised temporaries*;
* roots with the same local DAG shape are grouped into one ``Template`` whose
  per-instance differences (state indices, constants, temporary slots, output
  index) live in small tables;
* sub-expressions that are expensive and shared between different roots are
  materialised once into a scratch vector (``tmp``) in an earlier *stage*;
  stages are separated by a team barrier.

On the GPU a template instance is one lane's work item, a stage is one pass of
the team's warps over (template, 32-instance tile) pairs.  This plays the role of
CasADi's SX instruction tape (reference: CasadiFunctions.cpp:96-105, built by
base_solver.py:1830-2002) but is data-parallel by construction.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import ir

MERGE_INPUTS = bool(int(__import__("os").environ.get("PBGPU_MERGE_INPUTS", "1")))   # see extract(): input leaves are slots
MATERIALISE_COST = 6      # DFMA-equivalents: shared sub-expressions at least this costly get a tmp slot
SPLIT_COST = int(__import__("os").environ.get("PBGPU_SPLIT_COST", "120"))   # for functions with few roots: split big trees at nodes at least this costly
FEW_ROOTS = 64
BIG_SUM_COST = int(__import__("os").environ.get("PBGPU_BIG_SUM_COST", "1500"))   # expression trees at least this costly have their sums split ...
SPLIT_SUMS_ALWAYS = bool(int(__import__("os").environ.get("PBGPU_SPLIT_SUMS", "0")))   # A/B switch: also for the models that do not need it
_BIG_SUM = 10 ** 9
SUM_TERM_COST = 8                                                                # ... at terms at least this costly


@dataclass
class Template:
    key: tuple
    body: list                      # [(op, (ref, ...))]; ref = ('v', k) | ('s', slot) | ('c', slot) | ('m', slot) | ('i', j) | ('t',) | ('j',)
    result: tuple                   # ref of the result
    n_s: int
    n_c: int
    n_m: int
    n_i: int = 0                    # input-parameter slots ('i', slot): per instance an index into p[]
    out: list = field(default_factory=list)       # per instance: output index (into out vector or tmp)
    s_idx: list = field(default_factory=list)     # per instance: list of state indices
    c_val: list = field(default_factory=list)     # per instance: list of constants
    m_idx: list = field(default_factory=list)     # per instance: list of tmp slots
    i_idx: list = field(default_factory=list)     # per instance: list of input indices
    tops: list = field(default_factory=list)      # per instance: the DAG node this instance evaluates
    cost: int = 0

    @property
    def n_inst(self):
        return len(self.out)


@dataclass
class Stage:
    to_tmp: list      # templates writing tmp[]
    to_out: list      # templates writing out[]


@dataclass
class Program:
    name: str
    n_out: int
    n_tmp: int
    stages: list
    flops: int = 0            # DFMA-equivalent cost of one evaluation (OP_COST model)
    n_transcendental: int = 0
    tmp_nodes: frozenset = frozenset()   # DAG nodes that own a tmp slot


_SPLIT_OPS = {ir.EXP, ir.LOG, ir.TANH, ir.SINH, ir.COSH, ir.SIN, ir.COS, ir.ASINH, ir.ATAN, ir.ERF, ir.ERFC, ir.LOG10, ir.POW}
_TRANSCENDENTAL = {ir.EXP, ir.LOG, ir.TANH, ir.SINH, ir.COSH, ir.SIN, ir.COS, ir.ASINH, ir.ATAN,
                   ir.ERF, ir.ERFC, ir.LOG10, ir.POW, ir.SQRT, ir.HYPOT, ir.DIV, ir.RECIP}


# Splitting of heavy, narrow templates (``split=True``): the electrode-kinetics rows of the DFN are two templates of 20
# instances and ~500 DFMA-equivalents each -- fifteen transcendental calls in a row on 20 lanes of one warp while the
# other seven warps of the team wait at the stage barrier (a third of the residual's critical path).  A second pass
# materialises the transcendental nodes inside such templates: each becomes an instance of a small shared template
# (``tanh(c0 * s0 + c1)`` x 60) in an earlier stage, spread over all the warps.
SPLIT_TEMPLATE_COST = 100     # a template at least this costly ...
SPLIT_TEMPLATE_INST = 64      # ... with at most this many instances is split


def build_program(g: ir.Graph, roots, name="f", out_index=None, n_out=None, split=False, split_sums=False) -> Program:
    """roots: node ids; root k is written to out[out_index[k]] (default k).
    split_sums: the terms of very large sums become parallel template instances (BIG_SUM_COST)."""
    global _BIG_SUM
    _BIG_SUM = BIG_SUM_COST if (split_sums or SPLIT_SUMS_ALWAYS) else 10 ** 9
    prog = _build_program(g, roots, name, out_index, n_out, frozenset())
    if not split:
        return prog
    forced = set()
    for _ in range(4):    # materialising a node exposes the (now shorter) templates around it: iterate to a fixed point
        new = set()
        for st in prog.stages:
            for t in st.to_tmp + st.to_out:
                if t.cost >= SPLIT_TEMPLATE_COST and t.n_inst <= SPLIT_TEMPLATE_INST:
                    for top in t.tops:
                        new |= _heavy_nodes(g, top, prog.tmp_nodes)
        new -= forced
        if not new:
            break
        forced |= new
        prog = _build_program(g, roots, name, out_index, n_out, frozenset(forced))
    return prog


def _heavy_nodes(g, top, tmp_nodes):
    """Transcendental nodes strictly inside the local DAG of ``top`` (down to leaves / materialised nodes)."""
    out, seen, stack = set(), set(), [a for a in g.args(top)]
    while stack:
        x = stack.pop()
        if x in seen or x in tmp_nodes or g.op[x] in ir.LEAF_OPS:
            continue
        seen.add(x)
        if g.op[x] in _SPLIT_OPS:
            out.add(x)
        stack.extend(g.args(x))
    return out


def _build_program(g: ir.Graph, roots, name, out_index, n_out, forced) -> Program:
    roots = [int(r) for r in roots]
    if out_index is None:
        out_index = list(range(len(roots)))
    out_index = [int(i) for i in out_index]
    if n_out is None:
        n_out = (max(out_index) + 1) if out_index else 0
    if not roots:
        return Program(name, n_out, 0, [])

    ids = g.reachable(roots)
    op, a0, a1, a2 = g.op, g.a0, g.a1, g.a2

    # tree cost (capped) ------------------------------------------------------
    tc = {}
    for n in ids:
        n = int(n)
        c = ir.OP_COST[op[n]]
        for a in g.args(n):
            c += tc[a]
        tc[n] = min(c, 10 ** 6)

    # owners: top-down (ids are topological, parents have larger ids) -----------
    # owner[n] = the single consumer (a root slot or a materialised node) or MULTI
    MULTI = -2
    owner = {}

    def claim(a, tok):
        prev = owner.get(a)
        owner[a] = tok if (prev is None or prev == tok) else MULTI

    for k, r in enumerate(roots):
        claim(r, (0, k))
    materialised = set()
    few = len(roots) <= FEW_ROOTS
    sum_term = set()      # nodes that are a term of a sum (argument of an ADD / SUB and not one themselves)
    _SUM = (ir.ADD, ir.SUB)
    for n in ids[::-1]:
        n = int(n)
        o = owner[n]
        if op[n] in _SUM:
            for a in g.args(n):
                if op[a] not in _SUM:
                    sum_term.add(a)
        if op[n] not in ir.LEAF_OPS:
            if n in forced and not (o != MULTI and o[0] == 0 and roots[o[1]] == n):
                materialised.add(n)
                o = (1, n)
            elif o == MULTI:
                if tc[n] >= MATERIALISE_COST:
                    materialised.add(n)
                    o = (1, n)
            elif (n in sum_term and tc[n] >= SUM_TERM_COST and op[n] not in _SUM
                  and tc[roots[o[1]] if o[0] == 0 else o[1]] >= _BIG_SUM):
                # a term of a very large sum (the x-integral of the heat source in a lumped thermal model: one root, 3 000
                # operations on one lane while the rest of the team waits): the terms become instances of a shared template
                # in an earlier stage, the sum itself adds their slots
                materialised.add(n)
                o = (1, n)
            elif few and o[0] == 1 and tc[n] >= SPLIT_COST and op[n] in _TRANSCENDENTAL:
                materialised.add(n)
                o = (1, n)
            elif few and o[0] == 0 and roots[o[1]] != n and tc[n] >= SPLIT_COST and op[n] in _TRANSCENDENTAL:
                materialised.add(n)
                o = (1, n)
        for a in g.args(n):
            claim(a, o)

    mat_list = sorted(materialised)
    tmp_slot = {n: k for k, n in enumerate(mat_list)}

    # local DAG extraction ------------------------------------------------------
    shash = {}
    H_S, H_C, H_T, H_J, H_M, H_I = 101, 102, 103, 104, 105, 106

    def struct_hash(n):
        """Structure hash with leaves abstracted (deterministic: ints only)."""
        stack = [n]
        while stack:
            x = stack[-1]
            if x in shash:
                stack.pop()
                continue
            o = op[x]
            if x in tmp_slot:
                shash[x] = H_M
            elif o == ir.STATE:
                shash[x] = H_S
            elif o == ir.CONST:
                shash[x] = H_C
            elif o == ir.INPUT:
                shash[x] = H_I if MERGE_INPUTS else hash((200, int(g.val[x])))
            elif o == ir.TIME:
                shash[x] = H_T
            elif o == ir.CJ:
                shash[x] = H_J
            else:
                args = g.args(x)
                missing = [a for a in args if a not in shash]
                if missing:
                    stack.extend(missing)
                    continue
                hs = [shash[a] for a in args]
                if o in ir.COMMUTATIVE:
                    hs.sort()
                shash[x] = hash((o, tuple(hs)))
            stack.pop()
        return shash[n]

    def extract(top, define):
        """Local DAG of ``top`` down to leaves/temps -> (key, body, result, s, c, m).

        define=True: ``top`` is a materialised node whose definition is wanted
        (it is expanded even though it has a tmp slot)."""
        local = {}
        body = []
        s_list, c_list, m_list, i_list = [], [], [], []
        s_slot, c_slot, m_slot, i_slot = {}, {}, {}, {}

        def leaf_ref(x):
            o = op[x]
            if x in tmp_slot and not (define and x == top):
                if x not in m_slot:
                    m_slot[x] = len(m_list)
                    m_list.append(tmp_slot[x])
                return ("m", m_slot[x])
            if o == ir.STATE:
                i = int(g.val[x])
                if i not in s_slot:
                    s_slot[i] = len(s_list)
                    s_list.append(i)
                return ("s", s_slot[i])
            if o == ir.CONST:
                if x not in c_slot:
                    c_slot[x] = len(c_list)
                    c_list.append(g.val[x])
                return ("c", c_slot[x])
            if o == ir.INPUT:
                # a slot, not the index itself: the negative- and positive-electrode copies of an expression differ only in
                # which input (diffusivity, ...) they read, and share one template this way
                j = int(g.val[x])
                if not MERGE_INPUTS:
                    i_slot.setdefault(j, j)
                    while len(i_list) <= j:
                        i_list.append(len(i_list))
                    return ("i", j)
                if j not in i_slot:
                    i_slot[j] = len(i_list)
                    i_list.append(j)
                return ("i", i_slot[j])
            if o == ir.TIME:
                return ("t",)
            if o == ir.CJ:
                return ("j",)
            return None

        # iterative post-order with canonical child order
        stack = [(top, False)]
        while stack:
            x, expanded = stack.pop()
            if x in local:
                continue
            lr = leaf_ref(x)
            if lr is not None:
                local[x] = lr
                continue
            args = list(g.args(x))
            if op[x] in ir.COMMUTATIVE:
                args.sort(key=struct_hash)
            if not expanded:
                stack.append((x, True))
                for a in reversed(args):
                    if a not in local:
                        stack.append((a, False))
                continue
            refs = tuple(local[a] for a in args)
            local[x] = ("v", len(body))
            body.append((op[x], refs))
        result = local[top]
        key = (tuple(body), result)
        return key, body, result, s_list, c_list, m_list, i_list

    # levels --------------------------------------------------------------------
    templates_tmp: dict = {}
    templates_out: dict = {}
    level_of_tmp = {}

    def add(top, dest_index, to_tmp):
        key, body, result, s_list, c_list, m_list, i_list = extract(top, to_tmp)
        lvl = 0
        for slot in m_list:
            lvl = max(lvl, level_of_tmp[slot] + 1)
        store = templates_tmp if to_tmp else templates_out
        tk = (lvl, key)
        t = store.get(tk)
        if t is None:
            cost = sum(ir.OP_COST[o] for o, _ in body)
            t = Template(key, body, result, len(s_list), len(c_list), len(m_list), n_i=len(i_list), cost=cost)
            store[tk] = t
        t.out.append(dest_index)
        t.tops.append(top)
        t.s_idx.append(s_list)
        t.c_val.append(c_list)
        t.m_idx.append(m_list)
        t.i_idx.append(i_list)
        return lvl

    # materialised nodes in topological (id) order so dependencies have levels first
    for n in mat_list:
        level_of_tmp[tmp_slot[n]] = add(n, tmp_slot[n], True)
    for k, r in enumerate(roots):
        add(r, out_index[k], False)

    n_levels = 1 + max([lv for lv, _ in templates_tmp] + [lv for lv, _ in templates_out])
    stages = [Stage([], []) for _ in range(n_levels)]
    for (lv, _), t in templates_tmp.items():
        stages[lv].to_tmp.append(t)
    for (lv, _), t in templates_out.items():
        stages[lv].to_out.append(t)
    for st in stages:
        st.to_tmp.sort(key=lambda t: -t.cost * t.n_inst)
        st.to_out.sort(key=lambda t: -t.cost * t.n_inst)

    prog = Program(name, n_out, len(mat_list), stages)
    prog.tmp_nodes = frozenset(mat_list)
    for st in stages:
        for t in st.to_tmp + st.to_out:
            prog.flops += t.cost * t.n_inst
            prog.n_transcendental += t.n_inst * sum(1 for o, _ in t.body if o in _TRANSCENDENTAL)
    return prog


# ---------------------------------------------------------------------------
# reference interpreter of a Program (tests: Program == Graph.evaluate)
# ---------------------------------------------------------------------------
def run_program(prog: Program, t=0.0, y=None, p=None, cj=0.0):
    out = np.zeros(prog.n_out)
    tmp = np.zeros(max(prog.n_tmp, 1))

    def val(ref, vals, s, c, m, ii):
        k = ref[0]
        if k == "v":
            return vals[ref[1]]
        if k == "s":
            return y[s[ref[1]]]
        if k == "c":
            return c[ref[1]]
        if k == "m":
            return tmp[m[ref[1]]]
        if k == "i":
            return p[ii[ref[1]]]
        if k == "t":
            return t
        return cj

    with np.errstate(all="ignore"):
        for st in prog.stages:
            for dest, tl in ((tmp, st.to_tmp), (out, st.to_out)):
                for tp in tl:
                    for q in range(tp.n_inst):
                        s, c, m, ii = tp.s_idx[q], tp.c_val[q], tp.m_idx[q], tp.i_idx[q]
                        vals = []
                        for o, refs in tp.body:
                            a = [np.float64(val(r, vals, s, c, m, ii)) for r in refs]
                            if o == ir.SELECT:
                                v = a[1] if a[0] != 0.0 else a[2]
                            elif o in ir._UNARY_NUMPY:
                                v = ir._UNARY_NUMPY[o](a[0])
                            else:
                                v = ir._BINARY_NUMPY[o](a[0], a[1])
                            vals.append(np.float64(v))
                        dest[tp.out[q]] = val(tp.result, vals, s, c, m, ii)
    return out
