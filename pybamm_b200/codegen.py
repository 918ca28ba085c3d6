"""Emit model functions (residual, J - cj*M, events, initial conditions, output
variables) and the static sparse-LU tables as C / CUDA source.

Replaces the CasADi functions that ``IDAKLUSolver.set_up`` serialises for the native
solver (packages/pybamm/src/pybamm/solvers/idaklu_solver.py:402-628) and the optional
gcc AOT path (packages/pybamm/src/pybamm/codegen/compilation.py:61-196).

One generator, two dialects selected by preprocessor macros of the including file:

* ``cuda``: team-parallel device functions for sm_100a.  A *team* (one warp or one
  CTA) evaluates one system; every stage walks (template, 32-instance tile) pairs
  warp by warp; per-instance tables are slot-major so that lanes read consecutive
  entries; stages are separated by ``PB_SYNC()``.
* ``c``: the same programs as plain scalar C loops for the CPU oracle / CPU baseline
  (``oracle/``).  The Jacobian is written in the bundle's CSC entry order there, and
  into the LU slots of ``symbolic_lu`` in the CUDA dialect.
"""
from __future__ import annotations

import hashlib

import numpy as np

from . import ir
from .program import Program, build_program
from . import symbolic_lu as slu

_FMT = "%.17g"


def _dlit(v):
    v = float(v)
    if np.isnan(v):
        return "PB_NAN"
    if np.isinf(v):
        return "PB_INF" if v > 0 else "(-PB_INF)"
    s = _FMT % v
    if "." not in s and "e" not in s and "n" not in s:
        s += ".0"
    return s


_UN = {
    ir.NEG: "(-{a})", ir.ABS: "fabs({a})", ir.EXP: "PB_EXP({a})", ir.LOG: "PB_LOG({a})", ir.SQRT: "PB_SQRT({a})",
    ir.TANH: "PB_TANH({a})", ir.SINH: "PB_SINH({a})", ir.COSH: "PB_COSH({a})", ir.SIN: "sin({a})", ir.COS: "cos({a})",
    ir.ASINH: "PB_ASINH({a})", ir.ATAN: "atan({a})", ir.ERF: "erf({a})", ir.ERFC: "erfc({a})",
    ir.FLOOR: "floor({a})", ir.CEIL: "ceil({a})", ir.RECIP: "(1.0 / {a})", ir.SQR: "({a} * {a})",
    ir.LOG10: "log10({a})", ir.SIGN: "(({a} > 0.0) - ({a} < 0.0))",
}
_BIN = {
    ir.ADD: "({a} + {b})", ir.SUB: "({a} - {b})", ir.MUL: "({a} * {b})", ir.DIV: "({a} / {b})",
    ir.POW: "PB_POW({a}, {b})", ir.MIN: "fmin({a}, {b})", ir.MAX: "fmax({a}, {b})",
    ir.LT: "(({a} < {b}) ? 1.0 : 0.0)", ir.LE: "(({a} <= {b}) ? 1.0 : 0.0)",
    ir.MOD: "({a} - floor({a} / {b}) * {b})", ir.HYPOT: "hypot({a}, {b})",
}


# multi-warp teams without the tile dispatch (every warp walks the templates and takes its tiles): measured on the DFN, 5 236
# against 5 286 sims/s -- off
WARP_LOOPS = bool(int(__import__("os").environ.get("PBGPU_WARP_LOOPS", "0")))
# one-warp teams: the templates of a stage as plain loops instead of the tile dispatch (A/B switch)
ONE_WARP_LOOPS = bool(int(__import__("os").environ.get("PBGPU_ONE_WARP_LOOPS", "1")))

# ops that compile to a call of an out-of-line libm copy (pbgpu_team.cuh) and can be issued 2 or 4 at a time (PB_B2 / PB_B4)
_BATCH_UN = {ir.EXP: "exp", ir.LOG: "log", ir.TANH: "tanh", ir.SINH: "sinh", ir.COSH: "cosh", ir.ASINH: "asinh", ir.SQRT: "sqrt"}
# Measured and left off (default 1): libdevice's exp / log / tanh branch on their argument ranges, so the copies inside a
# 2- / 4-wide routine are NOT interleaved by the compiler and the latency stays (SPM events 8 735 against 8 610 cycles) while
# the code grows: SPM 92.6k (4-wide) / 120.2k (2-wide) against 124.8k sims/s, SPMe 17.2k / 19.1k / 19.7k, DFN 4 892 / 4 831 / 5 134.
BATCH_W = int(__import__("os").environ.get("PBGPU_BATCH_W", "1"))


class _Emitter:
    def __init__(self, prefix):
        self.prefix = prefix
        self.tables = []     # lines
        self.code = []
        self.table_bytes = 0
        self._ntab = 0
        self.batch = False

    def table(self, ctype, values, hint="t", qual="PB_TABLE"):
        name = f"{self.prefix}_{hint}{self._ntab}"
        self._ntab += 1
        vals = list(values)
        if ctype == "double":
            body = ", ".join(_dlit(v) for v in vals)
            self.table_bytes += 8 * len(vals)
        else:
            body = ", ".join(str(int(v)) for v in vals)
            self.table_bytes += (2 if "short" in ctype else 4) * len(vals)
        self.tables.append(f"static {qual} const {ctype} {name}[{max(len(vals), 1)}] = {{{body if vals else '0'}}};")
        return name


def _itype(maxval):
    return "unsigned short" if maxval < 65536 else "int"


# Per-instance tables of mesh models repeat: the stencil coefficients of a particle depend on the radial node only, the
# state indices advance by a fixed stride from one particle to the next.  Such columns are stored once per period
# (DFN residual: 50 KB of tables -> 14 KB, which is what decides whether they stay in the 28 KB of L1 left beside the
# shared-memory carve-out).  PBGPU_TABLE_COMPRESS=0 emits the plain per-instance tables.
TABLE_COMPRESS = bool(int(__import__("os").environ.get("PBGPU_TABLE_COMPRESS", "1")))
_MAX_PERIOD = 64


def _period(vals):
    """Smallest P >= 2 with vals[q + P] - vals[q] constant for all q (0 for float columns), or None."""
    n = len(vals)
    for P in range(2, min(n // 2, _MAX_PERIOD) + 1):
        d = vals[P:] - vals[:-P]
        if np.all(d == d[0]) and (d[0] == 0 or np.issubdtype(vals.dtype, np.integer)):
            return P, d[0]
    return None


def _blocked(vals):
    """Smallest B >= 2 dividing n with vals constant inside every block of B, or None."""
    n = len(vals)
    for B in range(2, n // 2 + 1):
        if n % B == 0 and n // B <= 4 * _MAX_PERIOD:
            m = vals.reshape(-1, B)
            if np.all(m == m[:, :1]):
                return B
    return None


def _slot_expr(em, values, q, hint):
    """Expression for per-instance integer ``values`` indexed by instance ``q``."""
    vals = np.asarray(values, dtype=np.int64)
    if len(vals) == 0:
        return "0"
    if np.all(vals == vals[0]):
        return str(int(vals[0]))
    d = vals[1] - vals[0]
    if len(vals) > 2 and np.all(np.diff(vals) == d):
        return f"({int(vals[0])} + {int(d)} * {q})"
    if len(vals) == 2:
        return f"({int(vals[0])} + {int(d)} * {q})"
    if TABLE_COMPRESS and len(vals) >= 8:
        pr = _period(vals)
        if pr is not None:
            P, stride = pr
            base = vals[:P]
            dd = np.diff(base)
            if np.all(dd == dd[0]):     # affine inside the period as well: no table at all
                return f"({int(base[0])} + {int(dd[0])} * ({q} % {P}) + {int(stride)} * ({q} / {P}))"
            name = em.table(_itype(int(base.max())), base, hint)
            return f"((int)PB_LD({name}[{q} % {P}]) + {int(stride)} * ({q} / {P}))"
    name = em.table(_itype(int(vals.max())), vals, hint)
    return f"(int)PB_LD({name}[{q}])"


def _const_expr(em, col, q):
    """Expression for a per-instance constant column (not all equal)."""
    if TABLE_COMPRESS and len(col) >= 8:
        pr = _period(col)
        if pr is not None:
            P = pr[0]
            name = em.table("double", col[:P], "c")
            return f"PB_LD({name}[{q} % {P}])"
        B = _blocked(col)
        if B is not None:
            name = em.table("double", col[::B], "c")
            return f"PB_LD({name}[{q} / {B}])"
    name = em.table("double", col, "c")
    return f"PB_LD({name}[{q}])"


def _emit_template(em: _Emitter, tp, dest, lines, ind):
    """Body of one template for instance index ``q`` (already in scope)."""
    n = tp.n_inst
    s_idx = np.asarray(tp.s_idx, dtype=np.int64).reshape(n, tp.n_s)
    c_val = np.asarray(tp.c_val, dtype=np.float64).reshape(n, tp.n_c)
    m_idx = np.asarray(tp.m_idx, dtype=np.int64).reshape(n, tp.n_m)
    refs = {}
    for k in range(tp.n_s):
        e = _slot_expr(em, s_idx[:, k], "q", "s")
        lines.append(f"{ind}const double s{k} = y[{e}];")
        refs[("s", k)] = f"s{k}"
    for k in range(tp.n_c):
        col = c_val[:, k]
        if np.all(col == col[0]) or (np.all(np.isnan(col))):
            refs[("c", k)] = _dlit(col[0])
        else:
            lines.append(f"{ind}const double c{k} = {_const_expr(em, col, 'q')};")
            refs[("c", k)] = f"c{k}"
    i_idx = np.asarray(tp.i_idx, dtype=np.int64).reshape(n, tp.n_i)
    for k in range(tp.n_i):
        col = i_idx[:, k]
        if np.all(col == col[0]):
            refs[("i", k)] = f"p[{int(col[0])}]"
        else:
            e = _slot_expr(em, col, "q", "i")
            lines.append(f"{ind}const double i{k} = p[{e}];")
            refs[("i", k)] = f"i{k}"
    for k in range(tp.n_m):
        e = _slot_expr(em, m_idx[:, k], "q", "m")
        lines.append(f"{ind}const double m{k} = tmp[{e}];")
        refs[("m", k)] = f"m{k}"

    def ref(r):
        kind = r[0]
        if kind == "v":
            return f"v{r[1]}"
        if kind == "t":
            return "t"
        if kind == "j":
            return "cj"
        return refs[r]

    def batch_name(o, args):
        """Short name of the out-of-line libm routine an op compiles to, or None."""
        if o == ir.POW and args[1][0] == "c" and np.all(c_val[:, args[1][1]] == -0.25):
            return "powm025"
        if o == ir.POW and args[1][0] == "c" and np.all(c_val[:, args[1][1]] == -1.25):
            return "powm125"
        return _BATCH_UN.get(o)

    def expr_of(o, args):
        a = [ref(r) for r in args]
        if o == ir.SELECT:
            return f"(({a[0]} != 0.0) ? {a[1]} : {a[2]})"
        if o == ir.POW and args[1][0] == "c" and np.all(c_val[:, args[1][1]] == -0.25):
            return f"PB_POWM025({a[0]})"     # reg_power's x^-1/4 and its derivative x^-5/4: two square roots instead of pow()
        if o == ir.POW and args[1][0] == "c" and np.all(c_val[:, args[1][1]] == -1.25):
            return f"PB_POWM125({a[0]})"
        if o == ir.DIV:
            return f"PB_DIV({a[0]}, {a[1]})"
        if o == ir.RECIP:
            return f"PB_DIV(1.0, {a[0]})"
        if o in _UN:
            return _UN[o].format(a=a[0])
        return _BIN[o].format(a=a[0], b=a[1])

    if BATCH_W < 2 or not em.batch:
        for k, (o, args) in enumerate(tp.body):
            lines.append(f"{ind}const double v{k} = {expr_of(o, args)};")
    else:
        # Independent calls of the same libm routine are issued as ONE call of a 2- or 4-wide copy whose chains the compiler
        # interleaves: a template is a serial FP64 dependency chain per lane (~8 cycles per instruction), and a terminal
        # voltage or a Butler-Volmer row is fifteen such routines in a row.  Level = number of routine calls on the longest
        # path to a node; equal level and name means independent.  Order: per level, the batched calls, then the rest.
        body = tp.body
        lvl = []
        for k, (o, args) in enumerate(body):
            base = max([lvl[r[1]] for r in args if r[0] == "v"], default=0)
            lvl.append(base + (1 if batch_name(o, args) else 0))
        for L in range(0, (max(lvl) if lvl else 0) + 1):
            groups = {}
            for k, (o, args) in enumerate(body):
                nm = batch_name(o, args)
                if nm and lvl[k] == L:
                    groups.setdefault(nm, []).append(k)
            for nm, ks in groups.items():
                while ks:
                    w = 4 if (len(ks) >= 4 and BATCH_W >= 4) else (2 if len(ks) >= 2 else 1)
                    chunk, ks = ks[:w], ks[w:]
                    if w == 1:
                        o, args = body[chunk[0]]
                        lines.append(f"{ind}const double v{chunk[0]} = {expr_of(o, args)};")
                    else:
                        outs = ", ".join(f"v{k}" for k in chunk)
                        ins = ", ".join(ref(body[k][1][0]) for k in chunk)
                        lines.append(f"{ind}PB_B{w}({nm}, {outs}, {ins})")
            for k, (o, args) in enumerate(body):
                if lvl[k] == L and not batch_name(o, args):
                    lines.append(f"{ind}const double v{k} = {expr_of(o, args)};")
    oe = _slot_expr(em, tp.out, "q", "o")
    lines.append(f"{ind}{dest}[{oe}] = {ref(tp.result)};")
    lines.append(f"{ind}PB_PAD")        # empty; the code-size sensitivity experiment defines it as a run of NOPs


def emit_program(prog: Program, fname: str, prefix: str, batch=False, multi=False, one_warp=False):
    """Return (tables_text, function_text).  batch: issue independent libm calls 2 / 4 at a time (PB_B2 / PB_B4).
    multi: the function evaluates ``npts`` points at once, (instance, point) pairs spread over the lanes -- for the output
    variables of one-warp models, whose long scalar expressions leave 31 lanes idle (pbgpu_ida.cuh, PB_VARS_BATCH):
    point k reads y from yb + k * PBM_N, its time from tb[k], writes outb + k * n_out and uses tmpb + k * n_tmp."""
    em = _Emitter(prefix)
    em.batch = batch
    L = []
    if multi:
        L.append(f"PB_FN void {fname}(const double* PB_RESTRICT yb, const double* PB_RESTRICT p, const double* PB_RESTRICT tb, "
                 f"const int npts, double* PB_RESTRICT outb, double* PB_RESTRICT tmpb) {{")
        L.append("  const double cj = 0.0; (void)cj; (void)p;")
        for si, st in enumerate(prog.stages):
            if si > 0:
                L.append("  PB_SYNC();")
            for t, dest in [(t, "tmp") for t in st.to_tmp] + [(t, "out") for t in st.to_out]:
                L.append(f"  for (int idx = PB_TID; idx < {t.n_inst} * npts; idx += PB_TEAM) {{")
                L.append(f"    const int q = idx % {t.n_inst}, pt = idx / {t.n_inst}; (void)q;")
                L.append(f"    const double* PB_RESTRICT y = yb + pt * PBM_N; double* PB_RESTRICT tmp = tmpb + pt * {max(prog.n_tmp, 1)}; "
                         f"double* PB_RESTRICT out = outb + pt * {prog.n_out}; const double t = tb[pt]; (void)y; (void)tmp; (void)out; (void)t;")
                _emit_template(em, t, dest, L, "    ")
                L.append("  }")
        L.append("}")
        return "\n".join(em.tables), "\n".join(L), em.table_bytes
    L.append(f"PB_FN void {fname}(const double* PB_RESTRICT y, const double* PB_RESTRICT p, const double t, const double cj, "
             f"double* PB_RESTRICT out, double* PB_RESTRICT tmp) {{")
    L.append("  (void)y; (void)p; (void)t; (void)cj; (void)out; (void)tmp;")
    for si, st in enumerate(prog.stages):
        tiles_tp, tiles_base = [], []
        tpls = [(t, "tmp") for t in st.to_tmp] + [(t, "out") for t in st.to_out]
        # tiles sorted by descending cost so that round-robin over warps balances
        order = sorted(range(len(tpls)), key=lambda k: -tpls[k][0].cost)
        for k in order:
            t = tpls[k][0]
            for base in range(0, t.n_inst, 32):
                tiles_tp.append(k)
                tiles_base.append(base)
        nt = len(tiles_tp)
        if si > 0:
            L.append("  PB_SYNC();")
        if nt == 0:
            continue
        if len(tpls) == 1 or one_warp:
            # one template, or a one-warp team (its warp would take every tile anyway): plain loops, no dispatch tables
            for k in order:
                t, dest = tpls[k]
                L.append(f"  for (int q = PB_TID; q < {t.n_inst}; q += PB_TEAM) {{")
                _emit_template(em, t, dest, L, "    ")
                L.append("  }")
            continue
        if WARP_LOOPS:
            # the same tile -> warp assignment (tile g goes to warp g % NWARPS, tiles in descending cost order) without the
            # dispatch: every warp walks the templates in order and takes its tiles of each (A/B switch PBGPU_WARP_LOOPS)
            g0 = 0
            for k in order:
                t, dest = tpls[k]
                n_tiles = (t.n_inst + 31) // 32
                L.append(f"  for (int tile = ((PB_WARP - {g0} % PB_NWARPS + PB_NWARPS) % PB_NWARPS); tile < {n_tiles}; tile += PB_NWARPS) {{")
                L.append(f"    const int tb = tile * 32, qe = (tb + 32 < {t.n_inst}) ? tb + 32 : {t.n_inst};")
                L.append("    for (int q = tb + PB_LANE; q < qe; q += PB_LANES) {")
                _emit_template(em, t, dest, L, "      ")
                L.append("    }")
                L.append("  }")
                g0 += n_tiles
            continue
        # tile dispatch tables: warp-uniform look-ups on the critical path of every tile -> constant memory
        ttp = em.table("unsigned char" if len(tpls) < 256 else "unsigned short", tiles_tp, "tt", qual="PB_CTABLE")
        tbs = em.table(_itype(max(tiles_base) if tiles_base else 0), tiles_base, "tb", qual="PB_CTABLE")
        L.append(f"  for (int tile = PB_WARP; tile < {nt}; tile += PB_NWARPS) {{")
        L.append(f"    const int tb = (int){tbs}[tile];")
        L.append(f"    switch ((int){ttp}[tile]) {{")
        for k, (t, dest) in enumerate(tpls):
            L.append(f"    case {k}: {{")
            L.append(f"      const int qe = (tb + 32 < {t.n_inst}) ? tb + 32 : {t.n_inst};")
            L.append("      for (int q = tb + PB_LANE; q < qe; q += PB_LANES) {")
            _emit_template(em, t, dest, L, "        ")
            L.append("      }")
            L.append("    } break;")
        L.append("    }")
        L.append("  }")
    L.append("}")
    return "\n".join(em.tables), "\n".join(L), em.table_bytes


def _emit_u32(lines, name, words):
    words = np.asarray(words, dtype=np.uint32).reshape(-1)
    body = ", ".join(f"{int(w)}u" for w in words) if len(words) else "0u"
    lines.append(f"static PB_TABLE PB_ALIGN16 const unsigned int {name}[{max(len(words), 1)}] = {{{body}}};")


# Events and output variables of the small models are one long scalar expression (terminal voltage: sixteen tanh / exp / log
# calls in a row on ONE lane, 8.7k cycles for the SPM against 2.5k for its whole residual; events + output points were half of
# the SPM / SPMe integration).  Splitting spreads the transcendental terms over lanes in earlier stages.
SPLIT_EV = bool(int(__import__("os").environ.get("PBGPU_SPLIT_EV", "0")))   # measured: SPM +1 %, SPMe -2 % (templates of a stage run one after the other on a one-warp team); batched calls below do the job
SPLIT_RES = bool(int(__import__("os").environ.get("PBGPU_SPLIT_RES", "0")))   # measured on the DFN: -6 % (two more stage barriers, 580 more tmp slots)

# shared memory the team may spend on the factor stream (bytes); larger streams stay in global memory / L1
FACTOR_STREAM_SMEM_BUDGET = 24 * 1024


def model_functions(bundle, sym: slu.SymbolicLU | None, dialect: str):
    """Build the Programs of a bundle. dialect 'c' -> Jacobian in CSC order; 'cuda' -> LU slots."""
    g = bundle.graph
    cjn = g.cj()
    jroots = [g.sub(int(n), cjn) if (r == c and r < bundle.n_diff) else int(n)
              for n, r, c in zip(bundle.jac_nodes, bundle.jac_rows, bundle.jac_cols)]
    progs = {}
    # models of several thousand states (lumped thermal DFN on fine meshes): the heat-source integral is ONE residual row of
    # 3 000 operations; its terms go to parallel lanes.  The smaller models keep the code they were measured with.
    big = dialect == "cuda" and bundle.n_states > 2048
    # the team-parallel dialect splits heavy narrow templates of the residual over the warps (program.build_program)
    progs["res"] = build_program(g, bundle.rhs_alg, "res", split=(dialect == "cuda" and SPLIT_RES), split_sums=big)
    if dialect == "cuda":
        progs["jac"] = build_program(g, jroots, "jac", out_index=sym.a_slot, n_out=sym.nnz_lu)
    else:
        progs["jac"] = build_program(g, jroots, "jac")
    progs["events"] = build_program(g, bundle.events, "events", split=(dialect == "cuda" and SPLIT_EV))
    progs["ic"] = build_program(g, bundle.ic, "ic")
    vroots = [int(n) for name in bundle.variables for n in bundle.variables[name]]
    progs["vars"] = build_program(g, vroots, "vars", split=(dialect == "cuda" and SPLIT_EV))
    return progs


def sample_matrices(bundle, p_nominal=None, n_samples=2, seed=0):
    """Values of J - cj*M (bundle CSC order) at a few states / cj for pivot-order selection."""
    g = bundle.graph
    rng = np.random.default_rng(seed)
    if p_nominal is None:
        p_nominal = bundle.meta.get("nominal_inputs")
    if p_nominal is None:
        raise ValueError("nominal inputs are needed to sample the Jacobian")
    p = np.asarray(p_nominal, dtype=np.float64)
    y0 = g.evaluate(bundle.ic, 0.0, None, p)
    diag = (bundle.jac_rows == bundle.jac_cols) & (bundle.jac_rows < bundle.n_diff)
    out = []
    for cj in (1.0, 1e-3, 1e3, 0.03, 30.0):
        for k in range(n_samples):
            y = y0 * (1.0 + (0.02 * rng.standard_normal(len(y0)) if k else 0.0))
            pk = p * (1.0 + (0.05 * rng.standard_normal(len(p)) if k else 0.0))
            v = g.evaluate(bundle.jac_nodes, 0.0, y, pk)
            if not np.all(np.isfinite(v)):
                continue
            out.append(v - cj * diag)
    # the consistent-initialisation matrix: identity on differential rows, J on algebraic rows
    if bundle.n_alg:
        v = g.evaluate(bundle.jac_nodes, 0.0, y0, p)
        drow = bundle.jac_rows < bundle.n_diff
        out.append(np.where(drow, diag.astype(float), v))
    return out


# One run-time-shaped PAR function for every record instead of one specialisation per shape.  Alone on an SM the
# specialised records are faster (15.9k against 22.5k cycles per DFN solve), but two teams share a 32 KB instruction
# cache: with the small footprint the whole integration runs 11 % faster (4.31k against 3.86k simulations/s).
PAR_GENERIC = bool(int(__import__("os").environ.get("PBGPU_PAR_GENERIC", "1")))


# Single-warp PAR records in a row that go through the software-pipelined loop pb_par_run instead of one call each.
# Measured on the DFN: no faster alone (745 against 750 cycles per record: the record is bound by instruction issue,
# not by the loads the pipeline hides) and 3 % slower inside the integration (larger instruction footprint): off.
PAR_RUN_MIN = int(__import__("os").environ.get("PBGPU_PAR_RUN_MIN", "99"))


# Measured on the DFN (5 920 systems): the table-driven loop halves pb_solve_program (11.5 -> 5.8 KB) but adds ~250 cycles of
# descriptor look-ups and dynamic dispatch to each of the 16 records (triangular solve 16.6k -> 20.7k cycles): 4 884 against
# 4 939 sims/s.  Off; the straight-line call sequence stays the default.
SOLVE_TABLE = bool(int(__import__("os").environ.get("PBGPU_SOLVE_TABLE", "0")))


def emit_solve_table(sym: slu.SymbolicLU):
    """The triangular solves as a TABLE walked by one small loop (pb_solve_program in pbgpu_ida.cuh) instead of forty
    straight-line call sites (10.7 KB of code for the DFN, fetched by every Newton iteration: the integration is bound by
    instruction fetch).  Per (record, warp) two words: offset | nt << 16 and G | mlg << 8 | mode << 16 | kind << 24 with
    kind 1 PAR, 2 forward CHAIN, 3 backward CHAIN, 0 nothing to do; per record the synchronisation that follows it
    (0 none, 1 warp, 2 team, 3: the record is the dense top block).  Returns None when the stream has a record kind the
    table does not cover (the straight-line program is emitted then)."""
    nw_max = 0
    rows = []          # (sync, {warp: (d0, d1)}, dense_off)
    for r in sym.solve_program:
        if r["kind"] == 3:
            rows.append((3, {}, r["offset"] + slu.SEQ_HDR + r["nblocks"]))
            continue
        if r["kind"] not in (1, 2):
            return None
        boff = r["offset"] + slu.SEQ_HDR + (1 if r.get("narrow") else r["nblocks"])
        ent = {}
        for v in range(r["nblocks"]):
            nt = r["desc"][v] & 0xFFFF
            wv = r["first_warp"] + v
            if r["kind"] == 1:
                G, mlg = r["param"] & 0xFF, r["param"] >> 8
                if boff >= 65536 or nt >= 65536 or mlg >= 256 or G >= 256:
                    return None
                ent[wv] = (boff | nt << 16, G | mlg << 8 | r["mode"] << 16 | 1 << 24)
                boff += (1 + mlg) * nt
            else:
                ent[wv] = (boff | nt << 16, (3 if (r["mode"] & 1) else 2) << 24)
                boff += 3 * nt
            nw_max = max(nw_max, wv + 1)
        rows.append((r["sync"], ent, 0))
    nw = max(nw_max, 1)
    flat, sync = [], []
    dense_off = 0
    for sy, ent, doff in rows:
        sync.append(sy)
        if sy == 3:
            dense_off = doff
        for wv in range(nw):
            d = ent.get(wv, (0, 0))
            flat += [d[0], d[1]]
    text = f"#define PBM_SOLVE_NREC {len(rows)}\n#define PBM_SOLVE_NW {nw}\n#define PBM_SOLVE_DENSE_OFF {dense_off}\n"
    text += f"static PB_CTABLE const unsigned int pbm_solve_desc[{len(flat)}] = {{" + ", ".join(f"{x}u" for x in flat) + "};\n"
    text += f"static PB_CTABLE const unsigned char pbm_solve_sync[{len(sync)}] = {{" + ", ".join(str(x) for x in sync) + "};\n"
    return text


def emit_solve_program(sym: slu.SymbolicLU) -> str:
    """The triangular solves as a call sequence: one call per record of the solve stream, expanded inside
    pb_solve_program (pbgpu_ida.cuh), which defines st, lu, w.  Runs of single-warp PAR records (the top of the
    elimination tree) become ONE call of the software-pipelined loop pb_par_run, with their shapes in the constant
    table pbm_par_run."""
    out, run_desc = [], []
    recs = sym.solve_program

    def call_of(r, v, boff):
        nt = r["desc"][v] & 0xFFFF
        if r["kind"] == 1:
            G, mlg = r["param"] & 0xFF, r["param"] >> 8
            call = (f"pb_par_generic(st + {boff}, {nt}, {G}, {mlg}, {r['mode']}u, lu, w);" if PAR_GENERIC else
                    f"pb_par_rec<{G}, {mlg}, {r['mode']}>(st + {boff}, {nt}, lu, w);")
            return call, boff + (1 + mlg) * nt
        if r["kind"] == 2:
            return f"pb_chain_rec<{r['mode'] & 1}>(st + {boff}, {nt}, lu, w);", boff + 3 * nt
        return f"pb_seq_rec<{r['mode'] & 1}>(st + {boff}, {nt}, lu, w);", boff + nt * (r["desc"][v] >> 16)

    def span(k):        # word range of record k in the stream
        q = recs[k]
        return q["offset"], slu.SEQ_HDR + len(q["desc"]) + len(q["body"])

    def prefetch(k):    # models whose solve stream stays in global memory: the next record's words are on their way to L1
        if k < len(recs):
            out.append("PB_SLV_PREFETCH({}, {})".format(*span(k)))

    prefetch(0)
    i = 0
    while i < len(recs):
        r = recs[i]
        if r.get("narrow"):
            j = i
            while recs[j]["sync"] == 1:     # the run ends with the record that carries the team barrier
                j += 1
            run = recs[i:j + 1]
            assert all(q.get("narrow") and q["nblocks"] == 1 and q["first_warp"] == 0 for q in run)
            for k in range(i + 1, j + 2):
                prefetch(k)
            out.append("if (PB_WARP == 0) {")
            k = 0
            while k < len(run):
                m = k
                while m < len(run) and run[m]["kind"] == 1:
                    m += 1
                if PAR_GENERIC and m - k >= PAR_RUN_MIN:
                    out.append(f"  pb_par_run(st, {len(run_desc)}, {m - k}, lu, w);")
                    for q in run[k:m]:
                        nt, G, mlg = q["desc"][0] & 0xFFFF, q["param"] & 0xFF, q["param"] >> 8
                        boff = q["offset"] + slu.SEQ_HDR + 1
                        assert boff < 65536 and nt <= 32 and mlg < 256
                        run_desc.append((boff | nt << 16, G | mlg << 8 | q["mode"] << 16))
                    out.append("  PB_REC_MARK")
                    k = m
                else:
                    q = run[k]
                    out.append("  " + call_of(q, 0, q["offset"] + slu.SEQ_HDR + 1)[0])
                    out.append("  PB_SYNCWARP();")
                    out.append("  PB_REC_MARK")
                    k += 1
            out.append("}")
            out.append("PB_SYNC();")
            out.append("PB_REC_MARK")
            i = j + 1
            continue
        prefetch(i + 1)
        boff = r["offset"] + slu.SEQ_HDR + r["nblocks"]
        if r["kind"] == 3:              # dense top block: all threads, barriers inside
            out += [f"PB_DENSE_REC({boff})", "PB_REC_MARK"]
            i += 1
            continue
        calls = []
        for v in range(r["nblocks"]):
            call, boff = call_of(r, v, boff)
            calls.append((r["first_warp"] + v, call))
        out.append(" else ".join(f"if (PB_WARP == {wv}) {{ {c} }}" for wv, c in calls))
        if r["sync"] == 2:
            out.append("PB_SYNC();")
        elif r["sync"] == 1:
            out.append("PB_SYNCWARP();")
        out.append("PB_REC_MARK")
        i += 1
    text = ""
    if run_desc:
        flat = [x for d in run_desc for x in d]
        text += f"#define PBM_PAR_RUN_N {len(run_desc)}\n"
        text += (f"static PB_CTABLE const unsigned int pbm_par_run[{len(flat)}] = {{" +
                 ", ".join(f"{x}u" for x in flat) + "};\n")
    return text + "#define PBM_SOLVE_PROGRAM \\\n  " + " \\\n  ".join(out) + "\n"


def default_team(n_states: int) -> int:
    """Threads per system: one warp for small models, a CTA of several warps for large ones.  PBGPU_TEAM overrides (A/B runs)."""
    env = __import__("os").environ.get("PBGPU_TEAM")
    if env:
        return int(env)
    if n_states <= 128:
        return 32
    if n_states <= 512:
        return 64
    if n_states <= 2048:
        return 256  # DFN-sized: 8 warps per team, two teams per SM (128 registers per thread)
    return 512      # several thousand states: one team per SM


def generate(bundle, dialect="cuda", sym: slu.SymbolicLU | None = None, extra_samples=None):
    """Return (source_text, info dict). The text is a header included by the integrator
    translation unit (CUDA) or the oracle model shim (C).
    extra_samples: value arrays of J - cj M (bundle entry order) the static pivot order must ALSO be valid for; they come
    first, so the row matching is taken from them (solver.BatchedIDASolver re-pivots with the matrices of a failed system)."""
    if dialect == "cuda" and sym is None:
        samples = [np.asarray(v, float) for v in (extra_samples or [])] + sample_matrices(bundle)
        sym = slu.analyse(bundle.n_states, bundle.jac_rows, bundle.jac_cols, samples,
                          team=default_team(bundle.n_states))
    progs = model_functions(bundle, sym, dialect)
    n_tmp = max(p.n_tmp for p in progs.values())
    n_var = progs["vars"].n_out
    out = []
    out.append(f"/* generated by pybamm_b200.codegen ({dialect}) -- do not edit */")
    # the long libm sequences go through PB_* so that the device build can route them to shared out-of-line copies
    # (instruction-cache footprint of the generated functions); everywhere else they are the plain libm calls
    out.append("#ifndef PB_CTABLE")
    out.append("#define PB_CTABLE")
    out.append("#endif")
    out.append("#ifndef PB_EXP")
    for fn in ("EXP", "LOG", "TANH", "SINH", "COSH", "ASINH"):
        out.append(f"#define PB_{fn}(x) {fn.lower()}(x)")
    out.append("#define PB_POW(x, y) pow(x, y)")
    out.append("#endif")
    out.append("#ifndef PB_POWM025")
    out.append("#define PB_POWM025(x) pow(x, -0.25)")
    out.append("#define PB_POWM125(x) pow(x, -1.25)")
    out.append("#endif")
    out.append("#ifndef PB_PAD")
    out.append("#define PB_PAD")
    out.append("#endif")
    out.append("#ifndef PB_SQRT")
    out.append("#define PB_SQRT(x) sqrt(x)")
    out.append("#endif")
    out.append("#ifndef PB_DIV")
    out.append("#define PB_DIV(a, b) ((a) / (b))")
    out.append("#endif")
    # 2- / 4-wide issue of independent libm calls: the device build defines PB_B2 / PB_B4 (pbgpu_team.cuh), any other
    # consumer of this header evaluates them one by one
    out.append("#ifndef PB_B2")
    for nm, mac in (("exp", "PB_EXP"), ("log", "PB_LOG"), ("tanh", "PB_TANH"), ("sinh", "PB_SINH"), ("cosh", "PB_COSH"),
                    ("asinh", "PB_ASINH"), ("sqrt", "PB_SQRT"), ("powm025", "PB_POWM025"), ("powm125", "PB_POWM125")):
        out.append(f"#define PB_B1_{nm}(x) {mac}(x)")
    out.append("#define PB_B2(f, r0, r1, x0, x1) const double r0 = PB_B1_##f(x0), r1 = PB_B1_##f(x1);")
    out.append("#define PB_B4(f, r0, r1, r2, r3, x0, x1, x2, x3) const double r0 = PB_B1_##f(x0), r1 = PB_B1_##f(x1), r2 = PB_B1_##f(x2), r3 = PB_B1_##f(x3);")
    out.append("#endif")
    out.append(f'#define PBM_NAME "{bundle.name}"')
    out.append(f"#define PBM_N {bundle.n_states}")
    out.append(f"#define PBM_NDIFF {bundle.n_diff}")
    out.append(f"#define PBM_NP {max(bundle.n_inputs, 1)}")
    out.append(f"#define PBM_NP_REAL {bundle.n_inputs}")
    out.append(f"#define PBM_NEV {bundle.n_events}")
    out.append(f"#define PBM_NNZ {len(bundle.jac_nodes)}")
    out.append(f"#define PBM_NTMP {max(n_tmp, 1)}")
    out.append(f"#define PBM_NTMP_JAC {max(progs['jac'].n_tmp, 1)}")
    out.append(f"#define PBM_NTMP_SMALL {max(max(p.n_tmp for k, p in progs.items() if k != 'jac'), 1)}")
    out.append(f"#define PBM_NVAR {n_var}")
    info = dict(n_tmp=n_tmp, n_var=n_var, table_bytes=0, programs={})
    for key in ("res", "jac", "events", "ic", "vars"):
        tabs, fn, tb = emit_program(progs[key], f"pbm_{key}", f"pbm_{key}", batch=(dialect == "cuda"),
                                    one_warp=(dialect == "cuda" and ONE_WARP_LOOPS and default_team(bundle.n_states) == 32))
        out.append(tabs)
        out.append(fn)
        info["table_bytes"] += tb
        pr = progs[key]
        info["programs"][key] = dict(stages=len(pr.stages), flops=pr.flops, transcendental=pr.n_transcendental,
                                     n_tmp=pr.n_tmp,
                                     templates=sum(len(s.to_tmp) + len(s.to_out) for s in pr.stages))
        out.append(f"#define PBM_FLOPS_{key.upper()} {pr.flops}")
    if dialect == "cuda" and n_var > 0 and default_team(bundle.n_states) == 32:
        tabs, fn, tb = emit_program(progs["vars"], "pbm_vars_multi", "pbm_varsm", multi=True)
        out.append(f"#define PBM_VARS_MULTI 1\n#define PBM_NTMP_VARS {max(progs['vars'].n_tmp, 1)}")
        out.append(tabs)
        out.append(fn)
        info["table_bytes"] += tb
    if dialect == "cuda":
        out.append(f"#define PBM_NNZ_LU {sym.nnz_lu}")
        lines = []
        fill = sorted(set(range(sym.nnz_lu)) - set(int(s) for s in sym.a_slot))
        it = _itype(sym.nnz_lu)
        lines.append(f"#define PBM_NFILL {len(fill)}")
        lines.append(f"static PB_TABLE const {it} pbm_fill[{max(len(fill), 1)}] = {{{', '.join(map(str, fill)) if fill else '0'}}};")
        # per structural entry: 0 algebraic row, 1 differential row off-diagonal, 2 differential diagonal
        kind = np.where(bundle.jac_rows < bundle.n_diff, np.where(bundle.jac_rows == bundle.jac_cols, 2, 1), 0)
        lines.append(f"static PB_TABLE const unsigned char pbm_ent_kind[{len(kind)}] = {{{', '.join(map(str, kind.tolist()))}}};")
        lines.append(f"static PB_TABLE const {it} pbm_ent_slot[{len(kind)}] = {{{', '.join(str(int(s)) for s in sym.a_slot)}}};")
        it_n = _itype(sym.n)
        lines.append(f"static PB_TABLE const {it_n} pbm_perm_r[{sym.n}] = {{{', '.join(str(int(x)) for x in sym.perm_r)}}};")
        lines.append(f"static PB_TABLE const {it_n} pbm_perm_c[{sym.n}] = {{{', '.join(str(int(x)) for x in sym.perm_c)}}};")
        # inverse permutations: where original row / unknown i sits in the permuted vector (the Newton loop writes its
        # right-hand side straight into the solve's w vector and reads the correction from it)
        ipr, ipc = np.empty(sym.n, np.int64), np.empty(sym.n, np.int64)
        ipr[np.asarray(sym.perm_r)] = np.arange(sym.n)
        ipc[np.asarray(sym.perm_c)] = np.arange(sym.n)
        lines.append(f"static PB_TABLE const {it_n} pbm_iperm_r[{sym.n}] = {{{', '.join(str(int(x)) for x in ipr)}}};")
        lines.append(f"static PB_TABLE const {it_n} pbm_iperm_c[{sym.n}] = {{{', '.join(str(int(x)) for x in ipc)}}};")
        lines.append(f"#define PBM_FAC_NLEV {sym.factor.n_levels}")
        lines.append(f"#define PBM_FWD_NLEV {sym.fwd.n_levels}")
        lines.append(f"#define PBM_BWD_NLEV {sym.bwd.n_levels}")
        if sym.dense_rows is not None:
            nT = len(sym.dense_rows)
            G = next(r["param"] & 0xFF for r in sym.solve_program if r["kind"] == 3)
            lines.append(f"#define PBM_DENSE_N {nT}")
            lines.append(f"#define PBM_DENSE_G {G}")
            lines.append(f"static PB_TABLE const {_itype(sym.nnz_lu + 1)} pbm_dense_slot[{nT * nT}] = {{" +
                         ", ".join(str(int(x)) for x in sym.dense_slots.reshape(-1)) + "};")
        lines.append(f"#define PBM_SLV_WORDS {sym.n_solve_words}")
        lines.append(f"#define PBM_SLV_ALLOC {len(sym.solve_stream)}")
        lines.append(f"#define PBM_SLV_NREC {sym.n_solve_records}")
        lines.append(f"#define PBM_LU_TEAM_MIN {sym.seq_info['team_min']}")
        lines.append(f"#define PBM_FAC_WORDS {len(sym.factor_stream)}")
        lines.append(f"#define PBM_FAC_NREC {sym.n_factor_records}")
        lines.append(f"#define PBM_FAC_SMEM {1 if 4 * len(sym.factor_stream) <= FACTOR_STREAM_SMEM_BUDGET else 0}")
        _emit_u32(lines, "pbm_slv_tab", sym.solve_stream)
        _emit_u32(lines, "pbm_fac_tab", sym.factor_stream)
        if 4 * len(sym.factor_stream) > FACTOR_STREAM_SMEM_BUDGET:
            # staging plan of pb_lu_factor: consecutive whole records packed into chunks that fit the solve's w scratch
            chunks = slu.factor_chunks(sym.factor_stream, 2 * sym.n)
            if chunks is not None:
                flat = [x for c in chunks for x in c] + [len(sym.factor_stream), 0]
                lines.append(f"#define PBM_FAC_NCHUNK {len(chunks)}")
                lines.append(f"#define PBM_FAC_CHUNK_WORDS {max(b - a for (a, _), (b, _) in zip(chunks, chunks[1:] + [(len(sym.factor_stream), 0)]))}")
                lines.append(f"static PB_CTABLE const unsigned int pbm_fac_chunk[{len(flat)}] = {{" + ", ".join(f"{x}u" for x in flat) + "};")
        lines.append(emit_solve_program(sym))
        table = emit_solve_table(sym) if SOLVE_TABLE else None
        if table is not None:
            lines.append(table)
        out.append("\n".join(lines))
        out.append(f"#define PBM_FLOPS_FACTOR {sym.flops_factor}")
        out.append(f"#define PBM_FLOPS_SOLVE {sym.flops_solve}")
        info["lu"] = dict(nnz_lu=sym.nnz_lu, ordering=sym.ordering, factor_levels=sym.factor.n_levels,
                          fwd_levels=sym.fwd.n_levels, bwd_levels=sym.bwd.n_levels,
                          solve_stream_words=len(sym.solve_stream), factor_stream_words=len(sym.factor_stream),
                          solve_records=sym.n_solve_records, factor_records=sym.n_factor_records,
                          est_cycles_solve=sym.est_cycles_solve, est_cycles_factor=sym.est_cycles_factor,
                          flops_factor=sym.flops_factor, flops_solve=sym.flops_solve,
                          seq_clusters=sym.seq_info["clusters"], seq_fwd_heights=sym.seq_info["fwd_heights"],
                          seq_bwd_heights=sym.seq_info["bwd_heights"], seq_team_min=sym.seq_info["team_min"],
                          perm_r=[int(x) for x in sym.perm_r], perm_c=[int(x) for x in sym.perm_c])
    else:
        rows = ", ".join(str(int(r)) for r in bundle.jac_rows)
        cols = ", ".join(str(int(c)) for c in bundle.jac_cols)
        out.append(f"static const int pbm_jac_rows[{len(bundle.jac_rows)}] = {{{rows}}};")
        out.append(f"static const int pbm_jac_cols[{len(bundle.jac_cols)}] = {{{cols}}};")
    text = "\n".join(out) + "\n"
    info["sha"] = hashlib.sha256(text.encode()).hexdigest()[:16]
    return text, info
