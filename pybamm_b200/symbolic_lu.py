"""Static-pattern sparse LU: ordering, symbolic factorisation and level schedules.

Role in the reference: SuiteSparse KLU behind ``SUNLinSol_KLU``
(idaklu_source/IDAKLUSolverOpenMP_solvers.hpp:29-33): a fill-reducing ordering and
a pivot sequence are fixed by the first factorisation and *reused* by every later
``klu_refactor``.  Here the same idea is taken to its end for a batch of systems
that share one sparsity pattern: the pivot order is chosen once, at code
generation time, from sample matrices ``J - cj*M``; the fill pattern, the
dependency levels and every index the numeric phase needs are emitted as tables,
so that the device kernels contain no searching, no pivoting and no indirection
beyond one table look-up per multiply-add.

Numeric phase the tables describe (all in one value array ``lu`` of length nnz_lu,
slot order chosen here):

factor:   entry-parallel left-looking LU.  The task for entry (i, j) of
          B = P A Q computes  v = b_ij - sum_{k<min(i,j)} l_ik u_kj ;
          U entries store v (the diagonal stores 1/v), L entries store v/u_jj.
          Tasks are grouped in dependency levels; inside a level they are
          sorted by length and their operand lists are stored jagged-diagonal
          (JDS) so that lanes read consecutive table entries.
solve:    row tasks of the unit-lower and upper triangular solves, again in
          levels with JDS operand lists; the row/column permutations are folded
          into the first gather and the last scatter.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla


@dataclass
class JDS:
    """Level-scheduled task tables, operands in ELL layout per level.

    Level l has nt tasks and maxlen operand rows; operand p of task q sits at op_base + p*nt + q, so
    lanes read consecutive entries and no address depends on another table look-up.  Short tasks are
    padded with a (zero slot, zero slot) operand pair."""
    n_levels: int
    lev_meta: np.ndarray     # (n_levels, 4) int32: task_base, nt, op_base, maxlen | (sync << 16)
    op_a: np.ndarray         # operand A (lu slot)
    op_b: np.ndarray         # operand B (lu slot for factor, w index for solves)
    task_out: np.ndarray     # per task: output slot / row
    task_aux: np.ndarray     # per task: aux (factor: diag slot or code; fwd: gather index; bwd: scatter index)
    n_real_ops: int

    @property
    def n_tasks(self):
        return len(self.task_out)

    @property
    def n_ops(self):
        return self.n_real_ops

    @property
    def sync(self):
        return (self.lev_meta[:, 3] >> 16) & 1


@dataclass
class SymbolicLU:
    n: int
    nnz_lu: int
    perm_r: np.ndarray       # B[i, :] = A[perm_r[i], :]
    perm_c: np.ndarray       # B[:, j] = A[:, perm_c[j]]
    slot_of: dict            # (i, j) in B coordinates -> slot
    a_slot: np.ndarray       # for each input entry k of A (rows[k], cols[k]) its slot
    diag_slot: np.ndarray    # (n,) slot of u_ii (holds 1/u_ii after factor)
    factor: JDS
    fwd: JDS
    bwd: JDS
    flops_factor: int
    flops_solve: int
    ordering: str
    solve_stream: np.ndarray = None    # uint32 words, see _solve_stream
    factor_stream: np.ndarray = None   # uint32 words, see _factor_stream
    n_solve_records: int = 0
    n_factor_records: int = 0
    est_cycles_solve: int = 0
    est_cycles_factor: int = 0

    def summary(self):
        return (f"n={self.n} nnz_lu={self.nnz_lu} ordering={self.ordering} streams: solve {len(self.solve_stream)} words/"
                f"{self.n_solve_records} rec (~{self.est_cycles_solve} cyc), factor {len(self.factor_stream)} words/"
                f"{self.n_factor_records} rec (~{self.est_cycles_factor} cyc); "
                f"factor: {self.factor.n_levels} levels/{self.factor.n_tasks} tasks/{self.factor.n_ops} ops ({len(self.factor.op_a)} padded) "
                f"(team syncs {int(self.factor.sync.sum())}); fwd {self.fwd.n_levels} lv, bwd {self.bwd.n_levels} lv "
                f"(team syncs {int(self.fwd.sync.sum())}+{int(self.bwd.sync.sum())})")


# ---------------------------------------------------------------------------
# orderings
# ---------------------------------------------------------------------------
def _nested_dissection(adj: sp.csr_matrix, leaf=8):
    """Simple recursive BFS-level-set nested dissection of a symmetric graph.

    Returns an elimination order (separators last).  Good for the chain- and
    thick-chain-like graphs of 1-D(+1-D) discretisations: elimination-tree height
    O(log n * separator size) instead of O(n)."""
    n = adj.shape[0]
    indptr, indices = adj.indptr, adj.indices
    order = []

    def bfs_levels(nodes_mask, start):
        level = {start: 0}
        frontier = [start]
        levels = [[start]]
        while frontier:
            nxt = []
            for u in frontier:
                for v in indices[indptr[u]:indptr[u + 1]]:
                    if nodes_mask[v] and v not in level:
                        level[v] = level[u] + 1
                        nxt.append(v)
            if nxt:
                levels.append(nxt)
            frontier = nxt
        return levels

    def components(nodes):
        mask = np.zeros(n, dtype=bool)
        mask[nodes] = True
        seen = set()
        comps = []
        for s in nodes:
            if s in seen:
                continue
            comp = [s]
            seen.add(s)
            stack = [s]
            while stack:
                u = stack.pop()
                for v in indices[indptr[u]:indptr[u + 1]]:
                    if mask[v] and v not in seen:
                        seen.add(v)
                        comp.append(v)
                        stack.append(v)
            comps.append(comp)
        return comps

    def dissect(nodes):
        """returns list of nodes in elimination order"""
        if len(nodes) <= leaf:
            return list(nodes)
        mask = np.zeros(n, dtype=bool)
        mask[nodes] = True
        # pseudo-peripheral start
        start = nodes[0]
        for _ in range(3):
            lv = bfs_levels(mask, start)
            start = min(lv[-1], key=lambda u: indptr[u + 1] - indptr[u])
        lv = bfs_levels(mask, start)
        if len(lv) < 3:
            return list(nodes)
        # separator = level closest to balancing the two halves, prefer small levels
        sizes = np.array([len(x) for x in lv])
        cum = np.cumsum(sizes)
        total = cum[-1]
        best, best_score = None, None
        for k in range(1, len(lv) - 1):
            left = cum[k - 1]
            right = total - cum[k]
            score = max(left, right) + 4 * sizes[k]
            if best_score is None or score < best_score:
                best, best_score = k, score
        sep = lv[best]
        rest = [u for k, l in enumerate(lv) if k != best for u in l]
        out = []
        for comp in components(rest):
            out.extend(dissect(comp))
        out.extend(sep)
        return out

    all_nodes = list(range(n))
    for comp in components(all_nodes):
        order.extend(dissect(comp))
    return np.asarray(order, dtype=np.int64)


def _peel_then_nd(sym: sp.csr_matrix, leaf=4):
    """Elimination order for "chains hanging off a core" graphs (particle radial chains off the through-cell
    variables in a DFN): repeatedly eliminate vertices of degree <= 1 (no fill, depth = chain length), then
    nested dissection of what is left."""
    n = sym.shape[0]
    adj = [set(sym.indices[sym.indptr[i]:sym.indptr[i + 1]].tolist()) for i in range(n)]
    alive = np.ones(n, bool)
    order = []
    stack = [i for i in range(n) if len(adj[i]) <= 1]
    while stack:
        u = stack.pop()
        if not alive[u] or len(adj[u]) > 1:
            continue
        alive[u] = False
        order.append(u)
        for v in adj[u]:
            adj[v].discard(u)
            if alive[v] and len(adj[v]) <= 1:
                stack.append(v)
        adj[u] = set()
    core = np.nonzero(alive)[0]
    if len(core):
        sub = sym[core][:, core].tocsr()
        q = _nested_dissection(sub, leaf=leaf)
        order.extend(core[q].tolist())
    return np.asarray(order, dtype=np.int64)


def _symbolic_fill(B: sp.csr_matrix):
    """Row patterns of L (strict lower) and U (upper incl. diagonal) of B without pivoting."""
    n = B.shape[0]
    B = B.tocsr()
    Lrows, Urows = [], []
    for i in range(n):
        cols = set(B.indices[B.indptr[i]:B.indptr[i + 1]].tolist())
        cols.add(i)
        # up-looking: for k in ascending order of current lower pattern, merge U row k
        lower = sorted(c for c in cols if c < i)
        done = set()
        import heapq

        heap = list(lower)
        heapq.heapify(heap)
        while heap:
            k = heapq.heappop(heap)
            if k in done:
                continue
            done.add(k)
            for c in Urows[k]:
                if c > k and c not in cols:
                    cols.add(c)
                    if c < i:
                        heapq.heappush(heap, c)
        Lrows.append(sorted(c for c in cols if c < i))
        Urows.append(sorted(c for c in cols if c >= i))
    return Lrows, Urows


def _jds(levels_tasks, pad_a, pad_b, warp=32):
    """levels_tasks: list (per level) of tasks (out, aux, [(a, b), ...])."""
    meta = []
    op_a, op_b = [], []
    task_out, task_aux = [], []
    n_real = 0
    for li, tasks in enumerate(levels_tasks):
        tasks = sorted(tasks, key=lambda t: -len(t[2]))
        nt = len(tasks)
        maxlen = len(tasks[0][2]) if tasks else 0
        nxt = len(levels_tasks[li + 1]) if li + 1 < len(levels_tasks) else 10 ** 9
        sync = 0 if (nt <= warp and nxt <= warp) else 1
        meta.append((len(task_out), nt, len(op_a), maxlen | (sync << 16)))
        for t in tasks:
            task_out.append(t[0])
            task_aux.append(t[1])
            n_real += len(t[2])
        for p in range(maxlen):
            for t in tasks:
                if len(t[2]) > p:
                    op_a.append(t[2][p][0])
                    op_b.append(t[2][p][1])
                else:
                    op_a.append(pad_a)
                    op_b.append(pad_b)
    i32 = np.int32
    return JDS(len(levels_tasks), np.asarray(meta, i32).reshape(-1, 4), np.asarray(op_a, i32), np.asarray(op_b, i32),
               np.asarray(task_out, i32), np.asarray(task_aux, i32), n_real)


AUX_U = 0xFFFF      # factor task writes an off-diagonal U entry
AUX_DIAG = 0xFFFE   # factor task writes a diagonal entry (stores the reciprocal)


def analyse(n, rows, cols, samples, ordering="auto", diag_pivot_thresh=0.01) -> SymbolicLU:
    """rows/cols: structural pattern of A (every entry the numeric phase may set).
    samples: list of value arrays (same order as rows/cols) used to pick the pivot order
    and to validate it."""
    rows = np.asarray(rows, np.int64)
    cols = np.asarray(cols, np.int64)
    candidates = []
    names = ["peel_nd", "nd", "COLAMD", "MMD_AT_PLUS_A"] if ordering == "auto" else [ordering]
    A0 = sp.csc_matrix((np.asarray(samples[0], float), (rows, cols)), shape=(n, n))
    pat = sp.csr_matrix((np.ones(len(rows)), (rows, cols)), shape=(n, n))
    for name in names:
        try:
            if name in ("nd", "peel_nd"):
                # row permutation first: a maximum transversal from a pivoting LU so that the
                # diagonal of A[perm_r] is structurally and numerically sound, then a symmetric ND
                lu0 = spla.splu(A0, permc_spec="NATURAL", diag_pivot_thresh=diag_pivot_thresh)
                # SuperLU: Pr A Pc = L U with perm_r[i] = new position of row i
                pr = np.empty(n, np.int64)
                pr[lu0.perm_r] = np.arange(n)
                Bp = pat[pr, :]
                sym = (Bp + Bp.T).tocsr()
                sym.setdiag(0)
                sym.eliminate_zeros()
                q = _nested_dissection(sym) if name == "nd" else _peel_then_nd(sym)
                perm_r = pr[q]
                perm_c = q
            else:
                lu0 = spla.splu(A0, permc_spec=name, diag_pivot_thresh=diag_pivot_thresh)
                perm_c = np.asarray(lu0.perm_c, np.int64)   # column j of A -> position perm_c[j]
                pc = np.empty(n, np.int64)
                pc[perm_c] = np.arange(n)
                pr = np.empty(n, np.int64)
                pr[lu0.perm_r] = np.arange(n)
                perm_r, perm_c = pr, pc
            candidates.append((name, perm_r, perm_c))
        except Exception as e:  # pragma: no cover
            print(f"[symbolic_lu] ordering {name} failed: {e}")
    best = None
    for name, perm_r, perm_c in candidates:
        s = _build(n, rows, cols, perm_r, perm_c, name)
        # cost proxy (cycles per factorisation interval): ~30 triangular solves per factorisation (measured on the
        # battery models); the L+U values and the solve stream must stay in shared memory (8 + ~6 bytes per entry)
        score = 30.0 * s.est_cycles_solve + s.est_cycles_factor + 30.0 * s.nnz_lu
        ok = all(_validate(s, rows, cols, np.asarray(v, float)) < 1e-6 for v in samples)
        if not ok:
            continue
        if best is None or score < best[0]:
            best = (score, s)
    if best is None:
        raise RuntimeError("no static pivot order passed validation on the sample matrices")
    return best[1]


def _build(n, rows, cols, perm_r, perm_c, name) -> SymbolicLU:
    inv_r = np.empty(n, np.int64)
    inv_r[perm_r] = np.arange(n)
    inv_c = np.empty(n, np.int64)
    inv_c[perm_c] = np.arange(n)
    bi, bj = inv_r[rows], inv_c[cols]
    B = sp.csr_matrix((np.ones(len(rows)), (bi, bj)), shape=(n, n))
    Lrows, Urows = _symbolic_fill(B)
    # slots: row-major over (L row, U row)
    slot_of = {}
    for i in range(n):
        for j in Lrows[i]:
            slot_of[(i, j)] = len(slot_of)
        for j in Urows[i]:
            slot_of[(i, j)] = len(slot_of)
    nnz_lu = len(slot_of)
    diag_slot = np.array([slot_of[(i, i)] for i in range(n)], np.int64)
    a_slot = np.array([slot_of[(int(i), int(j))] for i, j in zip(bi, bj)], np.int64)

    # column structure of U for pairing: Ucols[j] = {k: (k, j) in U}
    Lset = [set(r) for r in Lrows]
    Ucol = [[] for _ in range(n)]
    for k in range(n):
        for j in Urows[k]:
            Ucol[j].append(k)
    # factor tasks with levels
    level = {}
    tasks_by_level = {}
    flops = 0
    for i in range(n):
        row_entries = [(i, j) for j in Lrows[i]] + [(i, j) for j in Urows[i]]
        for (_, j) in row_entries:
            m = min(i, j)
            pairs = []
            lv = 0
            # k in L-pattern of row i with k < m and (k, j) in U
            for k in Lrows[i]:
                if k >= m:
                    break
                if (k, j) in slot_of:
                    pairs.append((slot_of[(i, k)], slot_of[(k, j)]))
                    lv = max(lv, level[(i, k)] + 1, level[(k, j)] + 1)
            if i > j:
                lv = max(lv, level[(j, j)] + 1)
                aux = slot_of[(j, j)]
            elif i == j:
                aux = AUX_DIAG
            else:
                aux = AUX_U
            level[(i, j)] = lv
            tasks_by_level.setdefault(lv, []).append((slot_of[(i, j)], aux, pairs))
            flops += 2 * len(pairs) + 1
    nlev = max(tasks_by_level) + 1
    factor = _jds([tasks_by_level.get(l, []) for l in range(nlev)], nnz_lu, nnz_lu)

    # forward solve: w[i] = b[perm_r[i]] - sum_k L[i,k] w[k]
    rl = np.zeros(n, np.int64)
    fl = {}
    sflops = 0
    for i in range(n):
        lv = 0
        ops = []
        for k in Lrows[i]:
            ops.append((slot_of[(i, k)], k))
            lv = max(lv, rl[k] + 1)
        rl[i] = lv
        fl.setdefault(lv, []).append((i, int(perm_r[i]), ops))
        sflops += 2 * len(ops)
    fwd = _jds([fl.get(l, []) for l in range(max(fl) + 1)], nnz_lu, n)
    # backward solve: w[i] = (w[i] - sum_{j>i} U[i,j] w[j]) * lu[diag_i]; x[perm_c[i]] = w[i]
    rb = np.zeros(n, np.int64)
    blv = {}
    for i in range(n - 1, -1, -1):
        lv = 0
        ops = []
        for j in Urows[i]:
            if j == i:
                continue
            ops.append((slot_of[(i, j)], j))
            lv = max(lv, rb[j] + 1)
        rb[i] = lv
        blv.setdefault(lv, []).append((i, int(perm_c[i]), ops))
        sflops += 2 * len(ops) + 1
    bwd = _jds([blv.get(l, []) for l in range(max(blv) + 1)], nnz_lu, n)
    return build_streams(SymbolicLU(n, nnz_lu, np.asarray(perm_r), np.asarray(perm_c), slot_of, a_slot, diag_slot,
                                    factor, fwd, bwd, flops, sflops, name))


# ---------------------------------------------------------------------------
# numpy emulation of the numeric phase (validation of tables; tests)
# ---------------------------------------------------------------------------
def numeric_factor(s: SymbolicLU, a_vals):
    """lu has nnz_lu + 1 entries: the last one is the constant-zero padding slot."""
    lu = np.zeros(s.nnz_lu + 1)
    np.add.at(lu, s.a_slot, a_vals)
    f = s.factor
    for lv in range(f.n_levels):
        t0, nt, ob, ml = (int(x) for x in f.lev_meta[lv])
        ml &= 0xFFFF
        out = f.task_out[t0:t0 + nt]
        aux = f.task_aux[t0:t0 + nt]
        acc = lu[out].copy()
        for p in range(ml):
            sl = slice(ob + p * nt, ob + (p + 1) * nt)
            acc -= lu[f.op_a[sl]] * lu[f.op_b[sl]]
        for q in range(nt):
            if aux[q] == AUX_U:
                lu[out[q]] = acc[q]
            elif aux[q] == AUX_DIAG:
                lu[out[q]] = 1.0 / acc[q]
            else:
                lu[out[q]] = acc[q] * lu[aux[q]]
    return lu


def numeric_solve(s: SymbolicLU, lu, b):
    w = np.zeros(s.n + 1)
    for jds, back in ((s.fwd, False), (s.bwd, True)):
        for lv in range(jds.n_levels):
            t0, nt, ob, ml = (int(x) for x in jds.lev_meta[lv])
            ml &= 0xFFFF
            rows_ = jds.task_out[t0:t0 + nt]
            acc = w[rows_].copy() if back else b[jds.task_aux[t0:t0 + nt]].copy()
            for p in range(ml):
                sl = slice(ob + p * nt, ob + (p + 1) * nt)
                acc -= lu[jds.op_a[sl]] * w[jds.op_b[sl]]
            if back:
                acc *= lu[s.diag_slot[rows_]]
            w[rows_] = acc
    x = np.empty(s.n)
    x[s.bwd.task_aux] = w[s.bwd.task_out]
    return x


def _validate(s: SymbolicLU, rows, cols, vals):
    A = sp.csc_matrix((vals, (rows, cols)), shape=(s.n, s.n))
    rng = np.random.default_rng(1)
    xref = rng.standard_normal(s.n)
    b = A @ xref
    with np.errstate(all="ignore"):
        lu = stream_factor(s, vals)
        x = stream_solve(s, lu, b)
    if not np.all(np.isfinite(x)):
        return np.inf
    return float(np.max(np.abs(x - xref)) / np.max(np.abs(xref)))


# ---------------------------------------------------------------------------
# packed 32-bit "streams" consumed by the device kernels (pbgpu_ida.cuh)
# ---------------------------------------------------------------------------
# A stream is a sequence of records  [meta][task word x nt][operand word x nt x ml]  (ELL inside a record:
# operand p of task q at  base + 1 + nt + p*nt + q).  The next record starts right after, so a kernel walks the
# stream with one running offset and no other table.  Tasks of one dependency level are split into up to three
# records by length (bounded ELL padding); only the last record of a level carries the END flag (barrier).
#   meta         = nt | ml << 16 | flags << 28      flags: 1 = END of level, 2 = backward-solve record (scale by 1/u_ii),
#                                                          4 = team barrier needed after the level (factor stream),
#                                                          8 = GROUP mode (32/nt lanes per task + shuffle reduction)
#   solve task   = row | diag_slot << 16            operand = lu_slot | w_index << 16
#   factor task  = out_slot | aux << 16             operand = lu_slot_a | lu_slot_b << 16
# Padding operands are (zero slot nnz_lu, zero index n / zero slot nnz_lu).
F_END, F_BWD, F_TEAM, F_GROUP = 1, 2, 4, 8
MAX_GROUPS = 3


def _split_groups(tasks, lane=32):
    """tasks: sorted by descending operand count. Returns a list of consecutive groups minimising padded size."""
    n = len(tasks)
    if n == 0:
        return []
    lens = [len(t[2]) for t in tasks]
    # candidate cut positions: where the operand count changes (tasks are sorted by it)
    cuts = sorted(set([0, n] + [i for i in range(1, n) if lens[i] != lens[i - 1]]))
    best = {(0, 0): (0, None)}   # (cut index, groups used) -> (cost, prev)
    for gi in range(1, MAX_GROUPS + 1):
        for ci, c in enumerate(cuts[1:], 1):
            cand = None
            for pi in range(ci):
                if (pi, gi - 1) not in best:
                    continue
                a = cuts[pi]
                cost = best[(pi, gi - 1)][0] + (c - a) * (1 + lens[a]) + 24
                if cand is None or cost < cand[0]:
                    cand = (cost, pi)
            if cand is not None:
                best[(ci, gi)] = cand
    last = len(cuts) - 1
    g = min((gi for gi in range(1, MAX_GROUPS + 1) if (last, gi) in best), key=lambda gi: best[(last, gi)][0])
    out, ci = [], last
    while ci > 0:
        pi = best[(ci, g)][1]
        out.append(tasks[cuts[pi]:cuts[ci]])
        ci, g = pi, g - 1
    return out[::-1]


def _emit_records(words, level_tasks, flags, pad_word, task_word, pad_task, allow_group=True):
    """Append the records of one dependency level. Returns (n_records, estimated cycles of the level for one warp).

    Records with few long tasks use GROUP mode: nt is padded to a power of two <= 16 and 32/nt lanes share one
    task (lane l takes operands l/nt, l/nt + 32/nt, ...: its words are base + nt + l + 32k, conflict-free),
    followed by an xor-shuffle reduction.  ml is padded so that nt*ml is a multiple of 32."""
    tasks = sorted(level_tasks, key=lambda t: -len(t[2]))
    groups = _split_groups(tasks)
    cyc = 0
    for gi, grp in enumerate(groups):
        nt, ml = len(grp), len(grp[0][2])
        fl = flags | (F_END if gi == len(groups) - 1 else 0)
        ntp = 1
        while ntp < nt:
            ntp *= 2
        if allow_group and ntp <= 16 and ml >= 3:
            g = 32 // ntp
            mlp = ((ml + g - 1) // g) * g
            fl |= F_GROUP
            words.append(ntp | (mlp << 16) | (fl << 28))
            for q in range(ntp):
                words.append(task_word(grp[q]) if q < nt else pad_task)
            for p in range(mlp):
                for q in range(ntp):
                    ops = grp[q][2] if q < nt else ()
                    words.append((ops[p][0] | (ops[p][1] << 16)) if len(ops) > p else pad_word)
            k = mlp // g
            sh = {1: 5, 2: 4, 4: 3, 8: 2, 16: 1}[ntp]
            cyc += 60 + 12 * k + 28 * sh
            continue
        assert nt < 65536 and ml < 4096
        words.append(nt | (ml << 16) | (fl << 28))
        for t in grp:
            words.append(task_word(t))
        for p in range(ml):
            for t in grp:
                words.append((t[2][p][0] | (t[2][p][1] << 16)) if len(t[2]) > p else pad_word)
        cyc += 35 * ml + 25 * ((nt + 31) // 32)
    return len(groups), cyc + 45


def build_streams(s: SymbolicLU, team=128):
    """Fill s.solve_stream / s.factor_stream (and the cycle estimates used to rank orderings)."""
    n, nz = s.n, s.nnz_lu
    assert nz + 3 < 0xFFFE and n + 1 < 0xFFFF, "16-bit packed indices: pattern too large"
    one_slot = nz + 1                     # lu[nnz_lu] = 0.0, lu[nnz_lu + 1] = 1.0 (set once by the kernel)

    def level_tasks(j: JDS, lv):
        t0, nt, ob, ml = (int(x) for x in j.lev_meta[lv])
        ml &= 0xFFFF
        out = []
        for q in range(nt):
            ops = []
            for p in range(ml):
                a, b = int(j.op_a[ob + p * nt + q]), int(j.op_b[ob + p * nt + q])
                if a != nz:               # drop ELL padding of the JDS tables
                    ops.append((a, b))
            out.append((int(j.task_out[t0 + q]), int(j.task_aux[t0 + q]), ops))
        return out

    words, nrec, cyc = [], 0, 0
    for lv in range(s.fwd.n_levels):
        tasks = [t for t in level_tasks(s.fwd, lv) if t[2]]   # rows without L entries: w = gathered b already
        if not tasks:
            continue
        r, c = _emit_records(words, tasks, 0, nz | (n << 16), lambda t: t[0] | (one_slot << 16), n | (nz << 16))
        nrec, cyc = nrec + r, cyc + c
    for lv in range(s.bwd.n_levels):
        tasks = level_tasks(s.bwd, lv)
        r, c = _emit_records(words, tasks, F_BWD, nz | (n << 16), lambda t: t[0] | (int(s.diag_slot[t[0]]) << 16),
                             n | (nz << 16))
        nrec, cyc = nrec + r, cyc + c
    s.solve_stream = np.asarray(words, dtype=np.uint32)
    s.n_solve_records, s.est_cycles_solve = nrec, cyc

    words, nrec, cyc = [], 0, 0
    nlev = s.factor.n_levels
    sizes = [int(s.factor.lev_meta[lv][1]) for lv in range(nlev)]
    for lv in range(nlev):
        tasks = level_tasks(s.factor, lv)
        nxt = sizes[lv + 1] if lv + 1 < nlev else 10 ** 9
        fl = 0 if (sizes[lv] <= 32 and nxt <= 32) else F_TEAM
        # factor pad task: out = the spare slot nnz_lu + 2 (scratch), written as a U entry
        r, c = _emit_records(words, tasks, fl, nz | (nz << 16), lambda t: t[0] | (t[1] << 16), (nz + 2) | (AUX_U << 16),
                             allow_group=(fl == 0))
        # wide levels are spread over the whole team
        nrec, cyc = nrec + r, cyc + (c if sizes[lv] <= 32 else 45 + (c - 45) * 32 // team + 40)
    s.factor_stream = np.asarray(words, dtype=np.uint32)
    s.n_factor_records, s.est_cycles_factor = nrec, cyc
    return s


def _walk(stream):
    off = 0
    while off < len(stream):
        meta = int(stream[off])
        nt, ml, fl = meta & 0xFFFF, (meta >> 16) & 0xFFF, meta >> 28
        yield off, nt, ml, fl
        off += 1 + nt * (1 + ml)


def stream_factor(s: SymbolicLU, a_vals):
    """numpy emulation of the device factorisation driven by the packed stream (validation of the tables)."""
    lu = np.zeros(s.nnz_lu + 3)
    lu[s.nnz_lu + 1] = 1.0
    np.add.at(lu, s.a_slot, a_vals)
    st = s.factor_stream
    for off, nt, ml, fl in _walk(st):
        tw = st[off + 1:off + 1 + nt].astype(np.int64)
        out, aux = tw & 0xFFFF, tw >> 16
        acc = lu[out].copy()
        for p in range(ml):
            ow = st[off + 1 + nt + p * nt:off + 1 + nt + (p + 1) * nt].astype(np.int64)
            acc -= lu[ow & 0xFFFF] * lu[ow >> 16]
        for q in range(nt):
            if aux[q] == AUX_U:
                lu[out[q]] = acc[q]
            elif aux[q] == AUX_DIAG:
                lu[out[q]] = 1.0 / acc[q]
            else:
                lu[out[q]] = acc[q] * lu[aux[q]]
    return lu


def stream_solve(s: SymbolicLU, lu, b):
    w = np.zeros(s.n + 1)
    w[:s.n] = b[s.perm_r]
    st = s.solve_stream
    for off, nt, ml, fl in _walk(st):
        tw = st[off + 1:off + 1 + nt].astype(np.int64)
        row, dg = tw & 0xFFFF, tw >> 16
        acc = w[row].copy()
        for p in range(ml):
            ow = st[off + 1 + nt + p * nt:off + 1 + nt + (p + 1) * nt].astype(np.int64)
            acc -= lu[ow & 0xFFFF] * w[ow >> 16]
        if fl & F_BWD:
            acc *= lu[dg]
        w[row] = acc
    x = np.empty(s.n)
    x[s.perm_c] = w[:s.n]
    return x
