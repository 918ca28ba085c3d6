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

import os

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
    Lrows: list = None                 # strict-lower pattern per row of B = P A Q
    Urows: list = None                 # upper pattern (incl. diagonal) per row
    seq_info: dict = None              # cluster statistics of the SEQ solve stream (see build_seq_solve_stream)
    n_solve_words: int = 0             # words of the solve stream without the prefetch padding at its end
    clusters: list = None              # rows of every elimination-tree cluster (_seq_clusters)
    cl: np.ndarray = None              # row -> cluster
    inv_cluster: np.ndarray = None     # cluster has explicit inverse diagonal blocks (slots ("Q", i, j) / ("P", i, j))
    cluster_kind: list = None          # "chain" | "inv" | "seq" per cluster
    solve_program: list = None         # the records of the solve stream as dicts (code generator input)
    s_ops: dict = None                 # row -> [(slot, w index)]: solve-form forward coupling  S = -Q L_C,desc
    t_ops: dict = None                 # row -> [(slot, w index)]: solve-form backward coupling T = -P U_C,anc
    dense_rows: list = None            # rows of the dense top block T (ascending), or None
    dense_slots: np.ndarray = None     # (nT, nT) slot of L/U entry (row, col) of the block, nnz_lu where structurally zero
    factor_chains: list = None         # (first row, rows, first L slot, first U slot) of the chains factorised by the chain record

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
        # hubs: vertices adjacent to a large part of the core (the lumped temperature of a thermal model couples to every
        # through-cell variable).  Inside the dissection a hub makes every breadth-first level structure two levels deep and
        # the separators useless; taken out and eliminated last it is the border of a bordered system.
        deg = np.diff(sub.indptr)
        hub = (deg > max(32, 0.25 * len(core))) if len(core) > 64 else np.zeros(len(core), bool)
        if hub.any():
            rest = np.nonzero(~hub)[0]
            q = _nested_dissection(sub[rest][:, rest].tocsr(), leaf=leaf)
            order.extend(core[rest[q]].tolist())
            order.extend(core[np.nonzero(hub)[0]].tolist())
        else:
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


DENSE_MAX = 32      # rows of the dense top block (one warp computes its inverse, one lane per column)
DENSE_TEAM_MIN = 128   # only multi-warp teams use it: the block is applied by rows x lanes spread over the team
DENSE_MIN_HEIGHTS = 3  # tree heights it must replace to be worth two records
INV_MAX = 8         # clusters of 2..INV_MAX rows get explicit inverse diagonal blocks
ONE = object()      # placeholder of the constant-one slot (nnz_lu + 1) while slots are still being handed out
AUX_U = 0xFFFF      # factor task writes an off-diagonal U entry
AUX_DIAG = 0xFFFE   # factor task writes a diagonal entry (stores the reciprocal)
AUX_NEG = 0xFFFD    # factor task writes -sum(pairs), ignoring the old value of the slot (solve-form entries)


def analyse(n, rows, cols, samples, ordering="auto", diag_pivot_thresh=0.01, team=128) -> SymbolicLU:
    """rows/cols: structural pattern of A (every entry the numeric phase may set).
    samples: list of value arrays (same order as rows/cols) used to pick the pivot order
    and to validate it."""
    rows = np.asarray(rows, np.int64)
    cols = np.asarray(cols, np.int64)
    candidates, late = [], []
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
                # The pivots SuperLU chose belong to ITS elimination sequence; after the symmetric re-ordering below an entry
                # it matched may sit on the diagonal before any fill reaches it, and such an entry can be a structural
                # non-zero that is numerically zero.  When that can happen the natural diagonal (J - cj M of a semi-explicit
                # DAE has a full one) is a second candidate -- the first one for models so large that every candidate
                # costs a minute of symbolic work.
                dg = [np.abs(sp.csc_matrix((np.asarray(v, float), (rows, cols)), shape=(n, n))[pr, :].diagonal()) for v in samples]
                variants = [pr]
                if any(float(d.min()) == 0.0 for d in dg):
                    variants = [np.arange(n), pr] if n > 2048 else [pr, np.arange(n)]
                for vi, pr in enumerate(variants[1:]):
                    Bp = pat[pr, :]
                    sym = (Bp + Bp.T).tocsr()
                    sym.setdiag(0)
                    sym.eliminate_zeros()
                    q = _nested_dissection(sym) if name == "nd" else _peel_then_nd(sym)
                    late.append((name + "_alt", pr[q], q))
                pr = variants[0]
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
    # the alternatives (natural diagonal) are only looked at when no regular candidate is valid: the orders of the models
    # that never needed them stay what they were
    for pool in (candidates, late):
        if best is not None:
            break
        for name, perm_r, perm_c in pool:
            try:
                s = _build(n, rows, cols, perm_r, perm_c, name, team=team)
            except AssertionError:           # unsymmetric fill the cluster schedule cannot express: pad the pattern and retry
                try:
                    s = _build(n, rows, cols, perm_r, perm_c, name, team=team, symmetrise=True)
                except AssertionError as e:  # still not expressible: next candidate
                    print(f"[symbolic_lu] ordering {name} rejected: {e}")
                    continue
            # cost proxy (cycles per factorisation interval): ~30 triangular solves per factorisation (measured on the
            # battery models); the L+U values and the solve stream must stay in shared memory (8 + ~6 bytes per entry)
            score = 30.0 * s.est_cycles_solve + s.est_cycles_factor + 30.0 * s.nnz_lu
            ok = all(_validate(s, rows, cols, np.asarray(v, float)) < 1e-6 for v in samples)
            if not ok:
                continue
            if best is None or score < best[0]:
                best = (score, s)
            if n > 2048:      # every candidate costs a minute of symbolic work at this size: the first valid one is taken
                break
    if best is None:
        raise RuntimeError("no static pivot order passed validation on the sample matrices")
    return best[1]


def _build(n, rows, cols, perm_r, perm_c, name, team=128, symmetrise=False) -> SymbolicLU:
    inv_r = np.empty(n, np.int64)
    inv_r[perm_r] = np.arange(n)
    inv_c = np.empty(n, np.int64)
    inv_c[perm_c] = np.arange(n)
    bi, bj = inv_r[rows], inv_c[cols]
    B = sp.csr_matrix((np.ones(len(rows)), (bi, bj)), shape=(n, n))
    if symmetrise:
        # structural zeros that make the permuted pattern symmetric: the cluster schedule below relies on the elimination
        # tree of a structurally symmetric matrix (every U entry of a row points at an ancestor); a lumped thermal state
        # that reads more variables than it feeds breaks that by a few dozen entries
        B = ((B + B.T) > 0).astype(float).tocsr()
    Lrows, Urows = _symbolic_fill(B)
    # slots: row-major over (L row, U row)
    # the diagonal of row i sits in slot i (the backward solve reads 1/u_ii as lu[row] without a table look-up)
    slot_of = {(i, i): i for i in range(n)}
    clusters, cl, kinds = _seq_clusters(n, Lrows, Urows)
    dense_rows = None
    if team >= DENSE_TEAM_MIN and os.environ.get("PBGPU_DENSE_TOP", "1") != "0":
        clusters, cl, kinds, dense_rows = _dense_top(n, Lrows, clusters, cl, kinds)
    for c, rws in enumerate(clusters):     # chains: L entries in consecutive slots, then U entries in consecutive slots
        if kinds[c] == "chain":
            for k in range(1, len(rws)):
                slot_of[(rws[k], rws[k - 1])] = len(slot_of)
            for i in rws:
                slot_of[(i, Urows[i][1])] = len(slot_of)
    for i in range(n):
        for j in Lrows[i]:
            slot_of.setdefault((i, j), len(slot_of))
        for j in Urows[i]:
            slot_of.setdefault((i, j), len(slot_of))
    # small clusters of the elimination tree (separators, leaves) get the explicit inverse of their diagonal block:
    # Q = L_CC^-1 (strict lower part) and P = U_CC^-1 (strict upper part) in extra slots, computed by extra factor
    # tasks, so that the rows of a cluster are solved side by side instead of one after the other (PAR records below)
    inv_cluster = np.array([kinds[c] == "inv" and len(clusters[c]) >= 2 for c in range(len(clusters))])
    q_pairs, p_pairs = {}, {}         # (i, j) -> operand pairs of the factor task
    for c, rws in enumerate(clusters):
        if not inv_cluster[c]:
            continue
        for jj, j in enumerate(rws):                  # Q[i][j] = -(L[i][j] + sum_{j<k<i} L[i][k] Q[k][j])
            for i in rws[jj + 1:]:
                pr = [(slot_of[(i, j)], ONE)] if (i, j) in slot_of else []
                pr += [(slot_of[(i, k)], slot_of[("Q", k, j)]) for k in rws[jj + 1:] if k < i and (i, k) in slot_of
                       and ("Q", k, j) in slot_of]
                if pr:
                    slot_of[("Q", i, j)] = len(slot_of)
                    q_pairs[(i, j)] = pr
        for jj, j in enumerate(rws):                  # P[i][j] = -1/u_ii * sum_{i<k<=j} U[i][k] P[k][j],  P[j][j] = 1/u_jj
            for i in reversed(rws[:jj]):
                pr = [(slot_of[(i, k)], slot_of[(j, j)] if k == j else slot_of[("P", k, j)]) for k in rws if i < k <= j
                      and (i, k) in slot_of and (k == j or ("P", k, j) in slot_of)]
                if pr:
                    slot_of[("P", i, j)] = len(slot_of)
                    p_pairs[(i, j)] = pr
    # solve form of the coupling blocks: with S = -Q_CC L_C,desc and T = -P_CC U_C,anc (P including 1/u_ii) a cluster
    # is solved in ONE step per sweep,  x_C = w_C + Q w_C + S z_desc   /   x_C = w_C / u + P w_C + T x_anc .
    # S and T overwrite L_C,desc and U_C,anc after the factorisation proper (extra slots only for their fill); the
    # tasks run in levels after all others, rows of a cluster ordered so that an original entry is read by the rows
    # that need it before its own row replaces it (forward: last row first; backward: first row first).
    s_tasks, t_tasks = [], []         # (out slot, pairs, rank within the cluster)
    s_ops, t_ops = {}, {}             # row -> [(slot, w index)] of the solve-form coupling operands
    for c, rws in enumerate(clusters):
        if kinds[c] != "inv":
            continue
        inside = set(rws)
        for rank, i in enumerate(reversed(rws)):
            contrib = [(i, ONE)] + [(j, slot_of[("Q", i, j)]) for j in rws if j < i and ("Q", i, j) in slot_of]
            cols = {}
            for j, coef in contrib:
                for k in Lrows[j]:
                    if k not in inside:
                        cols.setdefault(k, []).append((coef, slot_of[(j, k)]))
            for k in sorted(cols):
                out = slot_of.setdefault((i, k), len(slot_of)) if (i, k) in slot_of else slot_of.setdefault(("S", i, k), len(slot_of))
                s_tasks.append((out, cols[k], rank))
                s_ops.setdefault(i, []).append((out, k))
        for rank, i in enumerate(rws):
            contrib = [(i, slot_of[(i, i)])] + [(j, slot_of[("P", i, j)]) for j in rws if j > i and ("P", i, j) in slot_of]
            cols = {}
            for j, coef in contrib:
                for k in Urows[j]:
                    if k != j and k not in inside:
                        cols.setdefault(k, []).append((coef, slot_of[(j, k)]))
            for k in sorted(cols):
                out = slot_of[(i, k)] if (i, k) in slot_of else slot_of.setdefault(("T", i, k), len(slot_of))
                t_tasks.append((out, cols[k], rank))
                t_ops.setdefault(i, []).append((out, k))
    nnz_lu = len(slot_of)
    for d in (q_pairs, p_pairs):
        for key, pr in d.items():
            d[key] = [(a, nnz_lu + 1 if b is ONE else b) for a, b in pr]
    s_tasks = [(out, [(nnz_lu + 1 if a is ONE else a, b) for a, b in pr], rank) for out, pr, rank in s_tasks]
    diag_slot = np.array([slot_of[(i, i)] for i in range(n)], np.int64)
    a_slot = np.array([slot_of[(int(i), int(j))] for i, j in zip(bi, bj)], np.int64)

    # column structure of U for pairing: Ucols[j] = {k: (k, j) in U}
    Lset = [set(r) for r in Lrows]
    Ucol = [[] for _ in range(n)]
    for k in range(n):
        for j in Urows[k]:
            Ucol[j].append(k)
    # Bidiagonal chains (the radial particle chains: 40 x 18 rows in the DFN) are factorised by ONE record that runs before
    # every level, one lane per chain: d_0 = 1 / a_00, then l_k = a_{k,k-1} d_{k-1}, d_k = 1 / (a_kk - l_k u_{k-1,k}) down the
    # chain.  Level-scheduled they were two dependency levels per row -- a quarter of the ~200 records of a DFN
    # factorisation, each about 1 100 cycles of latency however few tasks it holds.  Same operations in the same order,
    # bit-identical values.  PBGPU_FACTOR_CHAINS=0 keeps them in the levels (A/B).
    chain_rows = {}
    factor_chains = []      # (first row, rows, slot of the first L entry, slot of the first U entry)
    if os.environ.get("PBGPU_FACTOR_CHAINS", "1") != "0":
        for c, rws in enumerate(clusters):
            if kinds[c] == "chain" and len(rws) >= 2:
                for k, i in enumerate(rws):
                    chain_rows[i] = k
                factor_chains.append((rws[0], len(rws), slot_of[(rws[1], rws[0])], slot_of[(rws[0], Urows[rws[0]][1])]))
    # factor tasks with levels
    level = {}
    tasks_by_level = {}
    flops = 0
    for i in range(n):
        row_entries = [(i, j) for j in Lrows[i]] + [(i, j) for j in Urows[i]]
        for (_, j) in row_entries:
            if i in chain_rows and j <= i:
                # done by the chain record, which runs before level 0 (the U entry of a chain row stays a plain entry)
                assert (j == i or (j == i - 1 and chain_rows[i] >= 1)) and (chain_rows[i] >= 1 or not Lrows[i])
                level[(i, j)] = -1
                flops += 2
                continue
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
    slot_level = {slot_of[k]: v for k, v in level.items()}
    for (i, j), pr in q_pairs.items():
        lv = 1 + max(slot_level.get(x, -1) for ab in pr for x in ab)
        slot_level[slot_of[("Q", i, j)]] = lv
        tasks_by_level.setdefault(lv, []).append((slot_of[("Q", i, j)], AUX_U, pr))
        flops += 2 * len(pr)
    for (i, j), pr in p_pairs.items():
        lv = 1 + max(max(slot_level.get(x, -1) for ab in pr for x in ab), slot_level[slot_of[(i, i)]])
        slot_level[slot_of[("P", i, j)]] = lv
        tasks_by_level.setdefault(lv, []).append((slot_of[("P", i, j)], slot_of[(i, i)], pr))
        flops += 2 * len(pr) + 1
    base = max(tasks_by_level) + 1
    for out, pr, rank in s_tasks + t_tasks:
        tasks_by_level.setdefault(base + rank, []).append((out, AUX_NEG, pr))
        flops += 2 * len(pr)
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
                                    factor, fwd, bwd, flops, sflops, name, Lrows=Lrows, Urows=Urows,
                                    clusters=clusters, cl=cl, inv_cluster=inv_cluster, cluster_kind=kinds,
                                    s_ops=s_ops, t_ops=t_ops, dense_rows=dense_rows, factor_chains=factor_chains,
                                    dense_slots=None if dense_rows is None else np.array(
                                        [[slot_of.get((i, k), nnz_lu) for k in dense_rows] for i in dense_rows], np.int64)),
                         team=team)


# ---------------------------------------------------------------------------
# numpy emulation of the numeric phase (validation of tables; tests)
# ---------------------------------------------------------------------------
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
F_CHAINS = F_BWD      # factor stream only (its records have no direction): the chain record, see _build
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
    assert nz + 3 < 0xFFFD and n + 1 < 0xFFFF, "16-bit packed indices: pattern too large"
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

    build_seq_solve_stream(s, team)

    words, nrec, cyc = [], 0, 0
    if s.factor_chains:
        # chain record: [nt | 1 << 16 | flags << 28] then nt words r0 | K << 16 and nt words L0 | U0 << 16 (one lane per chain)
        nt = len(s.factor_chains)
        words.append(nt | (1 << 16) | ((F_CHAINS | F_TEAM | F_END) << 28))
        words += [r0 | (K << 16) for r0, K, l0, u0 in s.factor_chains]
        words += [l0 | (u0 << 16) for r0, K, l0, u0 in s.factor_chains]
        nrec, cyc = 1, 45 + 90 * max(K for _, K, _, _ in s.factor_chains)   # same scale as two levels per row
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


# ---------------------------------------------------------------------------
# SEQ solve stream: thread-sequential clusters of the elimination tree
# ---------------------------------------------------------------------------
# Level scheduling pays one dependent round trip through shared memory *and* the interpreter per dependency level.
# The battery models' factors are made of long thin chains (particle radial meshes) hanging off a nested-dissection
# tree of small dense separators: their depth is close to their size, so a level holds a handful of rows.  Here the
# rows are instead grouped in *clusters* (maximal single-child paths of the elimination tree: a chain, a separator);
# one thread walks one cluster row by row with no synchronisation, its only loop-carried dependency being the
# w[] value it has just stored.  Clusters of equal height run side by side, one record per height:
#   record  = [nw | flags << 16] [words in the record] [skip] [nt_v | len_v << 16] * nw  + nw blocks, one per warp:
#             block v holds len_v * nt_v words, word j of its thread q (= lane) at block + j * nt_v + q
#   thread  = [ntask] task ... task [dummy task]      (dummy: target of the software prefetch of the last task)
#   task    = [row | ml << 16] [op] * ml              op = lu_slot | w_index << 16,  ml = 1 + 4k (zero-operand padded);
#             backward tasks finish with  w[row] *= lu[row]  (slot i holds 1/u_ii)
# flags: 1 backward, 2 WIDE (warp v of the team runs block v), 4 team barrier after the record.
# Narrow records are run by warp 0 alone; the other warps jump to `skip` (the end of the narrow run) and wait there.
SEQ_HDR = 4
SEQ_NARROW = 1 << 31
PAR_ADD, PAR_DIAG = 2, 4
PAR_BASE_CYC, PAR_OP_CYC, PAR_SHFL_CYC = 150, 25, 40
CHAIN_BASE_CYC, CHAIN_ROW_CYC = 200, 12
PAR_WIDE_GROUPS = False   # several lanes per row in multi-warp records: measured no faster, costs stream padding
PAR_MLG_SHAPES = (0, 1, 2, 3, 4, 6, 8, 10, 12, 16, 20, 24, 32)   # operand counts per lane the device code is compiled for
# cycle model of one task (measured B200 latencies: LDS 29, DFMA/DADD 8): chain-critical LDS + 2 DFMA levels + store,
# + 4 cycles per operand beyond the two prefetched ones (two accumulators) + one dependent LDS pair when ml > 2
SEQ_TASK_CYC, SEQ_CHUNK_CYC, SEQ_LONG_CYC, SEQ_REC_CYC = 90, 45, 30, 150


def _dense_top(n, Lrows, clusters, cl, kinds):
    """The top of the elimination tree as ONE dense block.  Every tree height costs a record of several hundred cycles
    however few rows it has, and the last heights hold a handful of rows each; the clusters above a height h0 (a set
    closed under ancestors, at most DENSE_MAX rows) are merged into one cluster of kind "dense": the factorisation
    leaves their L and U entries plain and forms W = (L_TT U_TT)^-1 explicitly, the solve replaces all their records
    of both sweeps by  r_T = b_T - L_T,desc z_desc  and  x_T = W r_T.  The merged cluster goes last (it depends on
    everything and only the backward sweep reads it)."""
    nc = len(clusters)
    hf = np.zeros(nc, np.int64)
    for c, rws in enumerate(clusters):
        for i in rws:
            for k in Lrows[i]:
                if cl[k] != c:
                    hf[c] = max(hf[c], hf[cl[k]] + 1)
    hmax = int(hf.max())
    def fits(h):      # rows above height h: at most DENSE_MAX, and the scratch for the inverse (column-major, odd
        m = sum(len(clusters[c]) for c in range(nc) if hf[c] >= h)   # leading dimension) is the solve's w vector
        return m <= DENSE_MAX and m * (m | 1) <= n

    h0 = hmax + 1
    while h0 > 0 and fits(h0 - 1):
        h0 -= 1
    top = [c for c in range(nc) if hf[c] >= h0]
    rows = sorted(i for c in top for i in clusters[c])
    if hmax + 1 - h0 < DENSE_MIN_HEIGHTS or any(kinds[c] == "chain" for c in top):
        return clusters, cl, kinds, None
    keep = [c for c in range(nc) if hf[c] < h0]
    clusters = [clusters[c] for c in keep] + [rows]
    kinds = [kinds[c] for c in keep] + ["dense"]
    cl = np.full(n, -1, np.int64)
    for c, rws in enumerate(clusters):
        cl[rws] = c
    return clusters, cl, kinds, rows


def _seq_clusters(n, Lrows, Urows):
    """Clusters = maximal paths i1 -> i2 -> ... of the elimination tree in which every node has exactly one child."""
    first_below = [None] * n          # first row i > k with L[i, k] != 0
    for i in range(n):
        for k in Lrows[i]:
            if first_below[k] is None:
                first_below[k] = i
    parent = np.full(n, -1, np.int64)
    for i in range(n):
        cand = [j for j in Urows[i] if j > i][:1]
        if first_below[i] is not None:
            cand.append(first_below[i])
        if cand:
            parent[i] = min(cand)
    nchild = np.zeros(n, np.int64)
    only_child = np.full(n, -1, np.int64)
    for i in range(n):
        if parent[i] >= 0:
            nchild[parent[i]] += 1
            only_child[parent[i]] = i
    cl = np.full(n, -1, np.int64)
    clusters = []
    for i in range(n):
        c = only_child[i] if nchild[i] == 1 else -1
        if c >= 0 and clusters[cl[c]][-1] == c:
            cl[i] = cl[c]
            clusters[cl[i]].append(i)
        else:
            cl[i] = len(clusters)
            clusters.append([i])
    # split the pure bidiagonal start of a long cluster off as a CHAIN (consecutive rows i0 .. i0+K-1 with
    # L row = {previous row} and U row = {diagonal, next row}): its recurrence needs no tables at all
    out, kinds = [], []
    for rws in clusters:
        K = 0
        if len(rws) > INV_MAX:
            while K < len(rws) - 1:
                i = rws[K]
                if i != rws[0] + K or Lrows[i] != ([rws[K - 1]] if K else []) or Urows[i] != [i, rws[K + 1]]:
                    break
                K += 1
        if K >= 4:
            out.append(rws[:K])
            kinds.append("chain")
            rws = rws[K:]
        if rws:
            out.append(rws)
            kinds.append("inv" if len(rws) <= INV_MAX else "seq")
    order = sorted(range(len(out)), key=lambda c: out[c][0])     # ascending first row = topological
    clusters = [out[c] for c in order]
    kinds = [kinds[c] for c in order]
    for c, rws in enumerate(clusters):
        cl[rws] = c
    return clusters, cl, kinds


def build_seq_solve_stream(s: SymbolicLU, team=128):
    """Solve stream = SEQ records (long chains: one thread per cluster, rows one after the other) and PAR records
    (clusters with explicit inverse diagonal blocks: one group of G lanes per row, every lane one partial dot product).

    record header (4 words): [kind | mode << 4 | sync << 8 | first_warp << 12 | nblocks << 16] [words in the record]
                             [skip] [G | mlg << 8]                 followed by one descriptor word per block
    kind 0 SEQ: descriptor nt | len << 16, block = len * nt words (layout in the comment block above)
    kind 1 PAR: descriptor nt,             block = (1 + mlg) * nt words: [row] then mlg operand words per lane
    mode: 1 backward SEQ; PAR: 2 ADD (base + sum, else base - sum), 4 DIAG (base = w[row] / u_ii, else w[row])
    sync: 0 none, 1 warp, 2 team.   Warp v runs block v - first_warp.  Warps other than 0 jump to `skip` from a
    narrow run (records of warp 0 alone) and meet warp 0 at the team barrier that ends it."""
    n, nz = s.n, s.nnz_lu
    Lrows, Urows, clusters, cl, inv = s.Lrows, s.Urows, s.clusters, s.cl, s.inv_cluster
    nc = len(clusters)
    kinds = s.cluster_kind
    par = np.array([kinds[c] == "inv" for c in range(nc)])
    # heights from the actual dependencies (not from the tree): always a valid schedule
    hf = np.zeros(nc, np.int64)
    for c, rws in enumerate(clusters):
        for i in rws:
            for k in Lrows[i]:
                if cl[k] != c:
                    hf[c] = max(hf[c], hf[cl[k]] + 1)
    # backward sweep: as late as possible, so that the long leaf chains all land in the last record instead of
    # stretching every record on the way down (a record lasts as long as its longest thread)
    deps_b = [set() for _ in range(nc)]
    for c, rws in enumerate(clusters):
        for i in rws:
            for j in Urows[i]:
                if j != i and cl[j] != c:
                    assert cl[j] > c, "upper dependency on an earlier cluster"
                    deps_b[c].add(int(cl[j]))
    tail = np.zeros(nc, np.int64)
    for c in range(nc):               # dependents of d are clusters c < d: ascending order sees them first
        for d in deps_b[c]:
            tail[d] = max(tail[d], tail[c] + 1)
    hb = int(tail.max()) - tail
    pad_op = nz | (n << 16)
    n_warps = max(team // 32, 1)

    def seq_task(i, back):
        ops = [(s.slot_of[(i, j)], j) for j in Urows[i] if j != i] if back else [(s.slot_of[(i, k)], k) for k in Lrows[i]]
        if not back and not ops:
            return None, 0            # w[i] is already the gathered right-hand side
        ml = 1 + 4 * ((max(len(ops), 1) + 2) // 4)      # one prefetched operand + chunks of four
        wd = [i | (ml << 16)] + [a | (b << 16) for a, b in ops] + [pad_op] * (ml - len(ops))
        cyc = SEQ_TASK_CYC + (8 if back else 0) + (SEQ_LONG_CYC + SEQ_CHUNK_CYC * ((ml - 1) // 4) if ml > 1 else 0)
        return wd, cyc

    dummy = [n | (1 << 16), pad_op]

    def fold(progs, nbins):
        """longest-processing-time folding of cluster programs [cyc, ntask, words] into at most nbins threads"""
        progs = sorted(progs, key=lambda t: -t[0])
        if len(progs) <= nbins:
            return progs
        bins = [list(t) for t in progs[:nbins]]
        for t in progs[nbins:]:
            bmin = min(bins, key=lambda x: x[0])
            bmin[0] += t[0]
            bmin[1] += t[1]
            bmin[2] = bmin[2] + t[2]
        return sorted(bins, key=lambda t: -t[0])

    def seq_record(progs, back, nw):
        progs = fold(progs, 32 * nw)
        nb = min(nw, (len(progs) + 31) // 32)
        blocks = [progs[v::nb] for v in range(nb)]
        desc, body = [], []
        for blk in blocks:
            streams = [[t[1]] + t[2] + dummy for t in blk]
            ln = max(len(x) for x in streams)
            m = np.zeros((ln, len(blk)), np.int64)
            for q, x in enumerate(streams):
                m[:len(x), q] = x
            desc.append(len(blk) | (ln << 16))
            body += m.reshape(-1).tolist()
        return dict(kind=0, mode=1 if back else 0, nblocks=nb, param=0, desc=desc, body=body, cyc=progs[0][0],
                    threads=[d & 0xFFFF for d in desc])

    def chain_records(cs, back, nw):
        """kind 2: one lane per chain, three words per lane: [first row | K << 16] [L slot of row 1 | U slot of row 0 << 16]
        [parent row of the last chain row].  One record per 32 * nw chains."""
        cs = sorted(cs, key=lambda c: -len(clusters[c]))
        out = []
        for at in range(0, len(cs), 32 * nw):
            part = cs[at:at + 32 * nw]
            nb = min(nw, (len(part) + 31) // 32)
            desc, body = [], []
            for v in range(nb):
                blk = part[v::nb]
                m = np.zeros((3, len(blk)), np.int64)
                for q, c in enumerate(blk):
                    rws = clusters[c]
                    m[0, q] = rws[0] | (len(rws) << 16)
                    m[1, q] = s.slot_of[(rws[1], rws[0])] | (s.slot_of[(rws[0], rws[1])] << 16)
                    m[2, q] = Urows[rws[-1]][1]
                desc.append(len(blk))
                body += m.reshape(-1).tolist()
            K = len(clusters[part[0]])
            out.append(dict(kind=2, mode=1 if back else 0, nblocks=nb, param=K, desc=desc, body=body,
                            cyc=CHAIN_BASE_CYC + CHAIN_ROW_CYC * K, threads=desc))
        return out

    def par_records(rows_ops, mode, nw, split_rows=False):
        """rows_ops: list of (row, [(slot, w index), ...]).  The rows of one cluster stay in one block (= one warp): they
        read each other's old values, and "every lane reads before any lane writes" only holds inside a warp.
        split_rows: the operands are never rows of the record itself (dense top block), rows go to any warp and get
        as many lanes as the team has."""
        units = {}
        for row, ops in rows_ops:
            units.setdefault(row if split_rows else int(cl[row]), []).append((row, ops))
        units = sorted(units.values(), key=lambda u: -len(u))
        # lanes per row: as many as the warps can hold (with slack for packing whole clusters into warps)
        maxml_all = max(len(t[1]) for t in rows_ops)
        Gw = 1
        while split_rows and 2 * Gw * len(rows_ops) <= 32 * nw and Gw < min(maxml_all, 32):
            Gw *= 2
        while (PAR_WIDE_GROUPS and 2 * Gw * len(rows_ops) <= 24 * nw and 2 * Gw * len(units[0]) <= 32 and Gw < maxml_all
               and nw > 1):
            Gw *= 2
        records, cur = [], []
        for u in units:                      # first fit into the blocks of the current record
            for blk in cur:
                if (len(blk) + len(u)) * Gw <= 32:
                    blk += u
                    break
            else:
                if len(cur) == nw:
                    records.append(cur)
                    cur = []
                cur.append(list(u))
        if cur:
            records.append(cur)
        out = []
        for blocks in records:
            nb = len(blocks)
            maxml = max(len(t[1]) for blk in blocks for t in blk)
            G = Gw
            if nb == 1:
                while 2 * G * len(blocks[0]) <= 32 and G < maxml:
                    G *= 2
            mlg = (maxml + G - 1) // G
            mlg = next(v for v in PAR_MLG_SHAPES if v >= mlg) if mlg <= PAR_MLG_SHAPES[-1] else mlg
            desc, body = [], []
            for blk in blocks:
                nt = len(blk) * G
                m = np.full((1 + mlg, nt), pad_op, np.int64)
                for g, (row, ops) in enumerate(blk):
                    m[0, g * G:(g + 1) * G] = row
                    for k, (a, b) in enumerate(ops):
                        m[1 + k // G, g * G + k % G] = a | (b << 16)
                desc.append(nt)
                body += m.reshape(-1).tolist()
            lg = G.bit_length() - 1
            out.append(dict(kind=1, mode=mode, nblocks=nb, param=G | (mlg << 8), desc=desc, body=body,
                            cyc=PAR_BASE_CYC + PAR_OP_CYC * mlg + PAR_SHFL_CYC * lg, threads=desc))
        return out

    recs = []                         # records in execution order, grouped by height: [(height key, [records])]
    for back, heights in ((False, hf), (True, hb)):
        for h in range(int(heights.max()) + 1):
            cs = [c for c in range(nc) if heights[c] == h]
            progs, rows_a, rows_b, chains = [], [], [], []
            dense_c = None
            for c in cs:
                rws = clusters[c][::-1] if back else clusters[c]
                if kinds[c] == "dense":       # both sweeps of the block happen in the forward one
                    if not back:
                        dense_c = c
                    continue
                if kinds[c] == "chain":
                    chains.append(c)
                    continue
                if not par[c]:
                    wd, cyc, nt_ = [], 0, 0
                    for i in rws:
                        tw, tc = seq_task(i, back)
                        if tw is not None:
                            wd += tw
                            cyc += tc
                            nt_ += 1
                    if nt_:
                        progs.append([cyc, nt_, wd])
                    continue
                for i in rws:
                    if back:     # x_i = w_i / u_ii + sum_j P[i][j] w_j + sum_k T[i][k] x_k
                        ops = [(s.slot_of[("P", i, j)], j) for j in clusters[c] if j > i and ("P", i, j) in s.slot_of]
                        rows_b.append((i, ops + s.t_ops.get(i, [])))
                    else:        # x_i = w_i + sum_j Q[i][j] w_j + sum_k S[i][k] z_k
                        ops = [(s.slot_of[("Q", i, j)], j) for j in clusters[c] if j < i and ("Q", i, j) in s.slot_of]
                        ops = ops + s.s_ops.get(i, [])
                        if ops:
                            rows_b.append((i, ops))
            group = []
            # chains on the first warps, the first PAR phase of the other clusters beside them on the remaining warps
            first_par = rows_a if rows_a else rows_b
            seq_w = n_warps
            side_by_side = False
            if (progs or chains) and first_par and n_warps > 1 and len(progs) + len(chains) > 8:
                need = (len(first_par) + 23) // 24      # blocks hold whole clusters: allow for packing slack
                if need < n_warps:
                    seq_w, side_by_side = n_warps - need, True
            if chains and progs:
                side_by_side = False          # not worth a third concurrent stream: one record after the other
                seq_w = n_warps
            if chains:
                if len(chains) > 32 * seq_w:
                    side_by_side = False      # several chain records: no PAR record beside them
                for r in chain_records(chains, back, seq_w if len(chains) > 32 or side_by_side else 1):
                    r["first_warp"] = 0
                    group.append(r)
            if progs:
                r = seq_record(progs, back, seq_w if len(progs) > 32 or side_by_side else 1)
                r["first_warp"] = 0
                group.append(r)
            for phase, (rows_x, mode) in enumerate(((rows_a, 0), (rows_b, PAR_ADD | (PAR_DIAG if back else 0)))):
                if not rows_x:
                    continue
                beside = side_by_side and rows_x is first_par
                nw = (n_warps - seq_w) if beside else (n_warps if len(rows_x) > 32 else 1)
                for r in par_records(rows_x, mode, nw):
                    r["first_warp"] = seq_w if beside else 0
                    r["beside"] = beside
                    group.append(r)
            if dense_c is not None:
                # r_T = b_T - L_T,desc z_desc (plain L entries, all lanes of the team), then x_T = W r_T
                T = clusters[dense_c]
                pre = [(i, [(s.slot_of[(i, k)], k) for k in Lrows[i] if cl[k] != dense_c]) for i in T]
                pre = [t for t in pre if t[1]]
                if pre:
                    for r in par_records(pre, 0, n_warps, split_rows=True):
                        r["first_warp"], r["beside"], r["pre"], r["wide"] = 0, False, True, True
                        group.append(r)
                GD = 1
                while 2 * GD * len(T) <= team and GD < 8:
                    GD *= 2
                group.append(dict(kind=3, mode=0, nblocks=1, param=GD | (len(T) << 8), desc=[len(T)], body=list(T),
                                  cyc=PAR_BASE_CYC + PAR_SHFL_CYC * 3, threads=[len(T) * GD], first_warp=0,
                                  beside=False, wide=True, team_need=len(T) * GD))
            for r in group:
                r["back"], r["height"] = back, h
            recs += group
    # sync codes: nothing between the chain record and the PAR record running beside it; otherwise a warp barrier
    # when this record and the next one involve warp 0 alone, a team barrier in every other case
    def narrow(r):
        return r["first_warp"] == 0 and r["nblocks"] == 1 and not r.get("wide")

    for k, r in enumerate(recs):
        nxt = recs[k + 1] if k + 1 < len(recs) else None
        if nxt is not None and nxt.get("beside"):
            r["sync"] = 0
        elif nxt is not None and narrow(r) and narrow(nxt) and not r.get("beside"):
            r["sync"] = 1
        else:
            r["sync"] = 2
    words, offs = [], []
    for r in recs:
        offs.append(len(words))
        w0 = r["kind"] | (r["mode"] << 4) | (r["sync"] << 8) | (r["first_warp"] << 12) | (r["nblocks"] << 16)
        total = SEQ_HDR + len(r["desc"]) + len(r["body"])
        words += [w0, total, 0, r["param"]] + r["desc"] + r["body"]
    end = len(words)
    # narrow runs: maximal sequences of narrow records joined by warp barriers; other warps skip to the end of the run
    k = 0
    while k < len(recs):
        if narrow(recs[k]) and recs[k]["sync"] != 0 and not recs[k].get("beside") and (k == 0 or recs[k - 1]["sync"] == 2):
            e = k
            while recs[e]["sync"] == 1:
                e += 1
            stop = offs[e + 1] if e + 1 < len(recs) else end
            for q in range(k, e + 1):
                words[offs[q]] |= SEQ_NARROW
                words[offs[q] + 2] = stop
                recs[q]["narrow"] = True
            k = e + 1
        else:
            k += 1
    s.n_solve_words = end
    words += [0] * (4 * 32)           # the operand prefetch of the very last SEQ task reads up to 4 * nt words ahead
    s.solve_stream = np.asarray(words, dtype=np.uint32)
    # critical path estimate: records run one after the other except a PAR record beside its chain record
    est, k = 0, 0
    while k < len(recs):
        c = recs[k]["cyc"]
        if recs[k]["sync"] == 0 and k + 1 < len(recs):
            c = max(c, recs[k + 1]["cyc"])
            k += 1
        est += c + SEQ_REC_CYC
        k += 1
    s.n_solve_records, s.est_cycles_solve = len(recs), int(est)
    for r, o in zip(recs, offs):
        r["offset"] = o
    s.solve_program = recs
    s.seq_info = dict(clusters=nc, fwd_heights=int(hf.max()) + 1, bwd_heights=int(hb.max()) + 1,
                      records=[dict(back=r["back"], height=r["height"], kind=("SEQ", "PAR", "CHAIN", "DENSE")[r["kind"]], mode=r["mode"],
                                    sync=r["sync"], first_warp=r["first_warp"], threads=r["threads"],
                                    G=r["param"] & 0xFF, mlg=r["param"] >> 8, words=len(r["body"]), cyc=r["cyc"])
                               for r in recs],
                      team_min=max(max(32 * (r["first_warp"] + r["nblocks"]), r.get("team_need", 0)) for r in recs))
    return s


def _walk(stream):
    off = 0
    while off < len(stream):
        meta = int(stream[off])
        nt, ml, fl = meta & 0xFFFF, (meta >> 16) & 0xFFF, meta >> 28
        yield off, nt, ml, fl
        off += 1 + nt * (1 + ml)


def factor_chunks(stream, max_words):
    """Pack consecutive whole records of the factor stream into chunks of at most ``max_words`` words.
    Returns [(word offset, number of records), ...] or None when a single record does not fit."""
    chunks, start, nrec, end = [], 0, 0, 0
    for off, nt, ml, fl in _walk(stream):
        size = 1 + nt * (1 + ml)
        if size > max_words:
            return None
        if off + size - start > max_words:
            chunks.append((start, nrec))
            start, nrec = off, 0
        nrec += 1
        end = off + size
    if nrec:
        chunks.append((start, nrec))
    assert end == len(stream)
    return chunks


def stream_factor(s: SymbolicLU, a_vals):
    """numpy emulation of the device factorisation driven by the packed stream (validation of the tables)."""
    lu = np.zeros(s.nnz_lu + 3)
    lu[s.nnz_lu + 1] = 1.0
    np.add.at(lu, s.a_slot, a_vals)
    st = s.factor_stream
    for off, nt, ml, fl in _walk(st):
        if fl & F_CHAINS:
            for q in range(nt):
                h, sl = int(st[off + 1 + q]), int(st[off + 1 + nt + q])
                r0, K, l0, u0 = h & 0xFFFF, h >> 16, sl & 0xFFFF, sl >> 16
                inv = 1.0 / lu[r0]
                lu[r0] = inv
                for k in range(1, K):
                    l = lu[l0 + k - 1] * inv
                    lu[l0 + k - 1] = l
                    inv = 1.0 / (lu[r0 + k] - l * lu[u0 + k - 1])
                    lu[r0 + k] = inv
            continue
        tw = st[off + 1:off + 1 + nt].astype(np.int64)
        out, aux = tw & 0xFFFF, tw >> 16
        acc = np.where(aux == AUX_NEG, 0.0, lu[out])
        for p in range(ml):
            ow = st[off + 1 + nt + p * nt:off + 1 + nt + (p + 1) * nt].astype(np.int64)
            acc -= lu[ow & 0xFFFF] * lu[ow >> 16]
        for q in range(nt):
            if aux[q] == AUX_U or aux[q] == AUX_NEG:
                lu[out[q]] = acc[q]
            elif aux[q] == AUX_DIAG:
                lu[out[q]] = 1.0 / acc[q]
            else:
                lu[out[q]] = acc[q] * lu[aux[q]]
    if s.dense_rows is not None:      # W = (L_TT U_TT)^-1 behind the slots (the device keeps it in global scratch)
        nT = len(s.dense_rows)
        D = lu[s.dense_slots]
        Lm = np.tril(D, -1) + np.eye(nT)
        Um = np.triu(D, 1) + np.diag(1.0 / np.diag(D))
        lu = np.concatenate([lu, (np.linalg.inv(Um) @ np.linalg.inv(Lm)).reshape(-1)])
    return lu


def stream_solve(s: SymbolicLU, lu, b):
    """numpy emulation of the device triangular solves driven by the solve stream (validation of the tables)."""
    w = np.zeros(s.n + 1)
    w[:s.n] = b[s.perm_r]
    st = s.solve_stream
    off = 0
    while off < s.n_solve_words:
        w0, total, param = int(st[off]), int(st[off + 1]), int(st[off + 3])
        kind, mode, nb = w0 & 0xF, (w0 >> 4) & 0xF, (w0 >> 16) & 0xFF
        base = off + SEQ_HDR + nb
        if kind == 0:
            for v in range(nb):
                nt, ln = int(st[off + SEQ_HDR + v]) & 0xFFFF, int(st[off + SEQ_HDR + v]) >> 16
                for q in range(nt):
                    p = base + q
                    ntask = int(st[p])
                    p += nt
                    for _ in range(ntask):
                        hdr = int(st[p])
                        row, ml = hdr & 0xFFFF, hdr >> 16
                        acc = w[row]
                        for j in range(ml):
                            o = int(st[p + (1 + j) * nt])
                            acc -= lu[o & 0xFFFF] * w[o >> 16]
                        if mode & 1:
                            acc *= lu[row]
                        w[row] = acc
                        p += (1 + ml) * nt
                base += nt * ln
        elif kind == 2:
            for v in range(nb):
                nt = int(st[off + SEQ_HDR + v])
                for q in range(nt):
                    r0, K = int(st[base + q]) & 0xFFFF, int(st[base + q]) >> 16
                    L0, U0 = int(st[base + nt + q]) & 0xFFFF, int(st[base + nt + q]) >> 16
                    if mode & 1:
                        nxt = w[int(st[base + 2 * nt + q])]
                        for k in range(K - 1, -1, -1):
                            nxt = (w[r0 + k] - lu[U0 + k] * nxt) * lu[r0 + k]
                            w[r0 + k] = nxt
                    else:
                        for k in range(1, K):
                            w[r0 + k] -= lu[L0 + k - 1] * w[r0 + k - 1]
                base += 3 * nt
        elif kind == 3:
            nT = param >> 8
            rows = st[base:base + nT].astype(np.int64)
            w[rows] = lu[s.nnz_lu + 3:s.nnz_lu + 3 + nT * nT].reshape(nT, nT) @ w[rows]
        else:
            G, mlg = param & 0xFF, param >> 8
            for v in range(nb):
                new = {}
                nt = int(st[off + SEQ_HDR + v])
                for g in range(nt // G):
                    row = int(st[base + g * G])
                    acc = 0.0
                    for l in range(g * G, (g + 1) * G):
                        for j in range(mlg):
                            o = int(st[base + (1 + j) * nt + l])
                            acc += lu[o & 0xFFFF] * w[o >> 16]
                    bse = w[row] * lu[row] if mode & PAR_DIAG else w[row]
                    new[row] = bse + acc if mode & PAR_ADD else bse - acc
                base += (1 + mlg) * nt
                for row, val in new.items():      # every lane of the warp reads before any lane writes
                    w[row] = val
        off += total
    x = np.empty(s.n)
    x[s.perm_c] = w[:s.n]
    return x
