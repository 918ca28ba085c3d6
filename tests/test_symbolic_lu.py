"""Static sparse LU: clusters, CHAIN / PAR records and the solve form of the coupling blocks, checked against scipy on
synthetic "chains hanging off a banded core" matrices (the shape of a DFN Jacobian) and on the frozen battery bundles.
CPU only: the numpy emulations interpret the very tables the device kernels read."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from pybamm_b200 import codegen, symbolic_lu as slu
from tests import util


def chain_core_matrix(n_chain, chain_len, core_per_node, rng, unsym=True):
    """n_chain radial chains (tridiagonal), each coupled through its last two nodes to a block of core unknowns; the
    core blocks form a block-tridiagonal line (what is left of a DFN after the particles)."""
    n_core = n_chain * core_per_node
    n = n_chain * chain_len + n_core
    rows, cols = [], []

    def add(i, j):
        rows.append(i)
        cols.append(j)

    for c in range(n_chain):
        base = c * chain_len
        for k in range(chain_len):
            add(base + k, base + k)
            if k:
                add(base + k, base + k - 1)
                add(base + k - 1, base + k)
        cb = n_chain * chain_len + c * core_per_node
        for u in range(core_per_node):
            for v in range(core_per_node):
                add(cb + u, cb + v)
            add(cb + u, base + chain_len - 1)
            add(base + chain_len - 1, cb + u)
            if u == 0:
                add(cb + u, base + chain_len - 2)
            if c + 1 < n_chain:
                nb = cb + core_per_node
                add(cb + u, nb + u)
                add(nb + u, cb + u)
                if unsym and u + 1 < core_per_node:
                    add(cb + u, nb + u + 1)
    rows, cols = np.asarray(rows), np.asarray(cols)
    key = rows * n + cols
    _, first = np.unique(key, return_index=True)
    rows, cols = rows[first], cols[first]
    vals = rng.standard_normal(len(rows)) * 0.3
    vals[rows == cols] = 4.0 + rng.random(int(np.sum(rows == cols)))
    return n, rows, cols, vals


@pytest.mark.parametrize("team", [32, 128, 256])
@pytest.mark.parametrize("shape", [(6, 12, 3), (40, 18, 5), (3, 30, 2)])
def test_factor_and_solve_streams_match_scipy(shape, team):
    rng = np.random.default_rng(sum(shape) + team)
    n, rows, cols, vals = chain_core_matrix(*shape, rng)
    samples = [vals, vals * (1 + 0.1 * rng.standard_normal(len(vals)))]
    s = slu.analyse(n, rows, cols, samples, team=team)
    assert "chain" in s.cluster_kind and "inv" in s.cluster_kind
    for v in samples + [vals + 0.05 * rng.standard_normal(len(vals)) * (rows != cols)]:
        A = sp.csc_matrix((v, (rows, cols)), shape=(n, n))
        lu = slu.stream_factor(s, v)
        for _ in range(3):
            b = rng.standard_normal(n)
            x = slu.stream_solve(s, lu, b)
            xr = spla.spsolve(A, b)
            assert np.max(np.abs(x - xr)) <= 1e-10 * max(1.0, np.max(np.abs(xr)))


def test_dense_top_block_replaces_the_last_tree_heights(monkeypatch):
    """Multi-warp teams: the clusters above a tree height become one dense block (W = (L_TT U_TT)^-1 behind the slots,
    two records for both sweeps).  Same solutions as without it, fewer records, block closed under ancestors."""
    rng = np.random.default_rng(7)
    n, rows, cols, vals = chain_core_matrix(40, 18, 5, rng)
    samples = [vals, vals * (1 + 0.1 * rng.standard_normal(len(vals)))]
    s = slu.analyse(n, rows, cols, samples, team=256)
    monkeypatch.setenv("PBGPU_DENSE_TOP", "0")
    s0 = slu.analyse(n, rows, cols, samples, team=256)
    assert s0.dense_rows is None and s.dense_rows is not None
    assert 8 <= len(s.dense_rows) <= slu.DENSE_MAX and s.cluster_kind[-1] == "dense"
    assert len(s.solve_program) < len(s0.solve_program)
    kinds = [r["kind"] for r in s.solve_program]
    assert kinds.count(3) == 1 and s.solve_program[kinds.index(3)]["back"] is False
    assert s.seq_info["team_min"] <= 256
    assert slu.analyse(n, rows, cols, samples, team=32).dense_rows is None      # single-warp teams keep the records
    for v in samples:
        A = sp.csc_matrix((v, (rows, cols)), shape=(n, n))
        lu, lu0 = slu.stream_factor(s, v), slu.stream_factor(s0, v)
        assert len(lu) == s.nnz_lu + 3 + len(s.dense_rows) ** 2
        for _ in range(3):
            b = rng.standard_normal(n)
            xr = spla.spsolve(A, b)
            for sym, f in ((s, lu), (s0, lu0)):
                x = slu.stream_solve(sym, f, b)
                assert np.max(np.abs(x - xr)) <= 1e-10 * max(1.0, np.max(np.abs(xr)))


def test_cluster_and_record_invariants_on_the_battery_models():
    for name in util.BATTERY:
        b = util.bundle(name)
        team = codegen.default_team(b.n_states)
        s = slu.analyse(b.n_states, b.jac_rows, b.jac_cols, codegen.sample_matrices(b), team=team)
        # clusters partition the rows, in topological (ascending first row) order
        allrows = sorted(i for c in s.clusters for i in c)
        assert allrows == list(range(s.n))
        ordered = s.clusters[:-1] if s.dense_rows is not None else s.clusters     # the dense top block goes last
        assert [c[0] for c in ordered] == sorted(c[0] for c in ordered)
        if s.dense_rows is not None:
            assert s.cluster_kind[-1] == "dense" and s.clusters[-1] == sorted(s.dense_rows)
            assert len(s.dense_rows) <= slu.DENSE_MAX
            top = set(s.dense_rows)       # closed under ancestors: no row outside it depends on it in the forward sweep
            assert all(k not in top for i in range(s.n) if i not in top for k in s.Lrows[i])
        for c, kind in zip(s.clusters, s.cluster_kind):
            if kind == "chain":     # consecutive rows, bidiagonal factors, consecutive slots
                assert c == list(range(c[0], c[0] + len(c))) and len(c) >= 4
                for k, i in enumerate(c):
                    assert s.Lrows[i] == ([c[k - 1]] if k else [])
                    assert len(s.Urows[i]) == 2
                l0, u0 = s.slot_of[(c[1], c[0])], s.slot_of[(c[0], c[1])]
                assert [s.slot_of[(c[k], c[k - 1])] for k in range(1, len(c))] == list(range(l0, l0 + len(c) - 1))
                assert [s.slot_of[(i, s.Urows[i][1])] for i in c] == list(range(u0, u0 + len(c)))
            elif kind == "inv":
                assert len(c) <= slu.INV_MAX
        # every PAR block holds whole clusters (its lanes read each other's old values), at most 32 lanes, and the
        # widest record fits the team
        assert s.seq_info["team_min"] <= team
        for r in s.solve_program:
            assert r["first_warp"] + r["nblocks"] <= max(team // 32, 1)
            if r["kind"] != 1 or r.get("pre"):     # (the rows of the dense block's first step never read each other)
                continue
            G, mlg = r["param"] & 0xFF, r["param"] >> 8
            off = 0
            seen = {}
            for v, nt in enumerate(r["desc"]):
                assert nt <= 32 and nt % G == 0
                blk = np.asarray(r["body"][off:off + (1 + mlg) * nt]).reshape(1 + mlg, nt)
                for row in set(blk[0].tolist()):
                    cl = int(s.cl[row])
                    assert seen.setdefault(cl, v) == v, "rows of one cluster in two warps"
                off += (1 + mlg) * nt
        # the diagonal of row i lives in slot i
        assert all(s.slot_of[(i, i)] == i for i in range(s.n))


def test_solve_form_overwrites_only_coupling_entries():
    """S and T replace L_C,desc / U_C,anc of the inverse-block clusters in place; chain entries and diagonals keep
    their LU values (the CHAIN records and the diagonal scaling read them)."""
    b = util.bundle("spme")
    s = slu.analyse(b.n_states, b.jac_rows, b.jac_cols, codegen.sample_matrices(b), team=32)
    v = np.asarray(codegen.sample_matrices(b)[0], float)
    A = sp.csc_matrix((v, (b.jac_rows, b.jac_cols)), shape=(s.n, s.n)).toarray()
    B = A[np.ix_(s.perm_r, s.perm_c)]
    lu = slu.stream_factor(s, v)
    # plain dense LU without pivoting of the permuted matrix
    L, U = np.eye(s.n), B.copy()
    for k in range(s.n):
        L[k + 1:, k] = U[k + 1:, k] / U[k, k]
        U[k + 1:] -= np.outer(L[k + 1:, k], U[k])
    assert np.allclose([lu[i] for i in range(s.n)], 1.0 / np.diag(U), rtol=1e-9)
    for c, kind in zip(s.clusters, s.cluster_kind):
        if kind == "chain":
            for k in range(1, len(c)):
                assert np.isclose(lu[s.slot_of[(c[k], c[k - 1])]], L[c[k], c[k - 1]], rtol=1e-9)
            for i in c:
                j = s.Urows[i][1]
                assert np.isclose(lu[s.slot_of[(i, j)]], U[i, j], rtol=1e-9, atol=1e-300)


def test_hub_vertex_is_eliminated_last_and_the_streams_still_match_scipy():
    """A lumped state coupled to every core unknown, reading more of them than it feeds (the temperature of a thermal DFN):
    taken out of the dissection and eliminated last, static pivots on the natural diagonal."""
    rng = np.random.default_rng(77)
    n0, rows, cols, vals = chain_core_matrix(30, 10, 4, rng)
    hub = n0
    core = np.arange(30 * 10, n0)
    feeds = core[::2]                                  # column of the hub: half of the core
    rows = np.concatenate([rows, np.full(len(core), hub), feeds, [hub]])
    cols = np.concatenate([cols, core, np.full(len(feeds), hub), [hub]])
    vals = np.concatenate([vals, 0.05 * rng.standard_normal(len(core)), 0.05 * rng.standard_normal(len(feeds)), [5.0]])
    n = n0 + 1
    samples = [vals, vals * (1 + 0.1 * rng.standard_normal(len(vals)))]
    s = slu.analyse(n, rows, cols, samples, ordering="peel_nd", team=256)
    assert int(s.perm_c[-1]) == hub and int(s.perm_r[-1]) == hub
    for v in samples:
        A = sp.csc_matrix((v, (rows, cols)), shape=(n, n))
        lu = slu.stream_factor(s, v)
        b = rng.standard_normal(n)
        x = slu.stream_solve(s, lu, b)
        xr = spla.spsolve(A, b)
        assert np.max(np.abs(x - xr)) <= 1e-10 * max(1.0, np.max(np.abs(xr)))


def test_config4_thermal_dfn_fits_the_16_bit_streams():
    """BASELINE config 4 (3 523 states, 13 031 structural entries): peel + dissection with the temperature as border
    gives 17.5k L+U slots (SuperLU's MMD: 16.9k), inside the 16-bit slot / row indices of the packed streams."""
    b = util.bundle("dfn_thermal40")
    samples = codegen.sample_matrices(b)
    s = slu.analyse(b.n_states, b.jac_rows, b.jac_cols, samples, team=512)
    assert s.nnz_lu < 20000 and s.n_solve_records <= 24
    t_row = b.n_diff - 1                                # the lumped temperature is the last differential state
    assert int(s.perm_c[-1]) == t_row
    rng = np.random.default_rng(5)
    for v in (samples[0], samples[4]):
        v = np.asarray(v, float)
        A = sp.csc_matrix((v, (b.jac_rows, b.jac_cols)), shape=(b.n_states,) * 2)
        lu = slu.stream_factor(s, v)
        rhs = rng.standard_normal(b.n_states)
        x = slu.stream_solve(s, lu, rhs)
        xr = spla.spsolve(A, rhs)
        assert np.max(np.abs(x - xr)) <= 1e-8 * max(1.0, np.max(np.abs(xr)))
