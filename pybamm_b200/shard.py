"""Batch sharding across GPUs: systems are independent, so the N_sys rows of
``inputs`` are split across devices / ranks with no collective inside the integration.

The reference splits the batch into contiguous static chunks, one per OpenMP solver, and solves the
remainder serially (idaklu_source/IDAKLUSolverGroup.cpp:99-160); step counts vary several-fold
across a C-rate sweep, so contiguous chunks load-balance badly.  Here rows are dealt round-robin
after sorting by a cost proxy (snake order), and every shard keeps its rows in ascending original
order so that gathers are monotone copies.
"""
from __future__ import annotations

import numpy as np


class CostModel:
    """Model-free predictor of the work of a system from its inputs: a least-squares fit of log(steps) on the logs
    of the input magnitudes, learned from the step counts of the previous solve of the same solver (``stats[:, 0]``).
    It only has to rank systems (launch order, shard balance), so a crude fit is enough: on the DFN bench the step
    count correlates +0.55 with the current alone (a fast discharge ends on the steep knee of the voltage curve)."""

    def __init__(self):
        self.coef = None

    @staticmethod
    def _features(inputs):
        x = np.asarray(inputs, dtype=np.float64)
        f = np.log(np.abs(x) + 1e-300)
        f = np.where(np.isfinite(f), f, 0.0)
        return np.column_stack([np.ones(len(x)), f])

    def fit(self, inputs, steps):
        steps = np.asarray(steps, dtype=np.float64).reshape(-1)
        inputs = np.asarray(inputs, dtype=np.float64)
        if inputs.ndim != 2 or len(steps) != len(inputs) or len(steps) < 4 * (inputs.shape[1] + 1) or inputs.shape[1] == 0:
            return
        A = self._features(inputs)
        ok = steps > 0
        if ok.sum() < A.shape[1] + 1:
            return
        self.coef, *_ = np.linalg.lstsq(A[ok], np.log(steps[ok]), rcond=None)
        self.n_inputs = inputs.shape[1]

    def predict(self, inputs):
        inputs = np.asarray(inputs, dtype=np.float64)
        if self.coef is None or inputs.ndim != 2 or inputs.shape[1] != self.n_inputs:
            return None
        return np.exp(np.clip(self._features(inputs) @ self.coef, -50.0, 50.0))


def cost_proxy(inputs: np.ndarray, model: CostModel | None = None) -> np.ndarray:
    """Expected work per system: the fitted cost model of previous solves when there is one, else uniform."""
    inputs = np.asarray(inputs, dtype=np.float64)
    if model is not None:
        c = model.predict(inputs)
        if c is not None:
            return c
    return np.ones(inputs.shape[0])


def launch_order(inputs: np.ndarray, model: CostModel | None = None) -> np.ndarray:
    """Row order for one device: most expensive systems first, so that the persistent kernel's teams do not end the
    batch waiting for one long straggler (the queue hands out rows in index order)."""
    return np.argsort(-cost_proxy(inputs, model), kind="stable")


def partition(n_sys: int, n_shards: int, cost=None):
    """Return ``n_shards`` sorted index arrays covering range(n_sys) exactly once.

    Sizes differ by at most one.  With ``cost`` the rows are dealt in snake order by descending cost
    (so that every shard gets a similar cost mix); without it the deal is plain round-robin."""
    if n_shards < 1:
        raise ValueError("n_shards must be positive")
    if cost is None:
        order = np.arange(n_sys)
    else:
        cost = np.asarray(cost, dtype=np.float64).reshape(-1)
        if len(cost) != n_sys:
            raise ValueError("cost has wrong length")
        order = np.argsort(-cost, kind="stable")
    shards = [[] for _ in range(n_shards)]
    for k, idx in enumerate(order):
        r, c = divmod(k, n_shards)
        s = c if (r % 2 == 0) else n_shards - 1 - c
        shards[s].append(int(idx))
    return [np.sort(np.asarray(s, dtype=np.int64)) for s in shards]


def gather_rows(parts, shards, n_sys):
    """Inverse of partition for per-system arrays: parts[k] has the rows of shards[k]."""
    first = next(p for p in parts if p is not None and len(p))
    out = np.empty((n_sys,) + first.shape[1:], dtype=first.dtype)
    for p, idx in zip(parts, shards):
        if len(idx):
            out[idx] = p
    return out
