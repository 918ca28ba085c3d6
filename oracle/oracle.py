"""ctypes driver of the CPU oracle (oracle/ida_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Never imported by the product package.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np


class Options(C.Structure):
    _fields_ = [("rtol", C.c_double), ("maxord", C.c_int), ("max_num_steps", C.c_long), ("maxnef", C.c_int),
                ("maxcor", C.c_int), ("maxncf", C.c_int), ("epcon", C.c_double), ("epsNewtIC", C.c_double),
                ("step_tol_ic", C.c_double), ("max_iter_ic", C.c_int), ("max_backtracks_ic", C.c_int),
                ("scalesol", C.c_int), ("dt_init", C.c_double), ("dt_max", C.c_double)]


_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_lp = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")

STAT_NAMES = ("nsteps", "nrevals", "nlinsetups", "netfails", "nniters", "nncfails", "njevals", "nge")


class Oracle:
    def __init__(self, so_path):
        self.lib = C.CDLL(os.path.abspath(so_path))
        sizes = np.zeros(6, np.int32)
        self.lib.pbo_sizes.argtypes = [_ip]
        self.lib.pbo_sizes(sizes)
        self.n, self.n_diff, self.n_p, self.n_ev, self.nnz, self.n_var = (int(x) for x in sizes)
        self.lib.pbo_eval.argtypes = [C.c_int, _dp, _dp, C.c_double, C.c_double, _dp]
        self.lib.pbo_default_options.argtypes = [C.POINTER(Options)]
        self.lib.pbo_solve.restype = C.c_int
        self.perm = None
        self.perm_r = self.perm_c = None

    def set_rcm(self, rows, cols):
        """Bandwidth-reducing symmetric permutation for the banded LU."""
        import scipy.sparse as sp
        from scipy.sparse.csgraph import reverse_cuthill_mckee

        A = sp.csr_matrix((np.ones(len(rows)), (rows, cols)), shape=(self.n, self.n))
        A = (A + A.T).tocsr()
        self.perm = np.ascontiguousarray(reverse_cuthill_mckee(A, symmetric_mode=True), dtype=np.int32)

    def set_sparse(self, perm_r, perm_c):
        """Row / column order of the static-pattern sparse LU (lu="sparse": the CPU-baseline leg; pybamm_b200.symbolic_lu)."""
        self.perm_r = np.ascontiguousarray(perm_r, dtype=np.int32)
        self.perm_c = np.ascontiguousarray(perm_c, dtype=np.int32)

    def eval(self, which, y, p, t=0.0, cj=0.0):
        idx = {"res": 0, "jac": 1, "events": 2, "ic": 3, "vars": 4}[which]
        n_out = {"res": self.n, "jac": self.nnz, "events": max(self.n_ev, 1), "ic": self.n, "vars": max(self.n_var, 1)}[which]
        out = np.zeros(n_out)
        y = np.ascontiguousarray(y if y is not None else np.zeros(self.n), dtype=np.float64)
        p = np.ascontiguousarray(p if len(p) else np.zeros(1), dtype=np.float64)
        self.lib.pbo_eval(idx, y, p, float(t), float(cj), out)
        return out

    def solve(self, t_eval, inputs, t_interp=None, rtol=1e-4, atol=1e-6, y0=None, save_vars=False, n_threads=1,
              n_max=None, options=None, lu="band"):
        """Returns dict(t [list of arrays], y [list of (n_t, n_ret)], flag, y_term, stats)."""
        t_eval = np.unique(np.asarray(t_eval, dtype=np.float64))
        t_interp = np.zeros(0) if t_interp is None else np.unique(np.asarray(t_interp, dtype=np.float64))
        t_interp = np.ascontiguousarray(np.setdiff1d(t_interp, t_eval))
        inputs = np.ascontiguousarray(np.atleast_2d(np.asarray(inputs, dtype=np.float64)))
        n_sys = inputs.shape[0]
        if inputs.shape[1] == 0:
            inputs = np.zeros((n_sys, 1))
        atol_v = np.ascontiguousarray(np.broadcast_to(np.asarray(atol, dtype=np.float64), (self.n,)).copy())
        opt = Options()
        self.lib.pbo_default_options(C.byref(opt))
        opt.rtol = rtol
        for k, v in (options or {}).items():
            setattr(opt, k, v)
        n_ret = self.n_var if save_vars else self.n
        if n_max is None:
            n_max = (len(t_interp) + len(t_eval) + 2) if len(t_interp) else 20000
        t_out = np.zeros((n_sys, n_max))
        y_out = np.zeros((n_sys, n_max, n_ret))
        n_out = np.zeros(n_sys, np.int32)
        flag = np.zeros(n_sys, np.int32)
        y_term = np.zeros((n_sys, self.n))
        stats = np.zeros((n_sys, 8), np.int64)
        vp = C.c_void_p
        self.lib.pbo_solve.argtypes = [C.c_int, _dp, C.c_int, vp, C.c_int, _dp, vp, _dp, C.POINTER(Options), vp,
                                       C.c_int, C.c_int, _dp, _dp, _ip, _ip, _dp, _lp, C.c_int, vp, C.c_int]
        sparse = lu == "sparse"
        if sparse and self.perm_r is None:
            raise ValueError("set_sparse(perm_r, perm_c) first")
        perm = self.perm_r if sparse else self.perm
        y0p = None
        if y0 is not None:
            y0 = np.ascontiguousarray(np.atleast_2d(np.asarray(y0, dtype=np.float64)))
            y0p = y0.ctypes.data_as(vp)
        rc = self.lib.pbo_solve(n_sys, t_eval, len(t_eval), t_interp.ctypes.data_as(vp) if len(t_interp) else None,
                                len(t_interp), inputs, y0p, atol_v, C.byref(opt),
                                perm.ctypes.data_as(vp) if perm is not None else None,
                                int(save_vars), n_max, t_out, y_out, n_out, flag, y_term, stats, n_threads,
                                self.perm_c.ctypes.data_as(vp) if sparse else None, int(sparse))
        if rc != 0:
            raise ValueError("t_eval must have at least 2 entries")
        return dict(t=[t_out[i, :n_out[i]] for i in range(n_sys)], y=[y_out[i, :n_out[i]] for i in range(n_sys)],
                    flag=flag, y_term=y_term, stats=stats, n_out=n_out)
